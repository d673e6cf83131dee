#!/usr/bin/env python
"""Random shapes / contents: separable 5 x 5 and 7 x 7 medians (b2d_median7.cu) against cv2.medianBlur, bit for bit."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import cv2
import numpy as np
import torch

from image_denoising_b200 import api

rng = np.random.RandomState(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
bad = 0
for case in range(int(sys.argv[2]) if len(sys.argv) > 2 else 80):
    h, w, n = int(rng.randint(1, 140)), int(rng.randint(1, 320)), int(rng.randint(1, 4))
    kind = case % 3
    if kind == 0:
        batch = rng.randint(0, 256, (n, h, w, 3)).astype(np.uint8)
    elif kind == 1:
        batch = rng.randint(0, 4, (n, h, w, 3)).astype(np.uint8) * 85          # many ties
    else:
        batch = (np.add.outer(np.arange(h), np.arange(w))[None, :, :, None] * np.ones((n, 1, 1, 3)) % 256).astype(np.uint8)
    x = torch.from_numpy(batch).cuda()
    for k in (5, 7):
        got = api.apply(x, "original_median", ksize=k).cpu().numpy()
        for i in range(n):
            want = cv2.medianBlur(batch[i], k)
            if not np.array_equal(got[i], want):
                bad += 1
                print("MISMATCH", (n, h, w), k, kind, int((got[i] != want).sum()))
print("cases done, mismatches:", bad)
sys.exit(1 if bad else 0)
