#!/usr/bin/env python
"""Times the device blob stage (b2d_blob_prep) and reports achieved HBM bandwidth:
algorithmic bytes = 3 read per source pixel + 12 written per destination pixel."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from image_denoising_b200 import blob

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=512)
ap.add_argument("--h", type=int, default=600)
ap.add_argument("--w", type=int, default=1000)
ap.add_argument("--target", type=int, default=600)
ap.add_argument("--max", type=int, default=1000)
ap.add_argument("--steps", type=int, default=10)
args = ap.parse_args()
means = np.array([[[102.9801, 115.9465, 122.7717]]])
peak = 6546.2
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
batch = torch.randint(0, 256, (args.n, args.h, args.w, 3), dtype=torch.uint8, device="cuda")
out, scale = blob.prep_blob_device(batch, means, args.target, args.max)
for _ in range(3):
    blob.prep_blob_device(batch, means, args.target, args.max, out=out)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
ev[0].record()
for i in range(args.steps):
    blob.prep_blob_device(batch, means, args.target, args.max, out=out)
    ev[i + 1].record()
torch.cuda.synchronize()
ms = float(np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]))
nbytes = batch.numel() + out.numel() * 4
print(json.dumps({"stage": "blob_prep", "shape": [args.n, args.h, args.w], "out": list(out.shape), "scale": scale,
                  "ms": round(ms, 4), "mpix_s": round(args.n * args.h * args.w / ms / 1e3, 1),
                  "algo_GBs": round(nbytes / ms / 1e6, 1), "hbm_frac": round(nbytes / ms / 1e6 / peak, 3)}))
