#!/usr/bin/env python
"""C3 building block for ncu: `speckle_bilateral_var1.0` (d = 9) on N x 1080 x 1920 (default N = 128, an eighth of
BASELINE config 3).  Prints ms per pass (CUDA events) for the whole spec and for the filter alone on clean input."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
batch = smooth_batch_torch(n, 1080, 1920, dev, 1003)
out = torch.empty_like(batch)
for spec in ("speckle_bilateral_var1.0", "original_bilateral", "speckle_var1.0"):
    plan = api.Plan(n, 1080, 1920, spec, 0)
    for _ in range(2):
        plan.run(batch, out, 3, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        plan.run(batch, out, 3, 0)
    e1.record()
    torch.cuda.synchronize()
    print("%-28s %d x 1080p: %.3f ms per pass" % (spec, n, e0.elapsed_time(e1) / reps))
    del plan
