#!/usr/bin/env python
"""C4 building block for ncu launch lists: `gaussian_wavelet_var1.5`, db1 / 3 levels, N x 2048 x 2048 (default N = 64,
a quarter of BASELINE config 4).  Prints ms per pass (CUDA events)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = torch.device("cuda", 0)
batch = smooth_batch_torch(n, 2048, 2048, dev, 1004)
out = torch.empty_like(batch)
plan = api.Plan(n, 2048, 2048, "gaussian_wavelet_var1.5", 0, wavelet="db1", levels=3)
for _ in range(2):
    plan.run(batch, out, 3, 0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    plan.run(batch, out, 3, 0)
e1.record()
torch.cuda.synchronize()
print("C4 block: %d x 2048 x 2048, %.3f ms per pass, %.1f Gpix/s" % (n, e0.elapsed_time(e1) / reps, n * 4.194304 / (e0.elapsed_time(e1) / reps)))
