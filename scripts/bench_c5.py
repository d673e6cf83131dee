#!/usr/bin/env python
"""C5 (mixed poisson / periodic / brownian / quant -> median 7 x 7) device timed on one GPU: the full 4096-image batch and
the first 512-image shard (what each rank owns when the batch is cut over 8 GPUs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import c5_specs, smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api  # noqa: E402

dev = torch.device("cuda", 0)
specs = c5_specs(4096)
for n in (4096, 512):
    batch = smooth_batch_torch(n, 600, 1000, dev, 1005)
    out = torch.empty_like(batch)
    plan = api.Plan(n, 600, 1000, specs[:n], 0, ksize=7)
    for _ in range(2):
        plan.run(batch, out, 3, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        plan.run(batch, out, 3, 0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print("C5 %d images: %.3f ms (%.3f ms per 512), %.1f Gpix/s" % (n, ms, ms * 512 / n, n * 0.6 / ms))
    del plan, batch, out
    api._PLANS.clear()
    torch.cuda.empty_cache()
