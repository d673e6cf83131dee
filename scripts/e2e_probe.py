#!/usr/bin/env python
"""Times b2d_run_host (pinned host buffers) for one spec on the C2 batch and reports the copy
rate it sustains; B2D_HOST_CHUNK_MB selects the pipeline chunk."""
import json
import sys
import time

import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from image_denoising_b200 import api

specs = sys.argv[1:] or ["sap_mean_var0.4"]
n, h, w = 512, 600, 1000
h_in = torch.randint(0, 256, (n, h, w, 3), dtype=torch.uint8).pin_memory()
h_out = torch.empty_like(h_in).pin_memory()
plans = [api.get_plan(n, h, w, s) for s in specs]
for p in plans:
    p.run_host(h_in, h_out)
best = 1e9
for _ in range(5):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for p in plans:
        p.run_host(h_in, h_out)
    best = min(best, (time.perf_counter() - t0) / len(plans))
print(json.dumps({"specs": specs, "ms_per_spec": round(best * 1e3, 2), "GBs_each_way": round(h_in.numel() / best / 1e9, 2),
                  "mpix_s": round(n * h * w / best / 1e6, 1)}))
