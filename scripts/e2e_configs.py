#!/usr/bin/env python
"""Does b2d_run_host overlap copies and kernels for the plans that have pre-passes / wavelet scratch (C4, C5)?
For each config: (a) kernels only (b2d_run, resident), (b) the plain pinned H2D + D2H of the same bytes on two streams,
(c) b2d_run_host.  Overlap = (c) close to max(a, b) instead of a + b (round 1 ran these plans as ONE chunk: a + b).
Sizes are a quarter of the BASELINE configs so that the pinned buffers stay small."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import c5_specs, smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api  # noqa: E402

dev = torch.device("cuda", 0)
CASES = [("C2/4 sap_mean", 128, 600, 1000, "sap_mean_var0.4", {}),
         ("C3/8 speckle_bilateral", 128, 1080, 1920, "speckle_bilateral_var1.0", {}),
         ("C4/4 gaussian_wavelet db1", 64, 2048, 2048, "gaussian_wavelet_var1.5", {"wavelet": "db1", "levels": 3}),
         ("C5/8 mixed -> median 7", 512, 600, 1000, None, {"ksize": 7})]
for name, n, h, w, spec, kw in CASES:
    batch = smooth_batch_torch(n, h, w, dev, 7)
    out = torch.empty_like(batch)
    plan = api.Plan(n, h, w, c5_specs(4096)[:n] if spec is None else spec, 0, **kw)
    h_in = batch.cpu().pin_memory()
    h_out = torch.empty_like(h_in).pin_memory()

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best * 1e3
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def copies():
        with torch.cuda.stream(s1):
            batch.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s2):
            h_out.copy_(out, non_blocking=True)
    k_ms = timed(lambda: plan.run(batch, out, 3, 0))
    c_ms = timed(copies)
    e_ms = timed(lambda: plan.run_host(h_in, h_out, 3, 0))
    print(json.dumps({"case": name, "images": n, "MB_each_way": round(batch.numel() / 1e6, 1), "kernels_ms": round(k_ms, 2),
                      "copies_ms": round(c_ms, 2), "run_host_ms": round(e_ms, 2), "serial_would_be_ms": round(k_ms + c_ms, 2),
                      "run_host_over_max": round(e_ms / max(k_ms, c_ms), 3)}), flush=True)
    del plan, batch, out, h_in, h_out
    api._PLANS.clear()
    torch.cuda.empty_cache()
