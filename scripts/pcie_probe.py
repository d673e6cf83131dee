#!/usr/bin/env python
"""Host<->device copy rates of the box (pinned memory), the ceiling of bench.py's `e2e` leg:
H2D alone, D2H alone, and both directions at once, whole buffer and in 32 MiB chunks."""
import json
import torch

n = 1843200000 // 2          # one direction of bench.py's C2 step
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        s1.synchronize(); s2.synchronize()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


def both_chunked(chunk=32 << 20):
    for o in range(0, n, chunk):
        with torch.cuda.stream(s1):
            d_in[o:o + chunk].copy_(h_in[o:o + chunk], non_blocking=True)
        with torch.cuda.stream(s2):
            h_out[o:o + chunk].copy_(d_out[o:o + chunk], non_blocking=True)


res = {}
for name, fn, nbytes in (("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n), ("both_32MiB_chunks", both_chunked, 2 * n)):
    ms = timed(fn)
    res[name] = {"ms": round(ms, 3), "GBs": round(nbytes / ms / 1e6, 2)}
res["e2e_ceiling_Mpix_s"] = round((n / 3) / (res["both"]["ms"] * 1e-3) / 1e6, 1)
print(json.dumps(res))
