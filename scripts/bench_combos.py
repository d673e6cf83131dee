"""Device-timed Mpix/s of a list of specs on one GPU (CUDA events, inputs resident).
  python scripts/bench_combos.py [--n 512 --h 600 --w 1000] spec [spec ...]   (spec may carry :k=v overrides)"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--h", type=int, default=600)
    ap.add_argument("--w", type=int, default=1000)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("specs", nargs="+")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    batch = smooth_batch_torch(a.n, a.h, a.w, dev, 1002)
    out = torch.empty_like(batch)
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
    for item in a.specs:
        spec, *kv = item.split(":")
        over = {}
        for t in kv:
            k, v = t.split("=")
            over[k] = v if k == "wavelet" else (float(v) if "." in v else int(v))
        plan = api.Plan(a.n, a.h, a.w, spec, 0, **over)
        for _ in range(3):
            plan.run(batch, out, 3, 0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(a.steps):
            plan.run(batch, out, 3, 0)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.steps
        mpix = a.n * a.h * a.w / 1e6
        gbs = mpix * 6e6 / (ms * 1e-3) / 1e9
        print(json.dumps({"spec": item, "ms": round(ms, 4), "mpix_s": round(mpix / (ms * 1e-3), 1),
                          "algo_GBs": round(gbs, 1), "hbm_frac": round(gbs / peaks["hbm_gbs"], 3)}), flush=True)
        del plan


if __name__ == "__main__":
    main()
