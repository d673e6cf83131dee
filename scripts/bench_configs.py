#!/usr/bin/env python
"""The five BASELINE.json configurations at their full sizes, device timed.

  python scripts/bench_configs.py                      # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
         --master-port P scripts/bench_configs.py      # N GPUs: the batch of every config is SHARDED
                                                       # (strong scaling, contiguous ranges, no collective)

One JSON line per config: whole-job Mpix/s (images of all ranks / max-over-ranks CUDA-event time),
the roofline the config is bound by and the achieved fraction (6 B/pixel against the measured HBM
peak for the bandwidth-bound C2; lane-instructions per pixel x pixels/s against 37.2 T lane-instr/s
for the instruction-bound C3 / C5).  Inputs: the `smooth` synthetic images, resident in
HBM before the timed region; 3 warm-ups; every config's batch is larger than the 126 MB L2 except C1."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import smooth_batch_torch  # noqa: E402
from image_denoising_b200 import api, shard  # noqa: E402

SEED = 3
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def c5_specs(n, seed=SEED):
    levels = {"periodic": ["3.14", "100", "size"], "brownian": ["0.9", "0.09", "0.009"], "quant": ["3", "7", "10"]}
    out = []
    for i in range(n):
        hsh = (i * 2654435761 + seed * 40503) & 0xffffffff
        noise = ["poisson", "periodic", "brownian", "quant"][(hsh >> 8) % 4]
        out.append(noise + "_median" + ("" if noise == "poisson" else "_var" + levels[noise][(hsh >> 16) % 3]))
    return out


CONFIGS = [
    ("C1", 16, 600, 1000, ["gaussian_median_var1.0"], {}, "hbm"),
    ("C2", 512, 600, 1000, ["sap_mean_var0.4", "gaussian_gaus_blur_var1.5"], {}, "hbm"),
    ("C3", 1024, 1080, 1920, ["speckle_bilateral_var1.0"], {}, "instr:539"),
    ("C4", 256, 2048, 2048, ["gaussian_wavelet_var1.5"], {"wavelet": "db1", "levels": 3}, "hbm9"),
    ("C5", 4096, 600, 1000, None, {"ksize": 7}, "instr"),
]


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    steps = int(os.environ.get("B2D_STEPS", "5"))
    for name, n, h, w, specs, over, bound in CONFIGS:
        lo, cnt = shard.my_range(n, world, rank)
        hi = lo + cnt
        batch = smooth_batch_torch(cnt, h, w, dev, 1000 + int(name[1]))
        out = torch.empty_like(batch)
        if specs is None:
            plans = [api.Plan(cnt, h, w, c5_specs(n)[lo:hi], local, **over)]
        else:
            plans = [api.Plan(cnt, h, w, s, local, **over) for s in specs]
        for _ in range(3):
            for p in plans:
                p.run(batch, out, SEED, lo)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            for p in plans:
                p.run(batch, out, SEED, lo)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms = ms.item()
        mpix = n * h * w * len(plans) / 1e6
        line = {"config": name, "n_gpus": world, "images": n, "shape": [h, w], "specs": specs or "mixed poisson/periodic/brownian/quant -> median 7x7",
                "ms": round(ms, 3), "mpix_s": round(mpix / (ms * 1e-3), 1), "scaling": "strong"}
        if bound.startswith("hbm"):
            bpp = 9.0 if bound == "hbm9" else 6.0
            gbs = mpix * bpp * 1e6 / (ms * 1e-3) / 1e9 / world
            line["roofline"] = {"bound": "hbm", "bytes_per_pixel": bpp, "achieved_GBs_per_gpu": round(gbs, 1),
                                "peak": peaks["hbm_gbs"], "frac": round(gbs / peaks["hbm_gbs"], 3)}
        elif bound == "instr:539":
            tins = mpix * 1e6 * 539 / (ms * 1e-3) / 1e12 / world
            line["roofline"] = {"bound": "fp32/int issue", "lane_instr_per_pixel": 539, "achieved_T_per_gpu": round(tins, 2),
                                "peak": 37.2, "frac": round(tins / 37.2, 3)}
        else:
            line["roofline"] = {"bound": "int issue (selection networks)", "note": "181 compare-exchanges per output byte: see DESIGN.md"}
        if rank == 0:
            print(json.dumps(line), flush=True)
        del plans, batch, out
        api._PLANS.clear()
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
