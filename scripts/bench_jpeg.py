"""Decode rate of the device JPEG stage against cv2.imdecode on the host cores.
  python scripts/bench_jpeg.py [--n 256 --h 600 --w 1000 --quality 90 --threads 0]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import cv2  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

from conftest import synth_image  # noqa: E402
from image_denoising_b200 import _lib, jpeg  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--h", type=int, default=600)
    ap.add_argument("--w", type=int, default=1000)
    ap.add_argument("--quality", type=int, default=90)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    a = ap.parse_args()
    base = [cv2.imencode(".jpg", synth_image(a.h, a.w, i, "smooth"), [cv2.IMWRITE_JPEG_QUALITY, a.quality])[1].tobytes() for i in range(8)]
    datas = [base[i % 8] for i in range(a.n)]
    mpix = a.n * a.h * a.w / 1e6
    dev = torch.device("cuda", 0)
    dec = jpeg.Decoder(a.n, a.h, a.w, dev, a.threads)
    out = torch.empty((a.n, a.h, a.w, 3), dtype=torch.uint8, device=dev)
    for _ in range(2):
        dec.decode(datas, out)
    torch.cuda.synchronize()
    L = _lib.lib()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        dec.decode(datas, out)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / a.steps
    # kernels alone: decode once more and time the stream with events around a second enqueue of the same work
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    th0 = time.perf_counter()
    e0.record()
    dec.decode(datas, out)
    e1.record()
    host_s = time.perf_counter() - th0
    torch.cuda.synchronize()
    want = cv2.imdecode(np.frombuffer(datas[3], np.uint8), cv2.IMREAD_COLOR)
    assert np.array_equal(out[3].cpu().numpy(), want)
    cv2.setNumThreads(1)
    t0 = time.perf_counter()
    for d in datas[:32]:
        cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR)
    cpu = (time.perf_counter() - t0) / 32
    print(json.dumps({"n": a.n, "shape": [a.h, a.w], "jpeg_bytes": len(datas[0]), "wall_ms_per_batch": round(dt * 1e3, 2),
                      "mpix_s": round(mpix / dt, 1), "images_s": round(a.n / dt, 1),
                      "host_part_ms": round(host_s * 1e3, 2), "device_after_host_ms": round(e0.elapsed_time(e1) - host_s * 1e3, 2),
                      "cv2_imdecode_ms_per_image_1thread": round(cpu * 1e3, 3), "cv2_mpix_s_1thread": round(a.h * a.w / 1e6 / cpu, 1),
                      "host_threads": a.threads or os.cpu_count(), "launches": int(L.b2d_launch_count())}))


if __name__ == "__main__":
    main()
