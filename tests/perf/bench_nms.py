"""b2d_nms at the reference's sizes: per-class test-time NMS (N <= 300, lib/model/test.py:1860) and RPN
proposal NMS (N = 6000 / 12000, lib/layer_utils/proposal_layer.py:46), against the NumPy oracle on one host core.
  python tests/perf/bench_nms.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from image_denoising_b200 import nms  # noqa: E402
from nms_cases import make_dets  # noqa: E402
from oracle import nms as O  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    for n, thresh in ((300, 0.3), (2000, 0.7), (6000, 0.7), (12000, 0.7)):
        dets = make_dets(n, seed=n, style="clustered")
        order = dets[:, 4].argsort()[::-1]
        srt = torch.from_numpy(np.ascontiguousarray(dets[order])).to(dev)
        for _ in range(3):
            keep = nms.nms_device(srt, thresh)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            keep = nms.nms_device(srt, thresh)
        e1.record()
        torch.cuda.synchronize()
        dev_ms = e0.elapsed_time(e1) / 20
        for _ in range(3):
            nms.nms(dets, thresh)
        t0 = time.perf_counter()
        for _ in range(20):
            got = nms.nms(dets, thresh)
        host_ms = (time.perf_counter() - t0) * 1e3 / 20
        t0 = time.perf_counter()
        want = O.nms(dets, thresh)
        cpu_ms = (time.perf_counter() - t0) * 1e3
        assert [int(k) for k in got] == [int(k) for k in want]
        print(json.dumps({"n": n, "thresh": thresh, "kept": len(got), "device_ms": round(dev_ms, 4), "host_call_ms": round(host_ms, 4),
                          "numpy_oracle_ms_1core": round(cpu_ms, 2)}), flush=True)


if __name__ == "__main__":
    main()
