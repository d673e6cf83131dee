"""Golden vectors for the NMS row, made by the reference's own pure-Python implementation
lib/nms/py_cpu_nms.py imported from /root/reference - run HERE:

  python tests/golden/make_golden_nms.py     ->  tests/golden/nms_cases.npz"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference/lib/nms")
sys.path.insert(0, os.path.dirname(HERE))

from nms_cases import make_dets  # noqa: E402


def main():
    from py_cpu_nms import py_cpu_nms
    out = {}
    for n in (1, 2, 7, 63, 64, 65, 300, 1000, 2500):
        for thresh in (0.3, 0.5, 0.7):
            for style in ("clustered", "spread"):
                dets = make_dets(n, seed=n * 10 + int(thresh * 10), style=style)
                keep = np.array(py_cpu_nms(dets, thresh), np.int64)
                key = "n%d_t%.1f_%s" % (n, thresh, style)
                out[key + "_keep"] = keep
    np.savez_compressed(os.path.join(HERE, "nms_cases.npz"), **out)
    print(len(out), "cases")


if __name__ == "__main__":
    main()
