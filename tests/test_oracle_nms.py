"""CPU: the NMS oracle (oracle/nms.py, a restatement of the CUDA `_nms`) against golden vectors made by the reference's
own lib/nms/py_cpu_nms.py (tests/golden/make_golden_nms.py), plus hand cases."""
import os

import numpy as np
import pytest

from nms_cases import make_dets
from oracle import nms as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "nms_cases.npz")


def golden_cases():
    z = np.load(GOLDEN)
    for key in sorted(z.files):
        n, t, style = key[:-5].split("_")
        yield int(n[1:]), float(t[1:]), style, z[key]


@pytest.mark.parametrize("n,thresh,style,keep", list(golden_cases()), ids=lambda v: str(v) if not isinstance(v, np.ndarray) else "")
def test_oracle_equals_reference_py_cpu_nms(n, thresh, style, keep):
    dets = make_dets(n, seed=n * 10 + int(thresh * 10), style=style)
    assert [int(k) for k in O.nms(dets, thresh)] == [int(k) for k in keep]


def test_hand_cases():
    a = np.array([[0, 0, 9, 9, 0.9], [0, 0, 9, 9, 0.8], [5, 0, 14, 9, 0.7], [100, 100, 109, 109, 0.6]], np.float32)
    # IoU(0,1) = 1; IoU(0,2) = 50 / 150 = 1/3
    assert O.nms(a, 0.5) == [0, 2, 3]
    assert O.nms(a, 0.3) == [0, 3]
    assert O.nms(a[:0], 0.5) == []
    # exactly at the threshold is kept (strict >): two 10x10 boxes shifted by 5 -> 1/3 in float32
    third = float(np.float32(50) / np.float32(150))
    assert O.nms(a[[0, 2]], third) == [0, 1]
