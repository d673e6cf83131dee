"""GPU: b2d_nms (the `_nms` drop-in) through the C ABI against the golden vectors of the reference's
lib/nms/py_cpu_nms.py and against the oracle at RPN sizes - identical index lists."""
import numpy as np
import pytest
import torch

from nms_cases import make_dets
from oracle import nms as O
from test_oracle_nms import golden_cases

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,thresh,style,keep", list(golden_cases()), ids=lambda v: str(v) if not isinstance(v, np.ndarray) else "")
def test_golden_cases(cuda, n, thresh, style, keep):
    from image_denoising_b200 import nms
    dets = make_dets(n, seed=n * 10 + int(thresh * 10), style=style)
    got = nms.nms(dets, thresh)
    assert [int(k) for k in got] == [int(k) for k in keep]


@pytest.mark.parametrize("n", [6000, 12000])                    # cfg.TEST / TRAIN.RPN_PRE_NMS_TOP_N (config.py:145,194)
@pytest.mark.parametrize("style", ["clustered", "spread"])
def test_rpn_sizes_against_oracle(cuda, n, style):
    from image_denoising_b200 import nms
    dets = make_dets(n, seed=n + 1, style=style)
    for thresh in (0.7, 0.3):
        got = nms.nms(dets, thresh)
        want = O.nms(dets, thresh)
        assert [int(k) for k in got] == [int(k) for k in want], (n, style, thresh, len(got), len(want))


def test_device_entry_and_edges(cuda):
    from image_denoising_b200 import nms
    assert nms.nms(np.zeros((0, 5), np.float32), 0.5) == []
    one = np.array([[1, 2, 30, 40, 0.5]], np.float32)
    assert [int(k) for k in nms.nms(one, 0.5)] == [0]
    a = np.array([[0, 0, 9, 9, 0.9], [0, 0, 9, 9, 0.8], [5, 0, 14, 9, 0.7], [100, 100, 109, 109, 0.6]], np.float32)
    third = float(np.float32(50) / np.float32(150))
    assert [int(k) for k in nms.nms(a, 0.5)] == [0, 2, 3]
    assert [int(k) for k in nms.nms(a, 0.3)] == [0, 3]
    assert [int(k) for k in nms.nms(a[[0, 2]], third)] == [0, 1]           # IoU == thresh is kept (strict >)
    wide = np.concatenate([a, np.arange(8, dtype=np.float32).reshape(4, 2)], 1)   # boxes_dim = 7: extra columns ignored
    assert [int(k) for k in nms.nms(wide, 0.5)] == [0, 2, 3]
    dets = make_dets(3000, seed=5, style="clustered")
    order = dets[:, 4].argsort()[::-1]
    srt = torch.from_numpy(np.ascontiguousarray(dets[order])).to(cuda)
    keep = nms.nms_device(srt, 0.7).cpu().numpy()
    assert list(keep) == O.nms_sorted(dets[order], 0.7)
    assert nms.nms_device(srt[:0].contiguous(), 0.7).numel() == 0
