"""GPU: b2d_nms_device_static - the allocation-free NMS entry point (static device scratch inside the library, no
counterpart in the reference) - through the C ABI against b2d_nms_device, the existing entry point with a
caller-owned workspace.  Index lists must be identical (the tolerance every NMS test here uses).

Named test_zz_* so that pytest collects it last: a failure here must not hide
any other test.  Passes alone in a fresh process and after the rest of the suite (it keeps no state but the library's
own static array, which every call rewrites before reading)."""
import ctypes

import numpy as np
import pytest
import torch

from nms_cases import make_dets

pytestmark = pytest.mark.gpu

MAX_BOXES = 4096          # B2D_NMS_STATIC_MAX_BOXES, include/b200denoise.h


def _sorted_boxes(n, seed, style, cuda):
    dets = make_dets(n, seed=seed, style=style)
    order = dets[:, 4].argsort()[::-1]
    return torch.from_numpy(np.ascontiguousarray(dets[order])).to(cuda)


def _static(boxes, thresh, keep, num):
    """One call of the new entry point on preallocated outputs (nothing is allocated between here and the launch)."""
    from image_denoising_b200 import _lib
    n, d = boxes.shape
    with torch.cuda.device(boxes.device):
        st = torch.cuda.current_stream(boxes.device).cuda_stream
        return _lib.lib().b2d_nms_device_static(ctypes.c_void_p(boxes.data_ptr()), n, d, float(thresh),
                                                ctypes.c_void_p(keep.data_ptr()), ctypes.c_void_p(num.data_ptr()),
                                                ctypes.c_void_p(st))


def test_static_scratch_nms_equals_workspace_nms(cuda):
    from image_denoising_b200 import _lib, nms
    keep = torch.full((MAX_BOXES,), -1, dtype=torch.int32, device=cuda)
    num = torch.full((1,), -1, dtype=torch.int32, device=cuda)
    # two calls in a row (and more): the large clustered set first, so that the smaller ones run on a scratch array
    # still holding its bits under another row pitch; the same input twice must give the same list twice
    cases = [(MAX_BOXES, 0.7, "clustered"), (700, 0.3, "clustered"), (700, 0.3, "clustered"), (3000, 0.5, "spread"),
             (65, 0.7, "clustered"), (64, 0.5, "spread"), (1, 0.5, "spread"), (MAX_BOXES, 0.3, "spread")]
    for i, (n, thresh, style) in enumerate(cases):
        boxes = _sorted_boxes(n, 40 + n, style, cuda)
        want = nms.nms_device(boxes, thresh).cpu().numpy()              # existing entry point, caller workspace
        keep.fill_(-1)
        num.fill_(-1)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(cuda)[0]
        _lib.check(_static(boxes, thresh, keep, num))
        torch.cuda.synchronize()
        if i > 0:                                                       # (the first call may load the kernels' module)
            assert torch.cuda.mem_get_info(cuda)[0] >= free_before, "the entry point must not take device memory"
        k = int(num.item())
        assert k == want.size, (n, thresh, style, k, want.size)
        got = keep[:k].cpu().numpy()
        assert np.array_equal(got, want), (n, thresh, style)
        assert bool((keep[k:] == -1).all()), "wrote past the kept count"
    # empty input: count 0, no launch
    num.fill_(-1)
    _lib.check(_static(_sorted_boxes(8, 1, "spread", cuda)[:0].contiguous(), 0.5, keep, num))
    assert int(num.item()) == 0


def test_static_scratch_nms_rejects_oversize_and_bad_arguments(cuda):
    from image_denoising_b200 import _lib
    L = _lib.lib()
    keep = torch.empty(MAX_BOXES + 1, dtype=torch.int32, device=cuda)
    num = torch.zeros(1, dtype=torch.int32, device=cuda)
    boxes = _sorted_boxes(MAX_BOXES + 1, 3, "spread", cuda)
    assert _static(boxes, 0.5, keep, num) == _lib.ERR_INVALID
    msg = L.b2d_last_error().decode()
    assert str(MAX_BOXES) in msg and "B2D_NMS_STATIC_MAX_BOXES" in msg and "b2d_nms_device" in msg, msg
    assert L.b2d_nms_device_static(None, 5, 5, 0.5, ctypes.c_void_p(keep.data_ptr()), ctypes.c_void_p(num.data_ptr()), None) == _lib.ERR_INVALID
    assert L.b2d_nms_device_static(ctypes.c_void_p(boxes.data_ptr()), 5, 3, 0.5, ctypes.c_void_p(keep.data_ptr()),
                                   ctypes.c_void_p(num.data_ptr()), None) == _lib.ERR_INVALID     # boxes_dim < 4
    torch.cuda.synchronize()
