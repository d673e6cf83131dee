"""Drop-in for lib/model/nms_wrapper.py::nms and lib/nms/gpu_nms.pyx::gpu_nms.

Same call, same result: ``keep = nms(dets, thresh)`` with ``dets`` an (N, 5) float32 array of
[x1, y1, x2, y2, score] rows returns the list of kept row indices in descending-score order
(lib/model/test.py:1860, lib/layer_utils/proposal_layer.py:46).  The score sort stays on the host exactly as in
gpu_nms.pyx:27-29 (``scores.argsort()[::-1]``, so ties break the same way); mask and sweep run on the GPU
(``b2d_nms``).  There is no CPU implementation here: ``force_cpu`` is accepted for signature compatibility
(the reference's cpu_nms returns the same indices) and ignored.
"""
import ctypes

import numpy as np

from . import _lib


def gpu_nms(dets, thresh, device_id=0):
    dets = np.ascontiguousarray(dets, dtype=np.float32)
    if dets.ndim != 2 or dets.shape[1] < 5:
        raise ValueError("dets must be (N, >= 5) [x1, y1, x2, y2, score]")
    n = dets.shape[0]
    if n == 0:
        return []
    order = dets[:, 4].argsort()[::-1]                       # gpu_nms.pyx:27-28
    sorted_dets = np.ascontiguousarray(dets[order, :])         # gpu_nms.pyx:29
    keep = np.zeros(n, dtype=np.int32)
    num = ctypes.c_int(0)
    L = _lib.lib()
    _lib.check(L.b2d_nms(ctypes.c_void_p(keep.ctypes.data), ctypes.byref(num), ctypes.c_void_p(sorted_dets.ctypes.data),
                         n, dets.shape[1], float(thresh), int(device_id)))
    return list(order[keep[:num.value]])                       # gpu_nms.pyx:30-31


def nms(dets, thresh, force_cpu=False):
    """lib/model/nms_wrapper.py:15-23."""
    if dets.shape[0] == 0:
        return []
    return gpu_nms(dets, thresh, device_id=0)


def nms_device(sorted_boxes, thresh):
    """(N, D >= 4) float32 CUDA tensor, rows already in descending-score order -> int32 CUDA tensor of kept rows."""
    import torch
    if not sorted_boxes.is_cuda or sorted_boxes.dtype != torch.float32 or sorted_boxes.dim() != 2 or not sorted_boxes.is_contiguous():
        raise ValueError("sorted_boxes must be a contiguous float32 CUDA tensor (N, D)")
    n, d = sorted_boxes.shape
    L = _lib.lib()
    need = ctypes.c_size_t(0)
    _lib.check(L.b2d_nms_workspace_bytes(n, ctypes.byref(need)))
    dev = sorted_boxes.device
    ws = torch.empty(max(need.value, 8), dtype=torch.uint8, device=dev)
    keep = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    num = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream(dev).cuda_stream
        _lib.check(L.b2d_nms_device(ctypes.c_void_p(sorted_boxes.data_ptr()), n, d, float(thresh), ctypes.c_void_p(keep.data_ptr()),
                                    ctypes.c_void_p(num.data_ptr()), ctypes.c_void_p(ws.data_ptr()), ctypes.c_void_p(st)))
    return keep[:int(num.item())]
