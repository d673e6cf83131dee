"""Oracle: greedy NMS as the reference's live path computes it (TEST INFRASTRUCTURE ONLY).

Restates ``_nms`` (lib/nms/nms_kernel.cu:21-144 behind lib/nms/gpu_nms.pyx:16-31): float32 IoU with the "+1"
convention in the kernel's operation order, suppression when IoU > thresh, greedy sweep in descending-score
order.  PINNED against the reference's own pure-Python implementation lib/nms/py_cpu_nms.py (imported from
/root/reference by tests/golden/make_golden_nms.py; inputs + outputs committed as tests/golden/nms_cases.npz):
the two agree on every box set whose areas are positive (py_cpu_nms drops NaN overlaps, the kernel keeps them).
"""
import numpy as np


def iou_row(box, others):
    """float32 IoU of one box against an (M, 4) array, operation order of devIoU (nms_kernel.cu:21-30)."""
    one = np.float32(1)
    left, right = np.maximum(box[0], others[:, 0]), np.minimum(box[2], others[:, 2])
    top, bottom = np.maximum(box[1], others[:, 1]), np.minimum(box[3], others[:, 3])
    width = np.maximum(right - left + one, np.float32(0))
    height = np.maximum(bottom - top + one, np.float32(0))
    inter = width * height
    sa = (box[2] - box[0] + one) * (box[3] - box[1] + one)
    sb = (others[:, 2] - others[:, 0] + one) * (others[:, 3] - others[:, 1] + one)
    with np.errstate(divide="ignore", invalid="ignore"):
        return inter / (sa + sb - inter)


def nms_sorted(boxes, thresh):
    """boxes: (N, >= 4) float32 rows in descending-score order -> kept row positions (ascending)."""
    b = np.ascontiguousarray(boxes[:, :4], dtype=np.float32)
    n = b.shape[0]
    removed = np.zeros(n, bool)
    keep = []
    t = np.float32(thresh)
    for i in range(n):
        if removed[i]:
            continue
        keep.append(i)
        if i + 1 < n:
            removed[i + 1:] |= iou_row(b[i], b[i + 1:]) > t
    return keep


def nms(dets, thresh):
    """lib/model/nms_wrapper.py::nms on (N, 5) [x1, y1, x2, y2, score]."""
    dets = np.asarray(dets, dtype=np.float32)
    if dets.shape[0] == 0:
        return []
    order = dets[:, 4].argsort()[::-1]
    return list(order[nms_sorted(dets[order], thresh)])
