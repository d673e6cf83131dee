"""Box sets of the NMS tests (deterministic; shared by tests/golden/make_golden_nms.py and the tests)."""
import numpy as np


def make_dets(n, seed, style="clustered", size=(600, 1000)):
    """(n, 5) float32 [x1, y1, x2, y2, score], x2 >= x1, y2 >= y1, distinct scores.  'clustered': boxes jittered
    around a few centres (heavy overlap, long suppression chains); 'spread': independent boxes."""
    rng = np.random.RandomState(seed)
    h, w = size
    if style == "clustered":
        k = max(1, n // 25)
        cx, cy = rng.uniform(0, w, k), rng.uniform(0, h, k)
        bw, bh = rng.uniform(20, 300, k), rng.uniform(20, 300, k)
        pick = rng.randint(0, k, n)
        x = cx[pick] + rng.normal(0, 12, n)
        y = cy[pick] + rng.normal(0, 12, n)
        ww = bw[pick] * rng.uniform(0.8, 1.25, n)
        hh = bh[pick] * rng.uniform(0.8, 1.25, n)
    else:
        x, y = rng.uniform(0, w, n), rng.uniform(0, h, n)
        ww, hh = rng.uniform(4, 400, n), rng.uniform(4, 400, n)
    x1 = np.clip(x - ww / 2, 0, w - 1)
    y1 = np.clip(y - hh / 2, 0, h - 1)
    x2 = np.clip(x + ww / 2, 0, w - 1)
    y2 = np.clip(y + hh / 2, 0, h - 1)
    scores = rng.permutation(n).astype(np.float64) / max(n, 1) * 0.98 + 0.01       # distinct
    if n >= 8:                                                                    # a few exact duplicates and integer boxes
        x1[3], y1[3], x2[3], y2[3] = x1[2], y1[2], x2[2], y2[2]
        x1[5], y1[5], x2[5], y2[5] = np.floor([x1[5], y1[5], x2[5], y2[5]])
    return np.stack([x1, y1, np.maximum(x2, x1), np.maximum(y2, y1), scores], 1).astype(np.float32)
