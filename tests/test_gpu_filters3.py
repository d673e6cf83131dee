"""GPU parity: the three 3x3 denoisers vs the reference's own cv2 calls
(bit-exact)."""
import numpy as np
import pytest
import torch

from conftest import synth_batch
from oracle import denoise as D

pytestmark = pytest.mark.gpu

SHAPES = [(64, 96), (600, 1000), (375, 500), (33, 344), (32, 256), (31, 260), (1, 4), (2, 8), (5, 1), (3, 2),
          (37, 53), (70, 342), (129, 1027)]
ORACLE = {"gaus_blur": D.gaus_blur_cv2, "mean": D.mean_cv2, "median": lambda u: D.median_cv2(u, 3)}


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("den", ["gaus_blur", "mean", "median"])
def test_filter3_bit_exact(cuda, shape, den):
    from image_denoising_b200 import api
    h, w = shape
    for kind in ("uniform", "smooth"):
        batch = synth_batch(3, h, w, seed=h * 7 + w, kind=kind)
        got = api.apply(torch.from_numpy(batch).to(cuda), "original_" + den).cpu().numpy()
        for i in range(len(batch)):
            want = ORACLE[den](batch[i])
            assert np.array_equal(got[i], want), (shape, den, kind, i, np.abs(got[i].astype(int) - want).max())


def test_mean_divider_exhaustive():
    s = np.arange(0, 9 * 255 + 1)
    assert np.array_equal(((s + 4) * 7282) >> 16, (2 * s + 9) // 18)
