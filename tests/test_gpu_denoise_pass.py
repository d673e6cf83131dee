"""GPU parity: median 5x5 / 7x7 (bit-exact vs cv2.medianBlur) and bilateral d=9
(<= 1 LSB vs cv2.bilateralFilter(u8, 9, 20, 100, BORDER_CONSTANT))."""
import numpy as np
import pytest
import torch

from conftest import synth_batch
from oracle import closures as C, denoise as D

pytestmark = pytest.mark.gpu

SHAPES = [(64, 96), (375, 500), (37, 53), (17, 344), (5, 7), (600, 1000)]


# the separable 7 x 7 kernel's tiles are 384 bytes x 48 rows, segments of 24 rows, steps of 2: sizes at and around them
MEDIAN_SHAPES = SHAPES + [(1, 1), (3, 2), (49, 129), (97, 256), (24, 128), (25, 131), (47, 127)]


@pytest.mark.parametrize("shape", MEDIAN_SHAPES)
@pytest.mark.parametrize("k", [5, 7])
def test_median_k_bit_exact(cuda, shape, k):
    from image_denoising_b200 import api
    h, w = shape
    for kind in ("uniform", "smooth"):
        batch = synth_batch(2, h, w, seed=h + w + k, kind=kind)
        got = api.apply(torch.from_numpy(batch).to(cuda), "original_median", ksize=k).cpu().numpy()
        for i in range(2):
            want = D.median_cv2(batch[i], k)
            assert np.array_equal(got[i], want), (shape, k, kind, int((got[i] != want).sum()))


@pytest.mark.parametrize("shape", SHAPES)
def test_bilateral_within_one_lsb(cuda, shape):
    from image_denoising_b200 import api
    h, w = shape
    for kind in ("uniform", "smooth"):
        batch = synth_batch(2, h, w, seed=h * 3 + w, kind=kind)
        got = api.apply(torch.from_numpy(batch).to(cuda), "original_bilateral").cpu().numpy()
        for i in range(2):
            want = D.bilateral_cv2(batch[i])
            d = np.abs(got[i].astype(int) - want.astype(int))
            assert d.max() <= 1 and (d > 0).mean() < 1e-3, (shape, kind, d.max(), (d > 0).mean())


def test_bilateral_other_diameter(cuda):
    import cv2
    from image_denoising_b200 import api
    batch = synth_batch(2, 40, 60, seed=8, kind="smooth")
    got = api.apply(torch.from_numpy(batch).to(cuda), "original_bilateral", d=5, sigma_color=30.0,
                    sigma_space=10.0).cpu().numpy()
    for i in range(2):
        want = cv2.bilateralFilter(batch[i], 5, 30, 10, borderType=cv2.BORDER_CONSTANT)
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3


@pytest.mark.parametrize("spec", ["speckle_bilateral_var1.0", "gaussian_bilateral_var0.1", "sap_bilateral_var0.4"])
def test_noise_then_bilateral_closure(cuda, spec):
    """Whole closure with injected fields: noise -> ->u8 -> bilateral."""
    from image_denoising_b200 import api
    from test_gpu_noise import draw_fields, to_dev
    noise, level = spec.split("_")[0], spec.split("var")[1]
    batch = synth_batch(2, 96, 128, seed=4, kind="smooth")
    rng = np.random.RandomState(3)
    fields = [draw_fields(noise, level, batch[0].shape, rng) for _ in batch]
    got = api.apply(torch.from_numpy(batch).to(cuda), spec, fields=to_dev(fields, noise, cuda)).cpu().numpy()
    for i in range(2):
        want = C.corrupt(batch[i], spec, fields=fields[i])
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3


def test_mixed_batch_per_image_specs(cuda):
    """One spec per image (BASELINE config 5 shape): groups are dispatched by image list."""
    from image_denoising_b200 import api
    batch = synth_batch(6, 64, 96, seed=12)
    specs = ["original_median", "periodic_median_var100", "original_gaus_blur", "periodic_median_var100",
             "original_median", "shader"]
    got = api.apply(torch.from_numpy(batch).to(cuda), specs, ksize=7).cpu().numpy()
    for i, s in enumerate(specs):
        want = C.corrupt(batch[i], s, median_k=7) if "original" not in s else \
            C.apply_denoise(batch[i], s.split("_", 1)[1], median_k=7)
        assert np.array_equal(got[i], want), (i, s)
