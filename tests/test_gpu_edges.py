"""GPU: degenerate shapes through the C ABI against the oracle (cv2 / NumPy on the same bytes).  The reference never
sees such images (VOC-sized JPEGs), but cv2's border rules are defined for them and the CUDA path must not read
outside an image: 1 x 1, single rows / columns, images smaller than the filter window and than a vector load, and
an empty batch."""
import numpy as np
import pytest
import torch

from oracle import closures as C

pytestmark = pytest.mark.gpu

TINY = [(1, 1), (1, 7), (5, 1), (2, 2), (3, 4), (8, 3), (2, 9), (7, 2), (4, 11)]


def _img(h, w, seed):
    return np.random.RandomState(seed).randint(0, 256, size=(3, h, w, 3)).astype(np.uint8)


@pytest.mark.parametrize("shape", TINY)
@pytest.mark.parametrize("den", ["gaus_blur", "mean", "median"])
def test_tiny_filters_bit_exact(cuda, shape, den):
    from image_denoising_b200 import api
    batch = _img(shape[0], shape[1], 7)
    got = api.apply(torch.from_numpy(batch).to(cuda), "original_" + den).cpu().numpy()
    for i in range(len(batch)):
        want = C.apply_denoise(batch[i], den)
        assert np.array_equal(got[i], want), (shape, den, i)


@pytest.mark.parametrize("shape", TINY)
@pytest.mark.parametrize("k", [5, 7])
def test_tiny_median_k(cuda, shape, k):
    from image_denoising_b200 import api
    batch = _img(shape[0], shape[1], 8)
    got = api.apply(torch.from_numpy(batch).to(cuda), "original_median", ksize=k).cpu().numpy()
    for i in range(len(batch)):
        assert np.array_equal(got[i], C.apply_denoise(batch[i], "median", median_k=k)), (shape, k, i)


@pytest.mark.parametrize("shape", TINY)
def test_tiny_bilateral(cuda, shape):
    from image_denoising_b200 import api
    batch = _img(shape[0], shape[1], 9)
    got = api.apply(torch.from_numpy(batch).to(cuda), "original_bilateral").cpu().numpy()
    for i in range(len(batch)):
        d = np.abs(got[i].astype(int) - C.apply_denoise(batch[i], "bilateral").astype(int))
        assert d.max() <= 1, (shape, i, int(d.max()))


@pytest.mark.parametrize("shape", TINY)
@pytest.mark.parametrize("spec", ["sap_var0.4", "sap_mean_var0.4", "shader", "periodic_var100", "periodic_median_var100"])
def test_tiny_noise_paths(cuda, shape, spec):
    from image_denoising_b200 import api
    batch = _img(shape[0], shape[1], 10)
    fields = None
    per_image = [None] * len(batch)
    if spec.startswith("sap"):
        rng = np.random.RandomState(3)
        u1, u2 = rng.random_sample(batch.shape), rng.random_sample(batch.shape)
        fields = (torch.from_numpy(u1).to(cuda), torch.from_numpy(u2).to(cuda))
        per_image = [{"u1": u1[i], "u2": u2[i]} for i in range(len(batch))]
    got = api.apply(torch.from_numpy(batch).to(cuda), spec, fields=fields).cpu().numpy()
    for i in range(len(batch)):
        want = C.corrupt(batch[i], spec, fields=per_image[i])
        assert np.array_equal(got[i], want), (shape, spec, i)


@pytest.mark.parametrize("spec", ["gaussian_median_var1.0", "poisson", "brownian_var0.09", "uniform_mean_var0.6", "quant_var3"])
def test_tiny_generator_paths_run_and_are_reproducible(cuda, spec):
    """Generator mode on degenerate shapes: no fault, same bytes on a second call, independent of the batch split."""
    from image_denoising_b200 import api
    for h, w in TINY:
        batch = torch.from_numpy(_img(h, w, 11)).to(cuda)
        a = api.apply(batch, spec, seed=5)
        b = api.apply(batch, spec, seed=5)
        assert torch.equal(a, b), (spec, h, w)
        tail = api.apply(batch[1:].contiguous(), spec, seed=5, first_index=1)
        assert torch.equal(a[1:], tail), (spec, h, w)
    torch.cuda.synchronize()


def test_empty_batch_is_a_no_op(cuda):
    from image_denoising_b200 import api
    empty = torch.empty((0, 16, 24, 3), dtype=torch.uint8, device=cuda)
    for spec in ("sap_mean_var0.4", "original_bilateral", "gaussian_wavelet_var1.5", "poisson_median"):
        out = api.apply(empty, spec)
        assert out.shape == empty.shape
    torch.cuda.synchronize()


@pytest.mark.parametrize("shape", [(2, 2), (3, 4), (8, 3), (12, 17), (9, 33)])
@pytest.mark.parametrize("wl", [("bior1.5", 0), ("db1", 0), ("db1", 2)])
def test_tiny_wavelet(cuda, shape, wl):
    """Images smaller than the 10-tap filter: pywt's symmetric extension folds more than once."""
    from image_denoising_b200 import api
    batch = _img(shape[0], shape[1], 12)
    got = api.apply(torch.from_numpy(batch).to(cuda), "original_wavelet", wavelet=wl[0], levels=wl[1]).cpu().numpy()
    for i in range(len(batch)):
        want = C.apply_denoise(batch[i], "wavelet", wavelet=wl[0], wavelet_levels=wl[1] or None)
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1, (shape, wl, i, int(d.max()))


def test_more_images_than_a_grid_dimension(cuda):
    """70 000 tiny images in one plan: the image index is a grid dimension (<= 65535) in every kernel, so the
    dispatcher, b2d_stats and the wavelet chunking split the launches; results must equal per-slice runs."""
    from image_denoising_b200 import api
    n, h, w = 70000, 8, 8
    rng = np.random.RandomState(4)
    batch = torch.from_numpy(rng.randint(0, 256, (n, h, w, 3)).astype(np.uint8)).to(cuda)
    for spec, kw in (("sap_mean_var0.4", {}), ("gaussian_median_var1.0", {"ksize": 5}), ("poisson_median", {}),
                     ("gaussian_wavelet_var1.0", {"wavelet": "db1", "levels": 3})):
        plan = api.get_plan(n, h, w, spec, cuda.index, **kw)
        full = plan.run(batch, seed=9, first_index=0)
        for lo in (0, 65530, 69990):
            part = api.apply(batch[lo:lo + 10].contiguous(), spec, seed=9, first_index=lo, **kw)
            assert torch.equal(part, full[lo:lo + 10]), (spec, lo)
        cs, sse = plan.stats(full, batch)
        cs2, sse2 = api.get_plan(10, h, w, spec, cuda.index, **kw).stats(full[69990:].contiguous(), batch[69990:].contiguous())
        assert torch.equal(cs[69990:], cs2) and torch.equal(sse[69990:], sse2), spec
    # mixed batch with a shared stand-alone denoiser: the single deferred denoise launch is split as well
    specs = ["poisson_median" if i % 2 else "sap_median_var0.4" for i in range(n)]
    full = api.get_plan(n, h, w, specs, cuda.index, ksize=5).run(batch, seed=9)
    for lo in (0, 65534, 69990):
        part = api.apply(batch[lo:lo + 10].contiguous(), specs[lo:lo + 10], seed=9, first_index=lo, ksize=5)
        assert torch.equal(part, full[lo:lo + 10]), lo
