"""GPU parity for the pixel-granular corruptions:
colour quantisation (Lab conversions bit-exact, assignment + palette bit-exact with injected
centres, fit by inertia) and bloom (Automold.add_sun_flare replay, committed golden fixture)."""
import os
import random

import numpy as np
import pytest
import torch

from conftest import synth_batch, synth_image
from oracle import closures as C, noise as N

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_lab_conversions_exhaustive(cuda):
    """All 2^24 colours through both conversions == cv2.cvtColor (8-bit fixed-point path)."""
    import cv2
    from image_denoising_b200 import api
    a, b, c = np.meshgrid(np.arange(256), np.arange(256), np.arange(256), indexing="ij")
    cube = np.stack([a, b, c], -1).astype(np.uint8).reshape(4096, 4096, 3)
    dev = torch.from_numpy(cube).to(cuda)
    for direction, code in ((0, cv2.COLOR_BGR2LAB), (1, cv2.COLOR_LAB2BGR)):
        got = api.lab_convert(dev, direction).cpu().numpy()
        want = cv2.cvtColor(cube, code)
        assert np.array_equal(got, want), (direction, int((got != want).any(-1).sum()))


@pytest.mark.parametrize("shape", [(64, 96), (375, 500), (37, 53)])
@pytest.mark.parametrize("k", [3, 7, 10])
@pytest.mark.parametrize("den", ["", "_median", "_mean"])
def test_quant_injected_centres_bit_exact(cuda, shape, k, den):
    from image_denoising_b200 import api
    h, w = shape
    batch = synth_batch(3, h, w, seed=k + h, kind="smooth")
    rng = np.random.RandomState(k)
    centres = [np.stack([rng.uniform(20, 235, k), rng.uniform(100, 160, k), rng.uniform(100, 160, k)], -1)
               for _ in batch]
    spec = "quant%s_var%d" % (den, k)
    f0 = torch.from_numpy(np.stack(centres)).to(cuda).contiguous()
    got = api.apply(torch.from_numpy(batch).to(cuda), spec, fields=(f0, None)).cpu().numpy()
    for i in range(len(batch)):
        want = C.corrupt(batch[i], spec, fields={"centers": centres[i]})
        assert np.array_equal(got[i], want), (spec, shape, i, int((got[i] != want).sum()))


@pytest.mark.parametrize("case", ["near_ties", "wild_centre"])
def test_quant_label_decision_paths(cuda, case):
    """The label is decided by fp32 scores unless the two best centres are within 0.25 of a tie (then the exact float64
    loop in NumPy's operation order runs) or a centre lies far outside the uint8 cube (then only the exact loop runs):
    centres that differ by 1e-9 / 1e-4 force the first case on every pixel, a centre at 600 the second."""
    from image_denoising_b200 import api
    batch = synth_batch(2, 96, 128, seed=31, kind="smooth")
    rng = np.random.RandomState(4)
    if case == "near_ties":                                               # k = 10: every pixel has two or three near-equal bests
        base = np.stack([rng.uniform(20, 235, 4), rng.uniform(100, 160, 4), rng.uniform(100, 160, 4)], -1)
        centres = np.concatenate([base + 1e-4, base, base[:2] - 1e-9], 0)
    else:                                                                 # k = 7
        base = np.stack([rng.uniform(20, 235, 6), rng.uniform(100, 160, 6), rng.uniform(100, 160, 6)], -1)
        centres = np.concatenate([base, [[600.0, 128.0, 128.0]]], 0)
    k = len(centres)
    spec = "quant_var%d" % k
    f0 = torch.from_numpy(np.stack([centres, centres[::-1].copy()])).to(cuda).contiguous()
    got = api.apply(torch.from_numpy(batch).to(cuda), spec, fields=(f0, None)).cpu().numpy()
    for i, cen in enumerate([centres, centres[::-1].copy()]):
        want = C.corrupt(batch[i], spec, fields={"centers": cen})
        assert np.array_equal(got[i], want), (case, i, int((got[i] != want).sum()))


def _inertia(img, out):
    import cv2
    lab = cv2.cvtColor(img, cv2.COLOR_BGR2LAB).reshape(-1, 3).astype(np.float64)
    q = cv2.cvtColor(out, cv2.COLOR_BGR2LAB).reshape(-1, 3).astype(np.float64)
    return float(((lab - q) ** 2).sum(-1).mean())


@pytest.mark.parametrize("k", [3, 7, 10])
def test_quant_generated_fit_by_inertia(cuda, k):
    """Generator mode: own k-means fit.  Palette size <= k and quantisation error comparable with
    sklearn's MiniBatchKMeans on the same image (distribution-level parity)."""
    from image_denoising_b200 import api
    img = synth_image(240, 320, 5, "smooth")
    img[:80] = (img[:80].astype(int) * 0.3).astype(np.uint8)          # some structure for the clusters
    img[160:, :, 2] = 255 - img[160:, :, 2]
    got = api.apply(torch.from_numpy(img[None]).to(cuda), "quant_var%d" % k, seed=9).cpu().numpy()[0]
    assert len(np.unique(got.reshape(-1, 3), axis=0)) <= k
    ref = C.corrupt(img, "quant_var%d" % k)
    e_got, e_ref = _inertia(img, got), _inertia(img, ref)
    assert e_got <= 1.35 * e_ref + 2.0, (k, e_got, e_ref)
    again = api.apply(torch.from_numpy(img[None]).to(cuda), "quant_var%d" % k, seed=9).cpu().numpy()[0]
    assert np.array_equal(got, again)                                  # counter based: reproducible


def _draw_tensor(draws_per_image, cuda):
    arr = np.array([[[a, cx, cy, r, c[0], c[1], c[2]] for (a, cx, cy, r, c) in d] for d in draws_per_image], np.float64)
    return torch.from_numpy(arr).to(cuda).contiguous()


def test_bloom_golden_fixture(cuda):
    """The committed output of the reference's own Automold.add_sun_flare (random.seed(7))."""
    from image_denoising_b200 import api
    want = np.load(os.path.join(GOLDEN, "bloom_seed7_320x480.npz"))["out"]
    img = synth_image(320, 480, 7, "smooth")
    draws = N.bloom_draws(img.shape, random.Random(7))
    got = api.apply(torch.from_numpy(img[None]).to(cuda), "bloom", fields=(_draw_tensor([draws], cuda), None))
    d = np.abs(got.cpu().numpy()[0].astype(int) - want.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3, (d.max(), (d > 0).mean())


@pytest.mark.parametrize("shape", [(600, 1000), (375, 500), (301, 333)])
def test_bloom_injected_draws_vs_cv2_replay(cuda, shape):
    from image_denoising_b200 import api
    h, w = shape
    batch = synth_batch(2, h, w, seed=3, kind="smooth")
    draws = [N.bloom_draws(batch[0].shape, random.Random(100 + i)) for i in range(2)]
    got = api.apply(torch.from_numpy(batch).to(cuda), "bloom", fields=(_draw_tensor(draws, cuda), None)).cpu().numpy()
    for i in range(2):
        want = N.bloom(batch[i], draws[i])
        d = np.abs(got[i].astype(int) - want.astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3, (shape, i, d.max(), (d > 0).mean())


def test_bloom_generated(cuda):
    """Generator mode: draws from Philox; output brighter than the input, reproducible, H < 300 rejected."""
    from image_denoising_b200 import api, _lib
    batch = torch.from_numpy(synth_batch(2, 320, 480, seed=4, kind="smooth")).to(cuda)
    a = api.apply(batch, "bloom", seed=5)
    assert torch.equal(a, api.apply(batch, "bloom", seed=5))
    assert not torch.equal(a, api.apply(batch, "bloom", seed=6))
    assert a.float().mean() > batch.float().mean()
    assert int(a[0, 100, 100].min()) >= 250                            # flare source centre is white
    with pytest.raises(_lib.B2DError):
        api.apply(torch.from_numpy(synth_batch(1, 200, 320)).to(cuda), "bloom")
