"""CPU: property tests that hold the skimage/pywt restatement (oracle/wavelet.py).
The real libraries cannot be installed here (parity unpinned, see the module
header), so the restatement is pinned by what any correct implementation of the
published algorithm must satisfy."""
import numpy as np
import pytest

from oracle import wavelet as W


@pytest.mark.parametrize("wl", ["db1", "bior1.5"])
@pytest.mark.parametrize("shape", [(64, 96), (37, 53), (75, 125), (33, 20)])
@pytest.mark.parametrize("level", [1, 2, 3])
def test_perfect_reconstruction(wl, shape, level):
    x = np.random.RandomState(1).rand(*shape)
    c = W.wavedec2(x, wl, level)
    r = W.waverec2(c, wl)[: shape[0], : shape[1]]
    assert np.abs(r - x).max() < 1e-12


def test_coefficient_lengths_follow_pywt():
    # floor((n + F - 1) / 2) per level
    for n in (37, 64, 600, 1000):
        for wl, F in (("db1", 2), ("bior1.5", 10)):
            a, d = W.dwt_axis(np.zeros(n), wl, 0)
            assert a.shape[0] == d.shape[0] == (n + F - 1) // 2


def test_haar_hand_computed():
    s = 2 ** -0.5
    a, d = W.dwt_axis(np.array([1.0, 2.0, 3.0, 4.0]), "db1", 0)
    assert np.allclose(a, [3 * s, 7 * s]) and np.allclose(d, [-s, -s])
    x = np.arange(16.0).reshape(4, 4)
    c = W.wavedec2(x, "db1", 1)
    assert np.allclose(c[0], [[5, 9], [21, 25]])            # 2x2 block sums / 2
    assert np.allclose(np.abs(c[1]["dd"]), 0)
    assert np.allclose(np.abs(c[1]["ad"]), 1) and np.allclose(np.abs(c[1]["da"]), 4)


def test_haar_parseval_even_sizes():
    x = np.random.RandomState(2).rand(64, 96)
    c = W.wavedec2(x, "db1", 3)
    e = (c[0] ** 2).sum() + sum((v ** 2).sum() for d in c[1:] for v in d.values())
    assert abs(e / (x ** 2).sum() - 1) < 1e-12


def test_bior15_filters_are_biorthogonal():
    dl, dh, rl, rh = (np.array(f) for f in W.WAVELETS["bior1.5"])
    assert abs(dl.sum() - 2 ** 0.5) < 1e-12 and abs(rl.sum() - 2 ** 0.5) < 1e-12
    assert abs(dh.sum()) < 1e-12 and abs(rh.sum()) < 1e-12
    # 5 vanishing moments on the decomposition side: rec_hi kills polynomials of degree < 5? no -
    # dec_hi (Haar-like) has 1 vanishing moment, rec_hi (long) has 5
    k = np.arange(10)
    for m in range(5):
        assert abs((rh * k ** m).sum()) < 1e-9 * 10 ** m
    assert abs((dh * k).sum()) > 0.1


def test_ycbcr_anchors_and_roundtrip():
    white = W.rgb2ycbcr(np.ones((1, 1, 3)))[0, 0]
    black = W.rgb2ycbcr(np.zeros((1, 1, 3)))[0, 0]
    assert np.allclose(white, [235, 128, 128]) and np.allclose(black, [16, 128, 128])
    x = np.random.RandomState(3).rand(8, 9, 3)
    assert np.abs(W.ycbcr2rgb(W.rgb2ycbcr(x)) - x).max() < 1e-12


def test_levels_rule():
    assert W.auto_levels((600, 1000), "bior1.5") == 3 and W.auto_levels((2048, 2048), "bior1.5") == 4
    assert W.auto_levels((375, 500), "bior1.5") == 2 and W.auto_levels((600, 1000), "db1") == 6


def test_soft_threshold_and_bayes():
    d = np.array([-3.0, -0.5, 0.0, 0.2, 2.0])
    assert np.allclose(W.soft_threshold(d, 1.0), [-2, 0, 0, 0, 1])
    det = np.random.RandomState(4).normal(0, 2.0, 10000)
    t = W.bayes_thresh(det, 1.0)
    assert abs(t - 1.0 / np.sqrt((det ** 2).mean() - 1.0)) < 1e-12
    assert abs(W.sigma_est_dwt(np.random.RandomState(5).normal(0, 0.3, 200000)) - 0.3) < 5e-3


def test_denoise_wavelet_reduces_noise_and_keeps_range():
    rng = np.random.RandomState(6)
    y, x = np.mgrid[0:128, 0:192]
    clean = np.stack([0.5 + 0.4 * np.sin(x / 17.0 + c) * np.cos(y / 23.0) for c in range(3)], -1)
    noisy = np.clip(clean + rng.normal(0, 0.1, clean.shape), 0, 1)
    for wl, lv in (("bior1.5", None), ("db1", 3)):
        out = W.denoise_wavelet(noisy, wl, lv)
        assert out.min() >= 0 and out.max() <= 1
        assert ((out - clean) ** 2).mean() < 0.5 * ((noisy - clean) ** 2).mean()


def test_published_pywavelets_examples():
    """Known answers printed in the PyWavelets documentation (pywavelets.readthedocs.io: `pywt.dwt`, `pywt.idwt`,
    `pywt.threshold`, `pywt.dwt_max_level`, signal extension modes) and the bior1.5 coefficient listing of
    wavelets.pybytes.com - restated from those pages; they anchor sign, phase and ordering conventions that the
    perfect-reconstruction properties above cannot see."""
    # >>> pywt.dwt([1, 2, 3, 4, 5, 6], 'db1') -> cA [2.12132034 4.94974747 7.77817459], cD [-0.70710678 x 3]
    x = np.array([[1., 2., 3., 4., 5., 6.]])
    a, d = W.dwt_axis(x, "db1", 1)
    assert np.allclose(a[0], [2.12132034, 4.94974747, 7.77817459], atol=5e-9)
    assert np.allclose(d[0], [-0.70710678] * 3, atol=5e-9)
    # >>> pywt.idwt(cA, cD, 'db1') -> [1. 2. 3. 4. 5. 6.]
    assert np.allclose(W.idwt_axis(a, d, "db1", 1)[0], x[0], atol=1e-12)
    # symmetric mode "... x2 x1 | x1 x2 ... xn | xn xn-1 ...": output length floor((n + F - 1) / 2); db2-sized check via bior1.5
    y = np.arange(1.0, 8.0)[None]                      # n = 7, F = 10 -> 8 coefficients
    a5, d5 = W.dwt_axis(y, "bior1.5", 1)
    assert a5.shape == (1, 8) and d5.shape == (1, 8)
    # >>> pywt.threshold(np.linspace(1, 4, 7), 2, 'soft') -> [0. 0. 0. 0.5 1. 1.5 2.]
    assert np.allclose(W.soft_threshold(np.linspace(1, 4, 7), 2), [0, 0, 0, 0.5, 1, 1.5, 2])
    # >>> pywt.dwt_max_level(data_len=1000, filter_len=10) -> 6   (the sym5 example: dec_len 10, as bior1.5)
    assert W.dwt_max_level(1000, 10) == 6
    # bior1.5 decomposition low-pass as listed: 0.0165728..., -0.0165728..., -0.1215339..., 0.1215339..., 0.7071..., ...
    lo = W.WAVELETS["bior1.5"][0]
    assert np.allclose(lo, [0.01657281518405971, -0.01657281518405971, -0.12153397801643787, 0.12153397801643787,
                            0.7071067811865476, 0.7071067811865476, 0.12153397801643787, -0.12153397801643787,
                            -0.01657281518405971, 0.01657281518405971], atol=0)
    hi = W.WAVELETS["bior1.5"][1]
    assert hi[4] == -0.7071067811865476 and hi[5] == 0.7071067811865476 and not any(hi[:4]) and not any(hi[6:])
