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


# ---------------------------------------------------------------------------------------------------------
# A second, independently derived transform: dense analysis / synthesis matrices built straight from the
# filter definitions and the half-sample-symmetric extension rule, no np.pad / slicing shared with the oracle.
# Perfect reconstruction cannot see a shifted or mirrored convention that is self-consistent; equality with
# these matrices on odd and even sizes can (VERDICT r01, next #8a).
# ---------------------------------------------------------------------------------------------------------
def _fold(i, n):
    """index of sample i of '... x1 x0 | x0 x1 ... x(n-1) | x(n-1) x(n-2) ...' (pywt mode 'symmetric')"""
    while i < 0 or i >= n:
        i = -i - 1 if i < 0 else 2 * n - i - 1
    return i


def _analysis_matrix(filt, n):
    """pywt dwt, one filter: out[k] = sum_j filt[j] * ext(x)[2 k + 1 - j], k = 0 .. floor((n + F - 1) / 2) - 1
    (= np.convolve(extended, filt)[F - 1 + 1 :: 2] read off the definition of full convolution + downsampling)."""
    F = len(filt)
    m = (n + F - 1) // 2
    A = np.zeros((m, n))
    for k in range(m):
        for j in range(F):
            A[k, _fold(2 * k + 1 - j, n)] += filt[j]
    return A


def _synthesis_matrix(filt, m):
    """pywt idwt, one branch: zero-stuff (u[2 i] = c[i]), full convolution with the reconstruction filter, keep
    samples F - 2 .. 2 m - 1: out[t] = sum_i c[i] * filt[t + F - 2 - 2 i], t = 0 .. 2 m - F + 1."""
    F = len(filt)
    S = np.zeros((2 * m - F + 2, m))
    for t in range(2 * m - F + 2):
        for i in range(m):
            j = t + F - 2 - 2 * i
            if 0 <= j < F:
                S[t, i] = filt[j]
    return S


@pytest.mark.parametrize("wl", ["db1", "bior1.5"])
@pytest.mark.parametrize("n", [2, 3, 7, 8, 9, 16, 33, 61, 64, 99])
def test_axis_transform_equals_dense_matrices(wl, n):
    dl, dh, rl, rh = W.WAVELETS[wl]
    x = np.random.RandomState(n).rand(n, 3)
    a, d = W.dwt_axis(x, wl, 0)
    Al, Ah = _analysis_matrix(dl, n), _analysis_matrix(dh, n)
    assert np.abs(a - Al @ x).max() < 1e-13 and np.abs(d - Ah @ x).max() < 1e-13
    m = a.shape[0]
    Sl, Sh = _synthesis_matrix(rl, m), _synthesis_matrix(rh, m)
    r = W.idwt_axis(a, d, wl, 0)
    assert np.abs(r - (Sl @ a + Sh @ d)).max() < 1e-13
    # the pair is a perfect-reconstruction filter bank on the first n samples (odd n: one extra sample)
    assert np.abs((Sl @ Al + Sh @ Ah)[:n] - np.eye(n)).max() < 1e-12


@pytest.mark.parametrize("wl", ["db1", "bior1.5"])
@pytest.mark.parametrize("shape", [(8, 8), (9, 14), (33, 20), (61, 99)])
@pytest.mark.parametrize("level", [1, 2, 3, 4])
def test_wavedec2_waverec2_equal_dense_matrices(wl, shape, level):
    """2-D multilevel: axis 0 then axis 1 per level, key letters = (axis 0, axis 1), approximation trimmed by one
    sample where the next detail band is shorter - all through the dense matrices."""
    dl, dh, rl, rh = W.WAVELETS[wl]
    x = np.random.RandomState(level).rand(*shape)
    c = W.wavedec2(x, wl, level)
    a, dets = x, []
    for _ in range(level):
        A0l, A0h = _analysis_matrix(dl, a.shape[0]), _analysis_matrix(dh, a.shape[0])
        A1l, A1h = _analysis_matrix(dl, a.shape[1]), _analysis_matrix(dh, a.shape[1])
        lo0, hi0 = A0l @ a, A0h @ a
        dets.append({"ad": lo0 @ A1h.T, "da": hi0 @ A1l.T, "dd": hi0 @ A1h.T})
        a = lo0 @ A1l.T
    assert np.abs(c[0] - a).max() < 1e-12
    for got, want in zip(c[1:], dets[::-1]):
        for k in ("ad", "da", "dd"):
            assert got[k].shape == want[k].shape and np.abs(got[k] - want[k]).max() < 1e-12, k
    rec = a
    for det in dets[::-1]:
        rec = rec[: det["dd"].shape[0], : det["dd"].shape[1]]
        S0l, S0h = _synthesis_matrix(rl, rec.shape[0]), _synthesis_matrix(rh, rec.shape[0])
        S1l, S1h = _synthesis_matrix(rl, rec.shape[1]), _synthesis_matrix(rh, rec.shape[1])
        lo0 = rec @ S1l.T + det["ad"] @ S1h.T
        hi0 = det["da"] @ S1l.T + det["dd"] @ S1h.T
        rec = S0l @ lo0 + S0h @ hi0
    assert np.abs(W.waverec2(c, wl) - rec).max() < 1e-12
    assert np.abs(rec[: shape[0], : shape[1]] - x).max() < 1e-11


def test_more_published_pywavelets_examples():
    """More known answers from the PyWavelets documentation pages (restated from the pages; pywt itself cannot be
    installed here): `wavedec` / `waverec`, `dwt2` / `wavedec2` on constant input, `dwt_coeff_len`, `threshold`
    with a substitute, and the full bior1.5 filter bank as `pywt.Wavelet('bior1.5').filter_bank` prints it."""
    s = 2 ** 0.5
    # >>> pywt.wavedec([1, 2, 3, 4, 5, 6, 7, 8], 'db1', level=2)
    # cA2 [ 5. 13.]  cD2 [-2. -2.]  cD1 [-0.70710678 -0.70710678 -0.70710678 -0.70710678]
    x = np.arange(1.0, 9.0)[None]
    a1, d1 = W.dwt_axis(x, "db1", 1)
    a2, d2 = W.dwt_axis(a1, "db1", 1)
    assert np.allclose(a2[0], [5, 13]) and np.allclose(d2[0], [-2, -2]) and np.allclose(d1[0], [-1 / s] * 4)
    # >>> pywt.waverec(coeffs, 'db1') -> [1. 2. 3. 4. 5. 6. 7. 8.]
    assert np.allclose(W.idwt_axis(W.idwt_axis(a2, d2, "db1", 1), d1, "db1", 1)[0], x[0])
    # >>> coeffs = pywt.wavedec2(np.ones((4, 4)), 'db1'); len(coeffs) - 1 -> 2; pywt.waverec2(coeffs, 'db1') -> ones
    c = W.wavedec2(np.ones((4, 4)), "db1", 2)
    assert c[0].shape == (1, 1) and np.allclose(c[0], 4.0)
    assert all(np.allclose(v, 0) for det in c[1:] for v in det.values())
    assert np.allclose(W.waverec2(c, "db1"), np.ones((4, 4)))
    # >>> cA, (cH, cV, cD) = pywt.dwt2([[1, 2], [3, 4]], 'db1') -> 5., -2 (horizontal detail = 'da'), -1 ('ad'), 0
    c = W.wavedec2(np.array([[1.0, 2.0], [3.0, 4.0]]), "db1", 1)
    assert np.allclose(c[0], 5) and np.allclose(c[1]["da"], -2) and np.allclose(c[1]["ad"], -1) and np.allclose(c[1]["dd"], 0)
    # >>> pywt.dwt_coeff_len(data_len=9, filter_len=4?, mode='symmetric') = floor((n + F - 1) / 2); the page's own
    # rule for every mode but periodization
    for n, F, want in ((9, 2, 5), (9, 10, 9), (8, 10, 8), (600, 10, 304), (1000, 10, 504), (2048, 2, 1024)):
        assert (n + F - 1) // 2 == want
    assert W.dwt_axis(np.zeros(600), "bior1.5", 0)[0].shape[0] == 304
    # >>> pywt.threshold(data, 2, 'soft') with data = np.linspace(1, 4, 7), and the sign convention on negatives
    assert np.allclose(W.soft_threshold(-np.linspace(1, 4, 7), 2), [0, 0, 0, -0.5, -1, -1.5, -2])
    # >>> pywt.Wavelet('bior1.5').filter_bank -> (dec_lo, dec_hi, rec_lo, rec_hi)
    c1, c2, h = 0.01657281518405971, 0.12153397801643787, 0.7071067811865476
    dl, dh, rl, rh = W.WAVELETS["bior1.5"]
    assert list(dl) == [c1, -c1, -c2, c2, h, h, c2, -c2, -c1, c1]
    assert list(dh) == [0, 0, 0, 0, -h, h, 0, 0, 0, 0]
    assert list(rl) == [0, 0, 0, 0, h, h, 0, 0, 0, 0]
    assert list(rh) == [c1, c1, -c2, -c2, h, -h, c2, c2, -c1, -c1]
    # the quadrature-mirror relations pywt builds a biorthogonal bank from: rec_lo[k] = (-1)^k dec_hi[k] up to
    # sign convention, rec_hi[k] = (-1)^(k+1) dec_lo[k]
    k = np.arange(10)
    assert np.allclose(np.array(rh), (-1.0) ** k * np.array(dl)) and np.allclose(np.array(rl), (-1.0) ** (k + 1) * np.array(dh))
    # db1: >>> pywt.Wavelet('db1').filter_bank
    assert [list(f) for f in W.WAVELETS["db1"]] == [[h, h], [-h, h], [h, h], [h, -h]]


def test_sigma_census_brackets_the_oracle():
    """oracle.wavelet.sigma_census (used by the GPU stage-wise check): the float64 oracle's own sigma is one of
    the admissible values; with continuous noise in front no coefficient is ambiguous and the range is a point."""
    import os
    rng = np.random.RandomState(8)
    y, x = np.mgrid[0:64, 0:96]
    smooth = np.clip(127.5 + 60 * np.cos(x / 9.0)[..., None] + 40 * np.sin(y / 7.0)[..., None] + rng.normal(0, 8, (64, 96, 3)), 0, 255).astype(np.uint8)
    binary = (rng.rand(61, 99, 3) < 0.5).astype(np.uint8) * 255
    inputs = [smooth, binary]
    nat = os.path.join(os.path.dirname(__file__), "golden", "natural_120x160.npz")
    if os.path.exists(nat):
        inputs.append(np.load(nat)["img"])
    for img in inputs:
        for wl in ("bior1.5", "db1"):
            stats = W.denoise_wavelet(img, wl, None, True)[1]
            for (smin, smax, n_sure, n_amb), st in zip(W.sigma_census(img, wl), stats):
                assert smin * (1 - 1e-12) <= st["sigma"] <= smax * (1 + 1e-12), (wl, smin, st["sigma"], smax, n_amb)
    noisy = np.clip(smooth / 255.0 + rng.normal(0, 0.1, smooth.shape), 0, 1)
    for smin, smax, n_sure, n_amb in W.sigma_census(noisy, "bior1.5"):
        assert n_amb == 0 and smin == smax
    # an injected sigma is what the thresholds are built from
    out_a = W.denoise_wavelet(noisy, "db1", 3)
    sig = [st["sigma"] for st in W.denoise_wavelet(noisy, "db1", 3, True)[1]]
    assert np.array_equal(out_a, W.denoise_wavelet(noisy, "db1", 3, sigmas=sig))
    assert not np.array_equal(out_a, W.denoise_wavelet(noisy, "db1", 3, sigmas=[2 * s for s in sig]))
