"""GPU: distribution checks of the Philox generators (generator mode has no
bit-level twin in the reference, whose draws come from NumPy's Mersenne
Twister; BASELINE.json asks for moment-level verification)."""
import numpy as np
import pytest
import torch

from conftest import synth_batch

pytestmark = pytest.mark.gpu

N, H, W = 4, 256, 384            # 1.18 M samples per field


def field(cuda, spec, seed=3):
    from image_denoising_b200 import api
    plan = api.get_plan(N, H, W, spec, 0)
    f0, f1 = plan.noise_field(seed=seed)
    return f0.double().cpu().numpy().ravel(), (f1.double().cpu().numpy().ravel() if f1 is not None else None)


@pytest.mark.parametrize("var", [0.1, 1.0, 1.5])
def test_gaussian_moments(cuda, var):
    x, _ = field(cuda, "gaussian_var%s" % var)
    n = x.size
    s = var ** 0.5
    assert abs(x.mean()) < 5 * s / np.sqrt(n)
    assert abs(x.std() / s - 1) < 5e-3
    z = x / s
    assert abs((z ** 3).mean()) < 0.02                 # skewness
    assert abs((z ** 4).mean() - 3.0) < 0.05           # kurtosis
    for q in (1.0, 2.0, 3.0):                          # tail mass vs the normal CDF
        from scipy.stats import norm
        assert abs((np.abs(z) > q).mean() - 2 * norm.sf(q)) < 4 * np.sqrt(2 * norm.sf(q) / n) + 1e-4
    # neighbouring samples are uncorrelated (pairs come from one Box-Muller draw)
    assert abs(np.corrcoef(z[0::2], z[1::2])[0, 1]) < 5e-3


@pytest.mark.parametrize("amount", [0.2, 0.4, 0.8])
def test_sap_rates(cuda, amount):
    u1, u2 = field(cuda, "sap_var%s" % amount)
    n = u1.size
    flip = u1 < amount
    assert abs(flip.mean() - amount) < 4 * np.sqrt(amount * (1 - amount) / n) + 2e-5
    salt = (u2 < 0.5)[flip]
    assert abs(salt.mean() - 0.5) < 4 * np.sqrt(0.25 / salt.size)


@pytest.mark.parametrize("high", [0.2, 0.6, 1.2])
def test_uniform_moments(cuda, high):
    x, _ = field(cuda, "uniform_var%s" % high)
    assert x.min() >= 0 and x.max() <= high
    assert abs(x.mean() / (high / 2) - 1) < 3e-3
    assert abs(x.var() / (high ** 2 / 12) - 1) < 6e-3


@pytest.mark.parametrize("scale", [0.1, 0.2, 0.3])
def test_rayleigh_moments(cuda, scale):
    x, _ = field(cuda, "rayleigh_var%s" % scale)
    assert x.min() >= 0
    assert abs(x.mean() / (scale * np.sqrt(np.pi / 2)) - 1) < 3e-3
    assert abs(x.var() / ((4 - np.pi) / 2 * scale ** 2) - 1) < 1e-2


@pytest.mark.parametrize("scale", [0.05, 0.1, 0.2])
def test_gamma_moments(cuda, scale):
    from scipy.stats import gamma, kstest
    x, _ = field(cuda, "gamma_var%s" % scale)
    a = 1.99
    assert x.min() >= 0
    assert abs(x.mean() / (a * scale) - 1) < 5e-3
    assert abs(x.var() / (a * scale ** 2) - 1) < 2e-2
    sub = x[::97]
    assert kstest(sub, gamma(a, scale=scale).cdf).statistic < 0.02


def test_poisson_distribution(cuda):
    """Per input grey level, the generator's output matches NumPy's Poisson pushed
    through the same expression (mean within 4 sigma, variance within 5 %)."""
    from image_denoising_b200 import api
    from oracle import noise as O
    rng = np.random.RandomState(0)
    for levels in ([0, 10, 100, 200, 255], list(range(0, 256, 1))):
        img = rng.choice(levels, size=(2, 384, 512, 3)).astype(np.uint8)
        got = api.apply(torch.from_numpy(img).to(cuda), "poisson", seed=5).cpu().numpy()
        vals = O.poisson_vals(img[0])
        for v in levels[:: max(1, len(levels) // 16)]:
            g = got[img == v].astype(np.float64)
            lam = v / 255.0 * vals
            ref = O.to_u8(np.clip(np.random.RandomState(1).poisson(lam, 400000) / vals, 0, 1)).astype(np.float64)
            se = ref.std() / np.sqrt(g.size) + ref.std() / np.sqrt(ref.size) + 1e-9
            assert abs(g.mean() - ref.mean()) < 5 * se + 1e-6, (v, vals, g.mean(), ref.mean())
            if ref.var() > 0:
                assert abs(g.var() / ref.var() - 1) < 0.06, (v, vals, g.var(), ref.var())


def test_brownian_increments(cuda):
    """Generator-mode brownian: the u8 field is a wrapped running sum whose
    increments have variance dt * 255^2 (checked mod 256 on small dt)."""
    from image_denoising_b200 import api
    zero = torch.zeros((2, 128, 256, 3), dtype=torch.uint8, device=cuda)
    out = api.apply(zero, "brownian_var0.009", seed=9).cpu().numpy().astype(np.int64)   # 0 + field, saturating
    for i in range(2):
        f = out[i].ravel()
        d = np.diff(f)
        d = (d + 128) % 256 - 128                       # unwrap single-step differences
        assert f[0] == 0
        assert abs(d.std() / (np.sqrt(0.009) * 255) - 1) < 0.03


def _corr(a, b):
    a = a - a.mean()
    b = b - b.mean()
    return float((a * b).mean() / np.sqrt((a * a).mean() * (b * b).mean()))


def test_generator_independence(cuda):
    """What the round count of the counter-based generator (Philox4x32-7, csrc/b2d_common.cuh) must deliver here:
    draws that are uncorrelated along the element index (lags 1..16 cross the 8-element group = one Philox block,
    lag 3 W = the pixel below), between consecutive images (counter words c2 / c3), between consecutive seeds
    (key), and between the two fields of one draw; every one of the 16 bits of a draw balanced and the bits
    pairwise independent; the 16-bit draws uniform (chi-square over 4096 cells)."""
    from image_denoising_b200 import api
    plan = api.get_plan(N, H, W, "uniform_var1.2", 0)
    n_img = H * W * 3
    x = plan.noise_field(seed=3)[0].double().cpu().numpy().ravel()
    k = np.floor(x / 1.2 * 65536.0).astype(np.int64)            # the 16-bit draws: x = high * (k + 0.5) / 65536
    assert k.min() >= 0 and k.max() <= 65535
    lim = 5.0 / np.sqrt(x.size)
    for lag in list(range(1, 17)) + [3 * W, 3 * W + 1, 3 * W - 1, n_img]:      # n_img: same element, next image
        assert abs(_corr(x[:-lag], x[lag:])) < lim, lag
    y = plan.noise_field(seed=4)[0].double().cpu().numpy().ravel()              # next seed
    assert abs(_corr(x, y)) < lim
    z = plan.noise_field(seed=3, first_index=1)[0].double().cpu().numpy().ravel()   # shifted image indices: same stream, moved
    assert np.array_equal(z[: 3 * n_img], x[n_img:])
    bits = ((k[:, None] >> np.arange(16)[None, :]) & 1).astype(np.float64)
    assert np.abs(bits.mean(0) - 0.5).max() < 5 * 0.5 / np.sqrt(k.size)
    cc = np.corrcoef(bits.T)
    assert np.abs(cc - np.eye(16)).max() < lim
    cells = np.bincount(k >> 4, minlength=4096).astype(np.float64)
    chi2 = ((cells - k.size / 4096.0) ** 2 / (k.size / 4096.0)).sum()
    assert abs(chi2 - 4095) < 6 * np.sqrt(2 * 4095), chi2                        # chi-square(4095): mean 4095, sd 90.5
    # Gaussian field: both members of a Box-Muller pair, neighbouring pairs, squares (no variance clustering)
    g = api.get_plan(N, H, W, "gaussian_var1.0", 0).noise_field(seed=3)[0].double().cpu().numpy().ravel()
    for lag in (1, 2, 3, 7, 8, 9, 3 * W, n_img):
        assert abs(_corr(g[:-lag], g[lag:])) < lim, lag
        assert abs(_corr(g[:-lag] ** 2, g[lag:] ** 2)) < lim, lag
    # s&p: the flip draw and the salt bit of one element, and of neighbours
    u1, u2 = api.get_plan(N, H, W, "sap_var0.4", 0).noise_field(seed=3)
    u1, u2 = u1.double().cpu().numpy().ravel(), u2.double().cpu().numpy().ravel()
    f, s = (u1 < 0.4).astype(np.float64), (u2 < 0.5).astype(np.float64)
    assert abs(_corr(f, s)) < lim
    for lag in (1, 2, 7, 8, 3 * W):
        assert abs(_corr(f[:-lag], f[lag:])) < lim and abs(_corr(s[:-lag], s[lag:])) < lim and abs(_corr(f[:-lag], s[lag:])) < lim, lag
