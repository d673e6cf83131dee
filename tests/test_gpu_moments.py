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
