"""CPU: the oracle against the reference's own libraries and known anchors."""
import random

import cv2
import numpy as np
import pytest

from conftest import synth_image
from oracle import closures as C, denoise as D, noise as N


@pytest.mark.parametrize("shape", [(64, 96), (37, 53), (5, 1), (1, 7)])
def test_filter_formulas_equal_cv2(shape):
    for kind in ("uniform", "smooth"):
        u8 = synth_image(shape[0], shape[1], 3, kind)
        assert np.array_equal(D.gaus_blur_cv2(u8), D.gaus_blur_formula(u8))
        assert np.array_equal(D.mean_cv2(u8), D.mean_formula(u8))
        for k in (3, 5, 7):
            if min(shape) >= 1:
                assert np.array_equal(D.median_cv2(u8, k), D.median_formula(u8, k))


def test_bilateral_formula_equals_cv2():
    for kind in ("uniform", "smooth"):
        u8 = synth_image(60, 90, 4, kind)
        d = np.abs(D.bilateral_cv2(u8).astype(int) - D.bilateral_formula(u8))
        assert d.max() <= 1 and (d > 0).mean() < 1e-3


def test_to_u8_wraps_like_numpy():
    x = np.array([0.0, 0.5, 1.0, 1.2, -0.00588, 3.9, 1e5, -1e5, 254.999 / 255])
    assert np.array_equal(N.to_u8(x), N.to_u8_explicit(x))
    assert list(N.to_u8(np.array([1.2, -1.5 / 255, 1000 / 255.0]))) == [50, 255, 232]
    v = np.arange(256, dtype=np.uint8)
    assert np.array_equal(N.to_u8(N.img_as_float(v)), v)          # ->u8 of an untouched pixel is exact


def test_random_noise_restatement_draw_order():
    """fields=None draws exactly what the reference's calls draw from np.random."""
    img = synth_image(40, 50, 1)
    np.random.seed(3)
    a = N.gaussian(img, 1.0)
    np.random.seed(3)
    n = np.random.normal(0.0, 1.0, img.shape)
    assert np.array_equal(a, np.clip(img / 255.0 + n, 0, 1))
    np.random.seed(3)
    s = N.sap(img, 0.4)
    np.random.seed(3)
    u1, u2 = np.random.random_sample(img.shape), np.random.random_sample(img.shape)
    assert np.array_equal(s, N.sap(img, 0.4, fields={"u1": u1, "u2": u2}))     # choice(p) == (u < p)
    np.random.seed(3)
    p = N.poisson(img)
    np.random.seed(3)
    k = np.random.poisson(img / 255.0 * N.poisson_vals(img))
    assert np.array_equal(p, N.poisson(img, fields={"k": k}))
    assert N.poisson_vals(np.array([0, 10, 100, 200, 255], np.uint8)) == 8.0


def test_periodic_and_brownian_fields():
    f = N.periodic_field((20, 30, 3), "pi")
    assert f.shape == (20, 30, 3) and f.dtype == np.uint8
    t = np.linspace(-np.pi, np.pi, 1800)
    assert f.ravel()[0] == np.uint8(int(np.sin(t[0]) * 255) & 0xFF)
    assert f.ravel()[1350] == int(np.sin(t[1350]) * 255)                      # positive half: plain trunc
    z = np.random.RandomState(0).normal(size=1799)
    b = N.brownian_field((20, 30, 3), 0.09, fields={"z": z}).ravel()
    ref = (np.concatenate(([0.0], np.cumsum(0.3 * z))) * 255).astype(np.int64) & 0xFF
    assert b[0] == 0 and np.array_equal(b, ref.astype(np.uint8))


def test_shader_is_min_255_3v_on_rgb():
    img = synth_image(16, 24, 2)
    assert np.array_equal(N.shader(img), np.minimum(255, 3 * img[:, :, ::-1].astype(int)).astype(np.uint8))


def test_quant_assignment_matches_sklearn_predict():
    from sklearn.cluster import MiniBatchKMeans
    img = synth_image(48, 64, 5, "smooth")
    lab = cv2.cvtColor(img, cv2.COLOR_BGR2LAB).reshape(-1, 3)
    clt = MiniBatchKMeans(n_clusters=7, random_state=0, n_init=3)
    labels = clt.fit_predict(lab)
    want = cv2.cvtColor(clt.cluster_centers_.astype("uint8")[labels].reshape(48, 64, 3), cv2.COLOR_LAB2BGR)
    got = N.quant(img, 7, fields={"centers": clt.cluster_centers_})
    assert (got != want).mean() < 1e-3        # ties between float32 (sklearn) and float64 distances only


def test_bloom_replay_matches_automold_draw_order():
    """oracle.bloom(draws) with draws taken from a seeded `random` equals what
    Automold.add_sun_flare computes with the same seed (golden fixture made by
    tests/golden/make_golden.py from /root/reference/tools/Automold.py)."""
    import os
    path = os.path.join(os.path.dirname(__file__), "golden", "bloom_seed7_320x480.npz")
    if not os.path.exists(path):
        pytest.skip("golden fixture missing")
    g = np.load(path)
    img = synth_image(320, 480, 7, "smooth")
    draws = N.bloom_draws(img.shape, random.Random(7))
    assert np.array_equal(N.bloom(img, draws), g["out"])


def test_ladder_precedence():
    assert C.ladder("gaussian_median_var1.0") == ("gaussian", "median", "1.0")
    assert C.ladder("sap_mean_var0.4") == ("sap", "mean", "0.4")
    assert C.ladder("brownian_wavelet_var0.009") == ("brownian", "wavelet", "0.009")
    assert C.ladder("periodic_gaus_blur_varsize") == ("periodic", "gaus_blur", "size")
    assert C.ladder("original") == (None, None, None)
    with pytest.raises(UnboundLocalError):
        C.corrupt(synth_image(8, 8), "gaussian_var9")
