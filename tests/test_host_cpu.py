"""CPU: host logic - the spec grammar, the C parser, the library's export list."""
import ctypes
import os
import random
import re

import numpy as np
import pytest

from image_denoising_b200 import _lib, spec
from oracle import closures as C

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NOISES = ["gaussian", "poisson", "sap", "speckle", "periodic", "brownian", "quant", "uniform", "gamma", "rayleigh",
          "bloom", "shader", "original"]
DENS = ["", "_wavelet", "_gaus_blur", "_mean", "_median", "_bilateral"]


def all_specs():
    for n in NOISES:
        for d in DENS:
            for tag, _ in spec.LEVELS.get(n, [("", 0)]):
                yield n + d + ("_var" + tag if tag else "")


def test_library_loads_and_exports_every_declared_symbol():
    L = _lib.lib()
    header = open(os.path.join(ROOT, "include", "b200denoise.h")).read()
    declared = set(re.findall(r"\b(b2d_[a-z_]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert b"sm_100a" in L.b2d_version()
    assert ctypes.sizeof(_lib.B2DSpec) == 48


@pytest.mark.parametrize("s", list(all_specs()))
def test_python_c_and_reference_ladder_agree(s):
    L = _lib.lib()
    cs = _lib.B2DSpec()
    assert L.b2d_parse_spec(s.encode(), ctypes.byref(cs)) == 0
    ps = spec.parse(s)
    pc = ps.to_c()
    assert (cs.noise, cs.denoise, cs.level, cs.ksize, cs.d, cs.wavelet, cs.levels) == \
           (pc.noise, pc.denoise, pc.level, pc.ksize, pc.d, pc.wavelet, pc.levels)
    noise, den, level = C.ladder(s)
    assert ps.noise == (noise or "none")
    if noise is not None:
        assert ps.denoise == (den or "none")
        if level is not None and level != "size" and noise != "periodic":
            assert ps.level == float(level)


def test_unknown_level_and_unsupported_specs():
    L = _lib.lib()
    cs = _lib.B2DSpec()
    assert L.b2d_parse_spec(b"gaussian_median_var9", ctypes.byref(cs)) == _lib.ERR_PARSE
    assert b"no known level" in L.b2d_last_error()
    assert L.b2d_parse_spec(b"noise_mix_var_low", ctypes.byref(cs)) == _lib.ERR_UNSUPPORTED
    assert L.b2d_parse_spec(b"curvelet", ctypes.byref(cs)) == _lib.ERR_UNSUPPORTED
    with pytest.raises(spec.SpecError):
        spec.parse("sap_mean_var0.5")
    with pytest.raises(spec.SpecError):
        spec.parse("noise_mix_var_low")


def test_substring_precedence_quirks():
    # 'var0.09' must not be read as 'var0.9'; 'var10' not as 'var1.0'; first noise name wins
    assert spec.parse("brownian_var0.09").level == 0.09 and spec.parse("brownian_var0.009").level == 0.009
    assert spec.parse("quant_var10").level == 10 and spec.parse("quant_median_var3").level == 3
    assert spec.parse("periodic_varsize").level == _lib.LEVEL_SIZE
    assert spec.parse("gaussian_poisson_var0.1").noise == "gaussian"
    assert spec.parse("speckle_gaussian_var1.0").noise == "gaussian"       # chain order, not string order
    assert spec.parse("sap_median_var0.4").denoise == "median"
    assert spec.parse("sap_var0.4_median").denoise == "none"               # needs the literal 'sap_median'
    assert spec.parse("original_median").denoise == "median"               # post-denoise pass on a clean image
    assert spec.parse("gaussian_median_var1.0", ksize=7).ksize == 7


def test_mix_resolution_uses_the_reference_lists():
    for side in ("test", "train"):
        lists = spec.MIX_LISTS[side]
        assert len(lists["var_low"]) == len(lists["var_medium"]) == len(lists["var_high"]) == 13
        for key, lst in lists.items():
            for s in lst:
                spec.parse(s)                                              # every entry is a valid concrete spec
            got = {spec.resolve_mix("noise_mix_" + key, side, random.Random(i)) for i in range(400)}
            assert got == set(lst)
    assert len(spec.MIX_LISTS["test"]["var_all"]) == 42 and len(spec.MIX_LISTS["train"]["var_all"]) == 38
    assert spec.MIX_LISTS["test"]["var_low"][9] == "quant_var3" and spec.MIX_LISTS["train"]["var_low"][9] == "quant_var10"
    assert spec.resolve_mix("sap_mean_var0.4") == "sap_mean_var0.4"


def test_mix_with_a_denoiser_keeps_the_denoiser():
    """`noise_mix_var_*_<denoise>`: the reference picks a noise-only closure and denoises in its post-denoise pass
    (minibatch.py:1636-1663; test.py:1787-1819 keeps only `wavelet` live, :1805-1812).  resolve() returns both parts."""
    r = random.Random(5)
    for den in ("gaus_blur", "mean", "median", "wavelet", "bilateral"):
        concrete, post = spec.resolve("noise_mix_var_low_" + den, "train", r)
        assert concrete in spec.MIX_LISTS["train"]["var_low"] and post == den
        assert spec.parse(concrete).denoise == "none"
        concrete, post = spec.resolve("noise_mix_var_low_" + den, "test", r)
        assert concrete in spec.MIX_LISTS["test"]["var_low"] and post == ("wavelet" if den == "wavelet" else None)
    assert spec.resolve("noise_mix_var_high", "train", r)[1] is None
    # elif order of the post pass: gaus_blur before mean before median before wavelet before bilateral
    assert spec.post_denoiser("noise_mix_var_low_wavelet_mean", "train") == "mean"
    assert spec.post_denoiser("noise_mix_var_low_wavelet_mean", "test") is None        # `mean` branch only prints
    # a concrete spec is one closure call: nothing is added
    assert spec.resolve("sap_mean_var0.4", "train") == ("sap_mean_var0.4", None)


def test_q3_gaussian_random_level_is_opt_in():
    """Quirk Q3 (minibatch.py:1575-1580, test.py:1678-1682): `gaussian` specs ignore their own level and denoiser."""
    assert spec.resolve("gaussian_median_var1.5", "train", random.Random(0)) == ("gaussian_median_var1.5", None)
    got = set()
    for i in range(200):
        concrete, post = spec.resolve("gaussian_median_var1.5", "train", random.Random(i), spec.COMPAT_GAUSSIAN_RANDOM_LEVEL)
        assert post == "median" and spec.parse(concrete).denoise == "none"
        got.add(concrete)
    assert got == {"gaussian_var0.1", "gaussian_var1.0", "gaussian_var1.5"}
    assert spec.resolve("gaussian_median_var1.5", "test", random.Random(0), spec.COMPAT_GAUSSIAN_RANDOM_LEVEL)[1] is None
    assert spec.resolve("sap_var0.4", "test", None, spec.COMPAT_GAUSSIAN_RANDOM_LEVEL) == ("sap_var0.4", None)
    assert _lib.COMPAT_GAUSSIAN_RANDOM_LEVEL == 16 and _lib.COMPAT_SAP_TRAIN06 == 2


def test_plan_create_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = _lib.lib()
    cs = spec.parse("sap_mean_var0.4").to_c()
    h = ctypes.c_void_p()
    assert L.b2d_plan_create(ctypes.byref(h), 0, 1, 8, 8, ctypes.byref(cs), 1) == _lib.ERR_NO_DEVICE
    assert b"no CPU fallback" in L.b2d_last_error()
    from image_denoising_b200 import api
    with pytest.raises(RuntimeError):
        api.Plan(1, 8, 8, "sap_mean_var0.4")


@pytest.mark.parametrize("log2_vals", [8, 5, 0])
def test_poisson_alias_table_represents_the_pmf(log2_vals):
    """b2d_poisson_table (host code of the library, no GPU): the alias table of input value v must
    reproduce the pmf of (255 * clip(Poisson(v / 255 * vals) / vals, 0, 1)).astype(uint8)
    (skimage random_noise 'poisson' + the reference's ->u8, minibatch.py:211-213) to 2^-16 per
    outcome plus the mass folded away by the rounding."""
    import ctypes
    from scipy.stats import poisson
    L = _lib.lib()
    tab = np.zeros(65536, np.uint16)
    _lib.check(L.b2d_poisson_table(log2_vals, tab.ctypes.data_as(ctypes.c_void_p)))
    tab = tab.reshape(256, 256)
    alias, thr = (tab & 0xff).astype(np.int64), (tab >> 8).astype(np.int64)
    vals = 1 << log2_vals
    k = np.arange(vals + 1)
    out_of_k = (255 * np.minimum(k, vals)) >> log2_vals
    for v in (0, 1, 7, 64, 128, 200, 255):
        lam = v / 255.0 * vals
        pk = poisson.pmf(k[:-1], lam) if lam > 0 else np.eye(1, vals, 0)[0]
        want = np.bincount(out_of_k[:-1], weights=pk, minlength=256)
        want[255] += max(0.0, 1.0 - pk.sum())
        # mass of outcome b = sum over slots of [own share] + [alias share], in units of 2^-16
        own = np.where(alias[v] == np.arange(256), 256, thr[v])
        got = np.bincount(np.arange(256), weights=own, minlength=256) + \
            np.bincount(alias[v], weights=256 - own, minlength=256)
        assert got.sum() == 65536
        assert np.abs(got / 65536.0 - want).max() <= 1.0 / 65536 + 1e-12, (log2_vals, v)
        mean_err = abs((got / 65536.0 * np.arange(256)).sum() - (want * np.arange(256)).sum())
        assert mean_err < 2e-3, (log2_vals, v, mean_err)
    with pytest.raises(_lib.B2DError):
        _lib.check(L.b2d_poisson_table(9, tab.ctypes.data_as(ctypes.c_void_p)))


@pytest.mark.parametrize("scale", [0.05, 0.1, 0.2, 0.7])
def test_gamma_alias_table_represents_the_folded_pmf(scale):
    """b2d_gamma_table (host code of the library): the alias table must reproduce the pmf of
    floor(255 * gamma.rvs(1.99, scale=S)) mod 256 (minibatch.py:1216-1217,1247: unclipped cv2.add,
    wrapping ->u8) to 2^-16 per outcome, and the fold table the conditional cdf of K // 256."""
    import ctypes
    from scipy.stats import gamma
    L = _lib.lib()
    alias_tab = np.zeros(256, np.uint16)
    fold = np.zeros(256 * 64, np.uint32)
    folds = ctypes.c_int(0)
    _lib.check(L.b2d_gamma_table(scale, alias_tab.ctypes.data_as(ctypes.c_void_p), fold.ctypes.data_as(ctypes.c_void_p),
                                 ctypes.byref(folds)))
    nf = folds.value
    assert 1 <= nf <= 64 and 256 * nf >= min(255 * scale * 40, 256 * 64 - 1)
    edges = np.arange(256 * nf + 1) / (255.0 * scale)
    cdf = gamma(1.99, scale=1.0).cdf(edges)
    cdf[-1] = 1.0
    pk = np.diff(cdf).reshape(nf, 256)
    want = pk.sum(axis=0)
    alias, thr = (alias_tab & 0xff).astype(np.int64), (alias_tab >> 8).astype(np.int64)
    own = np.where(alias == np.arange(256), 256, thr)
    got = np.bincount(np.arange(256), weights=own, minlength=256) + np.bincount(alias, weights=256 - own, minlength=256)
    assert got.sum() == 65536
    assert np.abs(got / 65536.0 - want).max() <= 1.0 / 65536 + 1e-10, scale
    f = fold[:256 * nf].reshape(256, nf).astype(np.float64) / 65536.0
    cond = np.cumsum(pk, axis=0).T / np.maximum(want, 1e-300)[:, None]
    assert np.abs(f - cond).max() <= 0.5 / 65536 + 1e-9
    assert np.all(f[:, -1] == 1.0)
    with pytest.raises(_lib.B2DError):
        _lib.check(L.b2d_gamma_table(0.0, alias_tab.ctypes.data_as(ctypes.c_void_p), None, None))


def test_blob_shape_matches_host_mirror():
    """b2d_blob_shape (host code of the library) == the scale / dsize of the reference's
    prep_im_for_blob through cv2.resize (lib/utils/blob.py:37-45), over many image sizes."""
    from image_denoising_b200 import blob
    rng = np.random.RandomState(4)
    shapes = [(600, 1000), (375, 500), (333, 500), (500, 375), (480, 640), (1080, 1920), (2048, 2048), (301, 333),
              (64, 96), (1000, 200)] + [(int(a), int(b)) for a, b in rng.randint(40, 1400, size=(60, 2))]
    means = np.array([[[102.9801, 115.9465, 122.7717]]])
    for h, w in shapes:
        for target, mx in ((600, 1000), (800, 1333), (224, 224)):
            im, s_ref = blob.prep_im_for_blob(np.zeros((h, w, 3), np.uint8), means, target, mx)
            s, oh, ow = blob.blob_shape(h, w, target, mx)
            assert s == s_ref and (oh, ow) == im.shape[:2], (h, w, target, mx, s, s_ref, (oh, ow), im.shape)
