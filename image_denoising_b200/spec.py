"""The ``{noise}_{denoise}_var{level}`` spec grammar of the reference.

The reference has no parser: the ``--noise`` string (tools/test_net.py:47-49,
tools/trainval_net.py:52-54) is probed with ``'substr' in spec`` chains
(lib/roi_data_layer/minibatch.py:1518-1673, lib/model/test.py:1612-1831 and,
inside every closure, e.g. minibatch.py:89-199).  ``parse`` reproduces that
precedence exactly - same order of tests, first match wins - and returns a
``Spec``; ``resolve_mix`` reproduces the ``mix`` branches (a ``random.choice``
over a hard-coded list).  Keyword overrides (ksize, wavelet, levels, ...)
extend the grammar for the BASELINE configs that need them (median 7x7,
db1 / 3 levels).
"""
import random as _random
from dataclasses import dataclass, replace

from . import _lib

# order of the top-level elif chain, minibatch.py:1575-1634
NOISE_ORDER = ["gaussian", "poisson", "sap", "speckle", "periodic", "brownian", "quant", "uniform", "gamma",
               "rayleigh", "bloom", "shader"]
# order of `<noise>_<denoise>` tests inside each closure
DENOISE_ORDER = ["wavelet", "gaus_blur", "mean", "median", "bilateral"]
# order of the post-denoise pass (minibatch.py:1636-1663), used for noise-free specs
POST_DENOISE_ORDER = ["gaus_blur", "mean", "median", "wavelet", "bilateral"]
# `'var<L>' in spec` tests per noise, in source order
LEVELS = {
    "gaussian": [("0.1", 0.1), ("1.0", 1.0), ("1.5", 1.5)],
    "sap": [("0.2", 0.2), ("0.4", 0.4), ("0.8", 0.8)],
    "speckle": [("0.5", 0.5), ("1.0", 1.0), ("2.0", 2.0)],
    "periodic": [("3.14", 3.141592653589793), ("100", 100.0), ("size", _lib.LEVEL_SIZE)],
    "brownian": [("0.9", 0.9), ("0.09", 0.09), ("0.009", 0.009)],
    "quant": [("3", 3.0), ("7", 7.0), ("10", 10.0)],
    "uniform": [("0.2", 0.2), ("0.6", 0.6), ("1.2", 1.2)],
    "gamma": [("0.05", 0.05), ("0.1", 0.1), ("0.2", 0.2)],
    "rayleigh": [("0.1", 0.1), ("0.2", 0.2), ("0.3", 0.3)],
}

# `mix` lists: lib/model/test.py:1613-1639 ("test") and minibatch.py:1519-1547 ("train")
_LOW = ['gaussian_var0.1', 'poisson', 'speckle_var0.5', 'sap_var0.2', 'uniform_var0.2', 'gamma_var0.05',
        'rayleigh_var0.1', 'periodic_var3.14', 'brownian_var0.9']
_MED = ['gaussian_var1.0', 'poisson', 'speckle_var1.0', 'sap_var0.4', 'uniform_var0.6', 'gamma_var0.1',
        'rayleigh_var0.2', 'periodic_var100', 'brownian_var0.09']
_HIGH = ['gaussian_var1.5', 'poisson', 'speckle_var2.0', 'sap_var0.8', 'uniform_var1.2', 'gamma_var0.2',
         'rayleigh_var0.3', 'periodic_varsize', 'brownian_var0.009']
MIX_LISTS = {
    "test": {
        "var_low": _LOW + ['quant_var3', 'original', 'bloom', 'shader'],
        "var_medium": _MED + ['quant_var7', 'original', 'shader', 'bloom'],
        "var_high": _HIGH + ['quant_var10', 'original', 'shader', 'bloom'],
        "var_all": ['gaussian_var0.1', 'poisson', 'speckle_var0.5', 'sap_var0.2', 'uniform_var0.2',
                    'gamma_var0.05', 'gamma_var0.05', 'rayleigh_var0.2', 'rayleigh_var0.1', 'periodic_var3.14',
                    'brownian_var0.9', 'quant_var3', 'gamma_var0.1', 'rayleigh_var0.1', 'gaussian_var1.0',
                    'poisson', 'speckle_var1.0', 'sap_var0.4', 'uniform_var0.6', 'gamma_var0.1', 'shader',
                    'original', 'shader', 'bloom', 'rayleigh_var0.2', 'periodic_var100', 'brownian_var0.09',
                    'quant_var7', 'gaussian_var1.5', 'poisson', 'speckle_var2.0', 'sap_var0.8', 'uniform_var1.2',
                    'gamma_var0.2', 'shader', 'original', 'rayleigh_var0.3', 'periodic_varsize',
                    'brownian_var0.009', 'quant_var10', 'original', 'shader'],
    },
    "train": {
        "var_low": _LOW + ['quant_var10', 'original', 'bloom', 'shader'],
        "var_medium": _MED + ['quant_var7', 'original', 'bloom', 'shader'],
        "var_high": _HIGH + ['quant_var3', 'original', 'bloom', 'shader'],
        "var_all": _LOW + ['quant_var3', 'shader', 'bloom'] + _MED[:6] + ['original', 'shader', 'bloom'] + _MED[6:]
                   + ['quant_var7'] + _HIGH + ['quant_var10', 'original', 'shader', 'bloom'],
    },
}


class SpecError(ValueError):
    """Unknown level / unusable spec (the reference raises UnboundLocalError in
    test.py, or silently yields ``im = []`` in minibatch.py:88)."""


@dataclass(frozen=True)
class Spec:
    noise: str = "none"
    denoise: str = "none"
    level: float = 0.0
    ksize: int = 3
    d: int = 9
    sigma_color: float = 20.0
    sigma_space: float = 100.0
    wavelet: str = "bior1.5"      # live files; backups / BASELINE config 4 use db1 with levels=3
    levels: int = 0               # 0 = skimage's automatic choice
    compat_flags: int = 0

    def to_c(self):
        return _lib.B2DSpec(_lib.NOISE[self.noise], _lib.DENOISE[self.denoise], float(self.level), int(self.ksize),
                            int(self.d), float(self.sigma_color), float(self.sigma_space),
                            _lib.WAVELET[self.wavelet], int(self.levels), int(self.compat_flags), 0)


def parse(spec, **overrides):
    """Parse a concrete (non-mix) spec string with the reference's precedence."""
    if isinstance(spec, Spec):
        return replace(spec, **overrides) if overrides else spec
    if "mix" in spec:
        raise SpecError("'mix' specs must go through resolve_mix() first: %r" % spec)
    if "curvelet" in spec:
        raise SpecError("curvelet denoising is out of scope (SURVEY.md 8(f) f3)")
    noise = next((n for n in NOISE_ORDER if n in spec), None)
    if noise is None:
        den = next((d for d in POST_DENOISE_ORDER if d in spec), "none")
        return Spec("none", den, **overrides)
    if noise in ("bloom", "shader"):
        return Spec(noise, "none", **overrides)
    den = next((d for d in DENOISE_ORDER if (noise + "_" + d) in spec), "none")
    level = 0.0
    if noise in LEVELS:
        for tag, val in LEVELS[noise]:
            if ("var" + tag) in spec:
                level = val
                break
        else:
            raise SpecError("spec %r: no known level for %s (the reference leaves `im` unbound)" % (spec, noise))
    return Spec(noise, den, level, **overrides)


COMPAT_SHADER_BGR = _lib.COMPAT_SHADER_BGR
COMPAT_SAP_TRAIN06 = _lib.COMPAT_SAP_TRAIN06
COMPAT_GAUSSIAN_RANDOM_LEVEL = _lib.COMPAT_GAUSSIAN_RANDOM_LEVEL     # Q3, host front-end flag
_Q3_LIST = ['gaussian_var0.1', 'gaussian_var1.0', 'gaussian_var1.5']   # minibatch.py:1575-1576, test.py:1678-1679


def resolve_mix(spec, side="test", rng=None):
    """The `mix` branch: pick a noise-only spec (test.py:1612-1639,
    minibatch.py:1518-1547).  ``rng``: a ``random.Random`` (the reference uses
    the never-seeded global ``random``).  Non-mix specs pass through.  A
    denoiser named in a mix spec is NOT part of the returned string: the
    reference applies it in its post-denoise pass - use ``resolve`` to get it."""
    if "mix" not in spec:
        return spec
    rng = rng or _random
    for key in ("var_low", "var_medium", "var_high", "var_all"):
        if key in spec:
            return rng.choice(MIX_LISTS[side][key])
    raise SpecError("mix spec %r names none of var_low/var_medium/var_high/var_all" % spec)


def post_denoiser(spec, side="train"):
    """The denoiser the reference's post-denoise pass applies to the closure's result for the
    top-level ``spec`` string (first match of the elif chain), or None.  Training side: all five are
    live (minibatch.py:1636-1663).  Test side: the gaus_blur / mean / median / bilateral branches
    only print (their cv2 calls are commented out, test.py:1787-1819); `wavelet` is live
    (test.py:1805-1812) and is reached only when none of gaus_blur / mean / median is in the spec."""
    if isinstance(spec, Spec):
        return None
    d = next((d for d in POST_DENOISE_ORDER if d in spec), None)
    if side == "test" and d != "wavelet":
        return None
    return d


def resolve(spec, side="test", rng=None, compat_flags=0):
    """-> (concrete spec, post) for one top-level ``--noise`` string.

    * `mix` specs: ``concrete`` = the noise-only string drawn from the reference's list, ``post`` = the
      denoiser its post-denoise pass runs on that closure's uint8 result (``noise_mix_var_low_mean``
      -> e.g. ``('sap_var0.2', 'mean')``).  This is the only way the reference denoises a mix.
    * ``compat_flags & COMPAT_GAUSSIAN_RANDOM_LEVEL`` (quirk Q3, minibatch.py:1575-1580,
      test.py:1678-1682): any spec containing `gaussian` ignores its own level and denoiser, draws one
      of the three noise-only gaussian specs with ``random.choice`` and leaves the denoiser to the
      post pass.  Off by default.
    * everything else: ``(spec, None)`` - one closure call, at most one denoise pass (the canonical
      semantics); the training-side second filtering (Q1) stays opt-in
      (``double_filter=True`` in minibatch.py)."""
    if isinstance(spec, Spec):
        return spec, None
    rng = rng or _random
    if "mix" in spec:
        return resolve_mix(spec, side, rng), post_denoiser(spec, side)
    if (compat_flags & COMPAT_GAUSSIAN_RANDOM_LEVEL) and "gaussian" in spec:
        return rng.choice(_Q3_LIST), post_denoiser(spec, side)
    return spec, None
