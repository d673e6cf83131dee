"""Oracle: the reference closures composed end to end (test infrastructure only).

``corrupt(img, spec, ...)`` evaluates one ``add_<noise>_noise(spec)`` closure of
``lib/model/test.py:193-1607`` (identical bodies in
``lib/roi_data_layer/minibatch.py:87-1516`` except that training returns the
float image on noise-only branches): noise -> [->u8 -> cv2 filter | wavelet ->
->u8 | ->u8].  Dispatch is by the reference's own substring tests, transcribed
as conditions (``ladder``), not through the product's parser.
"""
import numpy as np

from . import noise as N
from . import denoise as D
from . import wavelet as W

# order of the top-level `elif ('<name>' in spec)` chain, minibatch.py:1518-1634
NOISE_ORDER = ["gaussian", "poisson", "sap", "speckle", "periodic", "brownian",
               "quant", "uniform", "gamma", "rayleigh", "bloom", "shader"]
# order of the `<noise>_<denoise>` tests inside each closure, e.g. minibatch.py:89-184
DENOISE_ORDER = ["wavelet", "gaus_blur", "mean", "median", "bilateral"]
# order of the `'var<L>' in spec` tests per noise (first match wins)
LEVELS = {
    "gaussian": ["0.1", "1.0", "1.5"],
    "sap": ["0.2", "0.4", "0.8"],
    "speckle": ["0.5", "1.0", "2.0"],
    "quant": ["3", "7", "10"],
    "uniform": ["0.2", "0.6", "1.2"],
    "brownian": ["0.9", "0.09", "0.009"],
    "periodic": ["3.14", "100", "size"],
    "gamma": ["0.05", "0.1", "0.2"],
    "rayleigh": ["0.1", "0.2", "0.3"],
}


def ladder(spec):
    """(noise, denoise, level-string) the reference's `in` chains select for
    ``spec``; noise None = retain_original; level None = no/unknown level."""
    noise = next((n for n in NOISE_ORDER if n in spec), None)
    if noise is None:
        return None, None, None
    if noise in ("bloom", "shader"):
        return noise, None, None
    denoise = next((d for d in DENOISE_ORDER if (noise + "_" + d) in spec), None)
    level = None
    if noise in LEVELS:
        level = next((l for l in LEVELS[noise] if ("var" + l) in spec), None)
    return noise, denoise, level


def apply_noise(img, noise, level, fields=None, rng=np.random):
    """The noise expression of one closure; returns float64 (unclipped where the
    reference does not clip) or uint8 for the integer noises."""
    if noise == "gaussian":
        return N.gaussian(img, float(level), fields, rng)
    if noise == "speckle":
        return N.speckle(img, float(level), fields, rng)
    if noise == "sap":
        return N.sap(img, float(level), fields, rng)
    if noise == "poisson":
        return N.poisson(img, fields, rng)
    if noise == "uniform":
        return N.uniform(img, float(level), fields, rng)
    if noise == "gamma":
        return N.gamma(img, float(level), fields, rng if rng is not np.random else None)
    if noise == "rayleigh":
        return N.rayleigh(img, float(level), fields, rng if rng is not np.random else None)
    if noise == "brownian":
        return N.brownian(img, float(level), fields, rng)
    if noise == "periodic":
        return N.periodic(img, {"3.14": "pi", "100": 100, "size": "size"}[level])
    if noise == "quant":
        return N.quant(img, int(level), fields, None)
    if noise == "shader":
        return N.shader(img)
    if noise == "bloom":
        return N.bloom(img, fields["draws"])
    if noise is None:
        return img.copy()
    raise ValueError(noise)


def apply_denoise(x, denoise, wavelet="bior1.5", wavelet_levels=None, median_k=3):
    """x: the noise stage's result (float64 or uint8)."""
    if denoise == "wavelet":
        # float noises feed the float image; poisson / u8 noises quantise first
        return N.to_u8(W.denoise_wavelet(x, wavelet, wavelet_levels))
    u8 = x if x.dtype == np.uint8 else N.to_u8(x)
    if denoise is None:
        return u8
    if denoise == "gaus_blur":
        return D.gaus_blur_cv2(u8)
    if denoise == "mean":
        return D.mean_cv2(u8)
    if denoise == "median":
        return D.median_cv2(u8, median_k)
    if denoise == "bilateral":
        return D.bilateral_cv2(u8)
    raise ValueError(denoise)


def noise_stage(img, spec, fields=None, rng=np.random):
    """What the closure hands to its denoiser: float64 for the float noises, uint8 otherwise."""
    noise, denoise, level = ladder(spec)
    if noise in LEVELS and level is None:
        raise UnboundLocalError("unknown level in %r (test.py leaves `im` unbound)" % spec)
    x = apply_noise(img, noise, level, fields, rng)
    if noise == "poisson" and denoise == "wavelet":
        x = N.to_u8(x)                                  # minibatch.py:212-213
    return x


def corrupt(img, spec, fields=None, rng=np.random, wavelet="bior1.5",
            wavelet_levels=None, median_k=3):
    """One closure call: uint8 BGR in -> uint8 out (test.py semantics)."""
    denoise = ladder(spec)[1]
    return apply_denoise(noise_stage(img, spec, fields, rng), denoise, wavelet, wavelet_levels, median_k)
