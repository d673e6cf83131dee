"""Golden-vector cases ("golden vectors to create - none exist upstream").

One case = (input image, spec, injected fields drawn the way the reference's closure draws them,
from RandomState(3) = cfg.RNG_SEED, lib/model/config.py:255).  tests/golden/make_manifest.py evaluates
every case with the oracle (the reference's own cv2 / NumPy / SciPy / PIL calls + the restated
skimage / pywt functions) in the container that holds /root/reference and commits the SHA-256 of
each uint8 result with the library versions to tests/golden/manifest.json.  The CPU suite re-derives the digests (pins
the oracle), the GPU suite holds the CUDA path to them: digest-equal for the bit-exact branches, within the stated
tolerance of the digest-checked oracle image for the float branches."""
import hashlib
import os

import numpy as np

from conftest import synth_image

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
MANIFEST = os.path.join(GOLDEN, "manifest.json")
NATURAL = os.path.join(GOLDEN, "natural_120x160.npz")

LEVEL_MID = {"gaussian": "1.0", "speckle": "1.0", "sap": "0.4", "uniform": "0.6", "gamma": "0.1", "rayleigh": "0.2",
             "brownian": "0.09"}
OTHER_LEVELS = {"gaussian": ["0.1", "1.5"], "speckle": ["0.5", "2.0"], "sap": ["0.2", "0.8"], "uniform": ["0.2", "1.2"],
                "gamma": ["0.05", "0.2"], "rayleigh": ["0.1", "0.3"], "brownian": ["0.9", "0.009"]}
# tolerance class per branch (BASELINE.json north_star): "exact" = digest must match
FLOAT_DENOISERS = ("bilateral", "wavelet")


def inputs():
    """name -> uint8 BGR image.  natural_* is a crop of the reference's data/demo/000456.jpg as cv2.imread
    decodes it (made by make_manifest.py; the only natural-image statistics the reference ships)."""
    out = {"smooth_64x96": synth_image(64, 96, 5, "smooth"), "uniform_61x99": synth_image(61, 99, 6, "uniform")}
    if os.path.exists(NATURAL):
        out["natural_120x160"] = np.load(NATURAL)["img"]
    return out


def deterministic_specs():
    """retain_original (+ each post-denoise pass, minibatch.py:1636-1662) and shader; the periodic ladder is
    deterministic too but lives with the other noises in stochastic_specs()."""
    return ["original", "original_gaus_blur", "original_mean", "original_median", "original_bilateral", "original_wavelet",
            "shader"]


def stochastic_specs():
    """The WHOLE ladder of the reference: every noise closure x {noise only, wavelet, gaus_blur, mean, median,
    bilateral} x each of its three `'var<L>' in spec` levels (lib/roi_data_layer/minibatch.py:87-1484, twins in
    lib/model/test.py:193-1588) - 9 x 6 x 3 + poisson x 6 = 168 specs, no cell skipped."""
    from oracle import closures as C
    specs = []
    for noise in C.NOISE_ORDER:
        if noise in ("bloom", "shader"):
            continue                                    # no spec argument (minibatch.py:1485-1511): deterministic_specs / bloom fixture
        for lvl in C.LEVELS.get(noise, [None]):
            for den in ("", "_wavelet", "_gaus_blur", "_mean", "_median", "_bilateral"):
                specs.append(noise + den + ("_var" + lvl if lvl else ""))
    return specs


def noise_of(spec):
    from oracle import closures as C
    return C.ladder(spec)[0]


def draw_fields(spec, img, rng):
    """The draws of one closure call, same NumPy / SciPy calls and order as the reference
    (e.g. lib/roi_data_layer/minibatch.py:91, 259, 676, 822, 1216, 1356)."""
    from oracle import closures as C
    noise, _, level = C.ladder(spec)
    shape = img.shape
    if noise in ("gaussian", "speckle"):
        return {"n": rng.normal(0.0, float(level) ** 0.5, shape)}
    if noise == "sap":
        return {"u1": rng.random_sample(shape), "u2": rng.random_sample(shape)}
    if noise == "uniform":
        return {"u": rng.uniform(0.0, float(level), shape)}
    if noise == "gamma":
        from scipy.stats import gamma
        return {"g": gamma.rvs(1.99, loc=0.0, scale=float(level), size=shape, random_state=rng)}
    if noise == "rayleigh":
        from scipy.stats import rayleigh
        return {"r": rayleigh.rvs(loc=0.0, scale=float(level), size=shape, random_state=rng)}
    if noise == "poisson":
        from oracle import noise as N
        f = N.img_as_float(img)
        return {"k": rng.poisson(f * N.poisson_vals(f)).astype(np.float64)}
    if noise == "brownian":
        return {"z": rng.normal(size=(int(np.prod(shape)) - 1,))}
    if noise == "quant":
        k = int(spec.rsplit("var", 1)[1])
        return {"centers": np.stack([rng.uniform(20, 235, k), rng.uniform(100, 160, k), rng.uniform(100, 160, k)], -1)}
    return None


def cases():
    """Yields (case id, image, spec, fields) in a fixed order; the RNG restarts per case."""
    for name, img in inputs().items():
        for spec in deterministic_specs():
            yield "%s|%s" % (name, spec), img, spec, None
        for spec in stochastic_specs():
            yield "%s|%s" % (name, spec), img, spec, draw_fields(spec, img, np.random.RandomState(3))


def evaluate(img, spec, fields):
    """The oracle's result for one case.  `original_<denoise>` = retain_original followed by the post-denoise
    pass (lib/model/test.py:1602-1607, 1787-1831; minibatch.py:1636-1662); everything else is one closure."""
    from oracle import closures as C
    if spec == "original":
        return img.copy()
    if spec.startswith("original_"):
        return C.apply_denoise(img, spec[len("original_"):])
    return C.corrupt(img, spec, fields=fields)


def digest(arr):
    a = np.ascontiguousarray(arr)
    assert a.dtype == np.uint8
    return hashlib.sha256(("%s|" % (a.shape,)).encode() + a.tobytes()).hexdigest()


def versions():
    import cv2
    import scipy
    import PIL
    return {"numpy": np.__version__, "cv2": cv2.__version__, "scipy": scipy.__version__, "PIL": PIL.__version__}


def tolerance_class(spec):
    """'exact' | 'lsb' (<= 1 LSB on a small fraction) | 'brownian' (<= 1 LSB mod 256 on <= 1e-4) | 'periodic' |
    'brownian_lsb' (brownian field as above, then a float denoiser: checked stage by stage on the GPU)."""
    from oracle import closures as C
    noise, den, _ = C.ladder(spec)
    if noise == "brownian" and den in FLOAT_DENOISERS:
        return "brownian_lsb"
    if den in FLOAT_DENOISERS or (spec.startswith("original_") and spec[len("original_"):] in FLOAT_DENOISERS):
        return "lsb"
    if noise == "brownian":
        return "brownian"
    if noise == "periodic":
        return "periodic"
    return "exact"
