"""Oracle: the noise half of the reference closures (test infrastructure only).

Every function restates what one ``add_<noise>_noise`` closure computes BEFORE
the denoiser runs (``lib/roi_data_layer/minibatch.py`` line numbers cited per
function; ``lib/model/test.py`` holds the same expressions).  Each stochastic
function has two modes:

  * ``fields=None``  - draw from ``rng`` exactly as the reference draws from the
    global ``np.random`` (same calls, same order, same shapes);
  * ``fields={...}`` - consume pre-drawn noise fields so the CUDA path and the
    oracle can be fed identical randomness (bit-exact parity tests).

skimage 0.14.2 is not installable here, so ``random_noise`` / ``img_as_float``
are restated from the published source of that version (skimage/util/noise.py,
skimage/util/dtype.py).
"""
import numpy as np
import cv2


def img_as_float(img):
    """skimage.img_as_float: uint8 -> float64 / 255, float passthrough
    (call sites minibatch.py:207,673,1212,1353)."""
    img = np.asarray(img)
    if img.dtype == np.uint8:
        return img.astype(np.float64) / 255.0
    if img.dtype.kind == "f":
        return img.astype(np.float64, copy=False)
    raise TypeError("oracle.img_as_float: unsupported dtype %s" % img.dtype)


def to_u8(x):
    """The reference's ``(255 * x).astype(np.uint8)`` idiom (minibatch.py:118).
    Truncation toward zero through a signed 32-bit integer, low byte kept, so
    out-of-range values WRAP (1.2 -> 306 -> 50)."""
    with np.errstate(invalid="ignore"):
        return (255 * np.asarray(x)).astype(np.uint8)


def to_u8_explicit(x):
    """Library-independent restatement of :func:`to_u8` used by the CUDA path's
    documentation: cvttsd2si semantics (|v| >= 2**31 or NaN -> INT_MIN), & 0xff."""
    v = 255.0 * np.asarray(x, dtype=np.float64)
    bad = ~np.isfinite(v) | (np.abs(v) >= 2147483648.0)
    t = np.trunc(np.where(bad, 0.0, v)).astype(np.int64)
    t = np.where(bad, -2147483648, t)
    return (t & 0xFF).astype(np.uint8)


# --------------------------------------------------------------------------
# skimage.util.random_noise modes (minibatch.py:91,211,259,380)
# --------------------------------------------------------------------------
def gaussian(img, var, fields=None, rng=np.random):
    """random_noise(img, mode='gaussian', var=var) -> float64 in [0,1]."""
    f = img_as_float(img)
    n = fields["n"] if fields is not None else rng.normal(0.0, var ** 0.5, f.shape)
    return np.clip(f + n, 0.0, 1.0)


def speckle(img, var, fields=None, rng=np.random):
    """random_noise(img, mode='speckle', var=var): out = x + x*n."""
    f = img_as_float(img)
    n = fields["n"] if fields is not None else rng.normal(0.0, var ** 0.5, f.shape)
    return np.clip(f + f * n, 0.0, 1.0)


def sap(img, amount, fields=None, rng=np.random, salt_vs_pepper=0.5):
    """random_noise(img, mode='s&p', amount=amount).  skimage 0.14.2 draws
    ``flipped = choice([True, False], size, p=[amount, 1-amount])`` then
    ``salted = choice([True, False], size, p=[q, 1-q])``; numpy's choice with
    ``p`` is ``searchsorted(cumsum(p), random_sample(size), side='right')`` so
    True <=> u < p.  Injected fields are the two uniform draws u1, u2."""
    f = img_as_float(img)
    out = f.copy()
    if fields is not None:
        flipped = fields["u1"] < amount
        salted = fields["u2"] < salt_vs_pepper
    else:
        flipped = rng.choice([True, False], size=f.shape, p=[amount, 1 - amount])
        salted = rng.choice([True, False], size=f.shape,
                            p=[salt_vs_pepper, 1 - salt_vs_pepper])
    peppered = ~salted
    out[flipped & salted] = 1.0
    out[flipped & peppered] = 0.0
    return np.clip(out, 0.0, 1.0)


def poisson_vals(img):
    """vals = 2 ** ceil(log2(#unique values)) (skimage noise.py, 'poisson')."""
    n_unique = len(np.unique(np.asarray(img)))
    return float(2 ** np.ceil(np.log2(n_unique)))


def poisson(img, fields=None, rng=np.random):
    """random_noise(img_as_float(img), mode='poisson') (minibatch.py:207-211).
    Injected field: the integer draws k ~ Poisson(f * vals)."""
    f = img_as_float(img)
    vals = poisson_vals(f)
    k = fields["k"] if fields is not None else rng.poisson(f * vals)
    return np.clip(k / float(vals), 0.0, 1.0)


# --------------------------------------------------------------------------
# additive float noises that are NOT clipped (wrap on ->u8)
# --------------------------------------------------------------------------
def uniform(img, high, fields=None, rng=np.random):
    """cv2.add(img_as_float(img), np.random.uniform(0, high, shape))
    (minibatch.py:676-677).  cv2.add on float64 is a plain add."""
    f = img_as_float(img)
    u = fields["u"] if fields is not None else rng.uniform(low=0.0, high=high, size=f.shape)
    return cv2.add(f, u)


def gamma(img, scale, fields=None, rng=None, a=1.99):
    """cv2.add(image, scipy.stats.gamma.rvs(1.99, loc=0, scale=scale, size))
    (minibatch.py:1213-1217)."""
    from scipy.stats import gamma as _gamma
    f = img_as_float(img)
    if fields is not None:
        g = fields["g"]
    else:
        g = _gamma.rvs(a, loc=0.0, scale=scale, size=f.shape, random_state=rng)
    return cv2.add(f, g)


def rayleigh(img, scale, fields=None, rng=None):
    """cv2.add(image, scipy.stats.rayleigh.rvs(loc=0, scale=scale, size))
    (minibatch.py:1356-1357)."""
    from scipy.stats import rayleigh as _rayleigh
    f = img_as_float(img)
    if fields is not None:
        r = fields["r"]
    else:
        r = _rayleigh.rvs(loc=0.0, scale=scale, size=f.shape, random_state=rng)
    return cv2.add(f, r)


# --------------------------------------------------------------------------
# u8 additive fields with wrap-around casts (saturating cv2.add on uint8)
# --------------------------------------------------------------------------
def brownian_field(shape, dt, fields=None, rng=np.random):
    """B = [0, cumsum(sqrt(dt) * N(0,1)[n-1])]; (B*255).astype(uint8).reshape
    (minibatch.py:822-827).  Injected field: the n-1 standard normals."""
    h, w = shape[:2]
    n = h * w * 3
    z = fields["z"] if fields is not None else rng.normal(size=(n - 1,))
    dB = np.sqrt(dt) * z
    B = np.concatenate((np.zeros(shape=(1,)), np.cumsum(dB)))
    with np.errstate(invalid="ignore"):
        return (B * 255).astype(np.uint8).reshape(h, w, 3)


def brownian(img, dt, fields=None, rng=np.random):
    """cv2.add(img, brownian_noise) - saturating uint8 add (minibatch.py:828)."""
    return cv2.add(img, brownian_field(img.shape, dt, fields, rng))


def periodic_field(shape, level):
    """(sin(linspace(-L, L, size)) * 255).astype(uint8).reshape(h, w, 3)
    (minibatch.py:1042-1045).  ``level`` is 'pi' (var3.14), 100 or 'size'."""
    h, w = shape[:2]
    size = h * w * 3
    if level in ("pi", 3.14):
        lim = np.pi
    elif level == "size":
        lim = size
    else:
        lim = level
    time = np.linspace(-lim, lim, size)
    amplitude = np.sin(time)
    with np.errstate(invalid="ignore"):
        return (amplitude * 255).astype(np.uint8).reshape(h, w, 3)


def periodic(img, level):
    """cv2.add(img, periodic_noise) (minibatch.py:1046)."""
    return cv2.add(img, periodic_field(img.shape, level))


# --------------------------------------------------------------------------
# colour quantisation, brightness "shader", bloom
# --------------------------------------------------------------------------
def quant(img, k, fields=None, rng=None):
    """BGR->Lab, MiniBatchKMeans(k).fit_predict, palette = centres.astype(u8),
    Lab->BGR (minibatch.py:496-507).  Injected field: the float64 centres
    (k,3) in Lab; labels are then argmin squared L2 to those centres, which is
    what fit_predict returns for its own final centres."""
    lab = cv2.cvtColor(img, cv2.COLOR_BGR2LAB)
    h, w = lab.shape[:2]
    flat = lab.reshape((h * w, 3))
    if fields is not None:
        centers = np.asarray(fields["centers"], dtype=np.float64)
        x = flat.astype(np.float64)
        d = ((x[:, None, :] - centers[None, :, :]) ** 2).sum(-1)
        labels = np.argmin(d, axis=1)
    else:
        from sklearn.cluster import MiniBatchKMeans
        clt = MiniBatchKMeans(n_clusters=k, random_state=rng)
        labels = clt.fit_predict(flat)
        centers = clt.cluster_centers_
    q = centers.astype("uint8")[labels].reshape(h, w, 3)
    return cv2.cvtColor(q, cv2.COLOR_LAB2BGR)


def shader(img_bgr):
    """ImageEnhance.Brightness(PIL image).enhance(3) (minibatch.py:1505-1511).
    PIL decodes RGB, so the reference's result is RGB-ordered (quirk Q5); the
    oracle takes the BGR array the rest of the path uses and returns RGB."""
    from PIL import Image, ImageEnhance
    rgb = np.ascontiguousarray(img_bgr[:, :, ::-1])
    out = ImageEnhance.Brightness(Image.fromarray(rgb)).enhance(3)
    return np.array(out)


def bloom_draws(shape, pyrandom, no_of_flare_circles=8, flare_center=(100, 100),
                angle=-np.pi / 4, src_color=(255, 255, 255)):
    """The random draws add_sun_process makes, in its order
    (tools/Automold.py:575-584): per circle alpha, line index, radius, colour."""
    import math
    ang = angle % (2 * math.pi)
    xs, ys = [], []
    for rx in range(0, shape[1], 10):                 # add_sun_flare_line, Automold.py:565-573
        ry = math.tan(ang) * (rx - flare_center[0]) + flare_center[1]
        xs.append(rx)
        ys.append(2 * flare_center[1] - ry)
    draws = []
    for _ in range(no_of_flare_circles):
        alpha = pyrandom.uniform(0.05, 0.2)
        r = pyrandom.randint(0, len(xs) - 1)
        rad = pyrandom.randint(1, shape[0] // 100 - 2)
        col = tuple(pyrandom.randint(max(c - 50, 0), c) for c in src_color)
        draws.append((alpha, int(xs[r]), int(ys[r]), rad * rad * rad, col))
    return draws


def bloom(img, draws, flare_center=(100, 100), src_radius=400, src_color=(255, 255, 255)):
    """add_sun_process + flare_source replayed with cv2 itself from a list of
    pre-drawn circles (tools/Automold.py:553-563,575-586)."""
    overlay = img.copy()
    output = img.copy()
    for alpha, cx, cy, rad, col in draws:
        cv2.circle(overlay, (cx, cy), rad, col, -1)
        cv2.addWeighted(overlay, alpha, output, 1 - alpha, 0, output)
    overlay = output.copy()
    output = output.copy()
    num_times = src_radius // 10
    alpha = np.linspace(0.0, 1, num=num_times)
    rad = np.linspace(1, src_radius, num=num_times)
    for i in range(num_times):
        cv2.circle(overlay, (int(flare_center[0]), int(flare_center[1])), int(rad[i]), src_color, -1)
        alp = alpha[num_times - i - 1] * alpha[num_times - i - 1] * alpha[num_times - i - 1]
        cv2.addWeighted(overlay, alp, output, 1 - alp, 0, output)
    return output
