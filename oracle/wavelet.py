"""Oracle: NumPy restatement of skimage.restoration.denoise_wavelet (0.14.2)
and the PyWavelets (1.0.3) routines it reaches.  TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED against the real libraries: scikit-image 0.14.2 and
pywavelets 1.0.3 are pinned in the reference's requirements.txt:151,160 but
are neither vendored under /root/reference nor installable here (no network).
The algorithm below restates their published source:
  skimage/restoration/_denoise.py  (_wavelet_threshold, _bayes_thresh,
                                    _sigma_est_dwt, denoise_wavelet)
  skimage/color/colorconv.py       (rgb2ycbcr, ycbcr2rgb)
  pywt/_multilevel.py              (wavedecn, waverecn), pywt/_thresholding.py
and is anchored on the reference's call sites (minibatch.py:92-96 - bior1.5,
automatic levels; test.py_bkp:1913-1916 - db1, wavelet_levels=3).  It is held
by tests/test_oracle_wavelet.py: property tests, an independent dense-matrix
transform it must equal, and the known answers of the PyWavelets documentation.
`sigma_census` / the `sigmas=` argument exist for the stage-wise GPU parity check
(tests/wavelet_check.py); the reference's own behaviour is `denoise_wavelet(x, ...)`.
"""
import numpy as np

_S = 0.7071067811865476
_C1 = 0.01657281518405971     # sqrt(2) * 3 / 256
_C2 = 0.12153397801643787     # sqrt(2) * 11 / 128

WAVELETS = {
    # name: (dec_lo, dec_hi, rec_lo, rec_hi) as PyWavelets lists them
    "db1": ([_S, _S], [-_S, _S], [_S, _S], [_S, -_S]),
    "bior1.5": (
        [_C1, -_C1, -_C2, _C2, _S, _S, _C2, -_C2, -_C1, _C1],
        [0.0, 0.0, 0.0, 0.0, -_S, _S, 0.0, 0.0, 0.0, 0.0],
        [0.0, 0.0, 0.0, 0.0, _S, _S, 0.0, 0.0, 0.0, 0.0],
        [_C1, _C1, -_C2, -_C2, _S, -_S, _C2, _C2, -_C1, -_C1],
    ),
}
WAVELETS["haar"] = WAVELETS["db1"]


def dwt_max_level(data_len, filter_len):
    """pywt.dwt_max_level: floor(log2(data_len / (filter_len - 1)))."""
    if data_len < filter_len - 1:
        return 0
    return int(np.floor(np.log2(data_len / (filter_len - 1.0))))


def _take(a, axis, sl):
    idx = [slice(None)] * a.ndim
    idx[axis] = sl
    return a[tuple(idx)]


def dwt_axis(x, wavelet, axis):
    """Single-level DWT along ``axis``, mode='symmetric' (half-sample symmetric
    extension).  Equals np.convolve(np.pad(x, F-1, 'symmetric'), filt,
    'valid')[1::2] per line; output length floor((n + F - 1) / 2)."""
    dec_lo, dec_hi = WAVELETS[wavelet][0], WAVELETS[wavelet][1]
    F = len(dec_lo)
    n = x.shape[axis]
    pad = [(0, 0)] * x.ndim
    pad[axis] = (F - 1, F - 1)
    p = np.pad(x, pad, mode="symmetric")
    m = (n + F - 1) // 2
    shape = list(x.shape)
    shape[axis] = m
    a = np.zeros(shape, np.float64)
    d = np.zeros(shape, np.float64)
    for j in range(F):
        s = _take(p, axis, slice(F - j, F - j + 2 * m, 2))
        if dec_lo[j] != 0.0:
            a += dec_lo[j] * s
        if dec_hi[j] != 0.0:
            d += dec_hi[j] * s
    return a, d


def idwt_axis(a, d, wavelet, axis):
    """Single-level inverse along ``axis``: zero-upsample, convolve with the
    reconstruction filters, add, keep samples [F-2, 2m) (length 2m - F + 2)."""
    rec_lo, rec_hi = WAVELETS[wavelet][2], WAVELETS[wavelet][3]
    F = len(rec_lo)
    m = a.shape[axis]
    shape = list(a.shape)
    shape[axis] = 2 * m + F - 1
    y = np.zeros(shape, np.float64)
    for j in range(F):
        idx = [slice(None)] * a.ndim
        idx[axis] = slice(j, j + 2 * m, 2)
        if rec_lo[j] != 0.0:
            y[tuple(idx)] += rec_lo[j] * a
        if rec_hi[j] != 0.0:
            y[tuple(idx)] += rec_hi[j] * d
    return _take(y, axis, slice(F - 2, 2 * m))


def wavedec2(x, wavelet, level):
    """pywt.wavedecn for a 2-D array: [cA_n, {ad,da,dd}_n, ..., {ad,da,dd}_1];
    first key letter = axis 0."""
    a = np.asarray(x, np.float64)
    details = []
    for _ in range(level):
        lo0, hi0 = dwt_axis(a, wavelet, 0)
        aa, ad = dwt_axis(lo0, wavelet, 1)
        da, dd = dwt_axis(hi0, wavelet, 1)
        details.append({"ad": ad, "da": da, "dd": dd})
        a = aa
    return [a] + details[::-1]


def waverec2(coeffs, wavelet):
    """pywt.waverecn for 2-D; when the running approximation is one sample
    longer than the next detail band along an axis it is trimmed to fit."""
    a = coeffs[0]
    for det in coeffs[1:]:
        dshape = det["dd"].shape
        if a.shape != dshape:
            a = a[tuple(slice(s) for s in dshape)]
        lo0 = idwt_axis(a, det["ad"], wavelet, 1)
        hi0 = idwt_axis(det["da"], det["dd"], wavelet, 1)
        a = idwt_axis(lo0, hi0, wavelet, 0)
    return a


def soft_threshold(data, value):
    """pywt.threshold(data, value, 'soft') = data * max(1 - value/|data|, 0)."""
    mag = np.absolute(data)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = 1 - value / mag
        t = t.clip(min=0, max=None)
        return data * t


# --------------------------------------------------------------------------
# skimage.restoration._denoise (0.14.2)
# --------------------------------------------------------------------------
NORM_PPF_075 = 0.6744897501960817          # scipy.stats.norm.ppf(0.75)


def sigma_est_dwt(detail_coeffs):
    """_sigma_est_dwt: MAD of the non-zero finest-level 'dd' coefficients."""
    dc = detail_coeffs[np.nonzero(detail_coeffs)]
    return np.median(np.abs(dc)) / NORM_PPF_075


def bayes_thresh(details, var):
    """_bayes_thresh: var / sqrt(max(mean(d*d) - var, eps))."""
    dvar = np.mean(details * details)
    eps = np.finfo(details.dtype).eps
    return var / np.sqrt(max(dvar - var, eps))


def auto_levels(shape, wavelet):
    """wavelet_levels=None branch of _wavelet_threshold."""
    dlen = len(WAVELETS[wavelet][0])
    lv = min(dwt_max_level(s, dlen) for s in shape)
    return max(lv - 3, 1)


def wavelet_threshold(image, wavelet="db1", wavelet_levels=None, return_stats=False, sigma=None):
    """_wavelet_threshold(image, wavelet, method='BayesShrink', mode='soft').
    ``sigma`` (tests only): use this noise estimate instead of _sigma_est_dwt's - the stage-wise parity
    check hands the oracle the estimate the CUDA path derived (see sigma_census)."""
    original_extent = tuple(slice(s) for s in image.shape)
    if wavelet_levels is None:
        wavelet_levels = auto_levels(image.shape, wavelet)
    coeffs = wavedec2(image, wavelet, wavelet_levels)
    dcoeffs = coeffs[1:]
    if sigma is None:
        sigma = sigma_est_dwt(dcoeffs[-1]["dd"])
    var = sigma ** 2
    threshold = [{key: bayes_thresh(level[key], var) for key in level} for level in dcoeffs]
    denoised_detail = [{key: soft_threshold(level[key], thresh[key]) for key in level}
                       for thresh, level in zip(threshold, dcoeffs)]
    rec = waverec2([coeffs[0]] + denoised_detail, wavelet)[original_extent]
    if return_stats:
        return rec, {"sigma": sigma, "threshold": threshold, "levels": wavelet_levels}
    return rec


YCBCR_FROM_RGB = np.array([[65.481, 128.553, 24.966],
                           [-37.797, -74.203, 112.0],
                           [112.0, -93.786, -18.214]])
RGB_FROM_YCBCR = np.linalg.inv(YCBCR_FROM_RGB)


def rgb2ycbcr(rgb):
    """skimage.color.rgb2ycbcr on a float image in [0,1] (BT.601 studio swing)."""
    arr = np.dot(np.asarray(rgb, np.float64), YCBCR_FROM_RGB.T.copy())
    arr[..., 0] += 16
    arr[..., 1] += 128
    arr[..., 2] += 128
    return arr


def ycbcr2rgb(ycbcr):
    arr = ycbcr.copy()
    arr[..., 0] -= 16
    arr[..., 1] -= 128
    arr[..., 2] -= 128
    return np.dot(arr, RGB_FROM_YCBCR.T.copy())


def _hh2(x):
    """2 x 2 differences d - c - b + a of the symmetric-extended plane, sampled the way both wavelets'
    high-pass filters ([-s, s] at taps 4, 5 of bior1.5 = the two taps of db1) sample them: the integer
    skeleton of the finest 'dd' band.  Returns (a, b, c, d) planes of the band's shape."""
    F = 2
    p = np.pad(x, ((F - 1, F - 1), (F - 1, F - 1)), mode="symmetric")
    m0, m1 = (x.shape[0] + F - 1) // 2, (x.shape[1] + F - 1) // 2
    r0, r1 = p[1:1 + 2 * m0:2], p[2:2 + 2 * m0:2]
    return r0[:, 1:1 + 2 * m1:2], r0[:, 2:2 + 2 * m1:2], r1[:, 1:1 + 2 * m1:2], r1[:, 2:2 + 2 * m1:2]


def sigma_census(image, wavelet="db1", tiny=1e-6):
    """For each "YCbCr" channel of ``image``: what _sigma_est_dwt's result can be when the set of EXACTLY-zero
    finest 'dd' coefficients is decided by an arithmetic other than this float64 one.

    sigma = median(|dd_1|[dd_1 != 0]) / 0.6745 filters exact zeros.  dd_1 of both wavelets is a 2 x 2 difference
    (a - b) - (c - d) of channel samples.  When it cancels ALGEBRAICALLY but the four samples are not pairwise
    equal, the computed value is 0.0 or rounding noise (|d| ~ 1e-17 in float64, ~1e-8 in float32) depending on
    the last bit of the colour matrix product and the normalisation - BLAS build, float width, operation order:
    those coefficients are AMBIGUOUS, and counting k of them as non-zero shifts the rank of the median.
    uint8-valued images (s&p, poisson, quant, brownian, periodic, clean): decided exactly in integers
    (1000 x the BT.601 matrix is integral).  Float images: a coefficient is ambiguous when it is below ``tiny``
    of the channel range in the normalised or the un-normalised float64 evaluation.
    Returns a list of (sigma_min, sigma_max, n_sure, n_ambiguous): the estimates with all / none of the ambiguous
    coefficients counted.  A well-conditioned channel has n_ambiguous == 0 and sigma_min == sigma_max == the
    oracle's sigma."""
    if image.dtype == np.uint8:
        image = image.astype(np.float64) / 255.0
    image = np.asarray(image, np.float64)
    as_int = np.rint(image * 255.0)
    integral = bool(np.all(np.abs(image * 255.0 - as_int) < 1e-9))
    ycc = rgb2ycbcr(image)
    m1000 = np.rint(YCBCR_FROM_RGB * 1000.0).astype(np.int64)
    out = []
    for i in range(3):
        ch = ycc[..., i]
        mn, mx = ch.min(), ch.max()
        norm = (ch - mn) / (mx - mn)
        dd = np.abs(dwt_axis(dwt_axis(norm, wavelet, 0)[1], wavelet, 1)[1])
        F = len(WAVELETS[wavelet][0])
        ids = np.arange(ch.size, dtype=np.int64).reshape(ch.shape)
        m0, m1 = dd.shape

        def quad(plane):                                           # taps F/2-1 (-s) and F/2 (+s): bior1.5's 4, 5 = db1's 0, 1
            p = np.pad(plane, ((F - 1, F - 1), (F - 1, F - 1)), mode="symmetric")
            o4, o5 = F - (F // 2 - 1), F - F // 2
            ra, rb = p[o4:o4 + 2 * m0:2], p[o5:o5 + 2 * m0:2]
            return ra[:, o4:o4 + 2 * m1:2], rb[:, o4:o4 + 2 * m1:2], ra[:, o5:o5 + 2 * m1:2], rb[:, o5:o5 + 2 * m1:2]
        ia, ib, ic, id_ = quad(ids)
        # the same ARRAY ELEMENT on both sides (duplicated border samples): zero in every arithmetic
        zero = ((ia == ib) & (ic == id_)) | ((ia == ic) & (ib == id_))
        if integral:
            yi = as_int.astype(np.int64) @ m1000[i]                # exact, up to the common factor 1 / 255000
            a, b, c, d = quad(yi)
            amb = ((a - b - c + d) == 0) & ~zero
        else:
            # float image (clipped Gaussian / speckle ...): whatever is below `tiny` in the normalised or in the
            # un-normalised float64 evaluation - exact float64 zeros of distinct samples included, they are luck
            ddu = np.abs(dwt_axis(dwt_axis(ch, wavelet, 0)[1], wavelet, 1)[1]) / (mx - mn)
            amb = ~zero & (np.minimum(dd, ddu) <= tiny)
        sure = np.sort(dd[~amb & ~zero].ravel())
        k = int(amb.sum())
        if sure.size == 0:
            out.append((0.0, 0.0, 0, k))
            continue
        with_all = np.concatenate([np.zeros(k), sure])          # ambiguous ones counted (as ~0 values)
        out.append((float(np.median(with_all)) / NORM_PPF_075, float(np.median(sure)) / NORM_PPF_075, int(sure.size), k))
    return out


def denoise_wavelet(image, wavelet="db1", wavelet_levels=None, return_stats=False, sigmas=None):
    """denoise_wavelet(image, method='BayesShrink', mode='soft', wavelet=...,
    wavelet_levels=..., multichannel=True, convert2ycbcr=True)
    (minibatch.py:92-94).  ``image``: uint8 or float64 HxWx3 in the array's own
    channel order (the reference feeds BGR - quirk Q7 - so do the callers)."""
    if image.dtype == np.uint8:
        image = image.astype(np.float64) / 255.0
    image = np.asarray(image, np.float64)
    out = rgb2ycbcr(image)
    stats = []
    for i in range(3):
        mn, mx = out[..., i].min(), out[..., i].max()
        channel = out[..., i] - mn
        with np.errstate(divide="ignore", invalid="ignore"):
            channel /= mx - mn
        res = wavelet_threshold(channel, wavelet, wavelet_levels, return_stats,
                                None if sigmas is None else float(sigmas[i]))
        if return_stats:
            res, st = res
            st.update(min=mn, max=mx)
            stats.append(st)
        # the inner (single-channel) denoise_wavelet call clips to [0, 1]
        res = np.clip(res, 0, 1)
        out[..., i] = res * (mx - mn)
        out[..., i] += mn
    out = ycbcr2rgb(out)
    clip_range = (-1, 1) if image.min() < 0 else (0, 1)
    out = np.clip(out, *clip_range)
    if return_stats:
        return out, stats
    return out
