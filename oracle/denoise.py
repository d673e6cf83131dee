"""Oracle: the denoiser half of the reference closures (test infrastructure only).

``*_cv2`` functions are the reference's own calls, verbatim.  ``*_formula``
functions restate the arithmetic OpenCV performs for those calls in plain
NumPy integers / float32 so that (a) the formula the CUDA kernels implement is
written down and (b) tests can show formula == cv2 on this OpenCV build.
"""
import numpy as np
import cv2

# constants used at every call site of the reference
BILATERAL_D = 9
BILATERAL_SIGMA_COLOR = 20
BILATERAL_SIGMA_SPACE = 100


def gaus_blur_cv2(u8):
    """cv2.GaussianBlur(im_noise, (3, 3), 0)  (minibatch.py:119)."""
    return cv2.GaussianBlur(u8, (3, 3), 0)


def mean_cv2(u8):
    """cv2.blur(im_noise, (3, 3))  (minibatch.py:136)."""
    return cv2.blur(u8, (3, 3))


def median_cv2(u8, k=3):
    """cv2.medianBlur(im_noise, 3)  (minibatch.py:153); k=7 for BASELINE config 5."""
    return cv2.medianBlur(u8, k)


def bilateral_cv2(u8):
    """cv2.bilateralFilter(im_noise, 9, 20, 100, borderType=BORDER_CONSTANT)
    (minibatch.py:166-172)."""
    return cv2.bilateralFilter(u8, BILATERAL_D, BILATERAL_SIGMA_COLOR,
                               BILATERAL_SIGMA_SPACE, borderType=cv2.BORDER_CONSTANT)


# --------------------------------------------------------------------------
# restatements
# --------------------------------------------------------------------------
def _pad(u8, r, mode):
    return np.pad(u8, ((r, r), (r, r), (0, 0)), mode=mode)


def gaus_blur_formula(u8):
    """(sum([1 2 1] x [1 2 1] * p) + 8) >> 4, border REFLECT_101."""
    p = _pad(u8, 1, "reflect").astype(np.int32)
    h = p[:, :-2] + 2 * p[:, 1:-1] + p[:, 2:]
    v = h[:-2] + 2 * h[1:-1] + h[2:]
    return ((v + 8) >> 4).astype(np.uint8)


def mean_formula(u8):
    """round-half-up(S / 9) = (2 S + 9) // 18, border REFLECT_101."""
    p = _pad(u8, 1, "reflect").astype(np.int32)
    h = p[:, :-2] + p[:, 1:-1] + p[:, 2:]
    s = h[:-2] + h[1:-1] + h[2:]
    return ((2 * s + 9) // 18).astype(np.uint8)


def median_formula(u8, k=3):
    """per-channel median of the k*k neighbourhood, border REPLICATE."""
    r = k // 2
    p = _pad(u8, r, "edge")
    H, W = u8.shape[:2]
    stack = [p[dy:dy + H, dx:dx + W] for dy in range(k) for dx in range(k)]
    return np.sort(np.stack(stack, 0), axis=0)[(k * k) // 2]


def bilateral_tables(d=BILATERAL_D, sigma_color=BILATERAL_SIGMA_COLOR,
                     sigma_space=BILATERAL_SIGMA_SPACE):
    """OpenCV's bilateralFilter_8u tables: taps inside the radius-r disc with
    float32 space weights, and the 256*3-entry float32 colour LUT."""
    r = d // 2
    gs = -0.5 / (sigma_space * sigma_space)
    gc = -0.5 / (sigma_color * sigma_color)
    taps = []
    for i in range(-r, r + 1):
        for j in range(-r, r + 1):
            rr = np.sqrt(float(i) * i + float(j) * j)
            if rr > r:
                continue
            taps.append((i, j, np.float32(np.exp(rr * rr * gs))))
    lut = np.exp((np.arange(256 * 3, dtype=np.float64) ** 2) * gc).astype(np.float32)
    return taps, lut


def bilateral_formula(u8):
    """fp32 restatement of cv2.bilateralFilter(u8,9,20,100,BORDER_CONSTANT):
    zero padding that participates, w = space * LUT[|db|+|dg|+|dr|],
    out_c = rint(sum(w p_c) / sum(w)).  May differ from cv2 by 1 LSB where the
    SIMD summation order differs (see tests)."""
    taps, lut = bilateral_tables()
    r = BILATERAL_D // 2
    H, W = u8.shape[:2]
    p = np.pad(u8, ((r, r), (r, r), (0, 0)), mode="constant").astype(np.int32)
    c = u8.astype(np.int32)
    wsum = np.zeros((H, W), np.float32)
    acc = np.zeros((H, W, 3), np.float32)
    for i, j, sw in taps:
        q = p[r + i:r + i + H, r + j:r + j + W]
        w = sw * lut[np.abs(q - c).sum(-1)]
        wsum += w
        acc += w[..., None] * q.astype(np.float32)
    out = acc / wsum[..., None]
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)
