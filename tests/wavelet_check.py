"""Stage-wise parity check of the wavelet denoiser (GPU tests).

skimage's sigma = median(|dd_1|[dd_1 != 0]) / 0.6745 depends on WHICH finest-level coefficients are exactly zero,
and for algebraically cancelling 2 x 2 differences that is decided by the last bit of the float64 colour product
(BLAS blocking included): the reference's own value is not defined better than the range
oracle.wavelet.sigma_census() returns.  So the check is split in two, each half tight:
  1. the sigma the CUDA path derived (test hook b2d_wavelet_stats) lies inside the census range of the oracle
     (a single value, to fp32 precision, when no coefficient is ambiguous), and its channel min / max match;
  2. with THAT sigma handed to the oracle, the uint8 images agree to <= 1 LSB everywhere, and away from the
     integers (where the reference's truncation of its float64 result is decided by rounding noise no fp32 path
     reproduces) on all but a small fraction of the pixels.
"""
import numpy as np

from oracle import noise as N
from oracle import wavelet as W

SIGMA_RTOL = 5e-4          # fp32 planes and median vs float64
FRAC_AWAY = 0.02           # pixels farther than AWAY from an integer that may still differ by 1 LSB
AWAY = 0.05


def stagewise(got, x, gpu_stats, wavelet="bior1.5", levels=None, tag=""):
    """got: uint8 HxWx3 CUDA result; x: the oracle's noise-stage image (uint8 or float64 in [0, 1]);
    gpu_stats: (3, 4) {sigma, min, max, count} from Plan.wavelet_stats().  Returns a list of failure tuples."""
    fails = []
    xf = x.astype(np.float64) / 255.0 if x.dtype == np.uint8 else np.asarray(x, np.float64)
    ycc = W.rgb2ycbcr(xf)
    census = W.sigma_census(xf, wavelet)
    sig = []
    for c in range(3):
        s_gpu, lo, hi = float(gpu_stats[c, 0]), float(gpu_stats[c, 1]), float(gpu_stats[c, 2])
        mn, mx = ycc[..., c].min(), ycc[..., c].max()
        if abs(lo - mn) > 1e-4 * max(1.0, abs(mn)) or abs(hi - mx) > 1e-4 * max(1.0, abs(mx)):
            fails.append((tag, "channel range", c, lo, hi, float(mn), float(mx)))
        s_norm = s_gpu / (hi - lo) if hi > lo else 0.0
        smin, smax, n_sure, n_amb = census[c]
        if not (smin * (1 - SIGMA_RTOL) - 1e-9 <= s_norm <= smax * (1 + SIGMA_RTOL) + 1e-9):
            fails.append((tag, "sigma outside the census range", c, s_norm, smin, smax, n_sure, n_amb))
        if not (n_sure <= int(gpu_stats[c, 3]) <= n_sure + n_amb):
            fails.append((tag, "non-zero dd_1 count outside the census", c, int(gpu_stats[c, 3]), n_sure, n_amb))
        sig.append(s_norm)
    f = 255.0 * W.denoise_wavelet(xf, wavelet, levels, sigmas=sig)
    want = f.astype(np.int32).astype(np.uint8)          # the reference's ->u8 (values are in [0, 255])
    d = np.abs(got.astype(int) - want.astype(int))
    if d.max() > 1:
        fails.append((tag, "more than 1 LSB", int(d.max()), int((d > 1).sum())))
    away = np.abs(f - np.rint(f)) > AWAY
    if away.any() and (d[away] > 0).mean() > FRAC_AWAY:
        fails.append((tag, "1-LSB differences away from integers", float((d[away] > 0).mean())))
    return fails


def run_and_check(api, torch, cuda, img, spec, x, fields_dev=None, wavelet="bior1.5", levels=0, tag="", **kw):
    """One image through the C ABI (one plan, so the statistics hook sees this run), then `stagewise`."""
    h, w, _ = img.shape
    plan = api.get_plan(1, h, w, spec, cuda.index, wavelet=wavelet, levels=levels, **kw)
    got = plan.run(torch.from_numpy(np.ascontiguousarray(img[None])).to(cuda), fields=fields_dev).cpu().numpy()[0]
    st = plan.wavelet_stats(0)
    return got, stagewise(got, x, st, wavelet, levels or None, tag or spec)
