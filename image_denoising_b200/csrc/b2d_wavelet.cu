// b2d_wavelet.cu - skimage.restoration.denoise_wavelet(x, method='BayesShrink',
// mode='soft', wavelet in {'bior1.5' (live files), 'db1' (backups, BASELINE C4)},
// [wavelet_levels=3,] multichannel=True, convert2ycbcr=True) followed by the
// reference's (255 * out).astype(np.uint8)   (lib/roi_data_layer/minibatch.py:92-96,
// lib/model/test.py_bkp:1913-1916).  Algorithm restated in oracle/wavelet.py.
//
// General path (any wavelet / level count / odd sizes), fp32 planes in the
// caller's workspace, float64 only for the energy sums:
//   1. noise (float, unquantised where the reference keeps floats) -> "YCbCr" of the
//      array's own channel order (quirk Q7) -> 3 planes + per-channel min / max
//   2. L levels of separable DWT, pywt mode 'symmetric'
//   3. sigma = median(|dd_1| != 0) / 0.6745 by 3-pass radix select; per band energy
//   4. inverse DWT with soft thresholding applied while reading the detail bands
//   5. clip to the channel's [min, max], YCbCr -> RGB, clip [0, 1], * 255, truncate
// The per-channel normalisation to [0, 1] is affine and the DWT is linear, so
// it is folded into the thresholds (thr_un = var / sqrt(max(dvar - var, eps * s^2)),
// s = max - min) and the clip range instead of being a pass of its own.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
#include "b2d_wavelet.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace b2d {

// ---------------------------------------------------------------- geometry --
constexpr double kSd = 0.7071067811865476, kC1d = 0.01657281518405971, kC2d = 0.12153397801643787;
static const double kS = kSd, kC1 = kC1d, kC2 = kC2d;

void wavelet_geometry(WvGeom& g, int wavelet, int levels_req, int h, int w) {
  g.F = wavelet == B2D_WAVELET_DB1 ? 2 : 10;
  const double db1[4][2] = {{kS, kS}, {-kS, kS}, {kS, kS}, {kS, -kS}};
  const double bior[4][10] = {
      {kC1, -kC1, -kC2, kC2, kS, kS, kC2, -kC2, -kC1, kC1},
      {0, 0, 0, 0, -kS, kS, 0, 0, 0, 0},
      {0, 0, 0, 0, kS, kS, 0, 0, 0, 0},
      {kC1, kC1, -kC2, -kC2, kS, -kS, kC2, kC2, -kC1, -kC1}};
  for (int k = 0; k < 4; ++k)
    for (int j = 0; j < WV_MAXF; ++j) {
      const double v = j < g.F ? (g.F == 2 ? db1[k][j] : bior[k][j]) : 0.0;
      (k == 0 ? g.dec_lo : k == 1 ? g.dec_hi : k == 2 ? g.rec_lo : g.rec_hi)[j] = (float)v;
    }
  int L = levels_req;
  if (L <= 0) {   // skimage: max(min(dwt_max_level(s, F)) - 3, 1)
    auto maxlev = [&](int n) { return n < g.F - 1 ? 0 : (int)std::floor(std::log2((double)n / (g.F - 1.0))); };
    L = std::max(std::min(maxlev(h), maxlev(w)) - 3, 1);
  }
  g.L = std::min(L, WV_MAXL);
  g.h[0] = h; g.w[0] = w;
  for (int l = 1; l <= g.L; ++l) {
    g.h[l] = (g.h[l - 1] + g.F - 1) / 2;
    g.w[l] = (g.w[l - 1] + g.F - 1) / 2;
  }
  // per-channel float offsets
  const size_t HW = (size_t)h * w;
  const size_t T = g.L >= 1 ? std::max((size_t)h * g.w[1], (size_t)g.h[1] * w) : HW;   // row-pass temporaries, both directions
  size_t off = 0;
  g.off_plane = off; off += HW;
  g.off_tlo = off;   off += T;
  g.off_thi = off;   off += T;
  for (int l = 1; l <= g.L; ++l) {
    const size_t S = (size_t)g.h[l] * g.w[l];
    g.off_app[l] = off; off += S;
    for (int b = 0; b < 3; ++b) { g.off_det[l][b] = off; off += S; }
  }
  g.chan_floats = (off + 3) & ~(size_t)3;
}

// skimage.color.rgb2ycbcr matrix (BT.601 studio swing), applied to channels in array order
__constant__ float kYCC[3][3] = {{65.481f, 128.553f, 24.966f}, {-37.797f, -74.203f, 112.0f}, {112.0f, -93.786f, -18.214f}};

// ------------------------------------------------- 1. noise + colour + min/max
template <int KIND, int MODE>
__global__ void __launch_bounds__(256) k_wv_color(ImageSet s, NoiseArgs na, WvGeom g, float* ws, WvStats* stats,
                                                  int chunk_first) {
  const int z = blockIdx.z;                                   // image inside the chunk
  const int img_local = image_of(s, chunk_first + z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  float* base = ws + (size_t)z * 3 * g.chan_floats;
  float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
  // 24 bytes = 8 pixels = 3 noise groups per thread iteration
  const uint64_t npix = (uint64_t)s.h * s.w;
  for (uint64_t p8 = (uint64_t)blockIdx.x * 256 + threadIdx.x; p8 * 8 < npix; p8 += (uint64_t)gridDim.x * 256) {
    float xf[24];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const uint32_t grp = (uint32_t)(p8 * 3 + q);
      uint32_t v[8];
      float f[8];
      const uint64_t e0 = (uint64_t)grp * 8;
      if (e0 + 8 <= s.elems && (reinterpret_cast<uintptr_t>(img + e0) & 7u) == 0) {
        const uint2 w8 = __ldg(reinterpret_cast<const uint2*>(img + e0));
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          v[i] = (w8.x >> (8 * i)) & 0xffu;
          v[4 + i] = (w8.y >> (8 * i)) & 0xffu;
        }
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = e0 + i < s.elems ? (uint32_t)__ldg(img + e0 + i) : 0u;
      }
      GroupCtx gc{(uint32_t)img_local, grp, img};
      apply8_float<KIND, MODE>(na, gc, v, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) xf[8 * q + i] = f[i] * (1.0f / 255.0f);
    }
    // a thread owns 8 consecutive pixels: colour transform in registers, then two 16-byte stores per
    // plane (scalar stores touched 24 x 32 sectors per warp step and ran at 1.2 TB/s)
    float yv[3][8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float c0 = xf[3 * k], c1 = xf[3 * k + 1], c2 = xf[3 * k + 2];
#pragma unroll
      for (int c = 0; c < 3; ++c)
        yv[c][k] = kYCC[c][0] * c0 + kYCC[c][1] * c1 + kYCC[c][2] * c2 + (c == 0 ? 16.0f : 128.0f);
    }
    const uint64_t pix0 = p8 * 8;
    if (pix0 + 8 <= npix && (reinterpret_cast<uintptr_t>(base + g.off_plane) & 15u) == 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float4* dst = reinterpret_cast<float4*>(base + (size_t)c * g.chan_floats + g.off_plane + pix0);
        dst[0] = make_float4(yv[c][0], yv[c][1], yv[c][2], yv[c][3]);
        dst[1] = make_float4(yv[c][4], yv[c][5], yv[c][6], yv[c][7]);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          mn[c] = fminf(mn[c], yv[c][k]);
          mx[c] = fmaxf(mx[c], yv[c][k]);
        }
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (pix0 + k >= npix) break;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          base[(size_t)c * g.chan_floats + g.off_plane + pix0 + k] = yv[c][k];
          mn[c] = fminf(mn[c], yv[c][k]);
          mx[c] = fmaxf(mx[c], yv[c][k]);
        }
      }
    }
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&stats[z * 3 + c].mn_ord, f2ord(mn[c]));
      atomicMax(&stats[z * 3 + c].mx_ord, f2ord(mx[c]));
    }
  }
}

// ------------------------------------------------------------ 2. forward DWT
// Products and sums are rounded separately (no FMA): pywt / NumPy do the same, and it
// keeps coefficients that cancel exactly (duplicated border samples, flat regions)
// exactly zero - _sigma_est_dwt drops exact zeros before taking the median.
// The high-pass taps keep that rule.  The low-pass taps of the long filter (bior1.5: 10 taps, F > 2) accumulate
// with FMA - half the instructions, one rounding less per tap, and no output that must be an exact
// zero depends on them (dd, the band _sigma_est_dwt filters, is high-pass in both directions); db1 keeps
// separate roundings so that the general path and the Haar fast path (b2d_haar.cu) agree.
__device__ __forceinline__ float lo_acc(int F, float acc, float c, float x) {
  return F > 2 ? fmaf(c, x, acc) : __fadd_rn(acc, __fmul_rn(c, x));
}

// pywt 'symmetric' (half-sample) extension
__device__ __forceinline__ int sym_ext(int i, int n) {
  const int p = 2 * n;
  int m = i % p;
  if (m < 0) m += p;
  return m < n ? m : p - 1 - m;
}

// along x: in [h x w] -> lo, hi [h x wo];  out[i] = sum_j filt[j] * x[ext(2 i + 1 - j)]
__global__ void __launch_bounds__(256) k_dwt_rows(WvGeom g, float* ws, int level, size_t in_off) {
  const int h = g.h[level - 1], w = g.w[level - 1], wo = g.w[level];
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* in = base + in_off;
  float* lo = base + g.off_tlo;
  float* hi = base + g.off_thi;
  const int total = h * wo;
  for (int t = blockIdx.x * 256 + threadIdx.x; t < total; t += gridDim.x * 256) {
    const int y = t / wo, i = t - y * wo;
    const float* row = in + (size_t)y * w;
    float a = 0.f, d = 0.f;
    if (2 * i + 1 - (g.F - 1) >= 0 && 2 * i + 1 < w) {
#pragma unroll
      for (int j = 0; j < WV_MAXF; ++j)
        if (j < g.F) { const float x = row[2 * i + 1 - j]; a = lo_acc(g.F, a, g.dec_lo[j], x); d = __fadd_rn(d, __fmul_rn(g.dec_hi[j], x)); }
    } else {
      for (int j = 0; j < g.F; ++j) {
        const float x = row[sym_ext(2 * i + 1 - j, w)];
        a = lo_acc(g.F, a, g.dec_lo[j], x);
        d = __fadd_rn(d, __fmul_rn(g.dec_hi[j], x));
      }
    }
    lo[t] = a;
    hi[t] = d;
  }
}

// along y: lo, hi [h x wo] -> aa (app), da, ad, dd [ho x wo]
__global__ void __launch_bounds__(256) k_dwt_cols(WvGeom g, float* ws, int level) {
  const int h = g.h[level - 1], wo = g.w[level], ho = g.h[level];
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* lo = base + g.off_tlo;
  const float* hi = base + g.off_thi;
  float* aa = base + g.off_app[level];
  float* da = base + g.off_det[level][0];     // detail along axis 0 (y), approx along x
  float* ad = base + g.off_det[level][1];     // approx along y, detail along x
  float* dd = base + g.off_det[level][2];
  const int total = ho * wo;
  for (int t = blockIdx.x * 256 + threadIdx.x; t < total; t += gridDim.x * 256) {
    const int i = t / wo, x = t - i * wo;
    float s_aa = 0.f, s_da = 0.f, s_ad = 0.f, s_dd = 0.f;
    for (int j = 0; j < g.F; ++j) {
      const int y = sym_ext(2 * i + 1 - j, h);
      const float l = lo[(size_t)y * wo + x], hv = hi[(size_t)y * wo + x];
      s_aa = lo_acc(g.F, s_aa, g.dec_lo[j], l);
      s_da = __fadd_rn(s_da, __fmul_rn(g.dec_hi[j], l));
      s_ad = lo_acc(g.F, s_ad, g.dec_lo[j], hv);
      s_dd = __fadd_rn(s_dd, __fmul_rn(g.dec_hi[j], hv));
    }
    aa[t] = s_aa; da[t] = s_da; ad[t] = s_ad; dd[t] = s_dd;
  }
}

// ---- specialised versions (the ones that run): filter taps are compile-time constants, zero taps
// (8 of the 10 in bior1.5's dec_hi / rec_lo) cost nothing, a thread produces 4 consecutive outputs
// along the filtered axis from a register window (F + 6 inputs instead of 4 F), 2-D grids (no
// integer division).  Accumulation order per output is that of the generic kernels above, which
// stay as the executable definition (B2D_WV_GENERIC=1 runs them; the tests compare both).
template <int WAVE> struct WvCoef;
template <> struct WvCoef<B2D_WAVELET_DB1> {
  static constexpr int F = 2;
  static __host__ __device__ constexpr float dec_lo(int j) { return (float)kSd; }
  static __host__ __device__ constexpr float dec_hi(int j) { return j == 0 ? -(float)kSd : (float)kSd; }
  static __host__ __device__ constexpr float rec_lo(int j) { return (float)kSd; }
  static __host__ __device__ constexpr float rec_hi(int j) { return j == 0 ? (float)kSd : -(float)kSd; }
};
template <> struct WvCoef<B2D_WAVELET_BIOR15> {
  static constexpr int F = 10;
  static __host__ __device__ constexpr float dec_lo(int j) {
    return j == 0 ? (float)kC1d : j == 1 ? -(float)kC1d : j == 2 ? -(float)kC2d : j == 3 ? (float)kC2d : j == 4 ? (float)kSd
         : j == 5 ? (float)kSd : j == 6 ? (float)kC2d : j == 7 ? -(float)kC2d : j == 8 ? -(float)kC1d : (float)kC1d;
  }
  static __host__ __device__ constexpr float dec_hi(int j) { return j == 4 ? -(float)kSd : j == 5 ? (float)kSd : 0.0f; }
  static __host__ __device__ constexpr float rec_lo(int j) { return j == 4 || j == 5 ? (float)kSd : 0.0f; }
  static __host__ __device__ constexpr float rec_hi(int j) {
    return j == 0 ? (float)kC1d : j == 1 ? (float)kC1d : j == 2 ? -(float)kC2d : j == 3 ? -(float)kC2d : j == 4 ? (float)kSd
         : j == 5 ? -(float)kSd : j == 6 ? (float)kC2d : j == 7 ? (float)kC2d : j == 8 ? -(float)kC1d : -(float)kC1d;
  }
};
constexpr int WV_RO = 4;                                       // outputs per thread along the filtered axis

template <int WAVE>
__global__ void __launch_bounds__(256) k_dwt_rows_t(WvGeom g, float* ws, int level, size_t in_off) {
  using C = WvCoef<WAVE>;
  constexpr int F = C::F, NIN = F + 2 * (WV_RO - 1);
  const int w = g.w[level - 1], wo = g.w[level], y = blockIdx.y;
  const int i0 = (blockIdx.x * 256 + threadIdx.x) * WV_RO;
  if (i0 >= wo) return;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* row = base + in_off + (size_t)y * w;
  const int xlo = 2 * i0 + 2 - F;                              // x[2 i + 1 - j], j = 0 .. F-1, i = i0 .. i0 + 3
  float v[NIN];
  if (xlo >= 0 && xlo + NIN <= w) {
    if (NIN % 4 == 0 && (reinterpret_cast<uintptr_t>(row + xlo) & 15u) == 0) {    // 16 floats as 4 x 16-byte loads
#pragma unroll
      for (int e = 0; e < NIN / 4; ++e) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(row + xlo) + e);
        v[4 * e] = q.x; v[4 * e + 1] = q.y; v[4 * e + 2] = q.z; v[4 * e + 3] = q.w;
      }
    } else {
#pragma unroll
      for (int e = 0; e < NIN; ++e) v[e] = row[xlo + e];
    }
  } else {
#pragma unroll
    for (int e = 0; e < NIN; ++e) v[e] = row[sym_ext(xlo + e, w)];
  }
  float* lo = base + g.off_tlo + (size_t)y * wo;
  float* hi = base + g.off_thi + (size_t)y * wo;
#pragma unroll
  for (int o = 0; o < WV_RO; ++o) {
    if (i0 + o >= wo) break;
    float a = 0.f, d = 0.f;
#pragma unroll
    for (int j = 0; j < F; ++j) {
      const float x = v[2 * o + F - 1 - j];
      if (C::dec_lo(j) != 0.0f) a = lo_acc(F, a, C::dec_lo(j), x);
      if (C::dec_hi(j) != 0.0f) d = __fadd_rn(d, __fmul_rn(C::dec_hi(j), x));
    }
    lo[i0 + o] = a;
    hi[i0 + o] = d;
  }
}

template <int WAVE>
__global__ void __launch_bounds__(256) k_dwt_cols_t(WvGeom g, float* ws, int level) {
  using C = WvCoef<WAVE>;
  constexpr int F = C::F, NIN = F + 2 * (WV_RO - 1);
  const int h = g.h[level - 1], wo = g.w[level], ho = g.h[level];
  const int x = blockIdx.x * 256 + threadIdx.x, i0 = blockIdx.y * WV_RO;
  if (x >= wo) return;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* lo = base + g.off_tlo + x;
  const float* hi = base + g.off_thi + x;
  const int ylo = 2 * i0 + 2 - F;
  float vl[NIN], vh[NIN];
  const bool inside = ylo >= 0 && ylo + NIN <= h;
#pragma unroll
  for (int e = 0; e < NIN; ++e) {
    const int yy = inside ? ylo + e : sym_ext(ylo + e, h);
    vl[e] = lo[(size_t)yy * wo];
    vh[e] = hi[(size_t)yy * wo];
  }
  float* aa = base + g.off_app[level] + x;
  float* da = base + g.off_det[level][0] + x;
  float* ad = base + g.off_det[level][1] + x;
  float* dd = base + g.off_det[level][2] + x;
#pragma unroll
  for (int o = 0; o < WV_RO; ++o) {
    if (i0 + o >= ho) break;
    float s_aa = 0.f, s_da = 0.f, s_ad = 0.f, s_dd = 0.f;
#pragma unroll
    for (int j = 0; j < F; ++j) {
      const float l = vl[2 * o + F - 1 - j], hv = vh[2 * o + F - 1 - j];
      if (C::dec_lo(j) != 0.0f) { s_aa = lo_acc(F, s_aa, C::dec_lo(j), l); s_ad = lo_acc(F, s_ad, C::dec_lo(j), hv); }
      if (C::dec_hi(j) != 0.0f) { s_da = __fadd_rn(s_da, __fmul_rn(C::dec_hi(j), l)); s_dd = __fadd_rn(s_dd, __fmul_rn(C::dec_hi(j), hv)); }
    }
    const size_t r = (size_t)(i0 + o) * wo;
    aa[r] = s_aa; da[r] = s_da; ad[r] = s_ad; dd[r] = s_dd;
  }
}

// one axis of the inverse for outputs k0 .. k0 + 3 (k0 = 4 q): inputs i = 2 q + e, e = 0 .. F/2; output k0 + o uses
// taps j = (o & 1), +2, ... with e = (o + F - 2 - j) / 2; out-of-range inputs are loaded as 0 (the generic code skips them)
template <int WAVE>
__device__ __forceinline__ float idwt4(const float (&a)[WvCoef<WAVE>::F / 2 + 1], const float (&d)[WvCoef<WAVE>::F / 2 + 1], int o) {
  using C = WvCoef<WAVE>;
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < C::F; ++j) {
    if ((j & 1) != (o & 1)) continue;
    const int e = (o + C::F - 2 - j) / 2;
    if (C::rec_lo(j) != 0.0f) acc = fmaf(C::rec_lo(j), a[e], acc);
    if (C::rec_hi(j) != 0.0f) acc = fmaf(C::rec_hi(j), d[e], acc);
  }
  return acc;
}

template <int WAVE>
__global__ void __launch_bounds__(256) k_idwt_rows_t(WvGeom g, float* ws, const WvStats* stats, int level) {
  constexpr int NE = WvCoef<WAVE>::F / 2 + 1;
  const int wl = g.w[level], wr = g.w[level - 1], y = blockIdx.y;
  const int q = blockIdx.x * 256 + threadIdx.x, k0 = q * WV_RO;
  if (k0 >= wr) return;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const size_t r = (size_t)y * wl;
  const float* app = base + g.off_app[level] + r;
  const float* da = base + g.off_det[level][0] + r;
  const float* ad = base + g.off_det[level][1] + r;
  const float* dd = base + g.off_det[level][2] + r;
  const float t_da = stats[blockIdx.z].thr[level][0], t_ad = stats[blockIdx.z].thr[level][1],
              t_dd = stats[blockIdx.z].thr[level][2];
  float va[NE], vad[NE], vda[NE], vdd[NE];
#pragma unroll
  for (int e = 0; e < NE; ++e) {
    const int i = 2 * q + e;
    const bool ok = i < wl;
    va[e] = ok ? app[i] : 0.f;
    vad[e] = ok ? wv_soft(ad[i], t_ad) : 0.f;
    vda[e] = ok ? wv_soft(da[i], t_da) : 0.f;
    vdd[e] = ok ? wv_soft(dd[i], t_dd) : 0.f;
  }
  float* lo = base + g.off_tlo + (size_t)y * wr;
  float* hi = base + g.off_thi + (size_t)y * wr;
#pragma unroll
  for (int o = 0; o < WV_RO; ++o) {
    if (k0 + o >= wr) break;
    lo[k0 + o] = idwt4<WAVE>(va, vad, o);
    hi[k0 + o] = idwt4<WAVE>(vda, vdd, o);
  }
}

template <int WAVE>
__global__ void __launch_bounds__(256) k_idwt_cols_t(WvGeom g, float* ws, int level) {
  constexpr int NE = WvCoef<WAVE>::F / 2 + 1;
  const int hl = g.h[level], wr = g.w[level - 1], hr = g.h[level - 1];
  const int x = blockIdx.x * 256 + threadIdx.x, q = blockIdx.y, k0 = q * WV_RO;
  if (x >= wr) return;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* lo = base + g.off_tlo + x;
  const float* hi = base + g.off_thi + x;
  float vl[NE], vh[NE];
#pragma unroll
  for (int e = 0; e < NE; ++e) {
    const int i = 2 * q + e;
    const bool ok = i < hl;
    vl[e] = ok ? lo[(size_t)i * wr] : 0.f;
    vh[e] = ok ? hi[(size_t)i * wr] : 0.f;
  }
  float* out = base + (level == 1 ? g.off_plane : g.off_app[level - 1]) + x;
#pragma unroll
  for (int o = 0; o < WV_RO; ++o) {
    if (k0 + o >= hr) break;
    out[(size_t)(k0 + o) * wr] = idwt4<WAVE>(vl, vh, o);
  }
}

// ---- fused 2-D versions (the ones that run): rows and columns of one level in ONE kernel, the row
// pass result lives in shared memory instead of the tlo / thi planes (a level cost read 1 + write 1 +
// read 1 + write 1 planes + a read of the three detail bands for their energies; now read 1 + write 1).
// Arithmetic, operand order and rounding are those of k_dwt_rows_t / k_dwt_cols_t / k_idwt_rows_t /
// k_idwt_cols_t above (B2D_WV_FUSED=0 runs those; tests/test_gpu_wavelet.py compares the two paths).
constexpr int WV_FOY = 16, WV_FOX = 64;                        // forward: coefficient tile per CTA
constexpr int WV_IY = 32, WV_IX = 128;                         // inverse: output tile per CTA

template <int WAVE>
__global__ void __launch_bounds__(256) k_dwt2_t(WvGeom g, float* ws, WvStats* stats, int level, size_t in_off) {
  using C = WvCoef<WAVE>;
  constexpr int F = C::F, NIN = F + 2 * (WV_RO - 1), RIN = 2 * WV_FOY + F - 2;
  __shared__ __align__(16) float s_lo[RIN][WV_FOX];
  __shared__ __align__(16) float s_hi[RIN][WV_FOX];
  __shared__ double s_e[8][3];
  const int h = g.h[level - 1], w = g.w[level - 1], ho = g.h[level], wo = g.w[level];
  const int ox0 = blockIdx.x * WV_FOX, oy0 = blockIdx.y * WV_FOY;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* in = base + in_off;
  const int nout_y = min(WV_FOY, ho - oy0);
  const int rows = 2 * nout_y + F - 2;                         // input rows 2 oy0 + 2 - F ... of this tile
  const int ylo = 2 * oy0 + 2 - F;
  // 1. along x into shared memory: a thread keeps its group of 4 coefficients (q) and walks the input rows
  // r = tid / 16, + 16, ...; everything that depends on q only is computed once
  const int q = threadIdx.x % (WV_FOX / WV_RO), i0 = ox0 + q * WV_RO;
  const int xlo = 2 * i0 + 2 - F;
  const bool x_inside = xlo >= 0 && xlo + NIN <= w;
  const bool one_fold = h >= F;                                // |overshoot| <= F - 2 < h: a single reflection
  for (int r = threadIdx.x / (WV_FOX / WV_RO); r < rows && i0 < wo; r += 256 / (WV_FOX / WV_RO)) {
    int yy = ylo + r;
    if (one_fold) yy = yy < 0 ? -1 - yy : (yy >= h ? 2 * h - 1 - yy : yy);
    else yy = sym_ext(yy, h);
    const float* row = in + (unsigned)(yy * w);
    float v[NIN];
    if (x_inside) {
      if (NIN % 4 == 0 && (reinterpret_cast<uintptr_t>(row + xlo) & 15u) == 0) {
#pragma unroll
        for (int e = 0; e < NIN / 4; ++e) {
          const float4 t4 = __ldg(reinterpret_cast<const float4*>(row + xlo) + e);
          v[4 * e] = t4.x; v[4 * e + 1] = t4.y; v[4 * e + 2] = t4.z; v[4 * e + 3] = t4.w;
        }
      } else {
#pragma unroll
        for (int e = 0; e < NIN; ++e) v[e] = row[xlo + e];
      }
    } else {
#pragma unroll
      for (int e = 0; e < NIN; ++e) v[e] = row[sym_ext(xlo + e, w)];
    }
    float a[WV_RO], d[WV_RO];
#pragma unroll
    for (int o = 0; o < WV_RO; ++o) {
      a[o] = 0.f; d[o] = 0.f;
#pragma unroll
      for (int j = 0; j < F; ++j) {
        const float x = v[2 * o + F - 1 - j];
        if (C::dec_lo(j) != 0.0f) a[o] = lo_acc(F, a[o], C::dec_lo(j), x);
        if (C::dec_hi(j) != 0.0f) d[o] = __fadd_rn(d[o], __fmul_rn(C::dec_hi(j), x));
      }
    }
    *reinterpret_cast<float4*>(&s_lo[r][q * WV_RO]) = make_float4(a[0], a[1], a[2], a[3]);
    *reinterpret_cast<float4*>(&s_hi[r][q * WV_RO]) = make_float4(d[0], d[1], d[2], d[3]);
  }
  __syncthreads();
  // 2. along y out of shared memory: thread = (column, group of 4 coefficient rows)
  const int xl = threadIdx.x % WV_FOX, grp = threadIdx.x / WV_FOX, x = ox0 + xl;
  double e_da = 0.0, e_ad = 0.0, e_dd = 0.0;
  if (x < wo && grp * WV_RO < nout_y) {
    float vl[NIN], vh[NIN];
#pragma unroll
    for (int e = 0; e < NIN; ++e) {
      vl[e] = s_lo[2 * grp * WV_RO + e][xl];                   // rows past `rows` hold stale data and feed only
      vh[e] = s_hi[2 * grp * WV_RO + e][xl];                   // outputs i >= ho, which are dropped below
    }
    float* aa = base + g.off_app[level] + x;
    float* da = base + g.off_det[level][0] + x;
    float* ad = base + g.off_det[level][1] + x;
    float* dd = base + g.off_det[level][2] + x;
#pragma unroll
    for (int o = 0; o < WV_RO; ++o) {
      const int i = oy0 + grp * WV_RO + o;
      if (i >= ho) break;
      float s_aa = 0.f, s_da = 0.f, s_ad = 0.f, s_dd = 0.f;
#pragma unroll
      for (int j = 0; j < F; ++j) {
        const float l = vl[2 * o + F - 1 - j], hv = vh[2 * o + F - 1 - j];
        if (C::dec_lo(j) != 0.0f) { s_aa = lo_acc(F, s_aa, C::dec_lo(j), l); s_ad = lo_acc(F, s_ad, C::dec_lo(j), hv); }
        if (C::dec_hi(j) != 0.0f) { s_da = __fadd_rn(s_da, __fmul_rn(C::dec_hi(j), l)); s_dd = __fadd_rn(s_dd, __fmul_rn(C::dec_hi(j), hv)); }
      }
      const unsigned r = (unsigned)(i * wo);
      aa[r] = s_aa; da[r] = s_da; ad[r] = s_ad; dd[r] = s_dd;
      e_da += (double)s_da * s_da;
      e_ad += (double)s_ad * s_ad;
      e_dd += (double)s_dd * s_dd;
    }
  }
  // band energies (k_band_energy's sums, taken while the coefficients are in registers)
  for (int o = 16; o; o >>= 1) {
    e_da += __shfl_xor_sync(0xffffffffu, e_da, o);
    e_ad += __shfl_xor_sync(0xffffffffu, e_ad, o);
    e_dd += __shfl_xor_sync(0xffffffffu, e_dd, o);
  }
  if ((threadIdx.x & 31) == 0) { s_e[threadIdx.x >> 5][0] = e_da; s_e[threadIdx.x >> 5][1] = e_ad; s_e[threadIdx.x >> 5][2] = e_dd; }
  __syncthreads();
  if (threadIdx.x < 3) {
    double t = 0.0;
    for (int k = 0; k < 8; ++k) t += s_e[k][threadIdx.x];
    atomicAdd(&stats[blockIdx.z].energy[level][threadIdx.x], t);
  }
}

template <int WAVE>
__global__ void __launch_bounds__(256) k_idwt2_t(WvGeom g, float* ws, const WvStats* stats, int level) {
  constexpr int NE = WvCoef<WAVE>::F / 2 + 1, RC = WV_IY / 2 + NE - 2;
  __shared__ __align__(16) float s_lo[RC][WV_IX];
  __shared__ __align__(16) float s_hi[RC][WV_IX];
  const int hl = g.h[level], wl = g.w[level], hr = g.h[level - 1], wr = g.w[level - 1];
  const int kx0 = blockIdx.x * WV_IX, ky0 = blockIdx.y * WV_IY;
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float t_da = stats[blockIdx.z].thr[level][0], t_ad = stats[blockIdx.z].thr[level][1],
              t_dd = stats[blockIdx.z].thr[level][2];
  const int i_y0 = ky0 / 2;                                    // first coefficient row of the tile
  // 1. along x: (app, ad) -> lo, (da, dd) -> hi at the tile's coefficient rows
  for (int item = threadIdx.x; item < RC * (WV_IX / WV_RO); item += 256) {
    const int r = item / (WV_IX / WV_RO), ql = item - r * (WV_IX / WV_RO);
    const int k0 = kx0 + ql * WV_RO, y = i_y0 + r;
    float lo4[WV_RO] = {0.f, 0.f, 0.f, 0.f}, hi4[WV_RO] = {0.f, 0.f, 0.f, 0.f};
    if (k0 < wr && y < hl) {
      const int q = k0 / WV_RO;
      const size_t ro = (size_t)y * wl;
      const float* app = base + g.off_app[level] + ro;
      const float* da = base + g.off_det[level][0] + ro;
      const float* ad = base + g.off_det[level][1] + ro;
      const float* dd = base + g.off_det[level][2] + ro;
      float va[NE], vad[NE], vda[NE], vdd[NE];
#pragma unroll
      for (int e = 0; e < NE; ++e) {
        const int i = 2 * q + e;
        const bool ok = i < wl;
        va[e] = ok ? app[i] : 0.f;
        vad[e] = ok ? wv_soft(ad[i], t_ad) : 0.f;
        vda[e] = ok ? wv_soft(da[i], t_da) : 0.f;
        vdd[e] = ok ? wv_soft(dd[i], t_dd) : 0.f;
      }
#pragma unroll
      for (int o = 0; o < WV_RO; ++o) {
        lo4[o] = idwt4<WAVE>(va, vad, o);
        hi4[o] = idwt4<WAVE>(vda, vdd, o);
      }
    }
    *reinterpret_cast<float4*>(&s_lo[r][ql * WV_RO]) = make_float4(lo4[0], lo4[1], lo4[2], lo4[3]);
    *reinterpret_cast<float4*>(&s_hi[r][ql * WV_RO]) = make_float4(hi4[0], hi4[1], hi4[2], hi4[3]);
  }
  __syncthreads();
  // 2. along y: thread = (column, half of the tile's 8 groups of 4 output rows); rows y >= hl were stored as 0
  const int xl = threadIdx.x % WV_IX, half = threadIdx.x / WV_IX, x = kx0 + xl;
  if (x >= wr) return;
  float* out = base + (level == 1 ? g.off_plane : g.off_app[level - 1]) + x;
#pragma unroll
  for (int qq = 0; qq < WV_IY / WV_RO / 2; ++qq) {
    const int ql = half * (WV_IY / WV_RO / 2) + qq;            // group of 4 output rows inside the tile
    const int k0 = ky0 + ql * WV_RO;
    if (k0 >= hr) break;
    float vl[NE], vh[NE];
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      vl[e] = s_lo[2 * ql + e][xl];
      vh[e] = s_hi[2 * ql + e][xl];
    }
#pragma unroll
    for (int o = 0; o < WV_RO; ++o) {
      if (k0 + o >= hr) break;
      out[(size_t)(k0 + o) * wr] = idwt4<WAVE>(vl, vh, o);
    }
  }
}

// ------------------------------------------------------- 3. band statistics
__global__ void __launch_bounds__(256) k_band_energy(WvGeom g, const float* ws, WvStats* stats, int level) {
  const float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const int n = g.h[level] * g.w[level];
  double e[3] = {0, 0, 0};
  for (int t = blockIdx.x * 256 + threadIdx.x; t < n; t += gridDim.x * 256)
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      const float d = base[g.off_det[level][b] + t];
      e[b] += (double)d * d;
    }
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    for (int o = 16; o; o >>= 1) e[b] += __shfl_xor_sync(0xffffffffu, e[b], o);
    if ((threadIdx.x & 31) == 0) atomicAdd(&stats[blockIdx.z].energy[level][b], e[b]);
  }
}

// radix select over the bit patterns of |dd_1| (non-zero only): pass 0 bits 30..20,
// pass 1 bits 19..9, pass 2 bits 8..0; two ranks are tracked for even counts
__global__ void __launch_bounds__(256) k_select_hist(WvGeom g, const float* ws, const WvStats* stats,
                                                     unsigned int* hist, int pass) {
  __shared__ unsigned int sh[2][WV_BINS];
  for (int i = threadIdx.x; i < 2 * WV_BINS; i += 256) (&sh[0][0])[i] = 0u;
  __syncthreads();
  const float* dd = ws + (size_t)blockIdx.z * g.chan_floats + g.off_det[1][2];
  const int n = g.h[1] * g.w[1];
  const int shift = pass == 0 ? 20 : pass == 1 ? 9 : 0;
  const unsigned int mask = pass == 2 ? 0x1ffu : 0x7ffu;
  const unsigned int p0 = stats[blockIdx.z].prefix[0], p1 = stats[blockIdx.z].prefix[1];
  const int hshift = pass == 0 ? 31 : pass == 1 ? 20 : 9;     // bits above the current digit
  auto tally = [&](float v) {
    const unsigned int b = __float_as_uint(v) & 0x7fffffffu;
    if (b == 0u) return;
    const unsigned int digit = (b >> shift) & mask;
    const unsigned int high = pass == 0 ? 0u : (b >> hshift);
    if (pass == 0 || high == p0) atomicAdd(&sh[0][digit], 1u);
    if (pass != 0 && high == p1) atomicAdd(&sh[1][digit], 1u);
  };
  // 16-byte loads, two in flight per thread: the pass is a pure stream of the band
  const int head = min(n, (int)((16u - ((unsigned)reinterpret_cast<uintptr_t>(dd) & 15u)) & 15u) / 4);
  const int nvec = (n - head) / 4;
  const float4* dv = reinterpret_cast<const float4*>(dd + head);
  const int stride = gridDim.x * 256;
  int t = blockIdx.x * 256 + threadIdx.x;
  for (; t + stride < nvec; t += 2 * stride) {
    const float4 a = __ldg(dv + t), b = __ldg(dv + t + stride);
    tally(a.x); tally(a.y); tally(a.z); tally(a.w);
    tally(b.x); tally(b.y); tally(b.z); tally(b.w);
  }
  if (t < nvec) {
    const float4 a = __ldg(dv + t);
    tally(a.x); tally(a.y); tally(a.z); tally(a.w);
  }
  if (blockIdx.x == 0) {
    if ((int)threadIdx.x < head) tally(dd[threadIdx.x]);
    const int tail = head + nvec * 4 + threadIdx.x;
    if (tail < n) tally(dd[tail]);
  }
  __syncthreads();
  unsigned int* gh = hist + (size_t)blockIdx.z * 2 * WV_BINS;
  for (int i = threadIdx.x; i < 2 * WV_BINS; i += 256)
    if ((&sh[0][0])[i]) atomicAdd(&gh[i], (&sh[0][0])[i]);
}

// one CTA per plane: block-wide prefix sum over the digit histogram(s), pick the bin holding each rank
__global__ void __launch_bounds__(256) k_select_step(WvStats* stats, unsigned int* hist, int pass) {
  constexpr int PER = WV_BINS / 256;
  const int z = blockIdx.x, t = threadIdx.x, lane = t & 31, wid = t >> 5;
  WvStats& st = stats[z];
  unsigned int* h0 = hist + (size_t)z * 2 * WV_BINS;
  const int bins = pass == 2 ? 512 : WV_BINS;                  // digits of pass 2 are 9 bits wide: upper bins stay 0
  const int bits = pass == 2 ? 9 : 11;
  __shared__ unsigned long long s_warp[8], s_rank[2];
  __shared__ unsigned int s_sel[2];
  for (int r = 0; r < 2; ++r) {
    const unsigned int* hh = (pass == 0) ? h0 : h0 + r * WV_BINS;
    unsigned int loc[PER];
    unsigned long long sum = 0;
#pragma unroll
    for (int i = 0; i < PER; ++i) { loc[i] = hh[t * PER + i]; sum += loc[i]; }
    unsigned long long inc = sum;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += v;
    }
    __syncthreads();                                           // s_warp / s_sel of the previous rank are consumed
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    unsigned long long before = inc - sum, total = 0;
    for (int w = 0; w < 8; ++w) {
      if (w < wid) before += s_warp[w];
      total += s_warp[w];
    }
    unsigned long long k;
    if (pass == 0) k = r == 0 ? (total ? (total - 1) / 2 : 0) : total / 2;
    else k = st.rank[r];
    if (t == 0) { s_sel[r] = (unsigned int)(bins - 1); s_rank[r] = k - total; }   // no bin holds the rank (empty plane)
    __syncthreads();
    if (before <= k && k < before + sum) {                     // exactly one thread
      unsigned long long acc = before;
      int sel = 0;
#pragma unroll
      for (int i = 0; i < PER; ++i) {
        if (acc + loc[i] > k) { sel = t * PER + i; break; }
        acc += loc[i];
      }
      s_sel[r] = (unsigned int)sel;
      s_rank[r] = k - acc;
    }
    if (pass == 0 && r == 0 && t == 0) st.count = total;
  }
  __syncthreads();
  for (int i = t; i < 2 * WV_BINS; i += 256) h0[i] = 0u;       // ready for the next pass
  if (t == 0) {
    for (int r = 0; r < 2; ++r) {
      st.rank[r] = s_rank[r];
      st.prefix[r] = ((pass == 0 ? 0u : st.prefix[r]) << bits) | s_sel[r];
    }
    if (pass == 2) {
      const float v0 = __uint_as_float(st.prefix[0]), v1 = __uint_as_float(st.prefix[1]);
      const double med = st.count ? 0.5 * ((double)v0 + (double)v1) : 0.0;
      st.sigma = (float)(med / 0.6744897501960817);
    }
  }
}

// BayesShrink thresholds in un-normalised units (see header comment)
__global__ void k_thresholds(WvGeom g, WvStats* stats, int nplanes) {
  const int z = blockIdx.x * blockDim.x + threadIdx.x;
  if (z >= nplanes) return;
  WvStats& st = stats[z];
  st.lo = ord2f(st.mn_ord);
  st.hi = ord2f(st.mx_ord);
  const double s = (double)st.hi - (double)st.lo;
  const double var = (double)st.sigma * st.sigma;
  const double eps = 2.220446049250313e-16 * s * s;
  for (int l = 1; l <= g.L; ++l) {
    const double n = (double)g.h[l] * g.w[l];
    for (int b = 0; b < 3; ++b) {
      const double dvar = st.energy[l][b] / n;
      st.thr[l][b] = (float)(var / sqrt(fmax(dvar - var, eps)));
    }
  }
}

// ------------------------------------------------------------ 4. inverse DWT
__device__ __forceinline__ float soft(float d, float thr) { return wv_soft(d, thr); }

// one axis of pywt's idwt: out[k] = sum_i a[i] rec_lo[k + F - 2 - 2 i] + d[i] rec_hi[k + F - 2 - 2 i]
template <class LoadA, class LoadD>
__device__ __forceinline__ float idwt_tap(const WvGeom& g, int k, int m, LoadA la, LoadD ld) {
  float acc = 0.f;
  const int top = k + g.F - 2;
  for (int j = top & 1; j < g.F; j += 2) {
    const int i = (top - j) >> 1;
    if (i < 0 || i >= m) continue;
    acc = fmaf(g.rec_lo[j], la(i), acc);
    acc = fmaf(g.rec_hi[j], ld(i), acc);
  }
  return acc;
}

// along x: (app, ad) -> tlo, (da, dd) -> thi, each [h_l x wr], wr = w_{l-1}; details soft-thresholded on read
__global__ void __launch_bounds__(256) k_idwt_rows(WvGeom g, float* ws, const WvStats* stats, int level) {
  const int hl = g.h[level], wl = g.w[level], wr = g.w[level - 1];
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* app = base + g.off_app[level];
  const float* da = base + g.off_det[level][0];
  const float* ad = base + g.off_det[level][1];
  const float* dd = base + g.off_det[level][2];
  const float t_da = stats[blockIdx.z].thr[level][0], t_ad = stats[blockIdx.z].thr[level][1],
              t_dd = stats[blockIdx.z].thr[level][2];
  float* lo = base + g.off_tlo;
  float* hi = base + g.off_thi;
  const int total = hl * wr;
  for (int t = blockIdx.x * 256 + threadIdx.x; t < total; t += gridDim.x * 256) {
    const int y = t / wr, k = t - y * wr;
    const size_t r = (size_t)y * wl;
    lo[t] = idwt_tap(g, k, wl, [&](int i) { return app[r + i]; }, [&](int i) { return soft(ad[r + i], t_ad); });
    hi[t] = idwt_tap(g, k, wl, [&](int i) { return soft(da[r + i], t_da); }, [&](int i) { return soft(dd[r + i], t_dd); });
  }
}

// along y: tlo, thi [h_l x wr] -> out [hr x wr], hr = h_{l-1}; out = app[l-1] or the plane
__global__ void __launch_bounds__(256) k_idwt_cols(WvGeom g, float* ws, int level) {
  const int hl = g.h[level], wr = g.w[level - 1], hr = g.h[level - 1];
  float* base = ws + (size_t)blockIdx.z * g.chan_floats;
  const float* lo = base + g.off_tlo;
  const float* hi = base + g.off_thi;
  float* out = base + (level == 1 ? g.off_plane : g.off_app[level - 1]);
  const int total = hr * wr;
  for (int t = blockIdx.x * 256 + threadIdx.x; t < total; t += gridDim.x * 256) {
    const int k = t / wr, x = t - k * wr;
    out[t] = idwt_tap(g, k, hl, [&](int i) { return lo[(size_t)i * wr + x]; },
                      [&](int i) { return hi[(size_t)i * wr + x]; });
  }
}

// ----------------------------------------------------------- 5. back to uint8
struct WvFinalArgs { float inv[3][3]; };

__global__ void __launch_bounds__(256) k_wv_final(ImageSet s, WvGeom g, const float* ws, const WvStats* stats,
                                                  WvFinalArgs fa, int chunk_first) {
  const int z = blockIdx.z;
  const int img_local = image_of(s, chunk_first + z);
  const float* base = ws + (size_t)z * 3 * g.chan_floats;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  const float lo0 = stats[z * 3].lo, hi0 = stats[z * 3].hi, lo1 = stats[z * 3 + 1].lo, hi1 = stats[z * 3 + 1].hi,
              lo2 = stats[z * 3 + 2].lo, hi2 = stats[z * 3 + 2].hi;
  const uint64_t npix = (uint64_t)s.h * s.w;
  for (uint64_t p4 = (uint64_t)blockIdx.x * 256 + threadIdx.x; p4 * 4 < npix; p4 += (uint64_t)gridDim.x * 256) {
    uint32_t o[12];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint64_t pix = min(p4 * 4 + k, npix - 1);
      // inner (per-channel) denoise_wavelet clips to the normalised [0, 1] = [min, max] of the channel
      const float y = fminf(fmaxf(base[g.off_plane + pix], lo0), hi0) - 16.0f;
      const float cb = fminf(fmaxf(base[g.chan_floats + g.off_plane + pix], lo1), hi1) - 128.0f;
      const float cr = fminf(fmaxf(base[2 * g.chan_floats + g.off_plane + pix], lo2), hi2) - 128.0f;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float v = fa.inv[c][0] * y + fa.inv[c][1] * cb + fa.inv[c][2] * cr;
        v = fminf(fmaxf(v, 0.0f), 1.0f);
        o[3 * k + c] = (uint32_t)__float2int_rz(255.0f * v);
      }
    }
    uint8_t* op = out + p4 * 12;
    if (p4 * 4 + 4 <= npix && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
      uint32_t* ow = reinterpret_cast<uint32_t*>(op);
      ow[0] = pack4(o[0], o[1], o[2], o[3]);
      ow[1] = pack4(o[4], o[5], o[6], o[7]);
      ow[2] = pack4(o[8], o[9], o[10], o[11]);
    } else {
      for (int k = 0; k < 12 && p4 * 12 + k < s.elems; ++k) op[k] = (uint8_t)o[k];
    }
  }
}

__global__ void k_wv_init(WvStats* stats, int nplanes) {
  const int z = blockIdx.x * blockDim.x + threadIdx.x;
  if (z >= nplanes) return;
  WvStats& st = stats[z];
  st.mn_ord = 0x7fffffff;
  st.mx_ord = (int)0x80000000;
  for (int l = 0; l <= WV_MAXL; ++l)
    for (int b = 0; b < 3; ++b) { st.energy[l][b] = 0.0; st.thr[l][b] = 0.f; }
  st.prefix[0] = st.prefix[1] = 0u;
  st.rank[0] = st.rank[1] = 0ull;
  st.count = 0ull;
  st.sigma = 0.f;
}

// ------------------------------------------------------------------ host side
constexpr size_t WV_BUDGET = 6ull << 30;           // workspace cap: images are processed in chunks

static int wv_chunk(const WvGeom& g, int n) {
  const size_t per_img = 3 * g.chan_floats * sizeof(float);
  // (3 planes per image index grid.z: at most 21845 images per chunk)
  return (int)std::max<size_t>(1, std::min<size_t>(std::min<size_t>((size_t)n, 21845), WV_BUDGET / std::max<size_t>(per_img, 1)));
}

size_t wavelet_workspace_bytes(const b2d_plan& p) {
  size_t need = 0;
  for (const SpecGroup& grp : p.groups) {
    if (grp.spec.denoise != B2D_DENOISE_WAVELET) continue;
    WvGeom g;
    wavelet_geometry(g, grp.spec.wavelet, grp.spec.levels, p.h, p.w);
    const int chunk = wv_chunk(g, (int)grp.images.size());
    const size_t bytes = (size_t)chunk * 3 * (g.chan_floats * sizeof(float) + sizeof(WvStats) +
                                              2 * WV_BINS * sizeof(unsigned int)) + 1024;
    need = std::max(need, bytes);
  }
  return need;
}

void wavelet_stats_begin(WvStats* stats, unsigned int* hist, int planes, cudaStream_t st) {
  k_wv_init<<<(planes + 127) / 128, 128, 0, st>>>(stats, planes);
  cudaMemsetAsync(hist, 0, (size_t)planes * 2 * WV_BINS * sizeof(unsigned int), st);
  count_launch();
}

void wavelet_sigma_thresholds(const WvGeom& g, const float* ws, WvStats* stats, unsigned int* hist, int planes,
                              cudaStream_t st, int first_pass) {
  auto blocks = [](size_t n) { return (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8); };
  for (int pass = 0; pass < 3; ++pass) {
    if (pass >= first_pass) {                             // else: the caller's kernel has tallied this pass already
      k_select_hist<<<dim3(std::min(blocks((size_t)g.h[1] * g.w[1]), 32u), 1, planes), 256, 0, st>>>(g, ws, stats, hist, pass);
      count_launch();
    }
    k_select_step<<<planes, 256, 0, st>>>(stats, hist, pass);
    count_launch();
  }
  k_thresholds<<<(planes + 63) / 64, 64, 0, st>>>(g, stats, planes);
  count_launch();
}

template <int KIND>
static void launch_color(int mode, dim3 grid, cudaStream_t st, const ImageSet& s, const NoiseArgs& na, const WvGeom& g,
                         float* ws, WvStats* stats, int first) {
  if (mode == NM_INJECT) k_wv_color<KIND, NM_INJECT><<<grid, 256, 0, st>>>(s, na, g, ws, stats, first);
  else k_wv_color<KIND, NM_GEN><<<grid, 256, 0, st>>>(s, na, g, ws, stats, first);
}

static void invert3(const double m[3][3], double o[3][3]) {
  const double det = m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
                     m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      const int a = (j + 1) % 3, b = (j + 2) % 3, c = (i + 1) % 3, d = (i + 2) % 3;
      o[i][j] = (m[a][c] * m[b][d] - m[a][d] * m[b][c]) / det;
    }
}

int run_wavelet(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, int kind, int mode, const NoiseArgs& na,
                uint8_t* ws8, cudaStream_t st) {
  WvGeom g;
  wavelet_geometry(g, sp.wavelet, sp.levels, p.h, p.w);
  const int chunk = wv_chunk(g, s.count);
  uint8_t* wbase = ws8 + p.off_denoise + (p.ws_bytes - p.off_denoise - wavelet_workspace_bytes(p));
  float* ws = reinterpret_cast<float*>(wbase);
  WvStats* stats = reinterpret_cast<WvStats*>(wbase + (size_t)chunk * 3 * g.chan_floats * sizeof(float));
  unsigned int* hist = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(stats) + (size_t)chunk * 3 * sizeof(WvStats));
  const double M[3][3] = {{65.481, 128.553, 24.966}, {-37.797, -74.203, 112.0}, {112.0, -93.786, -18.214}};
  double inv[3][3];
  invert3(M, inv);
  WvFinalArgs fa;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) fa.inv[i][j] = (float)inv[i][j];
  auto blocks = [](size_t n) { return (unsigned)std::min<size_t>((n + 255) / 256, 148 * 8); };
  // the register-resident path loads / stores 8-byte words at img + multiples of 24: both batch bases must be 8-byte aligned
  const bool base8 = ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 7u) == 0;
  if (base8 && haar_supported(p, sp, mode)) return run_haar(p, s, kind, na, g, wbase, wavelet_workspace_bytes(p), fa.inv, st);

  for (int first = 0; first < s.count; first += chunk) {
    const int cnt = std::min(chunk, s.count - first);
    const int planes = cnt * 3;
    wavelet_stats_begin(stats, hist, planes, st);
    const dim3 cgrid(blocks(((size_t)p.h * p.w + 7) / 8), 1, cnt);
    switch (kind) {
      case NK_NONE: launch_color<NK_NONE>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_GAUSSIAN: launch_color<NK_GAUSSIAN>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_SPECKLE: launch_color<NK_SPECKLE>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_SAP: launch_color<NK_SAP>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_POISSON: launch_color<NK_POISSON>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_UNIFORM: launch_color<NK_UNIFORM>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_RAYLEIGH: launch_color<NK_RAYLEIGH>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_GAMMA: launch_color<NK_GAMMA>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_TABLE_ADD: launch_color<NK_TABLE_ADD>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      case NK_SHADER: launch_color<NK_SHADER>(mode, cgrid, st, s, na, g, ws, stats, first); break;
      default: return fail(B2D_ERR_INVALID, "wavelet: bad noise operator %d", kind);
    }
    count_launch();
    static const bool generic = [] { const char* e = knob("B2D_WV_GENERIC"); return e && e[0] == '1'; }();
    static const bool fused = [] { const char* e = knob("B2D_WV_FUSED"); return !(e && e[0] == '0'); }();
    const bool db1 = sp.wavelet == B2D_WAVELET_DB1;
    auto cdiv = [](int a, int b) { return (unsigned)((a + b - 1) / b); };
    for (int l = 1; l <= g.L; ++l) {
      const size_t in_off = l == 1 ? g.off_plane : g.off_app[l - 1];
      if (generic) {
        k_dwt_rows<<<dim3(blocks((size_t)g.h[l - 1] * g.w[l]), 1, planes), 256, 0, st>>>(g, ws, l, in_off);
        k_dwt_cols<<<dim3(blocks((size_t)g.h[l] * g.w[l]), 1, planes), 256, 0, st>>>(g, ws, l);
      } else if (fused) {
        const dim3 gf(cdiv(g.w[l], WV_FOX), cdiv(g.h[l], WV_FOY), planes);
        if (db1) k_dwt2_t<B2D_WAVELET_DB1><<<gf, 256, 0, st>>>(g, ws, stats, l, in_off);
        else k_dwt2_t<B2D_WAVELET_BIOR15><<<gf, 256, 0, st>>>(g, ws, stats, l, in_off);
        count_launch();
        continue;                                              // band energies are summed inside
      } else {
        const dim3 gr(cdiv(cdiv(g.w[l], WV_RO), 256), (unsigned)g.h[l - 1], planes);
        const dim3 gc(cdiv(g.w[l], 256), cdiv(g.h[l], WV_RO), planes);
        if (db1) {
          k_dwt_rows_t<B2D_WAVELET_DB1><<<gr, 256, 0, st>>>(g, ws, l, in_off);
          k_dwt_cols_t<B2D_WAVELET_DB1><<<gc, 256, 0, st>>>(g, ws, l);
        } else {
          k_dwt_rows_t<B2D_WAVELET_BIOR15><<<gr, 256, 0, st>>>(g, ws, l, in_off);
          k_dwt_cols_t<B2D_WAVELET_BIOR15><<<gc, 256, 0, st>>>(g, ws, l);
        }
      }
      k_band_energy<<<dim3(std::min(blocks((size_t)g.h[l] * g.w[l]), 64u), 1, planes), 256, 0, st>>>(g, ws, stats, l);
      count_launch(); count_launch(); count_launch();
    }
    wavelet_sigma_thresholds(g, ws, stats, hist, planes, st);
    for (int l = g.L; l >= 1; --l) {
      if (generic) {
        k_idwt_rows<<<dim3(blocks((size_t)g.h[l] * g.w[l - 1]), 1, planes), 256, 0, st>>>(g, ws, stats, l);
        k_idwt_cols<<<dim3(blocks((size_t)g.h[l - 1] * g.w[l - 1]), 1, planes), 256, 0, st>>>(g, ws, l);
      } else if (fused) {
        const dim3 gi(cdiv(g.w[l - 1], WV_IX), cdiv(g.h[l - 1], WV_IY), planes);
        if (db1) k_idwt2_t<B2D_WAVELET_DB1><<<gi, 256, 0, st>>>(g, ws, stats, l);
        else k_idwt2_t<B2D_WAVELET_BIOR15><<<gi, 256, 0, st>>>(g, ws, stats, l);
        count_launch();
        continue;
      } else {
        const dim3 gr(cdiv(cdiv(g.w[l - 1], WV_RO), 256), (unsigned)g.h[l], planes);
        const dim3 gc(cdiv(g.w[l - 1], 256), cdiv(g.h[l - 1], WV_RO), planes);
        if (db1) {
          k_idwt_rows_t<B2D_WAVELET_DB1><<<gr, 256, 0, st>>>(g, ws, stats, l);
          k_idwt_cols_t<B2D_WAVELET_DB1><<<gc, 256, 0, st>>>(g, ws, l);
        } else {
          k_idwt_rows_t<B2D_WAVELET_BIOR15><<<gr, 256, 0, st>>>(g, ws, stats, l);
          k_idwt_cols_t<B2D_WAVELET_BIOR15><<<gc, 256, 0, st>>>(g, ws, l);
        }
      }
      count_launch(); count_launch();
    }
    k_wv_final<<<dim3(blocks(((size_t)p.h * p.w + 3) / 4), 1, cnt), 256, 0, st>>>(s, g, ws, stats, fa, first);
    count_launch();
    p.wv_stats_dev = stats; p.wv_stats_first = first; p.wv_stats_count = cnt;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "wavelet launch");
  }
  return B2D_OK;
}

}  // namespace b2d

// Test hook: sigma / clip range / dd_1 threshold the last b2d_run left for one image of the last wavelet chunk.
extern "C" int b2d_wavelet_stats(const b2d_plan* plan, int image_in_group, float* h_out, void* cuda_stream) {
  using namespace b2d;
  if (!plan || !h_out) return fail(B2D_ERR_INVALID, "b2d_wavelet_stats: null argument");
  if (!plan->wv_stats_dev) return fail(B2D_ERR_INVALID, "b2d_wavelet_stats: no wavelet run on this plan yet");
  const int k = image_in_group - plan->wv_stats_first;
  if (k < 0 || k >= plan->wv_stats_count)
    return fail(B2D_ERR_INVALID, "b2d_wavelet_stats: image %d is not in the last chunk [%d, %d)", image_in_group,
                plan->wv_stats_first, plan->wv_stats_first + plan->wv_stats_count);
  DeviceGuard guard(plan->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  WvStats hs[3];
  cudaError_t e = cudaMemcpyAsync(hs, static_cast<const WvStats*>(plan->wv_stats_dev) + (size_t)k * 3, sizeof(hs),
                                  cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return fail_cuda(e, "b2d_wavelet_stats");
  for (int c = 0; c < 3; ++c) {
    h_out[4 * c + 0] = hs[c].sigma;
    h_out[4 * c + 1] = hs[c].lo;
    h_out[4 * c + 2] = hs[c].hi;
    h_out[4 * c + 3] = (float)hs[c].count;
  }
  return B2D_OK;
}

