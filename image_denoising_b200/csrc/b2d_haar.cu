// b2d_haar.cu - fast path of the BayesShrink wavelet denoiser for wavelet = db1 (Haar),
// wavelet_levels = 3 and H, W multiples of 8 (BASELINE config 4; the `db1, wavelet_levels=3`
// variant of the reference's backups, lib/model/test.py_bkp:1913-1916):
//     denoise_wavelet(x, method='BayesShrink', mode='soft', wavelet_levels=3, multichannel=True,
//                     convert2ycbcr=True);  im = (255 * out).astype(np.uint8)
// For these sizes every pywt coefficient depends on one 8 x 8 block only, so the whole
// transform -> threshold -> inverse chain stays in registers; nothing but the image, the finest
// diagonal band (for the median) and a few per-image statistics touches HBM:
//   pass A  k_haar_stats : noise (regenerated from the Philox counter) -> "YCbCr" -> 3-level Haar;
//                          per channel min / max, per band sum of squares, dd_1 band -> workspace
//   select               : 3-pass radix select of median(|dd_1| != 0) (kernels of b2d_wavelet.cu)
//   pass B  k_haar_apply : same forward transform, soft threshold, inverse, clip, colour back, ->u8
// HBM traffic: 3 B + 3 B (dd_1, fp32 at 1/4 resolution) in pass A, 3 x 3 B in the select, 3 + 3 B in
// pass B = 21 B / pixel against ~150 B / pixel of the general multi-pass path.
//
// Work split: a lane owns 8 pixels x 2 rows (two 24-byte units = 6 Philox groups); 4 lanes
// (lane & 3 = row pair) form one 8 x 8 block, a warp 8 blocks side by side (192 contiguous bytes
// per row).  Along x all three levels are thread-local; along y level 1 is local, level 2 / 3
// exchange partners with __shfl_xor(1) / (2).  Arithmetic mirrors the general kernels operation
// for operation (analysis: separate multiplies and adds; synthesis: fma), so both paths agree.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
#include "b2d_wavelet.cuh"

#include <algorithm>
#include <cstdlib>

namespace b2d {

constexpr float HS = 0.70710678118654752f;          // 1 / sqrt(2) as the float the general path uses

struct HaarCoef {                                     // one channel of one lane (8 px x 2 rows)
  float da[4], ad[4], dd[4];                          // level 1 (this lane's row of the 4 x 4 band)
  float l2[4];                                        // even lane: aa2[2], ad2[2]; odd lane: da2[2], dd2[2]
  float l3[2];                                        // r = 0: aa3, ad3;  r = 2: da3, dd3
};

// pywt dwt of the pair (even, odd): out = sum_j filt[j] * x[2 i + 1 - j], products and sums rounded separately
__device__ __forceinline__ void an(float e, float o, float& a, float& d) {
  a = __fadd_rn(__fmul_rn(HS, o), __fmul_rn(HS, e));
  d = __fadd_rn(__fmul_rn(-HS, o), __fmul_rn(HS, e));
}
// pywt idwt of (a, d) -> (even, odd) samples, accumulated with fma like idwt_tap
__device__ __forceinline__ void syn(float a, float d, float& e, float& o) {
  const float sa = __fmul_rn(HS, a);
  e = fmaf(HS, d, sa);
  o = fmaf(-HS, d, sa);
}

__device__ __forceinline__ void haar_forward(const float (&x0)[8], const float (&x1)[8], int r, HaarCoef& c) {
  float aa1[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float lo0, hi0, lo1, hi1;
    an(x0[2 * i], x0[2 * i + 1], lo0, hi0);
    an(x1[2 * i], x1[2 * i + 1], lo1, hi1);
    an(lo0, lo1, aa1[i], c.da[i]);
    an(hi0, hi1, c.ad[i], c.dd[i]);
  }
  float aa2[2];
  const bool odd = (r & 1) != 0;
#pragma unroll
  for (int j = 0; j < 2; ++j) {                       // level 2: rows r (even) and r + 1 (odd)
    float lo, hi;
    an(aa1[2 * j], aa1[2 * j + 1], lo, hi);
    const float plo = __shfl_xor_sync(0xffffffffu, lo, 1), phi = __shfl_xor_sync(0xffffffffu, hi, 1);
    float a, d, a2, d2;
    an(odd ? plo : lo, odd ? lo : plo, a, d);         // from the row-low band: aa / da
    an(odd ? phi : hi, odd ? hi : phi, a2, d2);       // from the row-high band: ad / dd
    aa2[j] = a;
    c.l2[j] = odd ? d : a;
    c.l2[2 + j] = odd ? d2 : a2;
  }
  {                                                   // level 3: rows r = 0 (even) and r = 2 (odd)
    float lo, hi;
    an(aa2[0], aa2[1], lo, hi);
    const float plo = __shfl_xor_sync(0xffffffffu, lo, 2), phi = __shfl_xor_sync(0xffffffffu, hi, 2);
    const bool low = (r & 2) != 0;
    float a, d, a2, d2;
    an(low ? plo : lo, low ? lo : plo, a, d);
    an(low ? phi : hi, low ? hi : phi, a2, d2);
    c.l3[0] = low ? d : a;
    c.l3[1] = low ? d2 : a2;
  }
}

__device__ __forceinline__ float soft_thr(float d, float thr) { return wv_soft(d, thr); }

// thr[level - 1][band], band 0 = da, 1 = ad, 2 = dd
__device__ __forceinline__ void haar_inverse(const HaarCoef& c, int r, const float (&thr)[3][3], float (&x0)[8],
                                             float (&x1)[8]) {
  const bool odd = (r & 1) != 0, low = (r & 2) != 0;
  float aa2[2];
  {                                                   // level 3 -> aa2 (held by even lanes)
    float t0, t1;                                     // r = 0: tlo = syn(aa3, ad3); r = 2: thi = syn(da3, dd3)
    syn(low ? soft_thr(c.l3[0], thr[2][0]) : c.l3[0], soft_thr(c.l3[1], low ? thr[2][2] : thr[2][1]), t0, t1);
    const float p0 = __shfl_xor_sync(0xffffffffu, t0, 2), p1 = __shfl_xor_sync(0xffffffffu, t1, 2);
    float e, o;
    syn(low ? p0 : t0, low ? t0 : p0, e, o);
    aa2[0] = low ? o : e;
    syn(low ? p1 : t1, low ? t1 : p1, e, o);
    aa2[1] = low ? o : e;
  }
  float aa1[4];
  {                                                   // level 2 -> aa1 (one row of 4 per lane)
    float t[4];
#pragma unroll
    for (int j = 0; j < 2; ++j)                       // even: syn(aa2, ad2) = tlo; odd: syn(da2, dd2) = thi
      syn(odd ? soft_thr(c.l2[j], thr[1][0]) : aa2[j], soft_thr(c.l2[2 + j], odd ? thr[1][2] : thr[1][1]), t[2 * j],
          t[2 * j + 1]);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float p = __shfl_xor_sync(0xffffffffu, t[k], 1);
      float e, o;
      syn(odd ? p : t[k], odd ? t[k] : p, e, o);
      aa1[k] = odd ? o : e;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {                       // level 1, local
    float tlo0, tlo1, thi0, thi1;
    syn(aa1[i], soft_thr(c.ad[i], thr[0][1]), tlo0, tlo1);
    syn(soft_thr(c.da[i], thr[0][0]), soft_thr(c.dd[i], thr[0][2]), thi0, thi1);
    syn(tlo0, thi0, x0[2 * i], x1[2 * i]);
    syn(tlo1, thi1, x0[2 * i + 1], x1[2 * i + 1]);
  }
}

// skimage.color.rgb2ycbcr matrix, same constants as the general path (b2d_wavelet.cu)
__constant__ float kHaarYCC[3][3] = {{65.481f, 128.553f, 24.966f}, {-37.797f, -74.203f, 112.0f}, {112.0f, -93.786f, -18.214f}};

struct HaarGeom {
  int bw, bh;                 // 8 x 8 blocks per row / column
  int strips_x;               // ceil(bw / 8): a warp covers 8 blocks
  int total_strips;           // strips_x * bh
  int w2;                     // W / 2 (row length of the dd_1 plane)
  size_t dd_floats;           // (H / 2) * (W / 2)
};

// Loads this lane's 8 px x 2 rows, applies the noise operator and converts to the three
// "YCbCr" channels: P[c][row][x].  Lanes right of the image (ragged last strip) shadow block
// bw - 1 so that the shuffles stay convergent; `live` tells whether the lane owns real pixels.
template <int KIND>
__device__ __forceinline__ void haar_load(const ImageSet& s, const NoiseArgs& na, const uint8_t* img, int img_local,
                                          int px0, int y0, float (&P)[3][2][8]) {
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const uint32_t byte0 = (uint32_t)(y0 + row) * (uint32_t)s.rowbytes + (uint32_t)px0 * 3u;   // multiple of 24
    const uint2* src = reinterpret_cast<const uint2*>(img + byte0);
    float xf[24];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const uint2 w = __ldg(src + q);
      uint32_t v[8];
      float f[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        v[i] = (w.x >> (8 * i)) & 0xffu;
        v[4 + i] = (w.y >> (8 * i)) & 0xffu;
      }
      GroupCtx gc{(uint32_t)img_local, byte0 / 8u + (uint32_t)q, img};
      apply8_float<KIND, NM_GEN>(na, gc, v, f);
#pragma unroll
      for (int i = 0; i < 8; ++i) xf[8 * q + i] = f[i] * (1.0f / 255.0f);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float c0 = xf[3 * k], c1 = xf[3 * k + 1], c2 = xf[3 * k + 2];
#pragma unroll
      for (int c = 0; c < 3; ++c)
        P[c][row][k] = kHaarYCC[c][0] * c0 + kHaarYCC[c][1] * c1 + kHaarYCC[c][2] * c2 + (c == 0 ? 16.0f : 128.0f);
    }
  }
}

// ------------------------------------------------------------------ pass A
template <int KIND>
__global__ void __launch_bounds__(256, 2) k_haar_stats(ImageSet s, NoiseArgs na, HaarGeom hg, float* dd1, WvStats* stats,
                                                        int chunk_first) {
  const int z = blockIdx.y;
  const int img_local = image_of(s, chunk_first + z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int lane = threadIdx.x & 31, r = lane & 3, bxw = lane >> 2;
  const int warp = blockIdx.x * 8 + (threadIdx.x >> 5), nwarps = gridDim.x * 8;
  float mn[3], mx[3], en[3][3][3];                    // en[channel][level - 1][band]
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    mn[c] = 3.0e38f;
    mx[c] = -3.0e38f;
#pragma unroll
    for (int l = 0; l < 3; ++l)
#pragma unroll
      for (int b = 0; b < 3; ++b) en[c][l][b] = 0.f;
  }
  const bool odd = (r & 1) != 0, low = (r & 2) != 0;
  for (int strip = warp; strip < hg.total_strips; strip += nwarps) {
    const int by = strip / hg.strips_x, sx = strip - by * hg.strips_x;
    const int bx_raw = sx * 8 + bxw;
    const bool live = bx_raw < hg.bw;
    const int bx = live ? bx_raw : hg.bw - 1;
    float P[3][2][8];
    haar_load<KIND>(s, na, img, img_local, bx * 8, by * 8 + 2 * r, P);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      HaarCoef k;
      haar_forward(P[c][0], P[c][1], r, k);
      if (!live) continue;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        mn[c] = fminf(mn[c], fminf(P[c][0][i], P[c][1][i]));
        mx[c] = fmaxf(mx[c], fmaxf(P[c][0][i], P[c][1][i]));
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        en[c][0][0] = fmaf(k.da[i], k.da[i], en[c][0][0]);
        en[c][0][1] = fmaf(k.ad[i], k.ad[i], en[c][0][1]);
        en[c][0][2] = fmaf(k.dd[i], k.dd[i], en[c][0][2]);
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {                   // even lanes: (aa2, ad2); odd lanes: (da2, dd2)
        const float a = k.l2[j] * k.l2[j], b = k.l2[2 + j] * k.l2[2 + j];
        en[c][1][0] += odd ? a : 0.f;
        en[c][1][1] += odd ? 0.f : b;
        en[c][1][2] += odd ? b : 0.f;
      }
      if (!odd) {                                     // r = 0: (aa3, ad3); r = 2: (da3, dd3)
        const float a = k.l3[0] * k.l3[0], b = k.l3[1] * k.l3[1];
        en[c][2][0] += low ? a : 0.f;
        en[c][2][1] += low ? 0.f : b;
        en[c][2][2] += low ? b : 0.f;
      }
      float4* dst = reinterpret_cast<float4*>(dd1 + ((size_t)z * 3 + c) * hg.dd_floats +
                                              (size_t)(by * 4 + r) * hg.w2 + (size_t)bx * 4);
      *dst = make_float4(k.dd[0], k.dd[1], k.dd[2], k.dd[3]);
    }
  }
  __shared__ double s_en[27];
  __shared__ int s_mn[3], s_mx[3];
  if (threadIdx.x < 27) s_en[threadIdx.x] = 0.0;
  if (threadIdx.x < 3) { s_mn[threadIdx.x] = 0x7fffffff; s_mx[threadIdx.x] = (int)0x80000000; }
  __syncthreads();
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o; o >>= 1) {
      mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
      mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
    }
    if (lane == 0) {
      atomicMin(&s_mn[c], f2ord(mn[c]));
      atomicMax(&s_mx[c], f2ord(mx[c]));
    }
#pragma unroll
    for (int l = 0; l < 3; ++l)
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        float e = en[c][l][b];
        for (int o = 16; o; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
        if (lane == 0) atomicAdd(&s_en[(c * 3 + l) * 3 + b], (double)e);
      }
  }
  __syncthreads();
  if (threadIdx.x < 27) {
    const int c = threadIdx.x / 9, l = (threadIdx.x / 3) % 3, b = threadIdx.x % 3;
    atomicAdd(&stats[z * 3 + c].energy[l + 1][b], s_en[threadIdx.x]);
  }
  if (threadIdx.x < 3) {
    atomicMin(&stats[z * 3 + threadIdx.x].mn_ord, s_mn[threadIdx.x]);
    atomicMax(&stats[z * 3 + threadIdx.x].mx_ord, s_mx[threadIdx.x]);
  }
}

// ------------------------------------------------------------------ pass B
__device__ __forceinline__ uint32_t pack_low_bytes(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

struct HaarFinal { float inv[3][3]; };

template <int KIND>
__global__ void __launch_bounds__(256, 2) k_haar_apply(ImageSet s, NoiseArgs na, HaarGeom hg, const WvStats* stats,
                                                        HaarFinal fa, int chunk_first) {
  const int z = blockIdx.y;
  const int img_local = image_of(s, chunk_first + z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  const int lane = threadIdx.x & 31, r = lane & 3, bxw = lane >> 2;
  const int warp = blockIdx.x * 8 + (threadIdx.x >> 5), nwarps = gridDim.x * 8;
  float thr[3][3][3], lo[3], hi[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const WvStats& st = stats[z * 3 + c];
    lo[c] = st.lo;
    hi[c] = st.hi;
#pragma unroll
    for (int l = 0; l < 3; ++l)
#pragma unroll
      for (int b = 0; b < 3; ++b) thr[c][l][b] = st.thr[l + 1][b];
  }
  for (int strip = warp; strip < hg.total_strips; strip += nwarps) {
    const int by = strip / hg.strips_x, sx = strip - by * hg.strips_x;
    const int bx_raw = sx * 8 + bxw;
    const bool live = bx_raw < hg.bw;
    const int bx = live ? bx_raw : hg.bw - 1;
    float P[3][2][8];
    haar_load<KIND>(s, na, img, img_local, bx * 8, by * 8 + 2 * r, P);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      HaarCoef k;
      haar_forward(P[c][0], P[c][1], r, k);
      haar_inverse(k, r, thr[c], P[c][0], P[c][1]);
    }
    if (!live) continue;
#pragma unroll
    for (int row = 0; row < 2; ++row) {
      uint32_t o[24];                                   // 2^23 + byte as fp32 bits: the low byte is the result
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        // the per-channel denoise_wavelet clips to the channel's own [min, max] (its normalised [0, 1])
        const float y = fminf(fmaxf(P[0][row][i], lo[0]), hi[0]) - 16.0f;
        const float cb = fminf(fmaxf(P[1][row][i], lo[1]), hi[1]) - 128.0f;
        const float cr = fminf(fmaxf(P[2][row][i], lo[2]), hi[2]) - 128.0f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float v = fa.inv[c][0] * y + fa.inv[c][1] * cb + fa.inv[c][2] * cr;
          v = fminf(fmaxf(v, 0.0f), 1.0f);
          // (uint8)(255 * v): product rounded as the general path does, then truncated by the magic add
          o[3 * i + c] = __float_as_uint(__fadd_rz(__fmul_rn(255.0f, v), 8388608.0f));
        }
      }
      const uint32_t byte0 = (uint32_t)(by * 8 + 2 * r + row) * (uint32_t)s.rowbytes + (uint32_t)bx * 24u;
      uint2* dst = reinterpret_cast<uint2*>(out + byte0);
#pragma unroll
      for (int q = 0; q < 3; ++q)
        dst[q] = make_uint2(pack_low_bytes(o[8 * q], o[8 * q + 1], o[8 * q + 2], o[8 * q + 3]),
                            pack_low_bytes(o[8 * q + 4], o[8 * q + 5], o[8 * q + 6], o[8 * q + 7]));
    }
  }
}

// ------------------------------------------------------------------ host side
bool haar_supported(const b2d_plan& p, const b2d_spec& sp, int mode) {
  static const bool off = [] { const char* e = knob("B2D_HAAR"); return e && e[0] == '0'; }();
  if (off || mode != NM_GEN) return false;
  if (sp.wavelet != B2D_WAVELET_DB1 || sp.levels != 3) return false;
  if (p.h % 8 || p.w % 8 || p.h < 8 || p.w < 8) return false;
  return (uint64_t)p.h * p.w * 3 < (1ull << 32);
}

template <int KIND>
static void launch_haar(int which, dim3 grid, cudaStream_t st, const ImageSet& s, const NoiseArgs& na, const HaarGeom& hg,
                        float* dd1, WvStats* stats, const HaarFinal& fa, int first) {
  if (which == 0) k_haar_stats<KIND><<<grid, 256, 0, st>>>(s, na, hg, dd1, stats, first);
  else k_haar_apply<KIND><<<grid, 256, 0, st>>>(s, na, hg, stats, fa, first);
}

static int dispatch_haar(int kind, int which, dim3 grid, cudaStream_t st, const ImageSet& s, const NoiseArgs& na,
                         const HaarGeom& hg, float* dd1, WvStats* stats, const HaarFinal& fa, int first) {
  switch (kind) {
#define B2D_HAAR_CASE(K) case K: launch_haar<K>(which, grid, st, s, na, hg, dd1, stats, fa, first); break;
    B2D_HAAR_CASE(NK_NONE) B2D_HAAR_CASE(NK_GAUSSIAN) B2D_HAAR_CASE(NK_SPECKLE) B2D_HAAR_CASE(NK_SAP)
    B2D_HAAR_CASE(NK_POISSON) B2D_HAAR_CASE(NK_UNIFORM) B2D_HAAR_CASE(NK_RAYLEIGH) B2D_HAAR_CASE(NK_GAMMA)
    B2D_HAAR_CASE(NK_TABLE_ADD) B2D_HAAR_CASE(NK_SHADER)
#undef B2D_HAAR_CASE
    default: return fail(B2D_ERR_INVALID, "wavelet: bad noise operator %d", kind);
  }
  count_launch();
  return B2D_OK;
}

// `wbase` / `avail`: the plan's wavelet workspace (sized for the general path, which is always larger
// per image than 3 dd_1 planes + statistics).
int run_haar(const b2d_plan& p, const ImageSet& s, int kind, const NoiseArgs& na, const WvGeom& g_full, uint8_t* wbase,
             size_t avail, const float (&inv)[3][3], cudaStream_t st) {
  HaarGeom hg;
  hg.bw = p.w / 8;
  hg.bh = p.h / 8;
  hg.strips_x = (hg.bw + 7) / 8;
  hg.total_strips = hg.strips_x * hg.bh;
  hg.w2 = p.w / 2;
  hg.dd_floats = (size_t)(p.h / 2) * (p.w / 2);
  const size_t per_img = 3 * (hg.dd_floats * sizeof(float) + sizeof(WvStats) + 2 * WV_BINS * sizeof(unsigned int));
  const int chunk = (int)std::min<size_t>((size_t)s.count, (avail - 1024) / per_img);
  if (chunk < 1) return fail(B2D_ERR_INVALID, "wavelet: workspace too small for the Haar path");
  float* dd1 = reinterpret_cast<float*>(wbase);
  WvStats* stats = reinterpret_cast<WvStats*>(wbase + (size_t)chunk * 3 * hg.dd_floats * sizeof(float));
  unsigned int* hist = reinterpret_cast<unsigned int*>(reinterpret_cast<uint8_t*>(stats) + (size_t)chunk * 3 * sizeof(WvStats));
  WvGeom g = g_full;                                  // the select / threshold kernels read these fields only
  g.chan_floats = hg.dd_floats;
  g.off_det[1][2] = 0;
  HaarFinal fa;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) fa.inv[i][j] = inv[i][j];
  const int sms = sm_count();
  for (int first = 0; first < s.count; first += chunk) {
    const int cnt = std::min(chunk, s.count - first);
    // two resident CTAs per SM: pick the CTA count per image that fills whole waves (cnt x per_image close to a
    // multiple of 2 x SMs), at least 2 waves' worth, at most one strip row of 8 warps per CTA
    const int max_pi = std::max(1, (hg.total_strips + 7) / 8);
    const int slots = 2 * sms;
    int per_image = std::min(max_pi, std::max(1, (2 * slots + cnt - 1) / cnt));
    {
      double best = -1.0;
      int best_pi = per_image;
      for (int c = per_image; c <= std::min(max_pi, per_image + 64); ++c) {
        const long total = (long)c * cnt;
        const double fill = (double)total / (double)(((total + slots - 1) / slots) * slots);
        if (fill > best + 0.02) { best = fill; best_pi = c; }
      }
      per_image = best_pi;
    }
    const dim3 grid((unsigned)per_image, (unsigned)cnt, 1);
    wavelet_stats_begin(stats, hist, cnt * 3, st);
    int rc = dispatch_haar(kind, 0, grid, st, s, na, hg, dd1, stats, fa, first);
    if (rc) return rc;
    wavelet_sigma_thresholds(g, dd1, stats, hist, cnt * 3, st);
    rc = dispatch_haar(kind, 1, grid, st, s, na, hg, dd1, stats, fa, first);
    if (rc) return rc;
    p.wv_stats_dev = stats; p.wv_stats_first = first; p.wv_stats_count = cnt;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "haar wavelet launch");
  }
  return B2D_OK;
}

}  // namespace b2d
