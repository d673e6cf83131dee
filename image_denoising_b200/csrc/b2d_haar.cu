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
// exchange partners with __shfl_xor(1) / (2).  Arithmetic: unnormalised butterflies on packed fp32 pairs
// (see below); agrees with the general kernels to <= 1 LSB on < 0.2 % of the pixels (tests/test_gpu_wavelet.py).
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
#include "b2d_wavelet.cuh"

#include <algorithm>
#include <cstdlib>

namespace b2d {

// ---------------------------------------------------------------------------------------------------------
// Arithmetic.  Two things keep the instruction count down (both kernels are issue bound):
//  * packed fp32 pairs (FADD2 / FMUL2 / FFMA2, one issue slot for two lanes' worth of IEEE fp32 operations): a pair
//    holds pixels (k, k + 4) of a row, which makes every butterfly of level 1 (pairs of blocks i, i + 2) and the
//    x-direction of level 2 (pairs j = 0, 1) a packed operation with no lane shuffling inside a register pair;
//  * the UNNORMALISED Haar butterfly: pywt's a = (e + o) / sqrt 2, d = (e - o) / sqrt 2 applied along x and along y
//    is (sums and differences of the 2 x 2 block) / 2, so level l coefficients are kept as 2^l times their value -
//    pure additions - and the exact powers of two are folded into the thresholds (x 2^l), the band energies
//    (x 4^-l), the stored dd_1 (x 1/2) and, on the way back, into one factor 1/4 per level and the colour matrix.
//    Exact zeros stay exact ((a - b) - (c - d) cancels in any scaling), which is what _sigma_est_dwt looks at.
// ---------------------------------------------------------------------------------------------------------
typedef unsigned long long f2;                        // two fp32 in one 64-bit register pair

__device__ __forceinline__ f2 pk(float lo, float hi) {
  f2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpk(f2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2 splat(float x) { return pk(x, x); }
__device__ __forceinline__ f2 add2(f2 a, f2 b) {
  f2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2 sub2(f2 a, f2 b) {
  f2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2 mul2(f2 a, f2 b) {
  f2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) {
  f2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ f2 add2_rz(f2 a, f2 b) {
  f2 r;
  asm("add.rz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ float min3f(float a, float b, float c) {
  float d;
  asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float max3f(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
__device__ __forceinline__ float shx(float v, int m) { return __shfl_xor_sync(0xffffffffu, v, m); }
__device__ __forceinline__ f2 shx2(f2 v, int m) {
  float lo, hi;
  unpk(v, lo, hi);
  return pk(shx(lo, m), shx(hi, m));
}
// pywt.threshold(d, t, 'soft') on both halves
__device__ __forceinline__ f2 soft2(f2 v, float t) {
  float lo, hi;
  unpk(v, lo, hi);
  return pk(wv_soft(lo, t), wv_soft(hi, t));
}

// The 8 values of one Philox group on the reference's float scale ([0, 1] = [0, 255]), BEFORE any ->u8 cast: what
// denoise_wavelet is fed (minibatch.py:91-94).  gaussian / speckle are the reference's own expressions in one
// FFMA.SAT each - clip(f + n, 0, 1), clip(f + f n, 0, 1) with f = v / 255; the other operators go through their
// general float view (b2d_noise.cuh).
template <int KIND>
__device__ __forceinline__ void unit8(const NoiseArgs& na, const GroupCtx& gc, uint2 w, float (&u)[8]) {
  if constexpr (KIND == NK_GAUSSIAN || KIND == NK_SPECKLE) {
    const uint4 r = group_random(na, gc, 0);
    float n[8];
    normal_pair(r.x, na.p0, n[0], n[1]);
    normal_pair(r.y, na.p0, n[2], n[3]);
    normal_pair(r.z, na.p0, n[4], n[5]);
    normal_pair(r.w, na.p0, n[6], n[7]);
    const float v[8] = {magic_byte<0>(w.x), magic_byte<1>(w.x), magic_byte<2>(w.x), magic_byte<3>(w.x),
                        magic_byte<0>(w.y), magic_byte<1>(w.y), magic_byte<2>(w.y), magic_byte<3>(w.y)};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float b = v[i] - 8388608.0f;              // the byte as a float, exact
      if (KIND == NK_GAUSSIAN) u[i] = __saturatef(fmaf(b, 1.0f / 255.0f, n[i]));
      else {
        const float f = b * (1.0f / 255.0f);
        u[i] = __saturatef(fmaf(f, n[i], f));
      }
    }
  } else {
    uint32_t v[8];
    float f[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] = (w.x >> (8 * i)) & 0xffu;
      v[4 + i] = (w.y >> (8 * i)) & 0xffu;
    }
    apply8_float<KIND, NM_GEN>(na, gc, v, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) u[i] = f[i] * (1.0f / 255.0f);
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

// Loads this lane's 8 px x 2 rows, applies the noise operator and converts to the three "YCbCr" channels:
// P[c][row][k] = (pixel k, pixel k + 4), k = 0..3.  Lanes right of the image (ragged last strip) shadow block bw - 1
// so that the shuffles stay convergent; `live` tells whether the lane owns real pixels.
template <int KIND>
__device__ __forceinline__ void haar_load(const ImageSet& s, const NoiseArgs& na, const uint8_t* img, int img_local,
                                          int px0, int y0, f2 (&P)[3][2][4]) {
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const uint32_t byte0 = (uint32_t)(y0 + row) * (uint32_t)s.rowbytes + (uint32_t)px0 * 3u;   // multiple of 24
    const uint2* src = reinterpret_cast<const uint2*>(img + byte0);
    float xf[24];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      float u[8];
      GroupCtx gc{(uint32_t)img_local, byte0 / 8u + (uint32_t)q, img};
      unit8<KIND>(na, gc, __ldg(src + q), u);
#pragma unroll
      for (int i = 0; i < 8; ++i) xf[8 * q + i] = u[i];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const f2 c0 = pk(xf[3 * k], xf[3 * k + 12]), c1 = pk(xf[3 * k + 1], xf[3 * k + 13]), c2 = pk(xf[3 * k + 2], xf[3 * k + 14]);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        P[c][row][k] = fma2(splat(kHaarYCC[c][2]), c2,
                            fma2(splat(kHaarYCC[c][1]), c1, fma2(splat(kHaarYCC[c][0]), c0, splat(c == 0 ? 16.0f : 128.0f))));
    }
  }
}

// One channel of one lane (8 px x 2 rows) after the forward transform, unnormalised (level l values x 2^l).
struct HaarCoef {
  f2 da[2], ad[2], dd[2];      // level 1: pair m holds blocks i = m and i = m + 2 of this lane's row of the 4 x 4 band
  f2 aa2, ad2, da2, dd2;       // level 2, pairs over j = 0, 1.  aa2: every lane; ad2: valid on even lanes; da2, dd2: on odd lanes
  float aa3, ad3, da3, dd3;    // level 3.  aa3, ad3: valid on lane r = 0; da3, dd3: on lane r = 2
};

__device__ __forceinline__ void haar_forward(const f2 (&P)[2][4], HaarCoef& c) {
  f2 a1[2];
#pragma unroll
  for (int m = 0; m < 2; ++m) {                       // x: pixel pairs (2 m, 2 m + 1) = blocks i = m (low half), m + 2 (high half)
    const f2 s0 = add2(P[0][2 * m], P[0][2 * m + 1]), d0 = sub2(P[0][2 * m], P[0][2 * m + 1]);
    const f2 s1 = add2(P[1][2 * m], P[1][2 * m + 1]), d1 = sub2(P[1][2 * m], P[1][2 * m + 1]);
    a1[m] = add2(s0, s1);                             // y: rows 0 / 1
    c.da[m] = sub2(s0, s1);
    c.ad[m] = add2(d0, d1);
    c.dd[m] = sub2(d0, d1);
  }
  // level 2: this lane's row of aa1 is (a1[0].lo, a1[1].lo, a1[0].hi, a1[1].hi); x pairs (2 j, 2 j + 1) -> j = 0, 1 packed.
  // Rows r (even lane) and r + 1 (odd lane): e = the even lane's value, o = the odd lane's; the difference bands are
  // written as (partner - own), which is e - o on the odd lanes - the ones that keep them.
  const f2 lo = add2(a1[0], a1[1]), hi = sub2(a1[0], a1[1]);
  const f2 plo = shx2(lo, 1), phi = shx2(hi, 1);
  c.aa2 = add2(lo, plo);
  c.da2 = sub2(plo, lo);
  c.ad2 = add2(hi, phi);
  c.dd2 = sub2(phi, hi);
  // level 3: rows r = 0 (e) and r = 2 (o) of aa2; same convention, difference bands valid on the r = 2 lane
  float x0, x1;
  unpk(c.aa2, x0, x1);
  const float lo3 = x0 + x1, hi3 = x0 - x1;
  const float plo3 = shx(lo3, 2), phi3 = shx(hi3, 2);
  c.aa3 = lo3 + plo3;
  c.da3 = plo3 - lo3;
  c.ad3 = hi3 + phi3;
  c.dd3 = phi3 - hi3;
}

// ------------------------------------------------------------------ pass A
template <int KIND>
__global__ void __launch_bounds__(256, 2) k_haar_stats(ImageSet s, NoiseArgs na, HaarGeom hg, float* dd1, WvStats* stats,
                                                        unsigned int* hist, int chunk_first) {
  // pass 0 of the radix select (bits 30..20 of |dd_1|, zeros skipped: k_select_hist's rule) is tallied here, while the
  // coefficients are in registers: one full read of the band less (wavelet_sigma_thresholds(..., first_pass = 1))
  __shared__ unsigned int s_hist[3][WV_BINS];
  for (int i = threadIdx.x; i < 3 * WV_BINS; i += 256) (&s_hist[0][0])[i] = 0u;
  __syncthreads();
  const int z = blockIdx.y;
  const int img_local = image_of(s, chunk_first + z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int lane = threadIdx.x & 31, r = lane & 3, bxw = lane >> 2;
  const int warp = blockIdx.x * 8 + (threadIdx.x >> 5), nwarps = gridDim.x * 8;
  // sums of squares of the UNNORMALISED coefficients, accumulated on every lane; the lanes that do not own a band
  // (see HaarCoef) are masked out once, after the loop
  float mn[3], mx[3], e3[3][3];
  f2 e1[3][3], e2[3][3];                              // [channel][band: 0 = da, 1 = ad, 2 = dd]
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    mn[c] = 3.0e38f;
    mx[c] = -3.0e38f;
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      e1[c][b] = e2[c][b] = splat(0.f);
      e3[c][b] = 0.f;
    }
  }
  for (int strip = warp; strip < hg.total_strips; strip += nwarps) {
    const int by = strip / hg.strips_x, sx = strip - by * hg.strips_x;
    const int bx_raw = sx * 8 + bxw;
    const bool live = bx_raw < hg.bw;
    const int bx = live ? bx_raw : hg.bw - 1;
    f2 P[3][2][4];
    haar_load<KIND>(s, na, img, img_local, bx * 8, by * 8 + 2 * r, P);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      HaarCoef k;
      haar_forward(P[c], k);
      if (!live) continue;
#pragma unroll
      for (int row = 0; row < 2; ++row)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float a, b;
          unpk(P[c][row][i], a, b);
          mn[c] = min3f(mn[c], a, b);
          mx[c] = max3f(mx[c], a, b);
        }
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        e1[c][0] = fma2(k.da[m], k.da[m], e1[c][0]);
        e1[c][1] = fma2(k.ad[m], k.ad[m], e1[c][1]);
        e1[c][2] = fma2(k.dd[m], k.dd[m], e1[c][2]);
      }
      e2[c][0] = fma2(k.da2, k.da2, e2[c][0]);
      e2[c][1] = fma2(k.ad2, k.ad2, e2[c][1]);
      e2[c][2] = fma2(k.dd2, k.dd2, e2[c][2]);
      e3[c][0] = fmaf(k.da3, k.da3, e3[c][0]);
      e3[c][1] = fmaf(k.ad3, k.ad3, e3[c][1]);
      e3[c][2] = fmaf(k.dd3, k.dd3, e3[c][2]);
      // dd_1 itself (x 1/2: exact) for the median: blocks i = 0..3 of this lane's row of the band
      float d0, d2, d1, d3;
      unpk(mul2(k.dd[0], splat(0.5f)), d0, d2);
      unpk(mul2(k.dd[1], splat(0.5f)), d1, d3);
      float4* dst = reinterpret_cast<float4*>(dd1 + ((size_t)z * 3 + c) * hg.dd_floats +
                                              (size_t)(by * 4 + r) * hg.w2 + (size_t)bx * 4);
      *dst = make_float4(d0, d1, d2, d3);
      const unsigned int b0 = __float_as_uint(d0) & 0x7fffffffu, b1 = __float_as_uint(d1) & 0x7fffffffu;
      const unsigned int b2 = __float_as_uint(d2) & 0x7fffffffu, b3 = __float_as_uint(d3) & 0x7fffffffu;
      if (b0) atomicAdd(&s_hist[c][b0 >> 20], 1u);
      if (b1) atomicAdd(&s_hist[c][b1 >> 20], 1u);
      if (b2) atomicAdd(&s_hist[c][b2 >> 20], 1u);
      if (b3) atomicAdd(&s_hist[c][b3 >> 20], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * WV_BINS; i += 256) {
    const unsigned int v = (&s_hist[0][0])[i];
    if (v) atomicAdd(&hist[((size_t)z * 3 + i / WV_BINS) * 2 * WV_BINS + (i % WV_BINS)], v);
  }
  __shared__ double s_en[27];
  __shared__ int s_mn[3], s_mx[3];
  if (threadIdx.x < 27) s_en[threadIdx.x] = 0.0;
  if (threadIdx.x < 3) { s_mn[threadIdx.x] = 0x7fffffff; s_mx[threadIdx.x] = (int)0x80000000; }
  __syncthreads();
  const bool odd = (r & 1) != 0;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o; o >>= 1) {
      mn[c] = fminf(mn[c], shx(mn[c], o));
      mx[c] = fmaxf(mx[c], shx(mx[c], o));
    }
    if (lane == 0) {
      atomicMin(&s_mn[c], f2ord(mn[c]));
      atomicMax(&s_mx[c], f2ord(mx[c]));
    }
#pragma unroll
    for (int b = 0; b < 3; ++b) {
      float a, bb;
      unpk(e1[c][b], a, bb);
      float v1 = 0.25f * (a + bb);                                   // level 1: every lane owns its row; (x / 2)^2
      unpk(e2[c][b], a, bb);
      float v2 = ((b == 1) != odd) ? 0.0625f * (a + bb) : 0.f;       // level 2: ad on even lanes, da / dd on odd lanes; (x / 4)^2
      float v3 = (r == (b == 1 ? 0 : 2)) ? 0.015625f * e3[c][b] : 0.f;   // level 3: ad on r = 0, da / dd on r = 2; (x / 8)^2
      for (int o = 16; o; o >>= 1) {
        v1 += shx(v1, o);
        v2 += shx(v2, o);
        v3 += shx(v3, o);
      }
      if (lane == 0) {
        atomicAdd(&s_en[(c * 3 + 0) * 3 + b], (double)v1);
        atomicAdd(&s_en[(c * 3 + 1) * 3 + b], (double)v2);
        atomicAdd(&s_en[(c * 3 + 2) * 3 + b], (double)v3);
      }
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
// T[level - 1][band] = threshold x 2^level (the coefficients are unnormalised); band 0 = da, 1 = ad, 2 = dd.
// Q = 4 x the reconstructed channel values, same pairing as P.
__device__ __forceinline__ void haar_inverse(const HaarCoef& c, bool odd, bool low, const float (&T)[3][3], f2 (&Q)[2][4]) {
  // level 3 -> aa2: lane r = 0 synthesises (aa3, ad3) along x, lane r = 2 (da3, dd3); then rows e / o along y
  const float first3 = low ? wv_soft(c.da3, T[2][0]) : c.aa3;
  const float second3 = wv_soft(low ? c.dd3 : c.ad3, low ? T[2][2] : T[2][1]);
  const f2 t3 = pk(first3 + second3, first3 - second3);
  const f2 aa2 = fma2(t3, splat(low ? -0.25f : 0.25f), mul2(shx2(t3, 2), splat(0.25f)));   // r = 0: (t + p) / 4; r = 2: (p - t) / 4
  // level 2 -> aa1: even lanes synthesise (aa2, ad2), odd lanes (da2, dd2)
  const f2 first2 = odd ? soft2(c.da2, T[1][0]) : aa2;
  const f2 second2 = soft2(odd ? c.dd2 : c.ad2, odd ? T[1][2] : T[1][1]);
  const f2 tE = add2(first2, second2), tO = sub2(first2, second2);
  const f2 sg = splat(odd ? -0.25f : 0.25f), q = splat(0.25f);
  f2 a1[2];
  a1[0] = fma2(tE, sg, mul2(shx2(tE, 1), q));          // blocks i = 0, 2
  a1[1] = fma2(tO, sg, mul2(shx2(tO, 1), q));          // blocks i = 1, 3
#pragma unroll
  for (int m = 0; m < 2; ++m) {                        // level 1, local
    const f2 ad = soft2(c.ad[m], T[0][1]), da = soft2(c.da[m], T[0][0]), dd = soft2(c.dd[m], T[0][2]);
    const f2 tlo0 = add2(a1[m], ad), tlo1 = sub2(a1[m], ad), thi0 = add2(da, dd), thi1 = sub2(da, dd);
    Q[0][2 * m] = add2(tlo0, thi0);
    Q[1][2 * m] = sub2(tlo0, thi0);
    Q[0][2 * m + 1] = add2(tlo1, thi1);
    Q[1][2 * m + 1] = sub2(tlo1, thi1);
  }
}

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
  const bool odd = (r & 1) != 0, low = (r & 2) != 0;
  float T[3][3][3], lo4[3], hi4[3], inv4[3][3], k0[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const WvStats& st = stats[z * 3 + c];
    lo4[c] = 4.0f * st.lo;                             // Q = 4 x value: the clip range and the offsets scale with it (exact)
    hi4[c] = 4.0f * st.hi;
#pragma unroll
    for (int l = 0; l < 3; ++l)
#pragma unroll
      for (int b = 0; b < 3; ++b) T[c][l][b] = st.thr[l + 1][b] * (float)(2 << l);
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int j = 0; j < 3; ++j) inv4[c][j] = 0.25f * fa.inv[c][j];
    // ycbcr2rgb subtracts (16, 128, 128) first; here the offsets ride in the chain's first addend
    k0[c] = -(inv4[c][0] * 64.0f + inv4[c][1] * 512.0f + inv4[c][2] * 512.0f);
  }
  for (int strip = warp; strip < hg.total_strips; strip += nwarps) {
    const int by = strip / hg.strips_x, sx = strip - by * hg.strips_x;
    const int bx_raw = sx * 8 + bxw;
    const bool live = bx_raw < hg.bw;
    const int bx = live ? bx_raw : hg.bw - 1;
    f2 P[3][2][4];
    haar_load<KIND>(s, na, img, img_local, bx * 8, by * 8 + 2 * r, P);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      HaarCoef k;
      haar_forward(P[c], k);
      haar_inverse(k, odd, low, T[c], P[c]);
      // the per-channel denoise_wavelet clips to the channel's own [min, max] (its normalised [0, 1])
#pragma unroll
      for (int row = 0; row < 2; ++row)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float a, b;
          unpk(P[c][row][i], a, b);
          P[c][row][i] = pk(fminf(fmaxf(a, lo4[c]), hi4[c]), fminf(fmaxf(b, lo4[c]), hi4[c]));
        }
    }
    if (!live) continue;
#pragma unroll
    for (int row = 0; row < 2; ++row) {
      uint32_t o[24];                                   // 2^23 + byte as fp32 bits: the low byte is the result
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const f2 v = fma2(splat(inv4[c][2]), P[2][row][i],
                            fma2(splat(inv4[c][1]), P[1][row][i], fma2(splat(inv4[c][0]), P[0][row][i], splat(k0[c]))));
          float a, b;
          unpk(v, a, b);
          // (uint8)(255 * clip(v, 0, 1)): product rounded as the general path does, then truncated by the magic add
          const f2 t = add2_rz(mul2(pk(__saturatef(a), __saturatef(b)), splat(255.0f)), splat(8388608.0f));
          unpk(t, a, b);
          o[3 * i + c] = __float_as_uint(a);            // pixel i
          o[3 * (i + 4) + c] = __float_as_uint(b);      // pixel i + 4
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
                        float* dd1, WvStats* stats, unsigned int* hist, const HaarFinal& fa, int first) {
  if (which == 0) k_haar_stats<KIND><<<grid, 256, 0, st>>>(s, na, hg, dd1, stats, hist, first);
  else k_haar_apply<KIND><<<grid, 256, 0, st>>>(s, na, hg, stats, fa, first);
}

static int dispatch_haar(int kind, int which, dim3 grid, cudaStream_t st, const ImageSet& s, const NoiseArgs& na,
                         const HaarGeom& hg, float* dd1, WvStats* stats, unsigned int* hist, const HaarFinal& fa, int first) {
  switch (kind) {
#define B2D_HAAR_CASE(K) case K: launch_haar<K>(which, grid, st, s, na, hg, dd1, stats, hist, fa, first); break;
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
  const int chunk = (int)std::min<size_t>(std::min<size_t>((size_t)s.count, 21845), (avail - 1024) / per_img);   // 3 planes per image in grid.z
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
    int rc = dispatch_haar(kind, 0, grid, st, s, na, hg, dd1, stats, hist, fa, first);
    if (rc) return rc;
    wavelet_sigma_thresholds(g, dd1, stats, hist, cnt * 3, st, 1);        // pass 0's histogram is already in `hist`
    rc = dispatch_haar(kind, 1, grid, st, s, na, hg, dd1, stats, hist, fa, first);
    if (rc) return rc;
    p.wv_stats_dev = stats; p.wv_stats_first = first; p.wv_stats_count = cnt;
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_cuda(e, "haar wavelet launch");
  }
  return B2D_OK;
}

}  // namespace b2d
