// b2d_quant.cu - `quant`: colour quantisation in Lab (lib/roi_data_layer/minibatch.py:496-507)
//     image = cv2.cvtColor(img, cv2.COLOR_BGR2LAB); image = image.reshape((h * w, 3))
//     clt = MiniBatchKMeans(n_clusters=k); labels = clt.fit_predict(image)
//     quant = clt.cluster_centers_.astype("uint8")[labels]
//     quant = cv2.cvtColor(quant.reshape((h, w, 3)), cv2.COLOR_LAB2BGR)
//
//   k_quant_fit     one CTA per image: k-means++ seeding + Lloyd iterations on a Philox-chosen
//                   sample of pixels (distribution-level stand-in for MiniBatchKMeans.fit, whose
//                   result depends on the sklearn version and its RNG - SURVEY 7.2 item 7);
//                   skipped when the float64 centres are injected;
//   k_quant_palette centres.astype(uint8) -> Lab -> BGR for the k palette entries;
//   k_quant_apply   BGR -> Lab (OpenCV's 8-bit fixed-point path, tables in b2d_lab_tables.h),
//                   label = argmin of the float64 squared distance in NumPy's operation order,
//                   write the palette colour: 3 B read + 3 B written per pixel.
// The two Lab conversions are bit-exact against cv2 over all 2^24 inputs (scripts/gen_lab_tables.py,
// tests/test_gpu_quant.py).
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <algorithm>

#define B2D_LAB_TABLE __device__ const
#include "b2d_lab_tables.h"

namespace b2d {

constexpr int QK_MAX = 32;            // validate_spec bounds k
constexpr int Q_SAMPLES = 4096;       // pixels sampled per image for the fit
constexpr int Q_ITERS = 16;

__device__ __forceinline__ int descale(int x, int n) { return (x + (1 << (n - 1))) >> n; }
__device__ __forceinline__ int clamp255(int v) { return min(max(v, 0), 255); }

// RGB2Lab_b (cv2, uint8): src in B, G, R order.  gam / cb may point to shared copies of the tables.
__device__ __forceinline__ uint32_t lab_forward(uint32_t b, uint32_t g, uint32_t r, const uint16_t* gam,
                                                const uint16_t* cb) {
  const int sb = gam[b], sg = gam[g], sr = gam[r];
  const int fx = cb[descale(sb * kLabFwd[0] + sg * kLabFwd[1] + sr * kLabFwd[2], 12)];
  const int fy = cb[descale(sb * kLabFwd[3] + sg * kLabFwd[4] + sr * kLabFwd[5], 12)];
  const int fz = cb[descale(sb * kLabFwd[6] + sg * kLabFwd[7] + sr * kLabFwd[8], 12)];
  const int L = clamp255(descale(296 * fy - 1336934, 15));
  const int a = clamp255(descale(500 * (fx - fy) + 128 * 32768, 15));
  const int bb = clamp255(descale(200 * (fy - fz) + 128 * 32768, 15));
  return (uint32_t)L | ((uint32_t)a << 8) | ((uint32_t)bb << 16);
}

// abToXZ_b as a function (C integer division truncates toward zero, like the table's initialiser)
__device__ __forceinline__ int ab_to_xz(int i) {
  if (i <= 3390) return i * 108 / 841 - (1 << 14) * 16 / 116 * 108 / 841;
  return i * i / (1 << 14) * i / (1 << 14);
}

// Lab2RGBinteger (cv2, uint8): returns B | G << 8 | R << 16
__device__ __forceinline__ uint32_t lab_inverse(int L, int a, int b) {
  const int y = kLabYF[2 * L], ify = kLabYF[2 * L + 1];
  const int adiv = ((5 * a * 53687 + (1 << 7)) >> 13) - 128 * (1 << 14) / 500;
  const int bdiv = ((b * 41943 + (1 << 4)) >> 9) - 128 * (1 << 14) / 200 + 1;
  const int x = ab_to_xz(ify + adiv), z = ab_to_xz(ify - bdiv);
  int c[3];
#pragma unroll
  for (int k = 0; k < 3; ++k)
    c[k] = min(max(descale(kLabInv[3 * k] * x + kLabInv[3 * k + 1] * y + kLabInv[3 * k + 2] * z, 14), 0), 4095);
  return (uint32_t)kLabInvGamma[c[2]] | ((uint32_t)kLabInvGamma[c[1]] << 8) | ((uint32_t)kLabInvGamma[c[0]] << 16);
}

__device__ __forceinline__ void stage_tables(uint16_t* s_gam, uint16_t* s_cb) {
  for (int i = threadIdx.x; i < 256; i += blockDim.x) s_gam[i] = kLabGamma[i];
  for (int i = threadIdx.x; i < 3072; i += blockDim.x) s_cb[i] = kLabCbrt[i];
  __syncthreads();
}

// ------------------------------------------------------------------ fit --
// One CTA of 512 threads per image (at 128 images per GPU - a C5 shard at N = 8 - the kernel's time is ONE CTA's latency).
constexpr int QF_THREADS = 512;
constexpr int QF_PER = Q_SAMPLES / QF_THREADS;   // samples per thread

__global__ void __launch_bounds__(QF_THREADS) k_quant_fit(ImageSet s, int k, uint32_t k0, uint32_t k1, uint64_t first_global,
                                                          double* centres /* [batch][QK_MAX][3] */) {
  __shared__ uint16_t s_gam[256], s_cb[3072];
  __shared__ uint32_t smp[Q_SAMPLES];            // L | a << 8 | b << 16
  __shared__ float mind[Q_SAMPLES];              // seeding: min distance^2; Lloyd: the warps' partial sums (wpart)
  __shared__ float cen[QK_MAX][3];
  __shared__ float4 sc4[QK_MAX];                 // (-2 c, |c|^2): argmin |x - c|^2 = argmin (|c|^2 - 2 x.c)
  __shared__ unsigned int sums[QK_MAX][4];       // integer sums: deterministic accumulation
  __shared__ float part[QF_THREADS];
  stage_tables(s_gam, s_cb);
  const int tid = threadIdx.x;
  const int img_local = image_of(s, blockIdx.x);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const uint64_t gi = first_global + (uint64_t)img_local;
  const uint32_t npix = (uint32_t)(s.h * s.w);
  for (int i = tid; i < Q_SAMPLES / 4; i += QF_THREADS) {
    const uint4 r = philox4x32((uint32_t)i, 0xC0u, (uint32_t)gi, (uint32_t)(gi >> 32), k0, k1);
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t p = (uint32_t)(((uint64_t)rr[j] * npix) >> 32);
      const uint8_t* px = img + (size_t)p * 3;
      smp[4 * i + j] = lab_forward(px[0], px[1], px[2], s_gam, s_cb);
    }
  }
  __syncthreads();
  auto dist2 = [&](uint32_t v, int c) {
    const float d0 = (float)(v & 255u) - cen[c][0], d1 = (float)((v >> 8) & 255u) - cen[c][1],
                d2 = (float)((v >> 16) & 255u) - cen[c][2];
    return d0 * d0 + d1 * d1 + d2 * d2;
  };
  // ---- k-means++ seeding: next centre drawn with probability proportional to min distance^2
  const uint4 ru = philox4x32(0u, 0xC1u, (uint32_t)gi, (uint32_t)(gi >> 32), k0, k1);
  if (tid == 0) {
    const uint32_t v = smp[ru.x % Q_SAMPLES];
    cen[0][0] = (float)(v & 255u); cen[0][1] = (float)((v >> 8) & 255u); cen[0][2] = (float)((v >> 16) & 255u);
  }
  for (int i = tid; i < Q_SAMPLES; i += QF_THREADS) mind[i] = 3.0e38f;
  __syncthreads();
  for (int c = 1; c < k; ++c) {
    // a run of 16 consecutive samples per partial sum (threads 0 .. 255; the draw below walks runs, then samples)
    if (tid < 256) {
      float loc = 0.f;
      for (int j = 0; j < Q_SAMPLES / 256; ++j) {
        const int i = tid * (Q_SAMPLES / 256) + j;
        const float d = fminf(mind[i], dist2(smp[i], c - 1));
        mind[i] = d;
        loc += d;
      }
      part[tid] = loc;
    }
    __syncthreads();
    if (tid == 0) {
      float tot = 0.f;
      for (int t = 0; t < 256; ++t) tot += part[t];
      const uint4 rc = philox4x32((uint32_t)c, 0xC2u, (uint32_t)gi, (uint32_t)(gi >> 32), k0, k1);
      float target = tot * ((float)(rc.x >> 8) * (1.0f / 16777216.0f));
      int t = 0;
      while (t < 255 && target >= part[t]) { target -= part[t]; ++t; }
      int i = t * (Q_SAMPLES / 256);
      const int end = i + Q_SAMPLES / 256 - 1;
      while (i < end && target >= mind[i]) { target -= mind[i]; ++i; }
      const uint32_t v = smp[i];
      cen[c][0] = (float)(v & 255u); cen[c][1] = (float)((v >> 8) & 255u); cen[c][2] = (float)((v >> 16) & 255u);
    }
    __syncthreads();
  }
  // ---- Lloyd iterations on the sample.  A thread labels its samples in registers by the fp32 score |c|^2 - 2 x.c (one
  // LDS.128 + 3 FFMA per centre); every centre's integer sums (deterministic in any order) are then reduced inside the warp
  // (REDUX) and across the warps through shared memory.  Round 1: 3 LDS + 3 I2F + 8 more per distance, and 4 shared
  // atomicAdds per sample on k x 4 addresses.
  unsigned int (*wpart)[QK_MAX][4] = reinterpret_cast<unsigned int (*)[QK_MAX][4]>(mind);   // [QF_THREADS / 32][QK_MAX][4]
  static_assert(sizeof(unsigned int) * (QF_THREADS / 32) * QK_MAX * 4 <= sizeof(float) * Q_SAMPLES, "wpart fits in mind");
  const int lane = tid & 31, wid = tid >> 5;
  for (int it = 0; it < Q_ITERS; ++it) {
    if (tid < k)
      sc4[tid] = make_float4(-2.0f * cen[tid][0], -2.0f * cen[tid][1], -2.0f * cen[tid][2],
                             cen[tid][0] * cen[tid][0] + cen[tid][1] * cen[tid][1] + cen[tid][2] * cen[tid][2]);
    __syncthreads();
    uint32_t lab[QF_PER];
    unsigned long long pk[QF_PER];                 // L | a << 12 | b << 24 | 1 << 36: one 64-bit add per sample and centre
#pragma unroll
    for (int j = 0; j < QF_PER; ++j) {
      const uint32_t v = smp[tid + QF_THREADS * j];
      const uint32_t L = v & 255u, a = (v >> 8) & 255u, b = (v >> 16) & 255u;
      const float f0 = (float)L, f1 = (float)a, f2 = (float)b;
      int best = 0;
      float bd = 3.0e38f;
      for (int c = 0; c < k; ++c) {
        const float4 m = sc4[c];
        const float d = fmaf(f0, m.x, fmaf(f1, m.y, fmaf(f2, m.z, m.w)));
        if (d < bd) { bd = d; best = c; }
      }
      lab[j] = (uint32_t)best;
      pk[j] = (unsigned long long)L | ((unsigned long long)a << 12) | ((unsigned long long)b << 24) | (1ull << 36);
    }
    for (int c = 0; c < k; ++c) {
      unsigned long long acc = 0ull;
#pragma unroll
      for (int j = 0; j < QF_PER; ++j)
        if (lab[j] == (uint32_t)c) acc += pk[j];
      const unsigned int a0 = __reduce_add_sync(0xffffffffu, (unsigned int)(acc & 0xfffu));
      const unsigned int a1 = __reduce_add_sync(0xffffffffu, (unsigned int)((acc >> 12) & 0xfffu));
      const unsigned int a2 = __reduce_add_sync(0xffffffffu, (unsigned int)((acc >> 24) & 0xfffu));
      const unsigned int n = __reduce_add_sync(0xffffffffu, (unsigned int)(acc >> 36));
      if (lane == 0) { wpart[wid][c][0] = a0; wpart[wid][c][1] = a1; wpart[wid][c][2] = a2; wpart[wid][c][3] = n; }
    }
    __syncthreads();
    if (tid < k * 4) {
      const int c = tid >> 2, x = tid & 3;
      unsigned int t = 0u;
#pragma unroll
      for (int w = 0; w < QF_THREADS / 32; ++w) t += wpart[w][c][x];
      sums[c][x] = t;
    }
    __syncthreads();
    if (tid < k && sums[tid][3]) {
      const float inv = 1.0f / (float)sums[tid][3];
      cen[tid][0] = (float)sums[tid][0] * inv; cen[tid][1] = (float)sums[tid][1] * inv; cen[tid][2] = (float)sums[tid][2] * inv;
    }
    __syncthreads();
  }
  if (tid < k) {
    double* o = centres + ((size_t)img_local * QK_MAX + tid) * 3;
    if (sums[tid][3]) {
      const double n = (double)sums[tid][3];
      o[0] = (double)sums[tid][0] / n; o[1] = (double)sums[tid][1] / n; o[2] = (double)sums[tid][2] / n;
    } else {
      o[0] = cen[tid][0]; o[1] = cen[tid][1]; o[2] = cen[tid][2];
    }
  }
}

// injected centres (fields.f0: [batch][k][3] float64) -> the padded workspace layout
__global__ void k_quant_take(ImageSet s, int k, const double* f0, double* centres) {
  const int img_local = image_of(s, blockIdx.x);
  for (int i = threadIdx.x; i < k * 3; i += blockDim.x)
    centres[(size_t)img_local * QK_MAX * 3 + i] = f0[(size_t)img_local * k * 3 + i];
}

// palette[c] = LAB2BGR(centres[c].astype(uint8)); astype on an in-range positive float truncates
__global__ void k_quant_palette(ImageSet s, int k, const double* centres, uint32_t* palette) {
  const int img_local = image_of(s, blockIdx.x);
  const int c = threadIdx.x;
  if (c >= k) return;
  const double* ce = centres + ((size_t)img_local * QK_MAX + c) * 3;
  palette[(size_t)img_local * QK_MAX + c] =
      lab_inverse((int)wrap_u8(ce[0]), (int)wrap_u8(ce[1]), (int)wrap_u8(ce[2]));
}

// ---------------------------------------------------------------- apply --
// label of one pixel (packed Lab) -> palette colour.  The label is the argmin of the float64 distances in NumPy's
// operation order - but 66 fp64 operations per pixel made the kernel fp64-pipe bound (2.6 ms per 512 images), and
// fp32 distances from shared memory still cost 14 instructions per centre.  Here fp32 SCORES decide first:
// |x - c|^2 - |x|^2 = |c|^2 - 2 x.c, three FFMAs on one LDS.128 per centre.  For x in [0, 255]^3 and centres inside
// [-512, 512]^3 a score is off by < 0.09 (rounding of the products / sums, of -2c and |c|^2 to fp32), so when the
// runner-up is more than 0.25 behind, the fp32 winner IS the float64 argmin; the pixels that sit nearly midway between
// two centres (< 1 %) take the exact loop.
__device__ __forceinline__ uint32_t quant_colour(uint32_t lab, int k, const float4* sc, const double (*cen)[3],
                                                 const uint32_t* pal, int wild) {
  const float f0 = (float)(lab & 255u), f1 = (float)((lab >> 8) & 255u), f2 = (float)((lab >> 16) & 255u);
  int best = 0;
  float b1 = 3.0e38f, b2 = 3.0e38f;                // smallest and second smallest score
  for (int c = 0; c < k; ++c) {
    const float4 m = sc[c];                        // (-2 c0, -2 c1, -2 c2, |c|^2)
    const float d = fmaf(f0, m.x, fmaf(f1, m.y, fmaf(f2, m.z, m.w)));
    if (d < b1) { b2 = b1; b1 = d; best = c; }
    else b2 = fminf(b2, d);
  }
  if (wild || !(b2 - b1 > 0.25f)) {
    const double x0 = (double)(lab & 255u), x1 = (double)((lab >> 8) & 255u), x2 = (double)((lab >> 16) & 255u);
    double bd = 0.0;
    best = 0;
    for (int c = 0; c < k; ++c) {
      // ((x - c) ** 2).sum(-1): products and sums rounded separately, left to right (no FMA)
      const double d0 = __dsub_rn(x0, cen[c][0]), d1 = __dsub_rn(x1, cen[c][1]), d2 = __dsub_rn(x2, cen[c][2]);
      const double d = __dadd_rn(__dadd_rn(__dmul_rn(d0, d0), __dmul_rn(d1, d1)), __dmul_rn(d2, d2));
      if (c == 0 || d < bd) { bd = d; best = c; }       // np.argmin: first minimum
    }
  }
  return pal[best];
}

__global__ void __launch_bounds__(256) k_quant_apply(ImageSet s, int k, const double* centres, const uint32_t* palette) {
  __shared__ uint16_t s_gam[256], s_cb[3072];
  __shared__ double cen[QK_MAX][3];
  __shared__ float4 sc[QK_MAX];
  __shared__ uint32_t pal[QK_MAX];
  __shared__ int s_wild;                           // an (injected) centre far outside the uint8 cube: exact loop only
  const int img_local = image_of(s, blockIdx.z);
  if (threadIdx.x == 0) s_wild = 0;
  __syncthreads();
  if (threadIdx.x < k * 3) {
    const double c = centres[(size_t)img_local * QK_MAX * 3 + threadIdx.x];
    (&cen[0][0])[threadIdx.x] = c;
    if (!(fabs(c) <= 512.0)) s_wild = 1;
  }
  if (threadIdx.x < k) {
    const double* c = centres + ((size_t)img_local * QK_MAX + threadIdx.x) * 3;
    sc[threadIdx.x] = make_float4((float)(-2.0 * c[0]), (float)(-2.0 * c[1]), (float)(-2.0 * c[2]),
                                  (float)(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]));
    pal[threadIdx.x] = palette[(size_t)img_local * QK_MAX + threadIdx.x];
  }
  stage_tables(s_gam, s_cb);
  const int wild = s_wild;
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  const int npix = s.h * s.w;
  auto one = [&](uint32_t b, uint32_t g, uint32_t r) {
    return quant_colour(lab_forward(b, g, r, s_gam, s_cb), k, sc, cen, pal, wild);
  };
  // four pixels = three words per thread and step when the image is word aligned
  const bool words = ((reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(out)) & 3u) == 0;
  const int nquad = words ? npix / 4 : 0;
  for (int q = blockIdx.x * 256 + threadIdx.x; q < nquad; q += gridDim.x * 256) {
    const uint32_t* pw = reinterpret_cast<const uint32_t*>(img) + (size_t)q * 3;
    const uint32_t w0 = __ldg(pw), w1 = __ldg(pw + 1), w2 = __ldg(pw + 2);
    const uint32_t q0 = one(w0 & 255u, (w0 >> 8) & 255u, (w0 >> 16) & 255u);
    const uint32_t q1 = one(w0 >> 24, w1 & 255u, (w1 >> 8) & 255u);
    const uint32_t q2 = one((w1 >> 16) & 255u, w1 >> 24, w2 & 255u);
    const uint32_t q3 = one((w2 >> 8) & 255u, (w2 >> 16) & 255u, w2 >> 24);
    uint32_t* ow = reinterpret_cast<uint32_t*>(out) + (size_t)q * 3;
    ow[0] = q0 | (q1 << 24);
    ow[1] = (q1 >> 8) | (q2 << 16);
    ow[2] = (q2 >> 16) | (q3 << 8);
  }
  for (int p = 4 * nquad + blockIdx.x * 256 + threadIdx.x; p < npix; p += gridDim.x * 256) {
    const uint8_t* px = img + (size_t)p * 3;
    const uint32_t q = one(px[0], px[1], px[2]);
    uint8_t* o = out + (size_t)p * 3;
    o[0] = (uint8_t)q; o[1] = (uint8_t)(q >> 8); o[2] = (uint8_t)(q >> 16);
  }
}

// test hook: the two colour conversions alone (dir 0: BGR -> Lab, 1: Lab -> BGR), n pixels
__global__ void k_lab_convert(const uint8_t* in, uint8_t* out, size_t n, int dir) {
  for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (size_t)gridDim.x * blockDim.x) {
    const uint8_t* px = in + p * 3;
    const uint32_t v = dir == 0 ? lab_forward(px[0], px[1], px[2], kLabGamma, kLabCbrt) : lab_inverse(px[0], px[1], px[2]);
    out[p * 3] = (uint8_t)v; out[p * 3 + 1] = (uint8_t)(v >> 8); out[p * 3 + 2] = (uint8_t)(v >> 16);
  }
}

cudaError_t launch_lab_convert(const uint8_t* in, uint8_t* out, size_t n, int dir, cudaStream_t st) {
  k_lab_convert<<<148 * 8, 256, 0, st>>>(in, out, n, dir);
  count_launch();
  return cudaGetLastError();
}

size_t quant_table_bytes(int n) { return (size_t)n * QK_MAX * (3 * sizeof(double) + sizeof(uint32_t)) + 256; }

int run_quant(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, uint8_t* tables, const b2d_fields* fields,
              uint64_t seed, uint64_t first_global, cudaStream_t st) {
  const int k = (int)sp.level;
  double* centres = reinterpret_cast<double*>(tables);
  uint32_t* palette = reinterpret_cast<uint32_t*>(tables + (size_t)p.n * QK_MAX * 3 * sizeof(double));
  if (fields && fields->f0) k_quant_take<<<s.count, 128, 0, st>>>(s, k, fields->f0, centres);
  else k_quant_fit<<<s.count, QF_THREADS, 0, st>>>(s, k, (uint32_t)seed, (uint32_t)(seed >> 32), first_global, centres);
  count_launch();
  k_quant_palette<<<s.count, QK_MAX, 0, st>>>(s, k, centres, palette);
  count_launch();
  // every CTA stages 6.6 KB of Lab tables first: as few CTAs per image as still fill the GPU (16 per SM over the batch)
  const size_t want = ((size_t)sm_count() * 16 + (size_t)s.count - 1) / (size_t)s.count;
  const unsigned bx = (unsigned)std::max<size_t>(1, std::min<size_t>(((size_t)p.h * p.w + 255) / 256, std::min<size_t>(want, 148 * 4)));
  k_quant_apply<<<dim3(bx, 1, s.count), 256, 0, st>>>(s, k, centres, palette);
  count_launch();
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? B2D_OK : fail_cuda(e, "quant launch");
}

}  // namespace b2d
