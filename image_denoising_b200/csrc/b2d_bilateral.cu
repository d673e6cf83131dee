// b2d_bilateral.cu - cv2.bilateralFilter(u8, d, sigmaColor, sigmaSpace,
// borderType=BORDER_CONSTANT) on BGR uint8 (call sites
// lib/roi_data_layer/minibatch.py:166-183: d=9, sigmaColor=20, sigmaSpace=100).
//
// OpenCV's formulation, reproduced operation for operation in fp32:
//   taps    = offsets (i, j) with sqrt(i^2 + j^2) <= r, r = d / 2, row-major order
//   border  = zero pixels that PARTICIPATE (BORDER_CONSTANT)
//   w       = space[i, j] * color_lut[|db| + |dg| + |dr|]          (lut: 766 floats)
//   out_c   = round_half_even(sum(w * p_c) * (1 / sum(w)))
//
// Kernel (radius 4): CTA = 16 rows x 64 pixels, thread = 4 adjacent pixels.  The first version kept
// b / g / r as fp32 planes in shared memory and was bound by shared-memory wavefronts (ncu: LSU
// data pipe 91 %, `mio_throttle` the top stall).  Now the tile holds only the packed pixel (4 B):
//   * a neighbour's three floats are rebuilt in registers (PRMT onto 2^23, one FADD each) and
//     shared by the up to 4 outputs of the thread that use it;
//   * the spatial weight is folded into the colour LUT: r^2 takes 10 values inside the disc, so
//     lut2[class][t] = space * lut[t] (the same fp32 product OpenCV forms per tap) costs one
//     shared-memory load and no multiply;
//   * the four accumulators of an output are two packed-fp32 pairs (b, g) and (r, weight sum),
//     updated by two FFMA2 (`fma.rn.f32x2`, IEEE per lane: results identical to scalar code).
// Per tap: VABSDIFF4, LEA, LDS, MOV, 2 x FFMA2 = 6 issue slots / 6 ALU-pipe cycles.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
#include "b2d_noise.cuh"

#include <cmath>

namespace b2d {

constexpr int BL_R = 4;
constexpr int BL_TW = 64, BL_TH = 16, BL_PX = 4;
constexpr int BL_SW = BL_TW + 2 * BL_R;          // 72
constexpr int BL_SH = BL_TH + 2 * BL_R;          // 24
constexpr int BL_PLANE = BL_SW * BL_SH;          // 1728
constexpr int BL_THREADS = (BL_TW / BL_PX) * BL_TH;   // 256
constexpr int BL_LUT = 256 * 3;
constexpr int BL_CLASSES = 10;                   // distinct i^2 + j^2 <= 16
constexpr int BL_TILE_ITERS = (BL_PLANE + BL_THREADS - 1) / BL_THREADS;   // 7

__host__ __device__ constexpr int bl_class(int i, int j) {
  const int r2 = i * i + j * j;
  return r2 == 0 ? 0 : r2 == 1 ? 1 : r2 == 2 ? 2 : r2 == 4 ? 3 : r2 == 5 ? 4 : r2 == 8 ? 5 : r2 == 9 ? 6
       : r2 == 10 ? 7 : r2 == 13 ? 8 : r2 == 16 ? 9 : -1;
}

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long v;
  asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi));
  return v;
}
__device__ __forceinline__ void fma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}

__global__ void __launch_bounds__(BL_THREADS, 4) k_bilateral_r4(ImageSet s, const float* lut2g) {
  extern __shared__ __align__(16) uint32_t smu[];
  uint32_t* pk = smu;
  float* lut2 = reinterpret_cast<float*>(smu + BL_PLANE);

  const int tid = threadIdx.x;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int x0 = blockIdx.x * BL_TW, y0 = blockIdx.y * BL_TH;

  for (int i = tid; i < BL_CLASSES * BL_LUT / 4; i += BL_THREADS)
    reinterpret_cast<float4*>(lut2)[i] = __ldg(reinterpret_cast<const float4*>(lut2g) + i);
  {                                               // tile + apron: all loads in flight before the first store
    uint32_t v[BL_TILE_ITERS][3];
#pragma unroll
    for (int k = 0; k < BL_TILE_ITERS; ++k) {
      const int i = tid + k * BL_THREADS;
      const int ty = i / BL_SW, tx = i - ty * BL_SW;
      const int y = y0 + ty - BL_R, x = x0 + tx - BL_R;
      const bool in = i < BL_PLANE && y >= 0 && y < s.h && x >= 0 && x < s.w;   // BORDER_CONSTANT: zeros take part
      const uint8_t* p = img + (in ? ((uint64_t)y * s.w + x) * 3 : 0);
#pragma unroll
      for (int c = 0; c < 3; ++c) v[k][c] = in ? (uint32_t)__ldg(p + c) : 0u;
    }
#pragma unroll
    for (int k = 0; k < BL_TILE_ITERS; ++k) {
      const int i = tid + k * BL_THREADS;
      if (i < BL_PLANE) pk[i] = v[k][0] | (v[k][1] << 8) | (v[k][2] << 16);
    }
  }
  __syncthreads();

  const int tx = tid & 15, ty = tid >> 4;
  const int base = (ty + BL_R) * BL_SW + BL_PX * tx + BL_R;      // centre of output 0
  uint32_t ck[BL_PX];
  unsigned long long acc_bg[BL_PX], acc_rw[BL_PX];               // (sum w b, sum w g), (sum w r, sum w)
#pragma unroll
  for (int p = 0; p < BL_PX; ++p) {
    ck[p] = pk[base + p];
    acc_bg[p] = acc_rw[p] = 0ull;
  }
#pragma unroll
  for (int i = -BL_R; i <= BL_R; ++i) {
    const int rowoff = base + i * BL_SW - BL_R;                  // neighbour column jj = 0 .. 11, 16-byte aligned
    uint32_t nk[12];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const uint4 vk = *reinterpret_cast<const uint4*>(pk + rowoff + 4 * q);
      nk[4 * q] = vk.x; nk[4 * q + 1] = vk.y; nk[4 * q + 2] = vk.z; nk[4 * q + 3] = vk.w;
    }
#pragma unroll
    for (int jj = 0; jj < 12; ++jj) {
      bool used = false;
#pragma unroll
      for (int p = 0; p < BL_PX; ++p) {
        const int j = jj - p - BL_R;
        used = used || (j >= -BL_R && j <= BL_R && bl_class(i, j) >= 0);
      }
      if (!used) continue;
      const unsigned long long n_bg = pack2(magic_byte<0>(nk[jj]) - 8388608.0f, magic_byte<1>(nk[jj]) - 8388608.0f);
      const unsigned long long n_r1 = pack2(magic_byte<2>(nk[jj]) - 8388608.0f, 1.0f);
#pragma unroll
      for (int p = 0; p < BL_PX; ++p) {
        const int j = jj - p - BL_R;                             // tap column offset for output p
        if (j < -BL_R || j > BL_R || bl_class(i, j) < 0) continue;
        const uint32_t t = __vsadu4(nk[jj], ck[p]);
        const float w = lut2[bl_class(i, j) * BL_LUT + t];
        const unsigned long long ww = pack2(w, w);
        fma2(acc_bg[p], n_bg, ww);
        fma2(acc_rw[p], n_r1, ww);
      }
    }
  }
  const int y = y0 + ty, x = x0 + BL_PX * tx;
  if (y >= s.h || x >= s.w) return;
  uint32_t o[12];
#pragma unroll
  for (int p = 0; p < BL_PX; ++p) {
    float ab, ag, ar, wsum;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(ab), "=f"(ag) : "l"(acc_bg[p]));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(ar), "=f"(wsum) : "l"(acc_rw[p]));
    const float inv = 1.0f / wsum;
    o[3 * p] = (uint32_t)__float2int_rn(ab * inv);
    o[3 * p + 1] = (uint32_t)__float2int_rn(ag * inv);
    o[3 * p + 2] = (uint32_t)__float2int_rn(ar * inv);
  }
  uint8_t* op = s.out + (uint64_t)img_local * s.elems + ((uint64_t)y * s.w + x) * 3;
  if (x + BL_PX <= s.w && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
    uint32_t* ow = reinterpret_cast<uint32_t*>(op);
    ow[0] = pack4(o[0], o[1], o[2], o[3]);
    ow[1] = pack4(o[4], o[5], o[6], o[7]);
    ow[2] = pack4(o[8], o[9], o[10], o[11]);
  } else {
    for (int p = 0; p < BL_PX && x + p < s.w; ++p)
      for (int c = 0; c < 3; ++c) op[3 * p + c] = (uint8_t)o[3 * p + c];
  }
}

// ---------------------------------------------------------------------------------------------------------
// Round 2: the same filter with a thread owning 4 x 2 outputs (tile 64 x 32 per CTA).
//   * 10 neighbour rows serve two output rows: a neighbour's floats are rebuilt 5 times per output row instead of 9
//     (the conversions were 22 % of the instructions), and the (b, g) pair comes out of ONE packed FADD2;
//   * the apron shrinks from 1.69 to 1.41 of the tile and the colour LUT (30 KB from L2 per CTA) is loaded once per
//     2048 outputs instead of 1024.
// Every output still sees its taps in row-major order (neighbour rows top to bottom, columns left to right), the order
// cv2 accumulates in - the results are bit-identical to the 4 x 1 kernel (B2D_BILATERAL=1 runs it: cross-check).
// ---------------------------------------------------------------------------------------------------------
constexpr int B2_TW = 64, B2_TH = 32, B2_PY = 2;
constexpr int B2_SW = B2_TW + 2 * BL_R;          // 72
constexpr int B2_SH = B2_TH + 2 * BL_R;          // 40
constexpr int B2_PLANE = B2_SW * B2_SH;          // 2880
constexpr int B2_THREADS = (B2_TW / BL_PX) * (B2_TH / B2_PY);   // 256
constexpr int B2_TILE_ITERS = (B2_PLANE + B2_THREADS - 1) / B2_THREADS;   // 12

__device__ __forceinline__ unsigned long long fadd2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// tmax: the first index from which every class table is exactly 0.0f (exp(-t^2 / (2 sigma_color^2)) underflows: t = 290
// for sigma_color = 20).  Clamping the colour distance there changes no bit of the result (those taps add w = 0) but makes
// every such lane read ONE shared-memory address - a broadcast.  On noised input (C3: speckle var 1.0) most taps are
// there; without the clamp their scattered indices cost 2.8 wavefronts per lookup and the shared-memory pipe, at 93 %,
// bounded the kernel (profiles/r02l).
// CLAMP is off for clean input (`original_bilateral`: colour distances are small, the min would only cost an ALU slot).
template <bool CLAMP>
__global__ void __launch_bounds__(B2_THREADS, 3) k_bilateral_r4x2(ImageSet s, const float* lut2g, uint32_t tmax) {
  extern __shared__ __align__(16) uint32_t smu[];
  uint32_t* pk = smu;
  float* lut2 = reinterpret_cast<float*>(smu + B2_PLANE);

  const int tid = threadIdx.x;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int x0 = blockIdx.x * B2_TW, y0 = blockIdx.y * B2_TH;

  for (int i = tid; i < BL_CLASSES * BL_LUT / 4; i += B2_THREADS)
    reinterpret_cast<float4*>(lut2)[i] = __ldg(reinterpret_cast<const float4*>(lut2g) + i);
#pragma unroll
  for (int half = 0; half < 2; ++half) {          // tile + apron, 6 pixels per thread in flight at a time
    uint32_t v[B2_TILE_ITERS / 2][3];
#pragma unroll
    for (int k = 0; k < B2_TILE_ITERS / 2; ++k) {
      const int i = tid + (half * (B2_TILE_ITERS / 2) + k) * B2_THREADS;
      const int ty = i / B2_SW, tx = i - ty * B2_SW;
      const int y = y0 + ty - BL_R, x = x0 + tx - BL_R;
      const bool in = i < B2_PLANE && y >= 0 && y < s.h && x >= 0 && x < s.w;   // BORDER_CONSTANT: zeros take part
      const uint8_t* p = img + (in ? ((uint64_t)y * s.w + x) * 3 : 0);
#pragma unroll
      for (int c = 0; c < 3; ++c) v[k][c] = in ? (uint32_t)__ldg(p + c) : 0u;
    }
#pragma unroll
    for (int k = 0; k < B2_TILE_ITERS / 2; ++k) {
      const int i = tid + (half * (B2_TILE_ITERS / 2) + k) * B2_THREADS;
      if (i < B2_PLANE) pk[i] = v[k][0] | (v[k][1] << 8) | (v[k][2] << 16);
    }
  }
  __syncthreads();

  const int tx = tid & 15, ty = tid >> 4;
  const int base = (B2_PY * ty + BL_R) * B2_SW + BL_PX * tx + BL_R;   // centre of output (row 0, pixel 0)
  uint32_t ck[B2_PY][BL_PX];
  unsigned long long acc_bg[B2_PY][BL_PX], acc_rw[B2_PY][BL_PX];     // (sum w b, sum w g), (sum w r, sum w)
#pragma unroll
  for (int ry = 0; ry < B2_PY; ++ry)
#pragma unroll
    for (int p = 0; p < BL_PX; ++p) {
      ck[ry][p] = pk[base + ry * B2_SW + p];
      acc_bg[ry][p] = acc_rw[ry][p] = 0ull;
    }
  const unsigned long long neg_magic = pack2(-8388608.0f, -8388608.0f);
#pragma unroll
  for (int ni = 0; ni < 2 * BL_R + B2_PY; ++ni) {                    // neighbour row = output row 0 + ni - 4
    const int rowoff = base + (ni - BL_R) * B2_SW - BL_R;             // neighbour column jj = 0 .. 11, 16-byte aligned
    uint32_t nk[12];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const uint4 vk = *reinterpret_cast<const uint4*>(pk + rowoff + 4 * q);
      nk[4 * q] = vk.x; nk[4 * q + 1] = vk.y; nk[4 * q + 2] = vk.z; nk[4 * q + 3] = vk.w;
    }
#pragma unroll
    for (int jj = 0; jj < 12; ++jj) {
      bool used = false;
#pragma unroll
      for (int ry = 0; ry < B2_PY; ++ry)
#pragma unroll
        for (int p = 0; p < BL_PX; ++p) {
          const int i = ni - BL_R - ry, j = jj - p - BL_R;
          used = used || (i >= -BL_R && i <= BL_R && j >= -BL_R && j <= BL_R && bl_class(i, j) >= 0);
        }
      if (!used) continue;
      const unsigned long long n_bg = fadd2(pack2(magic_byte<0>(nk[jj]), magic_byte<1>(nk[jj])), neg_magic);
      const unsigned long long n_r1 = pack2(magic_byte<2>(nk[jj]) - 8388608.0f, 1.0f);
#pragma unroll
      for (int ry = 0; ry < B2_PY; ++ry) {
        const int i = ni - BL_R - ry;                                 // tap row offset for output row ry
        if (i < -BL_R || i > BL_R) continue;
#pragma unroll
        for (int p = 0; p < BL_PX; ++p) {
          const int j = jj - p - BL_R;                                // tap column offset for output p
          if (j < -BL_R || j > BL_R || bl_class(i, j) < 0) continue;
          const uint32_t sad = __vsadu4(nk[jj], ck[ry][p]);
          const uint32_t t = CLAMP ? min(sad, tmax) : sad;
          const float w = lut2[bl_class(i, j) * BL_LUT + t];
          const unsigned long long ww = pack2(w, w);
          fma2(acc_bg[ry][p], n_bg, ww);
          fma2(acc_rw[ry][p], n_r1, ww);
        }
      }
    }
  }
  const int x = x0 + BL_PX * tx;
  if (x >= s.w) return;
#pragma unroll
  for (int ry = 0; ry < B2_PY; ++ry) {
    const int y = y0 + B2_PY * ty + ry;
    if (y >= s.h) break;
    uint32_t o[12];
#pragma unroll
    for (int p = 0; p < BL_PX; ++p) {
      float ab, ag, ar, wsum;
      asm("mov.b64 {%0, %1}, %2;" : "=f"(ab), "=f"(ag) : "l"(acc_bg[ry][p]));
      asm("mov.b64 {%0, %1}, %2;" : "=f"(ar), "=f"(wsum) : "l"(acc_rw[ry][p]));
      const float inv = 1.0f / wsum;
      o[3 * p] = (uint32_t)__float2int_rn(ab * inv);
      o[3 * p + 1] = (uint32_t)__float2int_rn(ag * inv);
      o[3 * p + 2] = (uint32_t)__float2int_rn(ar * inv);
    }
    uint8_t* op = s.out + (uint64_t)img_local * s.elems + ((uint64_t)y * s.w + x) * 3;
    if (x + BL_PX <= s.w && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
      uint32_t* ow = reinterpret_cast<uint32_t*>(op);
      ow[0] = pack4(o[0], o[1], o[2], o[3]);
      ow[1] = pack4(o[4], o[5], o[6], o[7]);
      ow[2] = pack4(o[8], o[9], o[10], o[11]);
    } else {
      for (int p = 0; p < BL_PX && x + p < s.w; ++p)
        for (int c = 0; c < 3; ++c) op[3 * p + c] = (uint8_t)o[3 * p + c];
    }
  }
}

// any diameter: one thread per pixel, taps from global memory (L1/L2 cached)
__global__ void __launch_bounds__(256) k_bilateral_any(ImageSet s, int radius, const float* space, const float* lut) {
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const uint64_t npix = (uint64_t)s.h * s.w;
  const int d = 2 * radius + 1;
  for (uint64_t pi = (uint64_t)blockIdx.x * 256 + threadIdx.x; pi < npix; pi += (uint64_t)gridDim.x * 256) {
    const int y = (int)(pi / s.w), x = (int)(pi - (uint64_t)y * s.w);
    const uint8_t* c = img + pi * 3;
    const int b0 = c[0], g0 = c[1], r0 = c[2];
    float wsum = 0.f, sb = 0.f, sg = 0.f, sr = 0.f;
    for (int i = -radius; i <= radius; ++i)
      for (int j = -radius; j <= radius; ++j) {
        const float sw = space[(i + radius) * d + j + radius];
        if (sw == 0.0f) continue;
        int b = 0, g = 0, r = 0;
        const int yy = y + i, xx = x + j;
        if (yy >= 0 && yy < s.h && xx >= 0 && xx < s.w) {
          const uint8_t* p = img + ((uint64_t)yy * s.w + xx) * 3;
          b = p[0]; g = p[1]; r = p[2];
        }
        const float w = sw * lut[abs(b - b0) + abs(g - g0) + abs(r - r0)];
        wsum += w;
        sb = fmaf((float)b, w, sb); sg = fmaf((float)g, w, sg); sr = fmaf((float)r, w, sr);
      }
    const float inv = 1.0f / wsum;
    uint8_t* o = s.out + (uint64_t)img_local * s.elems + pi * 3;
    o[0] = (uint8_t)__float2int_rn(sb * inv);
    o[1] = (uint8_t)__float2int_rn(sg * inv);
    o[2] = (uint8_t)__float2int_rn(sr * inv);
  }
}

// host: OpenCV's tables.  ws layout: [lut 768 floats][space (2r+1)^2 floats][pad][lut2 10 x 768 floats (r = 4)]
int run_bilateral(const b2d_spec& sp, const ImageSet& s, uint8_t* ws_tables, cudaStream_t st, bool noisy_input) {
  const int radius = sp.d / 2, d = 2 * radius + 1;
  const double gc = -0.5 / ((double)sp.sigma_color * sp.sigma_color);
  const double gs = -0.5 / ((double)sp.sigma_space * sp.sigma_space);
  constexpr int kLut2Off = BL_LUT + 15 * 15 + 31;                // floats; 16-byte aligned (1024)
  static_assert(kLut2Off % 4 == 0, "lut2 must be float4 aligned");
  float host[kLut2Off + BL_CLASSES * BL_LUT];
  for (int i = 0; i < kLut2Off + BL_CLASSES * BL_LUT; ++i) host[i] = 0.0f;
  for (int i = 0; i < BL_LUT; ++i) host[i] = (float)std::exp((double)i * i * gc);
  for (int i = -radius; i <= radius; ++i)
    for (int j = -radius; j <= radius; ++j) {
      const double r = std::sqrt((double)i * i + (double)j * j);
      host[BL_LUT + (i + radius) * d + j + radius] = r > radius ? 0.0f : (float)std::exp(r * r * gs);
    }
  float* dtab = reinterpret_cast<float*>(ws_tables);
  if (radius == BL_R)
    for (int i = -BL_R; i <= BL_R; ++i)
      for (int j = -BL_R; j <= BL_R; ++j) {
        const int c = bl_class(i, j);
        if (c < 0) continue;
        const float sw = host[BL_LUT + (i + radius) * d + j + radius];
        for (int t = 0; t < BL_LUT; ++t) host[kLut2Off + c * BL_LUT + t] = sw * host[t];   // fp32 product, as per tap
      }
  const size_t nfl = radius == BL_R ? (size_t)kLut2Off + BL_CLASSES * BL_LUT : (size_t)BL_LUT + d * d;
  cudaError_t e = cudaMemcpyAsync(dtab, host, sizeof(float) * nfl, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return fail_cuda(e, "bilateral tables");
  static const bool v1 = [] { const char* e = knob("B2D_BILATERAL"); return e && e[0] == '1'; }();   // the 4 x 1 kernel (cross-check)
  if (radius == BL_R && !v1) {
    const size_t smem = sizeof(uint32_t) * B2_PLANE + sizeof(float) * BL_CLASSES * BL_LUT;
    static bool attr_done[64] = {};             // per device: the attribute lives in the context
    const int dev = current_device() & 63;
    if (!attr_done[dev]) {
      cudaError_t ea = cudaFuncSetAttribute(k_bilateral_r4x2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (ea == cudaSuccess)
        ea = cudaFuncSetAttribute(k_bilateral_r4x2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (ea != cudaSuccess) return fail_cuda(ea, "bilateral shared-memory opt-in");
      attr_done[dev] = true;
    }
    uint32_t tmax = BL_LUT - 2;                   // entries 766, 767 are padding zeros: always a valid all-zero index
    while (tmax > 0) {
      bool zero = true;
      for (int c = 0; c < BL_CLASSES && zero; ++c) zero = host[kLut2Off + c * BL_LUT + tmax - 1] == 0.0f;
      if (!zero) break;
      --tmax;
    }
    const dim3 grid((s.w + B2_TW - 1) / B2_TW, (s.h + B2_TH - 1) / B2_TH, s.count);
    if (noisy_input) k_bilateral_r4x2<true><<<grid, B2_THREADS, smem, st>>>(s, dtab + kLut2Off, tmax);
    else k_bilateral_r4x2<false><<<grid, B2_THREADS, smem, st>>>(s, dtab + kLut2Off, tmax);
  } else if (radius == BL_R) {
    const size_t smem = sizeof(uint32_t) * BL_PLANE + sizeof(float) * BL_CLASSES * BL_LUT;
    static bool attr_done1[64] = {};
    const int dev = current_device() & 63;
    if (!attr_done1[dev]) {
      const cudaError_t ea = cudaFuncSetAttribute(k_bilateral_r4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (ea != cudaSuccess) return fail_cuda(ea, "bilateral shared-memory opt-in");
      attr_done1[dev] = true;
    }
    const dim3 grid((s.w + BL_TW - 1) / BL_TW, (s.h + BL_TH - 1) / BL_TH, s.count);
    k_bilateral_r4<<<grid, BL_THREADS, smem, st>>>(s, dtab + kLut2Off);
  } else {
    const uint64_t npix = (uint64_t)s.h * s.w;
    uint32_t bx = (uint32_t)std::min<uint64_t>((npix + 255) / 256, 148u * 16u);
    k_bilateral_any<<<dim3(bx, 1, s.count), 256, 0, st>>>(s, radius, dtab + BL_LUT, dtab);
  }
  count_launch();
  e = cudaGetLastError();
  return e == cudaSuccess ? B2D_OK : fail_cuda(e, "bilateral launch");
}

}  // namespace b2d
