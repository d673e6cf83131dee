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
// Kernel (radius 4): CTA = 16 rows x 64 pixels, thread = 4 adjacent pixels.
// The tile + apron lives in shared memory as four planes (b, g, r as fp32 and
// the packed pixel) so that a thread's 12 neighbours of one tap row arrive in
// 3 conflict-free LDS.128 per plane and are reused by its 4 outputs; the
// colour distance is one VABSDIFF4 (packed SAD); the LUT sits in shared memory.
// Compute bound: 49 taps x ~8 instructions per pixel (FP32/INT pipe roofline).
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <cmath>

namespace b2d {

constexpr int BL_R = 4;
constexpr int BL_TW = 64, BL_TH = 16, BL_PX = 4;
constexpr int BL_SW = BL_TW + 2 * BL_R;          // 72
constexpr int BL_SH = BL_TH + 2 * BL_R;          // 24
constexpr int BL_PLANE = BL_SW * BL_SH;          // 1728
constexpr int BL_THREADS = (BL_TW / BL_PX) * BL_TH;   // 256
constexpr int BL_LUT = 256 * 3;

struct BilateralArgs {
  float space[2 * BL_R + 1][2 * BL_R + 1];   // 0 outside the disc
  const float* lut;                           // device, BL_LUT floats
};

__global__ void __launch_bounds__(BL_THREADS, 2) k_bilateral_r4(ImageSet s, BilateralArgs ba) {
  extern __shared__ __align__(16) float sm[];
  float* pb = sm;
  float* pg = sm + BL_PLANE;
  float* pr = sm + 2 * BL_PLANE;
  uint32_t* pk = reinterpret_cast<uint32_t*>(sm + 3 * BL_PLANE);
  float* lut = sm + 4 * BL_PLANE;

  const int tid = threadIdx.x;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int x0 = blockIdx.x * BL_TW, y0 = blockIdx.y * BL_TH;

  for (int i = tid; i < BL_LUT; i += BL_THREADS) lut[i] = ba.lut[i];
  for (int i = tid; i < BL_PLANE; i += BL_THREADS) {
    const int ty = i / BL_SW, tx = i - ty * BL_SW;
    const int y = y0 + ty - BL_R, x = x0 + tx - BL_R;
    uint32_t b = 0, g = 0, r = 0;
    if (y >= 0 && y < s.h && x >= 0 && x < s.w) {
      const uint8_t* p = img + ((uint64_t)y * s.w + x) * 3;
      b = __ldg(p); g = __ldg(p + 1); r = __ldg(p + 2);
    }
    pb[i] = (float)b; pg[i] = (float)g; pr[i] = (float)r;
    pk[i] = b | (g << 8) | (r << 16);
  }
  __syncthreads();

  const int tx = tid & 15, ty = tid >> 4;
  const int base = (ty + BL_R) * BL_SW + BL_PX * tx + BL_R;      // centre of output 0
  uint32_t ck[BL_PX];
  float wsum[BL_PX], ab[BL_PX], ag[BL_PX], ar[BL_PX];
#pragma unroll
  for (int p = 0; p < BL_PX; ++p) {
    ck[p] = pk[base + p];
    wsum[p] = ab[p] = ag[p] = ar[p] = 0.0f;
  }
#pragma unroll
  for (int i = -BL_R; i <= BL_R; ++i) {
    const int rowoff = base + i * BL_SW - BL_R;                  // neighbour column jj = 0 .. 11
    float nb[12], ng[12], nr[12];
    uint32_t nk[12];
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      const float4 vb = *reinterpret_cast<const float4*>(pb + rowoff + 4 * q);
      const float4 vg = *reinterpret_cast<const float4*>(pg + rowoff + 4 * q);
      const float4 vr = *reinterpret_cast<const float4*>(pr + rowoff + 4 * q);
      const uint4 vk = *reinterpret_cast<const uint4*>(pk + rowoff + 4 * q);
      nb[4 * q] = vb.x; nb[4 * q + 1] = vb.y; nb[4 * q + 2] = vb.z; nb[4 * q + 3] = vb.w;
      ng[4 * q] = vg.x; ng[4 * q + 1] = vg.y; ng[4 * q + 2] = vg.z; ng[4 * q + 3] = vg.w;
      nr[4 * q] = vr.x; nr[4 * q + 1] = vr.y; nr[4 * q + 2] = vr.z; nr[4 * q + 3] = vr.w;
      nk[4 * q] = vk.x; nk[4 * q + 1] = vk.y; nk[4 * q + 2] = vk.z; nk[4 * q + 3] = vk.w;
    }
#pragma unroll
    for (int jj = 0; jj < 12; ++jj) {
#pragma unroll
      for (int p = 0; p < BL_PX; ++p) {
        const int j = jj - p - BL_R;                             // tap column offset for output p
        if (j < -BL_R || j > BL_R || i * i + j * j > BL_R * BL_R) continue;
        const uint32_t t = __vsadu4(nk[jj], ck[p]);
        const float w = ba.space[i + BL_R][j + BL_R] * lut[t];
        wsum[p] += w;
        ab[p] = fmaf(nb[jj], w, ab[p]);
        ag[p] = fmaf(ng[jj], w, ag[p]);
        ar[p] = fmaf(nr[jj], w, ar[p]);
      }
    }
  }
  const int y = y0 + ty, x = x0 + BL_PX * tx;
  if (y >= s.h || x >= s.w) return;
  uint32_t o[12];
#pragma unroll
  for (int p = 0; p < BL_PX; ++p) {
    const float inv = 1.0f / wsum[p];
    o[3 * p] = (uint32_t)__float2int_rn(ab[p] * inv);
    o[3 * p + 1] = (uint32_t)__float2int_rn(ag[p] * inv);
    o[3 * p + 2] = (uint32_t)__float2int_rn(ar[p] * inv);
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

// host: OpenCV's tables.  ws layout: [lut 768 floats][space (2r+1)^2 floats]
int run_bilateral(const b2d_spec& sp, const ImageSet& s, uint8_t* ws_tables, cudaStream_t st) {
  const int radius = sp.d / 2, d = 2 * radius + 1;
  const double gc = -0.5 / ((double)sp.sigma_color * sp.sigma_color);
  const double gs = -0.5 / ((double)sp.sigma_space * sp.sigma_space);
  float host[BL_LUT + 15 * 15];
  for (int i = 0; i < BL_LUT; ++i) host[i] = (float)std::exp((double)i * i * gc);
  for (int i = -radius; i <= radius; ++i)
    for (int j = -radius; j <= radius; ++j) {
      const double r = std::sqrt((double)i * i + (double)j * j);
      host[BL_LUT + (i + radius) * d + j + radius] = r > radius ? 0.0f : (float)std::exp(r * r * gs);
    }
  float* dtab = reinterpret_cast<float*>(ws_tables);
  cudaError_t e = cudaMemcpyAsync(dtab, host, sizeof(float) * (BL_LUT + d * d), cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return fail_cuda(e, "bilateral tables");
  if (radius == BL_R) {
    BilateralArgs ba;
    for (int i = 0; i < d; ++i)
      for (int j = 0; j < d; ++j) ba.space[i][j] = host[BL_LUT + i * d + j];
    ba.lut = dtab;
    const size_t smem = sizeof(float) * (4 * BL_PLANE + BL_LUT);
    static bool attr_set = false;
    if (!attr_set) {
      cudaFuncSetAttribute(k_bilateral_r4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      attr_set = true;
    }
    const dim3 grid((s.w + BL_TW - 1) / BL_TW, (s.h + BL_TH - 1) / BL_TH, s.count);
    k_bilateral_r4<<<grid, BL_THREADS, smem, st>>>(s, ba);
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
