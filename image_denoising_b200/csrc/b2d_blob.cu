// b2d_blob.cu - the step right after the corrupt -> denoise path, on the device:
//   prep_im_for_blob (lib/utils/blob.py:33-47): im.astype(float32) - PIXEL_MEANS, then
//       cv2.resize(im, None, None, fx=s, fy=s, interpolation=cv2.INTER_LINEAR)
//   im_list_to_blob  (lib/utils/blob.py:17-30): zero-padded (N, Hmax, Wmax, 3) float32
//   horizontal flip  (lib/roi_data_layer/minibatch.py:1676-1677): im[:, ::-1, :] before the resize
// so that the detector reads the network input straight from HBM.
//
// cv2.resize, INTER_LINEAR, float32, restated (bit-exact against cv2 4.13 with cv2.setUseOptimized(False) /
// cv2.ipp.setUseIPP(False) when B2D_BLOB_F32_COORDS is set; the IPP-backed wheels compute the source
// coordinate in higher precision and differ from that by up to 4e-5 relative on downscales: the default
// mode follows them to 2 ulp) (modules/imgproc/src/resize.cpp, 4.x):
//   dsize  = (cvRound(w * fx), cvRound(h * fy));  scale = 1 / fx (double)
//   per destination column dx: f = (float)((dx + 0.5) * scale - 0.5); sx = floor(f); f -= sx;
//       sx < 0 -> (sx, f) = (0, 0);  sx >= w - 1 -> (sx, f) = (w - 1, 0)   [second tap unused]
//       a0 = 1.f - f, a1 = f;  rows alike with b0, b1 and source rows clamped to [0, h - 1]
//       [rows: indices clipped to [0, h - 1] but the fraction is KEPT, so border rows blend a row with itself]
//   horizontal pass first: t(y) = S[y][sx] * a0 + S[y][sx + 1] * a1, then out = t(y0) * b0 + t(y1) * b1,
//   every product and sum rounded separately (no FMA in the library's baseline build)
// The mean subtraction `im -= pixel_means` (float32 array, float64 means) rounds ONCE from the
// exact double difference: a 3 x 256 table built on the host in double reproduces it bit for bit.
//
// Bound: HBM.  Algorithmic bytes per source pixel: 3 read + 12 * s^2 written.
#include "b2d_kernels.cuh"
#include "b2d_plan.h"

#include <algorithm>
#include <cmath>

namespace b2d {

struct BlobArgs {
  const uint8_t* in;      // (n, h, w, 3) uint8
  float* out;             // (n, blob_h, blob_w, 3) float32
  const uint8_t* flip;    // n flags (device) or nullptr
  int n, h, w, oh, ow, blob_h, blob_w;
  double scale_x, scale_y;    // 1 / fx, 1 / fy
  double mean[3];         // float32(v - mean_c) is tabulated per CTA
};
constexpr int BLOB_ROWS = 8;  // destination rows per CTA

// one thread = one destination pixel; the CTA's 256 x 3 floats go through shared memory so that every
// store instruction of a warp covers 128 contiguous bytes
// CLAMP_FRAC: the horizontal pass resets the fraction at the borders (second tap unused); the vertical pass
// only clips the row indices and keeps the fraction (both taps read the same row)
template <bool F32_COORDS, bool CLAMP_FRAC>
__device__ __forceinline__ void src_coord(int d, double scale, int n, int& s0, float& frac) {
  if (F32_COORDS) {                                   // OpenCV's generic C++ path: fraction from the float32 coordinate
    float f = (float)(((double)d + 0.5) * scale - 0.5);
    s0 = (int)floorf(f);
    frac = f - (float)s0;
  } else {                                            // double-precision coordinate (what IPP-backed builds return)
    const double f = ((double)d + 0.5) * scale - 0.5;
    const double fl = floor(f);
    s0 = (int)fl;
    frac = (float)(f - fl);
  }
  if (CLAMP_FRAC) {
    if (s0 < 0) { s0 = 0; frac = 0.0f; }
    if (s0 >= n - 1) { s0 = n - 1; frac = 0.0f; }
  }
}

template <bool F32_COORDS>
__global__ void __launch_bounds__(256) k_blob_prep(const __grid_constant__ BlobArgs a) {
  __shared__ float lut[3][256];
  __shared__ float stage[2][768];
  __shared__ int s_sy[BLOB_ROWS];
  __shared__ float s_fy[BLOB_ROWS];
#pragma unroll
  for (int c = 0; c < 3; ++c) lut[c][threadIdx.x] = (float)((double)threadIdx.x - a.mean[c]);   // one rounding, like numpy
  if (threadIdx.x < BLOB_ROWS)
    src_coord<F32_COORDS, false>(min((int)blockIdx.y * BLOB_ROWS + (int)threadIdx.x, a.oh - 1), a.scale_y, a.h,
                                 s_sy[threadIdx.x], s_fy[threadIdx.x]);
  __syncthreads();
  const int img = blockIdx.z;
  const int dx0 = blockIdx.x * 256, dx = dx0 + threadIdx.x;
  int sx = 0;
  float fx = 0.0f;
  src_coord<F32_COORDS, true>(min(dx, a.ow - 1), a.scale_x, a.w, sx, fx);
  const int dy_end = min((int)(blockIdx.y + 1) * BLOB_ROWS, a.blob_h);
  for (int dy = blockIdx.y * BLOB_ROWS; dy < dy_end; ++dy) {
    float v[3] = {0.0f, 0.0f, 0.0f};                    // zero padding of im_list_to_blob
    if (dy < a.oh && dx < a.ow) {
      int sy = s_sy[dy - blockIdx.y * BLOB_ROWS];
      const float fy = s_fy[dy - blockIdx.y * BLOB_ROWS];
      const int y1 = min(max(sy + 1, 0), a.h - 1), sx1 = min(sx + 1, a.w - 1);
      sy = min(max(sy, 0), a.h - 1);
      const float b0 = 1.0f - fy, b1 = fy, a0 = 1.0f - fx, a1 = fx;
      const bool mirror = a.flip != nullptr && a.flip[img] != 0;
      const int x0 = mirror ? a.w - 1 - sx : sx, x1 = mirror ? a.w - 1 - sx1 : sx1;
      const uint8_t* base = a.in + (size_t)img * a.h * a.w * 3;
      const uint8_t* r0 = base + (size_t)sy * a.w * 3;
      const uint8_t* r1 = base + (size_t)y1 * a.w * 3;
  #pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float p00 = lut[c][__ldg(r0 + 3 * x0 + c)], p01 = lut[c][__ldg(r0 + 3 * x1 + c)];
        const float p10 = lut[c][__ldg(r1 + 3 * x0 + c)], p11 = lut[c][__ldg(r1 + 3 * x1 + c)];
        const float t0 = __fadd_rn(__fmul_rn(p00, a0), __fmul_rn(p01, a1));
        const float t1 = __fadd_rn(__fmul_rn(p10, a0), __fmul_rn(p11, a1));
        v[c] = __fadd_rn(__fmul_rn(t0, b0), __fmul_rn(t1, b1));
      }
    }
    float* stg = stage[dy & 1];                          // double buffered: one barrier per row
    stg[3 * threadIdx.x] = v[0];
    stg[3 * threadIdx.x + 1] = v[1];
    stg[3 * threadIdx.x + 2] = v[2];
    __syncthreads();
    float* o = a.out + (((size_t)img * a.blob_h + dy) * a.blob_w + dx0) * 3;
    const int nfl = 3 * min(256, a.blob_w - dx0);
    for (int i = threadIdx.x; i < nfl; i += 256) o[i] = stg[i];
  }
}

// im_scale == 1 (600 x 1000 images with the default 600 / 1000 sizes), no padding, no flip: the resize is the
// identity (both fractions 0: p * 1 + q * 0 = p exactly), so the stage is a table lookup per byte.
// 16 input bytes -> 16 floats per thread; the channel of byte i of chunk k is (16 k + i) % 3 = (k + i) % 3.
template <int PH>
__device__ __forceinline__ void blob_lookup16(const float (&lut)[3][256], const uint32_t (&w)[4], float4* o) {
#pragma unroll
  for (int j = 0; j < 4; ++j)
    o[j] = make_float4(lut[(PH + 4 * j) % 3][w[j] & 0xffu], lut[(PH + 4 * j + 1) % 3][(w[j] >> 8) & 0xffu],
                       lut[(PH + 4 * j + 2) % 3][(w[j] >> 16) & 0xffu], lut[(PH + 4 * j + 3) % 3][w[j] >> 24]);
}

__global__ void __launch_bounds__(256) k_blob_scale1(const uint8_t* __restrict__ in, float* __restrict__ out,
                                                     size_t image_bytes, int n, double m0, double m1, double m2) {
  __shared__ float lut[3][256];
  lut[0][threadIdx.x] = (float)((double)threadIdx.x - m0);
  lut[1][threadIdx.x] = (float)((double)threadIdx.x - m1);
  lut[2][threadIdx.x] = (float)((double)threadIdx.x - m2);
  __syncthreads();
  const size_t chunks = image_bytes / 16;                       // caller guarantees image_bytes % 48 == 0
  const size_t total = chunks * (size_t)n;
  for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < total; k += (size_t)gridDim.x * 256) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(in) + k);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    const int ph = (int)(k % 3);                                // chunks % 3 == 0: the phase restarts with every image
    float4* o = reinterpret_cast<float4*>(out) + 4 * k;
    if (ph == 0) blob_lookup16<0>(lut, w, o);
    else if (ph == 1) blob_lookup16<1>(lut, w, o);
    else blob_lookup16<2>(lut, w, o);
  }
}

}  // namespace b2d

using namespace b2d;

// Output geometry of prep_im_for_blob for an h x w image (blob.py:37-45 + cv2.resize's dsize)
extern "C" int b2d_blob_shape(int h, int w, int target_size, int max_size, double* im_scale, int* out_h, int* out_w) {
  if (h <= 0 || w <= 0 || target_size <= 0 || max_size <= 0 || !im_scale || !out_h || !out_w)
    return fail(B2D_ERR_INVALID, "b2d_blob_shape: bad argument");
  const double mn = (double)std::min(h, w), mx = (double)std::max(h, w);
  double s = (double)target_size / mn;
  if (std::nearbyint(s * mx) > (double)max_size) s = (double)max_size / mx;      // np.round: half to even
  *im_scale = s;
  *out_h = (int)std::lrint((double)h * s);                                        // cvRound
  *out_w = (int)std::lrint((double)w * s);
  return B2D_OK;
}

extern "C" int b2d_blob_prep(const uint8_t* d_in, int n, int h, int w, float* d_out, int blob_h, int blob_w,
                             double im_scale, const double* pixel_means, const uint8_t* d_flip, int flags,
                             void* cuda_stream) {
  if (!d_in || !d_out || !pixel_means || n < 0 || h <= 0 || w <= 0 || !(im_scale > 0.0))
    return fail(B2D_ERR_INVALID, "b2d_blob_prep: bad argument");
  if (n == 0) return B2D_OK;
  BlobArgs a;
  a.in = d_in; a.out = d_out; a.flip = d_flip;
  a.n = n; a.h = h; a.w = w;
  a.oh = (int)std::lrint((double)h * im_scale);
  a.ow = (int)std::lrint((double)w * im_scale);
  a.blob_h = blob_h; a.blob_w = blob_w;
  if (a.oh <= 0 || a.ow <= 0 || blob_h < a.oh || blob_w < a.ow)
    return fail(B2D_ERR_INVALID, "b2d_blob_prep: blob %d x %d smaller than the resized image %d x %d", blob_h, blob_w, a.oh, a.ow);
  a.scale_x = a.scale_y = 1.0 / im_scale;
  for (int c = 0; c < 3; ++c) a.mean[c] = pixel_means[c];
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  if (n > 65535) return fail(B2D_ERR_INVALID, "b2d_blob_prep: n must be <= 65535");
  const size_t image_bytes = (size_t)h * w * 3;
  if (im_scale == 1.0 && !d_flip && blob_h == h && blob_w == w && image_bytes % 48 == 0 &&
      ((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15u) == 0) {
    const size_t total = image_bytes / 16 * (size_t)n;
    const unsigned blocks = (unsigned)std::min<size_t>((total + 255) / 256, 148u * 16u);
    k_blob_scale1<<<blocks, 256, 0, st>>>(d_in, d_out, image_bytes, n, pixel_means[0], pixel_means[1], pixel_means[2]);
    count_launch();
    const cudaError_t e1 = cudaGetLastError();
    return e1 == cudaSuccess ? B2D_OK : fail_cuda(e1, "b2d_blob_prep launch");
  }
  const dim3 grid((unsigned)((blob_w + 255) / 256), (unsigned)((blob_h + BLOB_ROWS - 1) / BLOB_ROWS), (unsigned)n);
  if (flags & B2D_BLOB_F32_COORDS) k_blob_prep<true><<<grid, 256, 0, st>>>(a);
  else k_blob_prep<false><<<grid, 256, 0, st>>>(a);
  count_launch();
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? B2D_OK : fail_cuda(e, "b2d_blob_prep launch");
}
