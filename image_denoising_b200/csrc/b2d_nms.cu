// b2d_nms.cu - SURVEY 8(f) f4: the reference's only native CUDA code, `_nms` (lib/nms/gpu_nms.hpp:1,
// lib/nms/nms_kernel.cu:34-144, reached through lib/nms/gpu_nms.pyx:16-31 and lib/model/nms_wrapper.py:15-23
// from lib/model/test.py:1860 and lib/layer_utils/proposal_layer.py:46).  Greedy non-maximum
// suppression of score-sorted boxes: box j is dropped when a kept box i < j has IoU(i, j) > thresh, IoU
// with the "+1" pixel convention in float32.
//
// Redesigned for one B200 instead of copied: the reference computes all (N/64)^2 tiles of the
// 64-bit suppression mask, copies the mask (18 MB at N = 12000) to the host and sweeps it there,
// with a cudaMalloc / cudaFree pair per call.  Here
//   k_nms_mask  : only the upper-triangular tiles (j > i is all the sweep reads), one CTA per tile,
//                 column boxes staged in shared memory, float32 operations in the reference's order
//                 with explicit round-to-nearest intrinsics (no FMA contraction): bit-identical IoU;
//   k_nms_sweep : the greedy sweep on the device, one CTA: per block of 64 boxes the 64 diagonal
//                 mask words are fetched in parallel, one thread resolves the block's internal
//                 chain out of shared memory, then every thread ORs the kept rows into its own
//                 word of the removal bitmap; writes the kept indices and their count.
// Nothing but the kept indices ever crosses PCIe, and memory comes from the stream-ordered pool.
#include "b2d_kernels.cuh"
#include "b2d_plan.h"

#include <mutex>

namespace b2d {

__device__ __forceinline__ float nms_iou(const float4 a, const float4 b) {
  const float left = fmaxf(a.x, b.x), right = fminf(a.z, b.z);
  const float top = fmaxf(a.y, b.y), bottom = fminf(a.w, b.w);
  const float width = fmaxf(__fadd_rn(__fsub_rn(right, left), 1.0f), 0.0f);
  const float height = fmaxf(__fadd_rn(__fsub_rn(bottom, top), 1.0f), 0.0f);
  const float inter = __fmul_rn(width, height);
  const float sa = __fmul_rn(__fadd_rn(__fsub_rn(a.z, a.x), 1.0f), __fadd_rn(__fsub_rn(a.w, a.y), 1.0f));
  const float sb = __fmul_rn(__fadd_rn(__fsub_rn(b.z, b.x), 1.0f), __fadd_rn(__fsub_rn(b.w, b.y), 1.0f));
  return __fdiv_rn(inter, __fsub_rn(__fadd_rn(sa, sb), inter));
}

// tile t of the upper triangle (row block r <= column block c), enumerated row by row
__global__ void __launch_bounds__(64) k_nms_mask(const float* boxes, int n, int dim, float thresh, int cb,
                                                 unsigned long long* mask) {
  __shared__ float4 s_col[64];
  // invert t = r * cb - r (r - 1) / 2 + (c - r)
  int t = blockIdx.x, r = 0;
  while (t >= cb - r) { t -= cb - r; ++r; }
  const int c = r + t;
  const int col0 = c * 64, row = r * 64 + threadIdx.x;
  const int ncol = min(64, n - col0);
  if ((int)threadIdx.x < ncol) {
    const float* p = boxes + (size_t)(col0 + threadIdx.x) * dim;
    s_col[threadIdx.x] = make_float4(p[0], p[1], p[2], p[3]);
  }
  __syncthreads();
  if (row >= n) return;
  const float* p = boxes + (size_t)row * dim;
  const float4 me = make_float4(p[0], p[1], p[2], p[3]);
  unsigned long long bits = 0ull;
  for (int j = (r == c ? (int)threadIdx.x + 1 : 0); j < ncol; ++j)
    if (nms_iou(me, s_col[j]) > thresh) bits |= 1ull << j;
  mask[(size_t)row * cb + c] = bits;
}

__global__ void __launch_bounds__(256) k_nms_sweep(const unsigned long long* mask, int n, int cb, int* keep, int* num_out) {
  extern __shared__ unsigned long long s_remv[];      // cb words
  __shared__ unsigned long long s_diag[64];
  __shared__ unsigned long long s_kept;
  __shared__ int s_count;
  for (int w = threadIdx.x; w < cb; w += blockDim.x) s_remv[w] = 0ull;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  for (int b = 0; b < cb; ++b) {
    const int nb = min(64, n - b * 64);
    if ((int)threadIdx.x < nb) s_diag[threadIdx.x] = mask[(size_t)(b * 64 + threadIdx.x) * cb + b];
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned long long removed = s_remv[b], kept = 0ull;
      int count = s_count;
      for (int i = 0; i < nb; ++i)
        if (!((removed >> i) & 1ull)) {
          kept |= 1ull << i;
          removed |= s_diag[i];
          keep[count++] = b * 64 + i;
        }
      s_kept = kept;
      s_count = count;
    }
    __syncthreads();
    const unsigned long long kept = s_kept;
    for (int w = b + 1 + threadIdx.x; w < cb; w += blockDim.x) {
      unsigned long long acc = s_remv[w];
      const unsigned long long* col = mask + (size_t)(b * 64) * cb + w;
#pragma unroll
      for (int half = 0; half < 2; ++half) {                 // 32 independent loads in flight, then the ORs
        unsigned long long v[32];
#pragma unroll
        for (int u = 0; u < 32; ++u) {
          const int i = half * 32 + u;
          v[u] = (i < nb && ((kept >> i) & 1ull)) ? col[(size_t)i * cb] : 0ull;
        }
#pragma unroll
        for (int u = 0; u < 32; ++u) acc |= v[u];
      }
      s_remv[w] = acc;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *num_out = s_count;
}

static int nms_launch(const float* d_boxes, int n, int dim, float thresh, int* d_keep, int* d_num, unsigned long long* d_mask,
                      cudaStream_t st) {
  const int cb = (n + 63) / 64;
  const long long tiles = (long long)cb * (cb + 1) / 2;
  k_nms_mask<<<(unsigned)tiles, 64, 0, st>>>(d_boxes, n, dim, thresh, cb, d_mask);
  k_nms_sweep<<<1, 256, (size_t)cb * sizeof(unsigned long long), st>>>(d_mask, n, cb, d_keep, d_num);
  count_launch(); count_launch();
  const cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? B2D_OK : fail_cuda(e, "nms kernels");
}

}  // namespace b2d

using namespace b2d;

#define CKN(x)                                          \
  do {                                                  \
    cudaError_t e__ = (x);                              \
    if (e__ != cudaSuccess) return fail_cuda(e__, #x);  \
  } while (0)

static const int kNmsMaxBoxes = 1 << 20;

extern "C" {

int b2d_nms_workspace_bytes(int boxes_num, size_t* bytes) {
  if (!bytes || boxes_num < 0 || boxes_num > kNmsMaxBoxes) return fail(B2D_ERR_INVALID, "b2d_nms_workspace_bytes: bad argument");
  const size_t cb = ((size_t)boxes_num + 63) / 64;
  *bytes = (size_t)boxes_num * cb * sizeof(unsigned long long);
  return B2D_OK;
}

int b2d_nms_device(const float* d_boxes, int boxes_num, int boxes_dim, float nms_overlap_thresh, int* d_keep, int* d_num,
                   void* d_ws, void* cuda_stream) {
  if (!d_num || boxes_num < 0 || boxes_num > kNmsMaxBoxes || boxes_dim < 4) return fail(B2D_ERR_INVALID, "b2d_nms_device: bad argument");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  if (boxes_num == 0) { CKN(cudaMemsetAsync(d_num, 0, sizeof(int), st)); return B2D_OK; }
  if (!d_boxes || !d_keep || !d_ws) return fail(B2D_ERR_INVALID, "b2d_nms_device: null buffer");
  return nms_launch(d_boxes, boxes_num, boxes_dim, nms_overlap_thresh, d_keep, d_num, reinterpret_cast<unsigned long long*>(d_ws), st);
}

int b2d_nms(int* keep_out, int* num_out, const float* boxes_host, int boxes_num, int boxes_dim, float nms_overlap_thresh,
            int device_id) {
  if (!num_out || boxes_num < 0 || boxes_num > kNmsMaxBoxes || boxes_dim < 4) return fail(B2D_ERR_INVALID, "b2d_nms: bad argument");
  *num_out = 0;
  if (boxes_num == 0) return B2D_OK;
  if (!keep_out || !boxes_host) return fail(B2D_ERR_INVALID, "b2d_nms: null buffer");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(B2D_ERR_NO_DEVICE, "no CUDA device: NMS has no CPU fallback");
  }
  if (device_id < 0 || device_id >= ndev) return fail(B2D_ERR_INVALID, "device %d out of range (%d devices)", device_id, ndev);
  DeviceGuard guard(device_id);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  cudaStream_t st = cudaStreamPerThread;
  const size_t cb = ((size_t)boxes_num + 63) / 64;
  const size_t box_bytes = (size_t)boxes_num * boxes_dim * sizeof(float), mask_bytes = (size_t)boxes_num * cb * sizeof(unsigned long long);
  uint8_t* d = nullptr;                                         // one stream-ordered allocation: boxes | keep | num | mask
  const size_t off_keep = (box_bytes + 255) & ~(size_t)255, off_num = off_keep + (((size_t)boxes_num * 4 + 255) & ~(size_t)255),
               off_mask = off_num + 256;
  CKN(cudaMallocAsync(&d, off_mask + mask_bytes, st));
  cudaError_t e = cudaMemcpyAsync(d, boxes_host, box_bytes, cudaMemcpyHostToDevice, st);
  int rc = B2D_OK;
  if (e == cudaSuccess)
    rc = nms_launch(reinterpret_cast<float*>(d), boxes_num, boxes_dim, nms_overlap_thresh, reinterpret_cast<int*>(d + off_keep),
                    reinterpret_cast<int*>(d + off_num), reinterpret_cast<unsigned long long*>(d + off_mask), st);
  if (e == cudaSuccess && rc == B2D_OK) e = cudaMemcpyAsync(num_out, d + off_num, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess && rc == B2D_OK) e = cudaStreamSynchronize(st);
  if (e == cudaSuccess && rc == B2D_OK && *num_out > 0)
    e = cudaMemcpyAsync(keep_out, d + off_keep, sizeof(int) * (size_t)*num_out, cudaMemcpyDeviceToHost, st);
  cudaFreeAsync(d, st);
  const cudaError_t e2 = cudaStreamSynchronize(st);
  if (rc != B2D_OK) return rc;
  if (e != cudaSuccess) return fail_cuda(e, "b2d_nms");
  if (e2 != cudaSuccess) return fail_cuda(e2, "b2d_nms");
  return B2D_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// b2d_nms_device_static: b2d_nms_device for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The suppression mask - the only scratch the two kernels need - lives in a file-scope device array
// sized for B2D_NMS_STATIC_MAX_BOXES boxes (4096 x 64 words = 2 MiB), held by every GPU that loads the library
// until the process exits.
//
// `g_nms_static_mask` is a HOST-side stand-in in host code, not a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, asked once per device with that device current (nms_static_mask below).
//
// Contents: k_nms_mask writes every word k_nms_sweep reads (the upper-triangular tiles of rows < n, in the row
// pitch cb of THIS call), so nothing has to start at zero and the previous call's bits - left in place, also
// under another pitch - are never read.  The array is shared by every call on a device: calls that overlap on
// the GPU (different streams, no ordering between them) race on it.
// ---------------------------------------------------------------------------------------------------------
namespace b2d {

constexpr int kNmsStaticMaxBoxes = B2D_NMS_STATIC_MAX_BOXES;
constexpr size_t kNmsStaticWords = (size_t)kNmsStaticMaxBoxes * ((kNmsStaticMaxBoxes + 63) / 64);

static __device__ __align__(256) unsigned long long g_nms_static_mask[kNmsStaticWords];

// device address of the array on the CURRENT device; set up once per device, under a lock
static int nms_static_mask(unsigned long long** out) {
  static unsigned long long* addr[64] = {};
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail_cuda(e, "cudaGetDevice");
  if (dev < 0 || dev >= 64) return fail(B2D_ERR_UNSUPPORTED, "b2d_nms_device_static: device index %d beyond 63", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_nms_static_mask);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_nms_static_mask)");
    if (!p) return fail(B2D_ERR_CUDA, "cudaGetSymbolAddress(g_nms_static_mask) returned a null address");
    addr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = addr[dev];
  return B2D_OK;
}

}  // namespace b2d

extern "C" int b2d_nms_device_static(const float* d_boxes, int boxes_num, int boxes_dim, float nms_overlap_thresh,
                                     int* d_keep, int* d_num, void* cuda_stream) {
  if (!d_num || boxes_num < 0 || boxes_dim < 4) return fail(B2D_ERR_INVALID, "b2d_nms_device_static: bad argument");
  if (boxes_num > kNmsStaticMaxBoxes)
    return fail(B2D_ERR_INVALID, "b2d_nms_device_static: %d boxes exceed the static scratch (B2D_NMS_STATIC_MAX_BOXES = %d); "
                                 "use b2d_nms_device with a caller-owned workspace", boxes_num, kNmsStaticMaxBoxes);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  if (boxes_num == 0) { CKN(cudaMemsetAsync(d_num, 0, sizeof(int), st)); return B2D_OK; }
  if (!d_boxes || !d_keep) return fail(B2D_ERR_INVALID, "b2d_nms_device_static: null buffer");
  unsigned long long* mask = nullptr;
  const int rc = nms_static_mask(&mask);
  if (rc != B2D_OK) return rc;
  return nms_launch(d_boxes, boxes_num, boxes_dim, nms_overlap_thresh, d_keep, d_num, mask, st);
}
