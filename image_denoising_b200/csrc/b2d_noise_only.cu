// b2d_noise_only.cu - the noise-only branches (`<noise>_var<L>` with no
// denoiser, e.g. lib/model/test.py:290-305) and the first half of two-pass
// paths (injected fields, bilateral / wavelet consumers that read uint8).
// One thread per 8-element group, grid-stride; any image size / alignment.
#include "b2d_kernels.cuh"

#include <algorithm>

namespace b2d {

constexpr int NO_THREADS = 256;

__device__ __forceinline__ void load_group_any(const uint8_t* img, uint64_t elems, uint32_t g, uint32_t (&v)[8]) {
  const uint64_t e = (uint64_t)g * 8u;
  const uint8_t* p = img + e;
  if ((reinterpret_cast<uintptr_t>(p) & 7u) == 0 && e + 8u <= elems) {
    const uint2 w = __ldg(reinterpret_cast<const uint2*>(p));
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] = byte_of(w.x, i);
      v[4 + i] = byte_of(w.y, i);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (e + i < elems) ? (uint32_t)__ldg(p + i) : 0u;
  }
}

__device__ __forceinline__ void store_group_any(uint8_t* img, uint64_t elems, uint32_t g, const uint32_t (&v)[8]) {
  const uint64_t e = (uint64_t)g * 8u;
  uint8_t* p = img + e;
  if ((reinterpret_cast<uintptr_t>(p) & 7u) == 0 && e + 8u <= elems) {
    *reinterpret_cast<uint2*>(p) =
        make_uint2(pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]));
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (e + i < elems) p[i] = (uint8_t)v[i];
  }
}

// generator mode keeps the group packed.  An image of odd size starts at any byte address: its groups
// are read as the three aligned words around them (funnel shift by the phase) and written as one
// aligned word plus 4 single bytes, instead of 8 + 8 byte accesses.
__device__ __forceinline__ uint2 load_group_words(const uint8_t* img, uint64_t elems, uint32_t g) {
  const uint64_t e = (uint64_t)g * 8u;
  const uint8_t* p = img + e;
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  if (e + 8u <= elems) {
    if ((a & 7u) == 0) return __ldg(reinterpret_cast<const uint2*>(p));
    const unsigned ph = (unsigned)(a & 3u);
    if (ph == 0) return make_uint2(__ldg(reinterpret_cast<const uint32_t*>(p)), __ldg(reinterpret_cast<const uint32_t*>(p + 4)));
    if (e >= 4u && e + 12u <= elems) {            // the three words lie inside the image
      const uint32_t* q = reinterpret_cast<const uint32_t*>(p - ph);
      const uint32_t w0 = __ldg(q), w1 = __ldg(q + 1), w2 = __ldg(q + 2);
      return make_uint2(__funnelshift_r(w0, w1, 8u * ph), __funnelshift_r(w1, w2, 8u * ph));
    }
  }
  uint2 w = make_uint2(0u, 0u);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (e + i < elems) w.x |= (uint32_t)__ldg(p + i) << (8 * i);
    if (e + 4 + i < elems) w.y |= (uint32_t)__ldg(p + 4 + i) << (8 * i);
  }
  return w;
}

__device__ __forceinline__ void store_group_words(uint8_t* img, uint64_t elems, uint32_t g, uint2 w) {
  const uint64_t e = (uint64_t)g * 8u;
  uint8_t* p = img + e;
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  if (e + 8u <= elems) {
    if ((a & 7u) == 0) { *reinterpret_cast<uint2*>(p) = w; return; }
    const unsigned ph = (unsigned)(a & 3u);
    if (ph == 0) {
      *reinterpret_cast<uint32_t*>(p) = w.x;
      *reinterpret_cast<uint32_t*>(p + 4) = w.y;
      return;
    }
    const unsigned hb = 4u - ph;                  // head bytes in front of the aligned word
#pragma unroll
    for (unsigned i = 0; i < 3; ++i)
      if (i < hb) p[i] = (uint8_t)(w.x >> (8u * i));
    *reinterpret_cast<uint32_t*>(p + hb) = __funnelshift_r(w.x, w.y, 8u * hb);
#pragma unroll
    for (unsigned i = 0; i < 3; ++i)
      if (i < ph) p[4u + hb + i] = (uint8_t)(w.y >> (8u * (hb + i)));
    return;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (e + i < elems) p[i] = (uint8_t)(w.x >> (8 * i));
    if (e + 4 + i < elems) p[4 + i] = (uint8_t)(w.y >> (8 * i));
  }
}

template <int KIND, int MODE>
__global__ void __launch_bounds__(NO_THREADS) k_noise_only(ImageSet s, NoiseArgs na) {
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out ? s.out + (uint64_t)img_local * s.elems : nullptr;
  const uint32_t groups = (uint32_t)((s.elems + 7u) >> 3);
  for (uint32_t g = blockIdx.x * NO_THREADS + threadIdx.x; g < groups; g += gridDim.x * NO_THREADS) {
    GroupCtx gc{(uint32_t)img_local, g, img};
    if (MODE == NM_GEN) {
      // generator mode runs the packed operators, the same code the fused kernels run
      const uint2 w = PackedNoise<KIND>::apply(na, gc, load_group_words(img, s.elems, g));
      store_group_words(out, s.elems, g, w);
      continue;
    }
    uint32_t v[8];
    load_group_any(img, s.elems, g, v);
    NoiseOp<KIND, MODE>::apply8(na, gc, v);
    if (MODE != NM_DUMP) store_group_any(out, s.elems, g, v);
  }
}

// generator mode on 16-byte aligned images: two groups per thread step, words never unpacked
template <int KIND>
__global__ void __launch_bounds__(NO_THREADS) k_noise_packed(ImageSet s, NoiseArgs na) {
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const uint4* in = reinterpret_cast<const uint4*>(img);
  uint4* out = reinterpret_cast<uint4*>(s.out + (uint64_t)img_local * s.elems);
  const uint32_t chunks = (uint32_t)(s.elems >> 4);
  for (uint32_t k = blockIdx.x * NO_THREADS + threadIdx.x; k < chunks; k += gridDim.x * NO_THREADS) {
    const uint4 q = __ldg(in + k);
    const uint2 a = PackedNoise<KIND>::apply(na, GroupCtx{(uint32_t)img_local, 2u * k, img}, make_uint2(q.x, q.y));
    const uint2 b = PackedNoise<KIND>::apply(na, GroupCtx{(uint32_t)img_local, 2u * k + 1u, img}, make_uint2(q.z, q.w));
    out[k] = make_uint4(a.x, a.y, b.x, b.y);
  }
}

template <int KIND, int MODE>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  if (MODE == NM_GEN && s.elems % 16 == 0 && s.out &&
      ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 15u) == 0) {
    const uint32_t chunks = (uint32_t)(s.elems >> 4);
    const uint32_t bx = std::min<uint32_t>((chunks + NO_THREADS - 1) / NO_THREADS, 148u * 8u);
    k_noise_packed<KIND><<<dim3(bx, 1, s.count), NO_THREADS, 0, st>>>(s, na);
    count_launch();
    return cudaGetLastError();
  }
  const uint32_t groups = (uint32_t)((s.elems + 7u) >> 3);
  uint32_t bx = (groups + NO_THREADS - 1) / NO_THREADS;
  if (bx > 148u * 8u) bx = 148u * 8u;             // grid-stride beyond 8 CTAs per SM
  const dim3 grid(bx, 1, s.count);
  k_noise_only<KIND, MODE><<<grid, NO_THREADS, 0, st>>>(s, na);
  count_launch();
  return cudaGetLastError();
}

template <int KIND>
static cudaError_t launch_mode(int mode, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (mode) {
    case NM_GEN: return launch_one<KIND, NM_GEN>(s, na, st);
    case NM_INJECT: return launch_one<KIND, NM_INJECT>(s, na, st);
    case NM_DUMP: return launch_one<KIND, NM_DUMP>(s, na, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_noise_only(int kind, int mode, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  switch (kind) {
    case NK_NONE: return launch_mode<NK_NONE>(mode, s, na, st);
    case NK_GAUSSIAN: return launch_mode<NK_GAUSSIAN>(mode, s, na, st);
    case NK_SPECKLE: return launch_mode<NK_SPECKLE>(mode, s, na, st);
    case NK_SAP: return launch_mode<NK_SAP>(mode, s, na, st);
    case NK_POISSON: return launch_mode<NK_POISSON>(mode, s, na, st);
    case NK_UNIFORM: return launch_mode<NK_UNIFORM>(mode, s, na, st);
    case NK_RAYLEIGH: return launch_mode<NK_RAYLEIGH>(mode, s, na, st);
    case NK_GAMMA: return launch_mode<NK_GAMMA>(mode, s, na, st);
    case NK_TABLE_ADD: return launch_mode<NK_TABLE_ADD>(mode, s, na, st);
    case NK_SHADER: return launch_mode<NK_SHADER>(mode, s, na, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2d
