// b2d_prepass.cu - small per-image pre-passes that produce parameters / u8
// noise fields consumed by the fused kernels.
//   poisson : vals = 2^ceil(log2(#unique values))         (skimage random_noise)
//   periodic: (sin(linspace(-L, L, n)) * 255) ->u8         (minibatch.py:1042-1045)
//   brownian: B = [0, cumsum(sqrt(dt) z)], (B * 255) ->u8  (minibatch.py:822-827)
#include "b2d_kernels.cuh"
#include "b2d_prepass.cuh"

namespace b2d {

// ---------------------------------------------------------------- poisson --
// 256-bit presence bitmap of the byte values of one image.  A CTA starts from what the other CTAs
// of the image have already published and stops as soon as the bitmap is full (every value seen):
// natural images saturate after a few KB, so the pass usually costs far less than one read of the
// batch; the worst case (few distinct values) streams the image once with 16-byte loads.
__global__ void __launch_bounds__(256) k_unique_bitmap(ImageSet s, uint32_t* bitmaps /* [n][8] */) {
  __shared__ __align__(16) uint32_t bm[8];
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint32_t* gbm = bitmaps + (size_t)img_local * 8;
  uint32_t seen = 0u;
  if (threadIdx.x < 8) bm[threadIdx.x] = seen = *reinterpret_cast<volatile uint32_t*>(gbm + threadIdx.x);
  __syncthreads();
  auto full = [&]() {
    const volatile uint32_t* v = bm;
    return (v[0] & v[1] & v[2] & v[3] & v[4] & v[5] & v[6] & v[7]) == 0xffffffffu;
  };
  auto tally = [&](uint32_t b) {
    const uint32_t bit = 1u << (b & 31u);
    if (!(bm[b >> 5] & bit)) atomicOr(&bm[b >> 5], bit);
  };
  const uint64_t head = min((uint64_t)((16u - ((uint32_t)reinterpret_cast<uintptr_t>(img) & 15u)) & 15u), s.elems);
  const uint64_t nvec = (s.elems - head) / 16;
  const uint4* vec = reinterpret_cast<const uint4*>(img + head);
  for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (uint64_t)gridDim.x * 256) {
    if (full()) break;
    const uint4 q = __ldg(vec + i);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tally(w[k] & 0xffu);
      tally((w[k] >> 8) & 0xffu);
      tally((w[k] >> 16) & 0xffu);
      tally(w[k] >> 24);
    }
  }
  if (blockIdx.x == 0) {                              // unaligned head and the < 16-byte tail
    if (threadIdx.x < head) tally(__ldg(img + threadIdx.x));
    const uint64_t t = head + nvec * 16 + threadIdx.x;
    if (t < s.elems) tally(__ldg(img + t));
  }
  __syncthreads();
  if (threadIdx.x < 8 && (bm[threadIdx.x] & ~seen)) atomicOr(gbm + threadIdx.x, bm[threadIdx.x]);
}

// the two small kernels touch only the images of this launch (b2d_run_host pipelines chunks of one batch
// through one workspace: rows of other chunks may be in use by the previous chunk's kernels)
__global__ void k_bitmap_clear(ImageSet s, uint32_t* bitmaps) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.count * 8) return;
  bitmaps[(size_t)image_of(s, i >> 3) * 8 + (i & 7)] = 0u;
}

__global__ void k_poisson_vals(ImageSet s, const uint32_t* bitmaps, float* vals) {
  const int z = blockIdx.x * blockDim.x + threadIdx.x;
  if (z >= s.count) return;
  const int i = image_of(s, z);
  int cnt = 0;
  for (int k = 0; k < 8; ++k) cnt += __popc(bitmaps[(size_t)i * 8 + k]);
  int p = 1;
  while (p < cnt) p <<= 1;          // 2 ** ceil(log2(cnt))
  vals[i] = (float)p;
}

cudaError_t launch_poisson_vals(const ImageSet& s, uint32_t* bitmaps, float* vals, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  k_bitmap_clear<<<(s.count * 8 + 255) / 256, 256, 0, st>>>(s, bitmaps);
  count_launch();
  uint32_t bx = (uint32_t)((s.elems + 256 * 16 * 8 - 1) / (256 * 16 * 8));   // 8 x 16-byte loads per thread
  if (bx < 1) bx = 1;
  k_unique_bitmap<<<dim3(bx, 1, s.count), 256, 0, st>>>(s, bitmaps);
  count_launch();
  k_poisson_vals<<<(s.count + 127) / 128, 128, 0, st>>>(s, bitmaps, vals);
  count_launch();
  return cudaGetLastError();
}

// --------------------------------------------------------------- periodic --
__global__ void __launch_bounds__(256) k_periodic_table(uint8_t* tab, uint64_t n, double start, double stop,
                                                        double step) {
  for (uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (uint64_t)gridDim.x * 256) {
    // np.linspace: y = arange(n) * step + start (two roundings); y[-1] = stop
    double t = __dadd_rn(__dmul_rn((double)i, step), start);
    if (i == n - 1 && n > 1) t = stop;
    tab[i] = (uint8_t)wrap_u8(__dmul_rn(sin(t), 255.0));
  }
}

cudaError_t launch_periodic_table(uint8_t* tab, uint64_t n, double lim, cudaStream_t st) {
  const double start = -lim, stop = lim;
  const double step = n > 1 ? (stop - start) / (double)(n - 1) : 0.0;
  uint32_t bx = (uint32_t)((n + 255) / 256);
  if (bx > 148u * 16u) bx = 148u * 16u;
  k_periodic_table<<<bx, 256, 0, st>>>(tab, n, start, stop, step);
  count_launch();
  return cudaGetLastError();
}

// --------------------------------------------------------------- brownian --
// dB[j] = sqrt(dt) * z[j], j in [0, n-1); B[i] = sum_{j<i} dB[j].  One thread
// per 8-element group, BR_CHUNK elements per CTA, float64 throughout.
constexpr int BR_THREADS = 256;
constexpr int BR_CHUNK = BR_THREADS * 8;

template <bool INJECT>
__device__ __forceinline__ void brownian_steps(const NoiseArgs& na, const BrownianArgs& ba, int img_local,
                                               uint32_t g, double (&d)[8]) {
  const uint64_t nz = na.elems - 1;
  if (INJECT) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint64_t j = (uint64_t)g * 8u + i;
      d[i] = j < nz ? __dmul_rn(ba.sqrt_dt, na.f0[(uint64_t)img_local * nz + j]) : 0.0;
    }
  } else {
    GroupCtx gc{(uint32_t)img_local, g, nullptr};
    const uint4 r = group_random(na, gc, 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float z0, z1;
      box_muller16(r16(r, 2 * k), r16(r, 2 * k + 1), z0, z1);
      const uint64_t j = (uint64_t)g * 8u + 2 * k;
      d[2 * k] = j < nz ? __dmul_rn(ba.sqrt_dt, (double)z0) : 0.0;
      d[2 * k + 1] = j + 1 < nz ? __dmul_rn(ba.sqrt_dt, (double)z1) : 0.0;
    }
  }
}

__device__ __forceinline__ double block_exclusive_scan(double x, double* total) {
  __shared__ double warp_sum[BR_THREADS / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double inc = x;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double y = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += y;
  }
  if (lane == 31) warp_sum[wid] = inc;
  __syncthreads();
  double base = 0.0, tot = 0.0;
#pragma unroll
  for (int k = 0; k < BR_THREADS / 32; ++k) {
    if (k < wid) base += warp_sum[k];
    tot += warp_sum[k];
  }
  __syncthreads();
  *total = tot;
  return base + inc - x;
}

template <bool INJECT>
__global__ void __launch_bounds__(BR_THREADS) k_brown_chunksum(ImageSet s, NoiseArgs na, BrownianArgs ba) {
  const int img_local = image_of(s, blockIdx.z);
  const uint32_t g = blockIdx.x * BR_THREADS + threadIdx.x;
  double d[8], sum = 0.0;
  brownian_steps<INJECT>(na, ba, img_local, g, d);
#pragma unroll
  for (int i = 0; i < 8; ++i) sum += d[i];
  double total;
  block_exclusive_scan(sum, &total);
  if (threadIdx.x == 0) ba.chunk_sums[(uint64_t)img_local * ba.nchunks + blockIdx.x] = total;
}

// exclusive scan of one image's chunk sums, in place, in SEQUENTIAL order (((t0 + t1) + t2) + ...: the order in which
// k_brown_march below accumulates them, so that both paths give the same bits).  One warp per image: 32 sums are loaded
// at once, every lane replays the 32 additions (round 1 ran one THREAD per image on 8 CTAs: 0.24 ms per 512 images).
__global__ void __launch_bounds__(BR_THREADS) k_brown_scan(ImageSet s, BrownianArgs ba) {
  const int z = blockIdx.x * (BR_THREADS / 32) + (threadIdx.x >> 5);
  if (z >= s.count) return;
  const int lane = threadIdx.x & 31;
  double* cs = ba.chunk_sums + (uint64_t)image_of(s, z) * ba.nchunks;
  double run = 0.0;
  for (uint32_t base = 0; base < ba.nchunks; base += 32) {
    const double v = base + lane < ba.nchunks ? cs[base + lane] : 0.0;
    double mine = 0.0;
#pragma unroll 8
    for (int j = 0; j < 32; ++j) {
      const double x = __shfl_sync(0xffffffffu, v, j);
      if (lane == j) mine = run;
      run = __dadd_rn(run, x);
    }
    if (base + lane < ba.nchunks) cs[base + lane] = mine;
  }
}

// one chunk of the field: the thread's 8 steps d (their sum already scanned inside the CTA: b = B at the thread's first
// element) -> (B * 255) -> u8, or cv2.add(image, that) when add_out is set
__device__ __forceinline__ void brownian_write(const ImageSet& s, const NoiseArgs& na, const BrownianArgs& ba, int img_local,
                                               uint32_t g, const double (&d)[8], double b) {
  uint8_t* f = (ba.add_out ? ba.add_out : ba.field) + (uint64_t)img_local * na.elems;
  const uint8_t* img = ba.add_out ? s.in + (uint64_t)img_local * s.elems : nullptr;
  uint32_t v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] = wrap_u8(__dmul_rn(b, 255.0));   // B[8g + i]
    b += d[i];
  }
  const uint64_t e = (uint64_t)g * 8u;
  if (e + 8u <= na.elems && ((reinterpret_cast<uintptr_t>(f + e) & 7u) == 0) &&
      (!img || (reinterpret_cast<uintptr_t>(img + e) & 7u) == 0)) {
    uint2 w = make_uint2(pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]));
    if (img) {                             // cv2.add(img, field): saturating, minibatch.py:828
      const uint2 x = __ldg(reinterpret_cast<const uint2*>(img + e));
      w = make_uint2(__vaddus4(x.x, w.x), __vaddus4(x.y, w.y));
    }
    *reinterpret_cast<uint2*>(f + e) = w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (e + i < na.elems) f[e + i] = (uint8_t)(img ? min(v[i] + (uint32_t)img[e + i], 255u) : v[i]);
  }
}

template <bool INJECT>
__global__ void __launch_bounds__(BR_THREADS) k_brown_field(ImageSet s, NoiseArgs na, BrownianArgs ba) {
  const int img_local = image_of(s, blockIdx.z);
  const uint32_t g = blockIdx.x * BR_THREADS + threadIdx.x;
  double d[8], sum = 0.0;
  brownian_steps<INJECT>(na, ba, img_local, g, d);
#pragma unroll
  for (int i = 0; i < 8; ++i) sum += d[i];
  double total;
  const double b = block_exclusive_scan(sum, &total) + ba.chunk_sums[(uint64_t)img_local * ba.nchunks + blockIdx.x];
  brownian_write(s, na, ba, img_local, g, d, b);
}

// Large batches: ONE CTA per image walks the image's chunks in order and carries the running sum - the normals are
// generated once (no chunk-sum pass, no scan).  Same arithmetic as the three kernels: per chunk the same block scan, across
// chunks the same sequential additions.  Used when the batch alone fills the GPU (>= 2 images per SM).
template <bool INJECT>
__global__ void __launch_bounds__(BR_THREADS) k_brown_march(ImageSet s, NoiseArgs na, BrownianArgs ba) {
  const int img_local = image_of(s, blockIdx.x);
  double run = 0.0;
  for (uint32_t chunk = 0; chunk < ba.nchunks; ++chunk) {
    const uint32_t g = chunk * BR_THREADS + threadIdx.x;
    double d[8], sum = 0.0;
    brownian_steps<INJECT>(na, ba, img_local, g, d);
#pragma unroll
    for (int i = 0; i < 8; ++i) sum += d[i];
    double total;
    const double b = block_exclusive_scan(sum, &total) + run;
    brownian_write(s, na, ba, img_local, g, d, b);
    run = __dadd_rn(run, total);
  }
}

cudaError_t launch_brownian_field(const ImageSet& s, const NoiseArgs& na, const BrownianArgs& ba, bool inject,
                                  cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  static const char* march_env = knob("B2D_BROWN_MARCH");           // 0 / 1 forces one of the two paths (cross-check)
  const bool march = march_env ? march_env[0] == '1' : s.count >= 2 * sm_count();
  if (march) {
    if (inject) k_brown_march<true><<<s.count, BR_THREADS, 0, st>>>(s, na, ba);
    else k_brown_march<false><<<s.count, BR_THREADS, 0, st>>>(s, na, ba);
    count_launch();
    return cudaGetLastError();
  }
  const dim3 grid(ba.nchunks, 1, s.count);
  if (inject) k_brown_chunksum<true><<<grid, BR_THREADS, 0, st>>>(s, na, ba);
  else k_brown_chunksum<false><<<grid, BR_THREADS, 0, st>>>(s, na, ba);
  count_launch();
  k_brown_scan<<<(s.count + BR_THREADS / 32 - 1) / (BR_THREADS / 32), BR_THREADS, 0, st>>>(s, ba);
  count_launch();
  if (inject) k_brown_field<true><<<grid, BR_THREADS, 0, st>>>(s, na, ba);
  else k_brown_field<false><<<grid, BR_THREADS, 0, st>>>(s, na, ba);
  count_launch();
  return cudaGetLastError();
}

uint32_t brownian_chunks(uint64_t elems) { return (uint32_t)((elems + BR_CHUNK - 1) / BR_CHUNK); }

}  // namespace b2d
