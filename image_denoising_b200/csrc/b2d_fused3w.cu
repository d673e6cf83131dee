// b2d_fused3w.cu - noise -> ->u8 -> 3x3 denoiser in ONE pass over HBM (the hot kernel).
//
// Replaces, per image, e.g.
//     gauss_array = random_noise(img, mode='gaussian', var=v)
//     im_noise    = (255 * gauss_array).astype(np.uint8)
//     im          = cv2.GaussianBlur(im_noise, (3, 3), 0)
// (lib/roi_data_layer/minibatch.py:117-119 and every `<noise>_{gaus_blur,mean,median}`
// branch) - three float64 temporaries and a filter pass on the CPU.
//
// Tile: FT_TH rows x FT_TWB bytes of one image per CTA (rows are 8-byte multiples).
//   prologue : every thread turns 16 input bytes (two 8-byte Philox groups) per step into
//              noised bytes with the PACKED operators (b2d_noise.cuh) and stores them to
//              shared memory; the 1-row / 8-byte apron is regenerated from the counter,
//              so it equals what the neighbouring CTA computes;
//   filter   : the tile is cut in horizontal bands; a thread owns KW adjacent words of a
//              band and slides down it, keeping two "row parts" in registers
//              (b2d_filter3w.cuh); 16-byte (8-byte for median) stores.
// Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_kernels.cuh"
#include "b2d_filter3w.cuh"

#include <cstdlib>

namespace b2d {

#ifndef B2D_FT_THREADS
#define B2D_FT_THREADS 256
#endif
#ifndef B2D_FT_MINB
#define B2D_FT_MINB 4
#endif
constexpr int FT_THREADS = B2D_FT_THREADS;
constexpr int FT_STEP = FT_THREADS / 64;   // tile rows per prologue step
constexpr int FT_TWB = 1024;               // tile width in bytes
#ifndef B2D_FT_TH
#define B2D_FT_TH 48
#endif
#ifndef B2D_FT_UNROLL
#define B2D_FT_UNROLL 4
#endif
constexpr int FT_UNROLL = B2D_FT_UNROLL;
constexpr int FT_TH = B2D_FT_TH;           // tile height in rows (48: 4 CTAs / SM by shared memory and registers)
constexpr int FT_APRON = 16;               // bytes each side in shared memory (8 are used)
constexpr int FT_SROW = FT_TWB + 2 * FT_APRON;
constexpr int FT_ROWS = FT_TH + 2;
constexpr size_t FT_SMEM = (size_t)FT_ROWS * FT_SROW;

template <int KIND>
__device__ __forceinline__ uint2 noise_group(const NoiseArgs& na, int img_local, const uint8_t* img, uint32_t g,
                                             uint2 w) {
  GroupCtx gc{(uint32_t)img_local, g, img};
  return PackedNoise<KIND>::apply(na, gc, w);
}

// A4: rows (and image bases) are only 4-byte aligned - rows of 4 (mod 8) bytes, W = 500.  Without a
// generator (KIND == NK_NONE) groups do not matter, so the same tile code runs with 32-bit accesses.
template <int KIND, int FILTER, int KW, bool A4 = false>
__global__ void __launch_bounds__(FT_THREADS, B2D_FT_MINB) k_fused3(ImageSet s, NoiseArgs na, int th) {
  static_assert(!A4 || KIND == NK_NONE, "the 4-byte aligned variant has no generator");
  using F = Filter3W<FILTER, KW>;
  constexpr int TPB = FT_TWB / (4 * KW);            // threads per band
  constexpr int BANDS = FT_THREADS / TPB;
  constexpr int BAND_H = FT_TH / BANDS;             // most rows of a band; this launch's tiles have th <= FT_TH rows
  extern __shared__ __align__(16) uint8_t smem[];

  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * FT_TWB;               // first byte column of the tile
  const int y0 = blockIdx.y * th;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int rows = min(th + 2, s.h - y0 + 2);       // tile row tr holds image row y0 - 1 + tr

  // ---- prologue: 16 bytes per thread and step, 4 tile rows per step ----------------
  {
    const int q = tid & 63, r0 = tid >> 6;
    const int c = c0 + 16 * q;
    const bool v0 = c < s.rowbytes, v1 = c + 8 < s.rowbytes;
    if (A4) {
      const int nwv = (s.rowbytes - c) >> 2;            // valid words of the thread's 16 bytes
      if (v0) {
#pragma unroll FT_UNROLL
        for (int tr = r0; tr < rows; tr += FT_STEP) {
          int ys = y0 - 1 + tr;
          if (ys < 0 || ys >= s.h) ys = border_map(ys, s.h, F::BORDER);
          const uint32_t* pw = reinterpret_cast<const uint32_t*>(img + (uint32_t)ys * (uint32_t)s.rowbytes + (uint32_t)c);
          uint4 v = make_uint4(__ldg(pw), 0u, 0u, 0u);
          if (nwv > 1) v.y = __ldg(pw + 1);
          if (nwv > 2) v.z = __ldg(pw + 2);
          if (nwv > 3) v.w = __ldg(pw + 3);
          *reinterpret_cast<uint4*>(smem + tr * FT_SROW + FT_APRON + 16 * q) = v;
        }
      }
      for (int it = tid; it < 2 * rows; it += FT_THREADS) {
        const int tr = it >> 1, side = it & 1;
        const int ca = side ? c0 + FT_TWB : c0 - 8;
        if (ca < 0 || ca >= s.rowbytes) continue;
        const int ys = border_map(y0 - 1 + tr, s.h, F::BORDER);
        const uint32_t* pw = reinterpret_cast<const uint32_t*>(img + (uint32_t)ys * (uint32_t)s.rowbytes + (uint32_t)ca);
        uint2 a = make_uint2(__ldg(pw), 0u);
        if (ca + 4 < s.rowbytes) a.y = __ldg(pw + 1);
        *reinterpret_cast<uint2*>(smem + tr * FT_SROW + (side ? FT_APRON + FT_TWB : FT_APRON - 8)) = a;
      }
    } else {
    if (v0) {
#pragma unroll FT_UNROLL
      for (int tr = r0; tr < rows; tr += FT_STEP) {
        int ys = y0 - 1 + tr;                             // only the first / last tile row can leave the image
        if (ys < 0 || ys >= s.h) ys = border_map(ys, s.h, F::BORDER);   // (warp-uniform: a warp shares tr)
        const uint32_t e = (uint32_t)ys * (uint32_t)s.rowbytes + (uint32_t)c;
        const uint2 a = __ldg(reinterpret_cast<const uint2*>(img + e));
        uint2 b = make_uint2(0u, 0u);
        if (v1) b = __ldg(reinterpret_cast<const uint2*>(img + e + 8));
        const uint2 na0 = noise_group<KIND>(na, img_local, img, e >> 3, a);
        uint2 nb0 = b;
        if (v1) nb0 = noise_group<KIND>(na, img_local, img, (e >> 3) + 1u, b);
        *reinterpret_cast<uint4*>(smem + tr * FT_SROW + FT_APRON + 16 * q) = make_uint4(na0.x, na0.y, nb0.x, nb0.y);
      }
    }
    // apron groups: 8 bytes left of the tile and 8 bytes right of it
    for (int it = tid; it < 2 * rows; it += FT_THREADS) {
      const int tr = it >> 1, side = it & 1;
      const int ca = side ? c0 + FT_TWB : c0 - 8;
      if (ca < 0 || ca >= s.rowbytes) continue;
      const int ys = border_map(y0 - 1 + tr, s.h, F::BORDER);
      const uint32_t e = (uint32_t)ys * (uint32_t)s.rowbytes + (uint32_t)ca;
      const uint2 a = noise_group<KIND>(na, img_local, img, e >> 3, __ldg(reinterpret_cast<const uint2*>(img + e)));
      *reinterpret_cast<uint2*>(smem + tr * FT_SROW + (side ? FT_APRON + FT_TWB : FT_APRON - 8)) = a;
    }
    }  // !A4
  }
  __syncthreads();

  // ---- left / right image border: one apron pixel ---------------------------------
  const bool left_edge = (c0 == 0);
  const bool right_edge = (s.rowbytes <= c0 + FT_TWB);
  if (left_edge || right_edge) {
    if (tid < rows) {                               // one thread per tile row, 3 bytes per side
      uint8_t* srow = smem + tid * FT_SROW + FT_APRON - c0;
      if (left_edge) {
        const int sp3 = 3 * border_map(-1, s.w, F::BORDER);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) srow[ch - 3] = srow[sp3 + ch];
      }
      if (right_edge) {
        const int sp3 = 3 * border_map(s.w, s.w, F::BORDER);
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) srow[s.rowbytes + ch] = srow[sp3 + ch];
      }
    }
    __syncthreads();
  }

  // ---- filter: a thread owns KW words of one band and slides down it -----------------
  const int band = tid / TPB, q = tid - band * TPB;
  const int c = c0 + 4 * KW * q;
  const int band_h = th / BANDS;
  const int o0 = band * band_h;                      // first output row (tile-relative) of the band
  const int nrows = min(band_h, s.h - y0 - o0);
  if (c >= s.rowbytes || nrows <= 0) return;
  const uint8_t* sp = smem + FT_APRON + 4 * KW * q + o0 * FT_SROW;
  auto load_row = [&](int tr, typename F::Part& p) {
    const uint8_t* rp = sp + tr * FT_SROW;
    uint32_t w[KW + 2];
    w[0] = *reinterpret_cast<const uint32_t*>(rp - 4);
    if (KW == 4) {
      const uint4 v = *reinterpret_cast<const uint4*>(rp);
      w[1] = v.x; w[2] = v.y; w[3] = v.z; w[KW] = v.w;
    } else {
      const uint2 v = *reinterpret_cast<const uint2*>(rp);
      w[1] = v.x; w[KW] = v.y;
    }
    w[KW + 1] = *reinterpret_cast<const uint32_t*>(rp + 4 * KW);
    RowLanes<KW> rl;
    rl.split(w);
    F::hpart(rl, p);
  };
  typename F::Part p0, p1, p2;
  load_row(0, p0);
  load_row(1, p1);
  uint8_t* optr = s.out + (uint64_t)img_local * s.elems + (uint64_t)(y0 + o0) * s.rowbytes + c;
  const bool full = c + 4 * KW <= s.rowbytes;        // rows are 8-byte multiples: else exactly 8 valid bytes
  const int nw4 = min(KW, (s.rowbytes - c) >> 2);    // A4: valid words of the thread's column
#pragma unroll
  for (int r = 0; r < BAND_H; ++r) {
    if (r >= nrows) break;
    typename F::Part& pt = (r % 3 == 0) ? p0 : (r % 3 == 1) ? p1 : p2;
    typename F::Part& pm = (r % 3 == 0) ? p1 : (r % 3 == 1) ? p2 : p0;
    typename F::Part& pb = (r % 3 == 0) ? p2 : (r % 3 == 1) ? p0 : p1;
    load_row(r + 2, pb);
    uint32_t o[KW];
    F::vcomb(pt, pm, pb, o);
    if (A4) {
#pragma unroll
      for (int k = 0; k < KW; ++k)
        if (k < nw4) reinterpret_cast<uint32_t*>(optr)[k] = o[k];
    } else {
      *reinterpret_cast<uint2*>(optr) = make_uint2(o[0], o[1]);
      if (KW == 4 && full) *reinterpret_cast<uint2*>(optr + 8) = make_uint2(o[KW - 2], o[KW - 1]);
    }
    optr += s.rowbytes;
  }
}

template <int KIND, int FILTER, bool A4 = false>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  constexpr int KW = FILTER == FK_MEDIAN3 ? 2 : 4;
  static bool attr_done[64] = {};               // per device: the attribute lives in the context
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    const cudaError_t e =
        cudaFuncSetAttribute(k_fused3<KIND, FILTER, KW, A4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FT_SMEM);
    if (e != cudaSuccess) return e;
    attr_done[dev & 63] = true;
  }
  // Tile height: FT_TH, except for batches of a few waves of CTAs, where the last, nearly empty wave decides (BASELINE
  // config 1: 16 images = 624 tiles of 48 rows on 592 slots = two waves for 1.05 waves of work).  Cost model: waves x
  // (rows generated per tile + 6 for the tile's fixed cost); multiples of 4 so that 2 or 4 bands divide the tile.  For
  // large batches this picks FT_TH (fewer apron rows), which is also what was measured best (profiles/r02r_tile_geometry.txt).
  const long slots = (long)sm_count() * B2D_FT_MINB;
  const long cols = (s.rowbytes + FT_TWB - 1) / FT_TWB;
  int th = FT_TH;
  if (cols * ((s.h + FT_TH - 1) / FT_TH) * s.count <= 6 * slots) {
    long best = -1;
    for (int t = FT_TH; t >= 16; t -= 4) {
      const long ctas = cols * ((s.h + t - 1) / t) * s.count;
      const long cost = ((ctas + slots - 1) / slots) * (t + 8);
      if (best < 0 || cost < best) { best = cost; th = t; }
    }
  }
  const dim3 grid((unsigned)cols, (s.h + th - 1) / th, s.count);
  k_fused3<KIND, FILTER, KW, A4><<<grid, FT_THREADS, FT_SMEM, st>>>(s, na, th);
  count_launch();
  return cudaGetLastError();
}

template <int KIND>
static cudaError_t launch_filter(int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (filter) {
    case FK_GAUS3: return launch_one<KIND, FK_GAUS3>(s, na, st);
    case FK_MEAN3: return launch_one<KIND, FK_MEAN3>(s, na, st);
    case FK_MEDIAN3: return launch_one<KIND, FK_MEDIAN3>(s, na, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_fused3_any(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st);
cudaError_t launch_march3(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st);
bool march3_supported(const ImageSet& s);
bool march3_a4_supported(const ImageSet& s);
cudaError_t launch_march3_a4(int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st);

cudaError_t launch_fused3(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  const bool fast = s.rowbytes % 8 == 0 && s.elems < (1ull << 31) &&
                    ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 7u) == 0;
  if (!fast) {
    // no generator, rows / bases only 4-byte aligned (W = 500): the tile kernel with 32-bit accesses
    const bool fast4 = kind == NK_NONE && s.rowbytes % 4 == 0 && s.elems < (1ull << 31) &&
                       ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 3u) == 0;
    static const char* a4_env = knob("B2D_A4");      // B2D_A4=0: the group-wise kernel (cross-check)
    if (fast4 && !(a4_env && a4_env[0] == '0')) {
      static const char* march4_env = knob("B2D_MARCH");   // as below: gaus_blur / mean march, median tiles
      if (filter != FK_MEDIAN3 && !(march4_env && march4_env[0] == '0') && march3_a4_supported(s))
        return launch_march3_a4(filter, s, na, st);
      switch (filter) {
        case FK_GAUS3: return launch_one<NK_NONE, FK_GAUS3, true>(s, na, st);
        case FK_MEAN3: return launch_one<NK_NONE, FK_MEAN3, true>(s, na, st);
        case FK_MEDIAN3: return launch_one<NK_NONE, FK_MEDIAN3, true>(s, na, st);
      }
      return cudaErrorInvalidValue;
    }
    return launch_fused3_any(kind, filter, s, na, st);
  }
  // Measured on B200 (scripts/bench_combos.py, 512 x 600 x 1000): without a generator the
  // row-marching persistent kernel (b2d_march3.cu) is 1.4-1.5x faster for gaus_blur / mean
  // (0.32 / 0.35 ms vs 0.48 / 0.50 ms: no apron rows, register prefetch); with a generator in the
  // prologue the tiles win (0.69 vs 0.78 ms for sap_mean: longer phases, fewer barriers), as they
  // do for the register-heavy median.  B2D_MARCH=0 / 1 forces one of the two (profiling).
  // (A third schedule - one warp per 512-byte column strip, taps by SHFL, no shared memory and no barrier - was
  // measured 13-15 % slower than the tiles: profiles/r02q_strip.txt, source in the commit before its removal.)
  static const char* march_env = knob("B2D_MARCH");
  const bool march = march_env ? march_env[0] == '1' : (kind == NK_NONE && filter != FK_MEDIAN3);
  if (march && march3_supported(s)) return launch_march3(kind, filter, s, na, st);
  switch (kind) {
    case NK_NONE: return launch_filter<NK_NONE>(filter, s, na, st);
    case NK_GAUSSIAN: return launch_filter<NK_GAUSSIAN>(filter, s, na, st);
    case NK_SPECKLE: return launch_filter<NK_SPECKLE>(filter, s, na, st);
    case NK_SAP: return launch_filter<NK_SAP>(filter, s, na, st);
    case NK_POISSON: return launch_filter<NK_POISSON>(filter, s, na, st);
    case NK_UNIFORM: return launch_filter<NK_UNIFORM>(filter, s, na, st);
    case NK_RAYLEIGH: return launch_filter<NK_RAYLEIGH>(filter, s, na, st);
    case NK_GAMMA: return launch_filter<NK_GAMMA>(filter, s, na, st);
    case NK_TABLE_ADD: return launch_filter<NK_TABLE_ADD>(filter, s, na, st);
    case NK_SHADER: return launch_filter<NK_SHADER>(filter, s, na, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2d
