// b2d_fused3.cu - noise -> ->u8 -> 3x3 denoiser in ONE pass over HBM: the GENERAL variant
// (rows that are a multiple of 4 but not of 8 bytes); the hot variant is b2d_fused3w.cu.
//
// Replaces, per image, e.g.
//     gauss_array = random_noise(img, mode='gaussian', var=v)
//     im_noise    = (255 * gauss_array).astype(np.uint8)
//     im          = cv2.GaussianBlur(im_noise, (3, 3), 0)
// (lib/roi_data_layer/minibatch.py:117-119 and every `<noise>_{gaus_blur,mean,
// median}` branch) - three float64 temporaries and a filter pass on the CPU.
//
// Tile: TILE_H rows x TILE_WB bytes of one image per CTA.  Prologue: the
// noised uint8 values of the tile + 1-row / 8-byte apron are generated
// straight into shared memory (Philox is counter based, so apron values equal
// what the neighbouring CTA generates).  Main loop: the packed band filter of the
// hot kernel (KW words per thread, two row-parts in registers).
// Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_kernels.cuh"
#include "b2d_filter3w.cuh"

namespace b2d {

constexpr int F3_THREADS = 256;
#ifndef B2D_F3ANY_UNROLL
#define B2D_F3ANY_UNROLL 1
#endif
constexpr int F3_UNROLL = B2D_F3ANY_UNROLL;       // prologue steps unrolled (Philox chains of several groups in flight)
constexpr int F3_APRON = 16;                      // shared bytes each side of the tile (8 are generated, 4 are read)
// Tile = TWB bytes x 49152 / TWB rows (1024 x 48 as the hot kernel, or 512 x 96: VOC's 1500-byte rows fill 3 x 512
// to 98 % but 2 x 1024 to 73 %); the band height of the filter phase is 12 rows (24 for the median) either way.
template <int TWB>
struct F3Geom {
  static constexpr int TILE_WB = TWB;
  static constexpr int TILE_H = 49152 / TWB;
  static constexpr int SROW = TWB + 2 * F3_APRON;
  static constexpr int ROWS = TILE_H + 2;
  static constexpr int GROUPS_PER_ROW = (TWB + 16) / 8 + 1;   // +1: rows whose start is 4 (mod 8)
  static constexpr size_t SMEM = (size_t)ROWS * SROW;
};

// load the 8 bytes of group g of an image (4-byte aligned base, elems % 4 == 0)
__device__ __forceinline__ uint2 load_group(const uint8_t* img, uint64_t elems, uint32_t g) {
  const uint64_t e = (uint64_t)g * 8u;
  const uint8_t* p = img + e;
  uint2 w = make_uint2(0u, 0u);
  if ((reinterpret_cast<uintptr_t>(p) & 7u) == 0 && e + 8u <= elems) {
    w = __ldg(reinterpret_cast<const uint2*>(p));
  } else {
    if (e + 4u <= elems) w.x = __ldg(reinterpret_cast<const uint32_t*>(p));
    if (e + 8u <= elems) w.y = __ldg(reinterpret_cast<const uint32_t*>(p + 4));
  }
  return w;
}

// Prologue: one 8-byte Philox group per thread and step, wherever it falls in the row (a row
// that starts at 4 (mod 8) begins and ends in the middle of a group; the group is generated
// whole and the half that belongs to the row is kept) - the PACKED operators of the hot kernel.
// Filter: the band code of k_fused3 (b2d_fused3w.cu), 32-bit stores.
template <int KIND, int FILTER, int KW, int TWB>
__global__ void __launch_bounds__(F3_THREADS, 4) k_fused3_any(ImageSet s, NoiseArgs na) {
  using F = Filter3W<FILTER, KW>;
  using G = F3Geom<TWB>;
  constexpr int F3_TILE_WB = G::TILE_WB, F3_TILE_H = G::TILE_H, F3_SROW = G::SROW, F3_ROWS = G::ROWS;
  constexpr int F3_GROUPS_PER_ROW = G::GROUPS_PER_ROW;
  constexpr int TPB = F3_TILE_WB / (4 * KW);        // threads per band
  constexpr int BANDS = F3_THREADS / TPB;
  constexpr int BAND_H = F3_TILE_H / BANDS;
  extern __shared__ __align__(16) uint8_t smem[];

  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * F3_TILE_WB;          // first byte column of the tile
  const int y0 = blockIdx.y * F3_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int cs = c0 - F3_APRON;                    // byte column held at smem row offset 0
  const int col_lo = max(c0 - 8, 0), col_hi = min(c0 + F3_TILE_WB + 8, s.rowbytes);
  const int rows_needed = min(F3_ROWS, s.h - y0 + 2);

  // ---- prologue: noised bytes -> shared memory -----------------------------
  // per tile row: first group and the byte column it starts at (64-bit arithmetic once per row)
  __shared__ int2 s_row[F3_ROWS];
  if (tid < rows_needed) {
    const int ys = border_map(y0 - 1 + tid, s.h, F::BORDER);
    const uint64_t fs = (uint64_t)ys * (uint64_t)s.rowbytes;     // flat index of the row start
    const uint32_t g0 = (uint32_t)((fs + col_lo) >> 3);
    s_row[tid] = make_int2((int)g0, (int)((int64_t)((uint64_t)g0 * 8u) - (int64_t)fs));
  }
  __syncthreads();
  const uint32_t g_full = (uint32_t)(s.elems >> 3);              // groups that lie completely inside the image
  const bool al8 = (reinterpret_cast<uintptr_t>(img) & 7u) == 0;
#pragma unroll F3_UNROLL
  for (int it = tid; it < rows_needed * F3_GROUPS_PER_ROW; it += F3_THREADS) {
    const int tr = it / F3_GROUPS_PER_ROW;
    const int gi = it - tr * F3_GROUPS_PER_ROW;
    const int2 ri = s_row[tr];
    const uint32_t g = (uint32_t)ri.x + (uint32_t)gi;
    const int cg = ri.y + 8 * gi;                                 // byte column of the group start
    if (cg >= col_hi) continue;
    uint2 raw;
    if (al8 && g < g_full) raw = __ldg(reinterpret_cast<const uint2*>(img + (uint64_t)g * 8u));
    else raw = load_group(img, s.elems, g);
    GroupCtx gc{(uint32_t)img_local, g, img};
    const uint2 w = PackedNoise<KIND>::apply(na, gc, raw);
    uint8_t* sdst = smem + tr * F3_SROW + (cg - cs);
    if (cg >= col_lo) *reinterpret_cast<uint32_t*>(sdst) = w.x;             // (cg < col_hi: tested above)
    if (cg + 4 >= col_lo && cg + 4 < col_hi) *reinterpret_cast<uint32_t*>(sdst + 4) = w.y;
  }
  __syncthreads();

  // ---- left / right image border: one apron pixel -------------------------
  const bool left_edge = (c0 == 0);
  const bool right_edge = (s.rowbytes <= c0 + F3_TILE_WB);
  if (left_edge || right_edge) {
    if (tid < rows_needed) {                        // one thread per tile row, 3 bytes per side
      uint8_t* srow = smem + tid * F3_SROW - cs;
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
  const int o0 = band * BAND_H;                      // first output row (tile-relative) of the band
  const int nrows = min(BAND_H, s.h - y0 - o0);
  if (c >= s.rowbytes || nrows <= 0) return;
  const uint8_t* sp = smem + F3_APRON + 4 * KW * q + o0 * F3_SROW;
  auto load_row = [&](int tr, typename F::Part& p) {
    const uint8_t* rp = sp + tr * F3_SROW;
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
  const int nw = min(KW, (s.rowbytes - c) >> 2);     // valid words of the thread's column (rows are word multiples)
#pragma unroll
  for (int r = 0; r < BAND_H; ++r) {
    if (r >= nrows) break;
    typename F::Part& pt = (r % 3 == 0) ? p0 : (r % 3 == 1) ? p1 : p2;
    typename F::Part& pm = (r % 3 == 0) ? p1 : (r % 3 == 1) ? p2 : p0;
    typename F::Part& pb = (r % 3 == 0) ? p2 : (r % 3 == 1) ? p0 : p1;
    load_row(r + 2, pb);
    uint32_t o[KW];
    F::vcomb(pt, pm, pb, o);
#pragma unroll
    for (int k = 0; k < KW; ++k)
      if (k < nw) reinterpret_cast<uint32_t*>(optr)[k] = o[k];
    optr += s.rowbytes;
  }
}

template <int KIND, int FILTER, int TWB>
static cudaError_t launch_width(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  constexpr int KW = FILTER == FK_MEDIAN3 ? 2 : 4;
  using G = F3Geom<TWB>;
  static bool attr_done[64] = {};               // per device: the attribute lives in the context
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(k_fused3_any<KIND, FILTER, KW, TWB>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM);
    if (e != cudaSuccess) return e;
    attr_done[dev & 63] = true;
  }
  const dim3 grid((s.rowbytes + G::TILE_WB - 1) / G::TILE_WB, (s.h + G::TILE_H - 1) / G::TILE_H, s.count);
  k_fused3_any<KIND, FILTER, KW, TWB><<<grid, F3_THREADS, G::SMEM, st>>>(s, na);
  count_launch();
  return cudaGetLastError();
}

template <int KIND, int FILTER>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  // the width that pads the rows least (ties: the wider tile)
  const int pad1024 = (s.rowbytes + 1023) / 1024 * 1024, pad512 = (s.rowbytes + 511) / 512 * 512;
  if (pad512 < pad1024) return launch_width<KIND, FILTER, 512>(s, na, st);
  return launch_width<KIND, FILTER, 1024>(s, na, st);
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

cudaError_t launch_fused3_any(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
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
