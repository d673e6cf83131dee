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
// what the neighbouring CTA generates).  Main loop: each thread owns a 4-byte
// column word and slides down the tile keeping two row-parts in registers.
// Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_kernels.cuh"

namespace b2d {

constexpr int F3_THREADS = 256;
constexpr int F3_TILE_WB = F3_THREADS * 4;       // 1024 bytes wide
constexpr int F3_TILE_H = 32;
constexpr int F3_APRON = 8;                       // bytes each side (one noise group)
constexpr int F3_SROW = F3_TILE_WB + 2 * F3_APRON;
constexpr int F3_ROWS = F3_TILE_H + 2;
constexpr int F3_GROUPS_PER_ROW = F3_SROW / 8 + 1;  // +1: rows whose start is 4 (mod 8)

__device__ __forceinline__ void unpack8(uint2 w, uint32_t (&v)[8]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[i] = byte_of(w.x, i);
    v[4 + i] = byte_of(w.y, i);
  }
}
__device__ __forceinline__ uint2 pack8(const uint32_t (&v)[8]) {
  return make_uint2(pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]));
}

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

template <int KIND, int FILTER>
__global__ void __launch_bounds__(F3_THREADS) k_fused3_any(ImageSet s, NoiseArgs na) {
  using F = Filter3<FILTER>;
  using Nz = NoiseOp<KIND, NM_GEN>;
  extern __shared__ __align__(16) uint8_t smem[];

  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * F3_TILE_WB;          // first byte column of the tile
  const int y0 = blockIdx.y * F3_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int cs = c0 - F3_APRON;                    // byte column held at smem row offset 0
  const int col_lo = max(cs, 0), col_hi = min(c0 + F3_TILE_WB + F3_APRON, s.rowbytes);
  const int rows_needed = min(F3_ROWS, s.h - y0 + 2);

  // ---- prologue: noised bytes -> shared memory -----------------------------
  for (int it = tid; it < rows_needed * F3_GROUPS_PER_ROW; it += F3_THREADS) {
    const int tr = it / F3_GROUPS_PER_ROW;
    const int gi = it - tr * F3_GROUPS_PER_ROW;
    const int ys = border_map(y0 - 1 + tr, s.h, F::BORDER);
    const uint64_t fs = (uint64_t)ys * (uint64_t)s.rowbytes;     // flat index of the row start
    const uint32_t g = (uint32_t)((fs + col_lo) >> 3) + gi;
    const int64_t cg = (int64_t)((uint64_t)g * 8u) - (int64_t)fs;  // byte column of the group start
    if (cg >= col_hi) continue;
    uint32_t v[8];
    unpack8(load_group(img, s.elems, g), v);
    GroupCtx gc{(uint32_t)img_local, g, img};
    Nz::apply8(na, gc, v);
    const uint2 w = pack8(v);
    uint8_t* srow = smem + tr * F3_SROW;
    const int o0 = (int)cg - cs, o1 = o0 + 4;
    if (cg >= col_lo && cg + 4 <= col_hi) *reinterpret_cast<uint32_t*>(srow + o0) = w.x;
    if (cg + 4 >= col_lo && cg + 8 <= col_hi) *reinterpret_cast<uint32_t*>(srow + o1) = w.y;
  }
  __syncthreads();

  // ---- left / right image border: one apron pixel -------------------------
  const bool left_edge = (c0 == 0);
  const bool right_edge = (s.rowbytes <= c0 + F3_TILE_WB);
  if (left_edge || right_edge) {
    for (int it = tid; it < rows_needed * 6; it += F3_THREADS) {
      const int tr = it / 6, k = it - tr * 6, side = k / 3, ch = k - side * 3;
      if ((side == 0 && !left_edge) || (side == 1 && !right_edge)) continue;
      const int px = side == 0 ? -1 : s.w;
      const int sp = border_map(px, s.w, F::BORDER);
      uint8_t* srow = smem + tr * F3_SROW;
      srow[px * 3 + ch - cs] = srow[sp * 3 + ch - cs];
    }
    __syncthreads();
  }

  // ---- main loop: sliding 3-row window down the tile -----------------------
  const int c = c0 + 4 * tid;
  if (c >= s.rowbytes) return;
  const uint8_t* sp = smem + F3_APRON + 4 * tid;
  auto row_part = [&](int tr) {
    const uint32_t wl = *reinterpret_cast<const uint32_t*>(sp + tr * F3_SROW - 4);
    const uint32_t wc = *reinterpret_cast<const uint32_t*>(sp + tr * F3_SROW);
    const uint32_t wr = *reinterpret_cast<const uint32_t*>(sp + tr * F3_SROW + 4);
    return F::hpart(word_left3(wl, wc), wc, word_right3(wc, wr));
  };
  typename F::Part p0 = row_part(0), p1 = row_part(1);
  uint8_t* optr = s.out + (uint64_t)img_local * s.elems + (uint64_t)y0 * s.rowbytes + c;
  const int nrows = min(F3_TILE_H, s.h - y0);
#pragma unroll 4
  for (int r = 0; r < nrows; ++r) {
    const typename F::Part p2 = row_part(r + 2);
    *reinterpret_cast<uint32_t*>(optr) = F::vcomb(p0, p1, p2);
    optr += s.rowbytes;
    p0 = p1;
    p1 = p2;
  }
}

template <int KIND, int FILTER>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  const dim3 grid((s.rowbytes + F3_TILE_WB - 1) / F3_TILE_WB, (s.h + F3_TILE_H - 1) / F3_TILE_H, s.count);
  const size_t smem = (size_t)F3_ROWS * F3_SROW;
  k_fused3_any<KIND, FILTER><<<grid, F3_THREADS, smem, st>>>(s, na);
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
