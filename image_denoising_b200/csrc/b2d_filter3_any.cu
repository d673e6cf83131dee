// b2d_filter3_any.cu - the 3x3 denoisers for images whose rows are NOT word aligned
// (W % 4 != 0, e.g. 499 or 333 pixels wide: rowbytes = 3 W is odd or 2 (mod 4)).
//
// Replaces cv2.GaussianBlur(im, (3, 3), 0) / cv2.blur(im, (3, 3)) / cv2.medianBlur(im, 3) of
// lib/roi_data_layer/minibatch.py:117-170 for those widths.  The first version of this path
// (b2d_filter3_bytes.cu) ran one thread per output byte, 9 byte loads each: 27x slower per pixel
// than the aligned tile kernel.  Here the rows are RE-ALIGNED on the way into shared memory, so the
// filter phase is the packed-lane code of the hot kernel (b2d_filter3w.cuh) unchanged:
//
//   load   : every row of the tile starts at an arbitrary byte address; a warp takes a row and
//            copies the ALIGNED words that cover it into shared memory with 4-byte cp.async (all
//            rows of the tile in flight, no registers);
//   filter : as k_fused3 (bands, KW words per thread), except that a row's words are funnel-shifted
//            by the row's phase (address & 3) as they are read; results into a second shared tile.
//   store  : the mirror image - words that lie completely inside the row are rebuilt with a
//            funnel shift from the result tile and written as aligned 32-bit stores, the 0-3
//            bytes at either end of the row segment as byte stores.
// Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_denoise.cuh"
#include "b2d_filter3w.cuh"

#include <cstdlib>

namespace b2d {

constexpr int FA_THREADS = 256;
constexpr int FA_TWB = 1024;               // tile width in bytes (one word per thread and row)
constexpr int FA_TH = 32;                  // tile height in rows
constexpr int FA_APRON = 16;               // shared bytes left of the tile (8 are loaded, 3 are used)
constexpr int FA_SROW = FA_TWB + 32;
constexpr int FA_ROWS = FA_TH + 2;
constexpr int FA_XW = 3;                   // words per row past the first 256: [c0 - 8, c0 + 1028) is loaded
constexpr size_t FA_SMEM = (size_t)FA_ROWS * FA_SROW + (size_t)FA_TH * FA_TWB;

// the aligned word at p, bytes outside [lo, hi) read as 0
__device__ __forceinline__ uint32_t guarded_word(const uint8_t* p, const uint8_t* lo, const uint8_t* hi) {
  if (p >= lo && p + 4 <= hi) return __ldg(reinterpret_cast<const uint32_t*>(p));
  uint32_t v = 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (p + i >= lo && p + i < hi) v |= (uint32_t)__ldg(p + i) << (8 * i);
  return v;
}

template <int FILTER, int KW>
__global__ void __launch_bounds__(FA_THREADS, 3) k_filter3_tile_any(ImageSet s) {
  using F = Filter3W<FILTER, KW>;
  constexpr int TPB = FA_TWB / (4 * KW);            // threads per band
  constexpr int BANDS = FA_THREADS / TPB;
  constexpr int BAND_H = FA_TH / BANDS;
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* s_out = smem + FA_ROWS * FA_SROW;

  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * FA_TWB;               // first byte column of the tile
  const int y0 = blockIdx.y * FA_TH;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const uint8_t* img_end = img + s.elems;
  const int rows = min(FA_ROWS, s.h - y0 + 2);      // tile row tr holds image row y0 - 1 + tr

  // ---- load: bytes [c0 - 8, c0 + 1028) of every tile row, re-aligned to the shared row ----
  auto row_src = [&](int tr) {
    int ys = y0 - 1 + tr;
    if (ys < 0 || ys >= s.h) ys = border_map(ys, s.h, F::BORDER);
    return img + (int64_t)ys * s.rowbytes + c0 - 8;                          // lands at shared offset APRON - 8
  };
  // phase (address & 3) of tile row tr, in 32-bit arithmetic: (base + ys * rowbytes) mod 4
  const unsigned ph_base = (unsigned)((reinterpret_cast<uintptr_t>(img) + (unsigned)c0 - 8u) & 3u);
  const unsigned rb3 = (unsigned)s.rowbytes & 3u;
  auto row_phase = [&](int tr) {
    int ys = y0 - 1 + tr;
    if (ys < 0 || ys >= s.h) ys = border_map(ys, s.h, F::BORDER);
    return (ph_base + ((unsigned)ys & 3u) * rb3) & 3u;
  };
  // A WARP owns a tile row (rows warp, warp + 8, ...): the aligned words that cover it go to shared
  // memory RAW with 4-byte cp.async - every row of the tile is in flight at once, nothing waits in
  // registers (ncu r01i: the register version sat 15 of 25 cycles per instruction on the loads at
  // 37 % occupancy).  The row's phase (address & 3) is applied by the readers below.
  constexpr int NW = FA_TWB / 4 + FA_XW + 1;        // aligned words per tile row (one more than re-aligned words)
  const int warp = tid >> 5, lane = tid & 31;
  const int nwl = min(NW, ((s.rowbytes - c0 + 11) >> 2) + 1);   // words that hold bytes of the row
#pragma unroll 1
  for (int tr = warp; tr < rows; tr += FA_THREADS / 32) {
    const uint8_t* src = row_src(tr);
    const uint8_t* base = src - (reinterpret_cast<uintptr_t>(src) & 3u);
    uint32_t* sdst = reinterpret_cast<uint32_t*>(smem + tr * FA_SROW + FA_APRON - 8);
    if (base >= img && base + 4 * NW <= img_end) {
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(sdst);
#pragma unroll
      for (int i = 0; i < (NW + 31) / 32; ++i) {
        const int j = lane + 32 * i;
        if (j < nwl)
          asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + 4 * j), "l"(base + 4 * j) : "memory");
      }
    } else {                                        // first / last rows of an image: bytes outside it read as 0
      for (int j = lane; j < nwl; j += 32) sdst[j] = guarded_word(base + 4 * j, img, img_end);
    }
  }
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  // ---- left / right image border: one apron pixel.  The row may end 1-3 bytes into the NEXT
  // tile (then this tile still needs the border pixel behind it) or 1-3 bytes into THIS one
  // (then its source pixel lies left of c0: that is why the load starts at c0 - 8). ----
  const bool left_edge = (c0 == 0);
  const bool right_edge = (s.rowbytes <= c0 + FA_TWB + 3);
  if (left_edge || right_edge) {
    if (tid < rows) {
      uint8_t* srow = smem + tid * FA_SROW + FA_APRON - c0 + row_phase(tid);
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

  // ---- filter: a thread owns KW words of one band and slides down it ------------------
  {
    const int band = tid / TPB, q = tid - band * TPB;
    const int c = c0 + 4 * KW * q;
    const int o0 = band * BAND_H;
    const int nrows = min(BAND_H, s.h - y0 - o0);
    if (c < s.rowbytes && nrows > 0) {
      const uint8_t* sp = smem + FA_APRON + 4 * KW * q + o0 * FA_SROW;
      auto load_row = [&](int tr, typename F::Part& p) {
        const uint8_t* rp = sp + tr * FA_SROW;
        const unsigned sh = 8u * row_phase(o0 + tr);
        uint32_t R[KW + 3];                             // raw aligned words; tile byte t sits `phase` bytes further
        R[0] = *reinterpret_cast<const uint32_t*>(rp - 4);
        if (KW == 4) {
          const uint4 v = *reinterpret_cast<const uint4*>(rp);
          R[1] = v.x; R[2] = v.y; R[3] = v.z; R[4] = v.w;
        } else {
          const uint2 v = *reinterpret_cast<const uint2*>(rp);
          R[1] = v.x; R[2] = v.y;
        }
        const uint2 t = *reinterpret_cast<const uint2*>(rp + 4 * KW);
        R[KW + 1] = t.x; R[KW + 2] = t.y;
        uint32_t w[KW + 2];
#pragma unroll
        for (int i = 0; i < KW + 2; ++i) w[i] = __funnelshift_r(R[i], R[i + 1], sh);   // (shift 0 returns R[i])
        RowLanes<KW> rl;
        rl.split(w);
        F::hpart(rl, p);
      };
      typename F::Part p0, p1, p2;
      load_row(0, p0);
      load_row(1, p1);
      uint8_t* optr = s_out + o0 * FA_TWB + 4 * KW * q;
#pragma unroll
      for (int r = 0; r < BAND_H; ++r) {
        if (r >= nrows) break;
        typename F::Part& pt = (r % 3 == 0) ? p0 : (r % 3 == 1) ? p1 : p2;
        typename F::Part& pm = (r % 3 == 0) ? p1 : (r % 3 == 1) ? p2 : p0;
        typename F::Part& pb = (r % 3 == 0) ? p2 : (r % 3 == 1) ? p0 : p1;
        load_row(r + 2, pb);
        uint32_t o[KW];
        F::vcomb(pt, pm, pb, o);
        if (KW == 4) *reinterpret_cast<uint4*>(optr) = make_uint4(o[0], o[1], o[KW - 2], o[KW - 1]);
        else *reinterpret_cast<uint2*>(optr) = make_uint2(o[0], o[1]);
        optr += FA_TWB;
      }
    }
  }
  __syncthreads();

  // ---- store: aligned words of the destination row segment, byte stores at its two ends ----
  const int out_rows = min(FA_TH, s.h - y0);
  const int len = min(FA_TWB, s.rowbytes - c0);     // bytes of a row that belong to this tile
  uint8_t* out_img = s.out + (uint64_t)img_local * s.elems;
#pragma unroll 1
  for (int r = warp; r < out_rows; r += FA_THREADS / 32) {      // a warp owns a row here as well
    uint8_t* dst = out_img + (uint64_t)(y0 + r) * s.rowbytes + c0;
    const int po = (int)(reinterpret_cast<uintptr_t>(dst) & 3u);
    const int head = min((4 - po) & 3, len);        // bytes in front of the first aligned word
    const int nfull = (len - head) >> 2;            // aligned words that lie completely inside the segment
    const int t0 = head + 4 * nfull, tail = len - t0;
    const uint8_t* srow = s_out + r * FA_TWB;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(srow);
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst + head);
    if (head) {                                     // (nfull > 0 implies head == (4 - po) & 3; sw[m + 1] <= sw[255])
      const unsigned sh = 8u * (unsigned)head;
#pragma unroll
      for (int i = 0; i < FA_TWB / 128; ++i) {
        const int m = lane + 32 * i;
        if (m < nfull) dw[m] = __funnelshift_r(sw[m], sw[m + 1], sh);
      }
    } else {
#pragma unroll
      for (int i = 0; i < FA_TWB / 128; ++i) {
        const int m = lane + 32 * i;
        if (m < nfull) dw[m] = sw[m];
      }
    }
    if (lane < head) dst[lane] = srow[lane];
    if (lane >= 4 && lane - 4 < tail) dst[t0 + lane - 4] = srow[t0 + lane - 4];
  }
}

template <int FILTER>
static cudaError_t launch_tile_any(const ImageSet& s, cudaStream_t st) {
  constexpr int KW = FILTER == FK_MEDIAN3 ? 2 : 4;
  static bool attr_done[64] = {};               // per device: the attribute lives in the context
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(k_filter3_tile_any<FILTER, KW>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FA_SMEM);
    if (e != cudaSuccess) return e;
    attr_done[dev & 63] = true;
  }
  const dim3 grid((s.rowbytes + FA_TWB - 1) / FA_TWB, (s.h + FA_TH - 1) / FA_TH, s.count);
  k_filter3_tile_any<FILTER, KW><<<grid, FA_THREADS, FA_SMEM, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_filter3_any(int filter, const ImageSet& s, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  // B2D_FILTER_BYTES=1 forces the one-thread-per-byte kernel (the cross-check of the two in the tests)
  static const char* bytes_env = knob("B2D_FILTER_BYTES");
  if ((bytes_env && bytes_env[0] == '1') || s.elems >= (1ull << 31)) return launch_filter3_bytes(filter, s, st);
  switch (filter) {
    case FK_GAUS3: return launch_tile_any<FK_GAUS3>(s, st);
    case FK_MEAN3: return launch_tile_any<FK_MEAN3>(s, st);
    case FK_MEDIAN3: return launch_tile_any<FK_MEDIAN3>(s, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2d
