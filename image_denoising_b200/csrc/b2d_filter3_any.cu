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
// Tile geometry: TWB bytes x (32 KiB / TWB) rows - 1024 x 32, 512 x 64 or 256 x 128.  Same shared memory, same 8-row
// (median: 16-row) bands for all three; the launcher takes the width that pads a row least: 1125-byte rows (W = 375, the
// reference's portrait VOC images) fill 5 tiles of 256 to 88 %, but 2 tiles of 1024 only to 55 %.
template <int TWB>
struct FaGeom {
  static constexpr int TH = 32768 / TWB;
  static constexpr int SROW = TWB + 32;
  static constexpr int ROWS = TH + 2;
  static constexpr size_t SMEM = (size_t)ROWS * SROW + (size_t)TH * TWB;
};

// the aligned word at p, bytes outside [lo, hi) read as 0
__device__ __forceinline__ uint32_t guarded_word(const uint8_t* p, const uint8_t* lo, const uint8_t* hi) {
  if (p >= lo && p + 4 <= hi) return __ldg(reinterpret_cast<const uint32_t*>(p));
  uint32_t v = 0u;
#pragma unroll
  for (int i = 0; i < 4; ++i)
    if (p + i >= lo && p + i < hi) v |= (uint32_t)__ldg(p + i) << (8 * i);
  return v;
}

template <int FILTER, int KW, int FA_TWB>
__global__ void __launch_bounds__(FA_THREADS, 3) k_filter3_tile_any(ImageSet s) {
  using F = Filter3W<FILTER, KW>;
  constexpr int FA_TH = FaGeom<FA_TWB>::TH, FA_SROW = FaGeom<FA_TWB>::SROW, FA_ROWS = FaGeom<FA_TWB>::ROWS;
  static_assert(FA_ROWS <= FA_THREADS, "one thread per tile row builds the border pixels");
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

  // ---- load: bytes [c0 - 8, c0 + TWB + 4) of every tile row, copied RAW in aligned 16-byte chunks ----
  auto row_y = [&](int tr) {
    int ys = y0 - 1 + tr;
    if (ys < 0 || ys >= s.h) ys = border_map(ys, s.h, F::BORDER);
    return ys;
  };
  // (address & 15) of the first loaded byte of tile row tr, in 32-bit arithmetic: (base + ys * rowbytes) mod 16
  const unsigned ph_base = (unsigned)((reinterpret_cast<uintptr_t>(img) + (unsigned)c0 - 8u) & 15u);
  const unsigned rb16 = (unsigned)s.rowbytes & 15u;
  auto row_ph16 = [&](int tr) { return (ph_base + ((unsigned)row_y(tr) & 15u) * rb16) & 15u; };
  // Tile byte t of row tr (t = -8 .. TWB + 3) sits at shared offset tr * SROW + 8 + ph16 + t: the readers below
  // split ph16 into a word offset (ph16 & 12, folded into their addresses) and a byte phase (ph16 & 3, a funnel shift).
  // A WARP owns a tile row (rows warp, warp + 8, ...): the aligned 16-byte chunks that cover it go to shared memory
  // with cp.async - every row of the tile in flight at once, nothing waits in registers.  (Round 1 copied 4-byte
  // words: ~300 LDGSTS per tile at ~8 LSU cycles each, as long as the tile's whole HBM time.)
  constexpr int N16 = FA_TWB / 16 + 2;              // chunks per tile row: 15 + TWB + 12 bytes at most
  const int warp = tid >> 5, lane = tid & 31;
  const int span = min(FA_TWB + 12, s.rowbytes - c0 + 12);     // bytes from c0 - 8 that hold row data (+ slack)
#pragma unroll 1
  for (int tr = warp; tr < rows; tr += FA_THREADS / 32) {
    const uint8_t* src = img + (int64_t)row_y(tr) * s.rowbytes + c0 - 8;
    const unsigned ph = (unsigned)(reinterpret_cast<uintptr_t>(src) & 15u);
    const uint8_t* base = src - ph;
    const int n16 = min(N16, (int)(ph + (unsigned)span + 15u) >> 4);
    uint8_t* sdst = smem + tr * FA_SROW;
    if (base >= img && base + 16 * n16 <= img_end) {
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(sdst);
#pragma unroll
      for (int i = 0; i < (N16 + 31) / 32; ++i) {
        const int j = lane + 32 * i;
        if (j < n16)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa + 16 * j), "l"(base + 16 * j) : "memory");
      }
    } else {                                        // first / last rows of a batch: bytes outside it read as 0
      uint32_t* sw = reinterpret_cast<uint32_t*>(sdst);
      for (int j = lane; j < 4 * n16; j += 32) sw[j] = guarded_word(base + 4 * j, img, img_end);
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
      uint8_t* srow = smem + tid * FA_SROW + 8 + row_ph16(tid) - c0;
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
      const uint8_t* sp = smem + 8 + 4 * KW * q + o0 * FA_SROW;
      auto load_row = [&](int tr, typename F::Part& p) {
        const unsigned ph = row_ph16(o0 + tr);
        const uint32_t* rp = reinterpret_cast<const uint32_t*>(sp + tr * FA_SROW + (ph & 12u));   // word offset of the row
        const unsigned sh = 8u * (ph & 3u);
        uint32_t R[KW + 3];                             // raw aligned words; tile byte t sits (ph & 3) bytes further
#pragma unroll
        for (int i = 0; i < KW + 3; ++i) R[i] = rp[i - 1];   // (the word offset rules out vector loads)
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

template <int FILTER, int TWB>
static cudaError_t launch_tile_any_w(const ImageSet& s, cudaStream_t st) {
  constexpr int KW = FILTER == FK_MEDIAN3 ? 2 : 4;
  static bool attr_done[64] = {};               // per device: the attribute lives in the context
  const int dev = current_device() & 63;
  if (!attr_done[dev]) {
    const cudaError_t e = cudaFuncSetAttribute(k_filter3_tile_any<FILTER, KW, TWB>,
                                               cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FaGeom<TWB>::SMEM);
    if (e != cudaSuccess) return e;
    attr_done[dev] = true;
  }
  const dim3 grid((s.rowbytes + TWB - 1) / TWB, (s.h + FaGeom<TWB>::TH - 1) / FaGeom<TWB>::TH, s.count);
  k_filter3_tile_any<FILTER, KW, TWB><<<grid, FA_THREADS, FaGeom<TWB>::SMEM, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

template <int FILTER>
static cudaError_t launch_tile_any(const ImageSet& s, cudaStream_t st) {
  auto padded = [&](int twb) { return (s.rowbytes + twb - 1) / twb * twb; };
  int best = 1024;                              // least padding; ties go to the wider tile (fewer apron words)
  if (padded(512) < padded(best)) best = 512;
  if (padded(256) < padded(best)) best = 256;
  if (best == 1024) return launch_tile_any_w<FILTER, 1024>(s, st);
  if (best == 512) return launch_tile_any_w<FILTER, 512>(s, st);
  return launch_tile_any_w<FILTER, 256>(s, st);
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
