// b2d_median7.cu - cv2.medianBlur(u8, 7) (BASELINE config 5) and (u8, 5), SEPARABLE versions: sorted rows are merged, not
// re-sorted.  The 7 x 7 kernel is described here, the 5 x 5 one (same scheme, smaller lists) further down.
//
// The selection-network kernels of b2d_median.cu ("forgetful selection", 2 x 2 output blocks) start every block from
// 36 unsorted values: 181 compare-exchanges = 362 min / max instructions per output byte, and the ALU pipe that runs them is
// 93 % busy.  Here a thread keeps SORTED lists alive while it marches down the image, so most of that work is merging:
//
//   block   : outputs (x, y), (x+1, y), (x, y+1), (x+1, y+1) of one channel.  Their four 7 x 7 windows share the 6 x 6
//             CORE (rows y-2 .. y+3, columns x-2 .. x+3); the left / right outputs also share the OUTER COLUMN x-3 / x+4
//             (core rows), and every output has its own PRIVATE ROW (y-3 or y+4 over its 7 columns).
//   level 1 : each core row is sorted once (6 values, 12 compare-exchanges), two rows are merged to a 12-list, and the
//             12-lists of the row pairs (y-2, y-1), (y, y+1), (y+2, y+3) to the middle 14 of the 36 core values: the 11
//             smallest / largest have >= 25 core values on one side and are on that side of all four medians.  Marching
//             down by 2 rows keeps two of the three 12-lists; the merge of two of them (24-list) serves two blocks.
//   level 2 : survivors + the sorted outer column (20 values): the 6 smallest / largest go the same way -> 8 survivors.
//   level 3 : survivors + the sorted private row (its 6 core columns, reused as the next block's new core row) = 14
//             values, whose two middle ones m6 <= m7 bracket the median of the 15: min(max(corner, m6), m7).
// 463 min / max operations per block = 116 per output byte instead of 362 (443 instructions with the 3-input sorts
// below; + the operand fetches).  All programs are
// generated and PROVEN (0-1 principle over every sorted input) by scripts/gen_median7_networks.py
// (b2d_median7_nets.inc); data independent, so they run on the two u16 lanes of a register: lane 0 = the block at byte
// column c, lane 1 = the block 6 bytes (2 pixels) to the right.  Border: REPLICATE, as OpenCV's medianBlur.
#include "b2d_kernels.cuh"

namespace b2d {

#define VMIN(a, b) __vminu2(a, b)
#define VMAX(a, b) __vmaxu2(a, b)
#include "b2d_median7_nets.inc"

// Sorting 6 byte values (per u16 lane): two 3-sorts by the 3-input minimum / maximum and "middle = sum - min - max" (byte
// sums cannot carry between the lanes; IADD3 issues at twice the rate of VIMNMX), then a proven 3 + 3 merge: 20
// instructions and 36 ALU-pipe cycles instead of the 12 compare-exchanges' 24 and 48.
__device__ __forceinline__ void sort3_lanes(uint32_t a, uint32_t b, uint32_t c, uint32_t& lo, uint32_t& mid, uint32_t& hi) {
  lo = __vimin3_u16x2(a, b, c);
  hi = __vimax3_u16x2(a, b, c);
  mid = (a + b + c) - (lo + hi);
}
__device__ __forceinline__ void sort6_lanes(const uint32_t (&v)[6], uint32_t (&o)[6]) {
  uint32_t t[6];
  sort3_lanes(v[0], v[1], v[2], t[0], t[1], t[2]);
  sort3_lanes(v[3], v[4], v[5], t[3], t[4], t[5]);
  B2D_NET_MERGE33(t, o);
}

constexpr int S7_XU = 32;                          // 12-byte units (4 pixels: the two lanes' blocks) per tile row
constexpr int S7_SEGS = 2;                         // vertical segments per tile, one thread each per unit and channel
constexpr int S7_SEG_H = 24;                       // rows per segment = 12 steps of the march
constexpr int S7_THREADS = S7_XU * 3 * S7_SEGS;    // 192
constexpr int S7_TILE_WB = S7_XU * 12;             // 384 bytes
constexpr int S7_TILE_H = S7_SEGS * S7_SEG_H;      // 48 rows
constexpr int S7_APRON = 12;                       // >= 9 bytes to the left, 9 beyond the unit to the right; multiple of 4
constexpr int S7_SROW = S7_TILE_WB + 2 * S7_APRON;
constexpr int S7_ROWS = S7_TILE_H + 7;             // rows y0 - 3 .. y0 + TILE_H + 3 (the last block reads row y + 4)
constexpr size_t S7_TILE_BYTES = (size_t)S7_ROWS * S7_SROW;
// The lists a step hands to later steps wait in shared memory ([36][threads] words), not in registers: with them the
// levels 2 / 3 of a step need ~100 registers, and 4 CTAs / SM (80 registers) run 8 % faster than 3.
//   words 0 .. 23 : Q, the 24-list an even step merges for itself and the odd step after it; once that step has read it,
//                   words 0 .. 11 take the odd step's new 12-list (the next even step's second list)
//   words 24 .. 35: the even step's new 12-list (the next even step's first list)
constexpr int S7_PARK = 36;
constexpr size_t S7_SMEM = S7_TILE_BYTES + (size_t)S7_PARK * S7_THREADS * 4;

// tile + apron -> shared memory, REPLICATE at the image border (tile row tr = image row ytop + tr); the caller syncs
__device__ __forceinline__ void load_tile(uint8_t* smem, const ImageSet& s, const uint8_t* img, int c0, int ytop, int rows) {
  const int tid = threadIdx.x;
  for (int it = tid; it < rows * (S7_SROW / 4); it += S7_THREADS) {
    const int tr = it / (S7_SROW / 4), wc = it - tr * (S7_SROW / 4);
    const int y = min(max(ytop + tr, 0), s.h - 1);
    const uint8_t* row = img + (uint64_t)y * s.rowbytes;
    const int c = c0 - S7_APRON + 4 * wc;
    uint8_t* dst = smem + tr * S7_SROW + 4 * wc;
    if (c >= 0 && c + 4 <= s.rowbytes && ((reinterpret_cast<uintptr_t>(row + c) & 3u) == 0)) {
      // straight to shared memory: every word of the tile in flight at once, nothing waits in registers
      asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(row + c)
                   : "memory");
    } else {                                        // clamp the pixel, keep the channel
      uint32_t b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cc = c + k;
        int px = cc >= 0 ? cc / 3 : -((-cc + 2) / 3);
        const int ch = cc - 3 * px;
        px = min(max(px, 0), s.w - 1);
        b[k] = __ldg(row + px * 3 + ch);
      }
      *reinterpret_cast<uint32_t*>(dst) = pack4(b[0], b[1], b[2], b[3]);
    }
  }
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

__global__ void __launch_bounds__(S7_THREADS, 4) k_median7_sep(ImageSet s) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * S7_TILE_WB, y0 = blockIdx.y * S7_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;

  load_tile(smem, s, img, c0, y0 - 3, S7_ROWS);
  __syncthreads();

  // thread = (segment, channel, unit): a warp shares segment and channel, its lanes are 32 units (12-byte stride: the
  // byte reads below hit 32 different banks)
  const int tx = tid & 31, ch = (tid >> 5) % 3, seg = tid / 96;
  const int c = c0 + 12 * tx + ch;                  // byte column of lane 0's left outputs; lane 1: c + 6
  if (12 * tx + c0 >= s.rowbytes) return;
  const int ys = y0 + seg * S7_SEG_H;               // first output row of the segment
  if (ys >= s.h) return;
  // value (lane 0, lane 1) at window row r (tile row of image row ys - 3 is seg * SEG_H) and column offset d pixels
  const uint8_t* q0 = smem + (seg * S7_SEG_H) * S7_SROW + S7_APRON + 12 * tx + ch;
  auto at = [&](int r, int d) {                     // r: image row ys - 3 + r
    const uint8_t* q = q0 + r * S7_SROW + 3 * d;
    uint32_t v;                                     // lane 1 * 2^16 + lane 0 as an IMAD: the FMA pipe idles, the ALU pipe
    asm("mad.lo.u32 %0, %1, 65536, %2;" : "=r"(v) : "r"((uint32_t)q[6]), "r"((uint32_t)q[0]));   // is the one that is full
    return v;
  };
  auto sorted_row = [&](int r, uint32_t (&o)[6]) {  // the 6 core columns d = -2 .. 3 of a row, sorted
    const uint32_t v[6] = {at(r, -2), at(r, -1), at(r, 0), at(r, 1), at(r, 2), at(r, 3)};
    sort6_lanes(v, o);
  };
  auto pair_list = [&](const uint32_t (&a)[6], const uint32_t (&b)[6], uint32_t (&o)[12]) {
    uint32_t in[12];
#pragma unroll
    for (int i = 0; i < 6; ++i) { in[i] = a[i]; in[6 + i] = b[i]; }
    B2D_NET_PAIR(in, o);
  };

  // warm-up: L0 = rows (ys - 2, ys - 1) = window rows 1, 2 -> words 24..35; L1 = rows (ys, ys + 1) = window rows 3, 4 ->
  // words 0..11; and the sorted row ys + 2 (window row 5), which the first step pairs with row ys + 3
  uint32_t nxt[6];
  uint32_t* const sq = reinterpret_cast<uint32_t*>(smem + S7_TILE_BYTES) + tid;   // this thread's column of parked words
  {
    uint32_t r1[6], r2[6], L[12];
    sorted_row(1, r1); sorted_row(2, r2); pair_list(r1, r2, L);
#pragma unroll
    for (int i = 0; i < 12; ++i) sq[(24 + i) * S7_THREADS] = L[i];
    sorted_row(3, r1); sorted_row(4, r2); pair_list(r1, r2, L);
#pragma unroll
    for (int i = 0; i < 12; ++i) sq[i * S7_THREADS] = L[i];
    sorted_row(5, nxt);
  }
  uint8_t* out = s.out + (uint64_t)img_local * s.elems + (uint64_t)ys * s.rowbytes + c;
  const int nrows = min(S7_SEG_H, s.h - ys);
  const bool v0 = c < s.rowbytes, v1 = c + 3 < s.rowbytes, v2 = c + 6 < s.rowbytes, v3 = c + 9 < s.rowbytes;

  // one step = one block (x 2 lanes): rows (ys + 2k, ys + 2k + 1); window rows wr .. wr + 7 with wr = 2k
  auto step = [&](int k, bool even) {
    const int wr = 2 * k;
    // level 1
    uint32_t Ln[12], core[14];
    {
      uint32_t r6[6];
      sorted_row(wr + 6, r6);                       // row y + 3
      pair_list(nxt, r6, Ln);                       // rows (y + 2, y + 3)
    }
    if (even) {
      // lists L_k (words 24..35), L_{k+1} (words 0..11) and the new L_{k+2}; Q = merge(L_{k+1}, L_{k+2}) serves this
      // block and the next one
      uint32_t in[24], Q[24];
#pragma unroll
      for (int i = 0; i < 12; ++i) { in[i] = sq[i * S7_THREADS]; in[12 + i] = Ln[i]; }
      B2D_NET_QUAD(in, Q);
#pragma unroll
      for (int i = 0; i < 24; ++i) sq[i * S7_THREADS] = Q[i];
      uint32_t in2[36];
#pragma unroll
      for (int i = 0; i < 12; ++i) in2[i] = sq[(24 + i) * S7_THREADS];
#pragma unroll
      for (int i = 0; i < 24; ++i) in2[12 + i] = Q[i];
      B2D_NET_CORE(in2, core);
#pragma unroll
      for (int i = 0; i < 12; ++i) sq[(24 + i) * S7_THREADS] = Ln[i];   // L_{k+2}: first list of the next even step
    } else {
      // Q = merge(L_k, L_{k+1}) is parked; the new list completes the core
      uint32_t in2[36];
#pragma unroll
      for (int i = 0; i < 12; ++i) in2[i] = Ln[i];
#pragma unroll
      for (int i = 0; i < 24; ++i) in2[12 + i] = sq[i * S7_THREADS];
      B2D_NET_CORE(in2, core);
#pragma unroll
      for (int i = 0; i < 12; ++i) sq[i * S7_THREADS] = Ln[i];          // L_{k+2}: second list of the next even step
    }
    // level 2: outer columns d = -3 (left outputs) and d = 4 (right outputs) over the core rows wr + 1 .. wr + 6
    uint32_t sl[8], sr[8];
    {
      const uint32_t cl[6] = {at(wr + 1, -3), at(wr + 2, -3), at(wr + 3, -3), at(wr + 4, -3), at(wr + 5, -3), at(wr + 6, -3)};
      uint32_t in[20];
#pragma unroll
      for (int i = 0; i < 14; ++i) in[i] = core[i];
      uint32_t so[6];
      sort6_lanes(cl, so);
#pragma unroll
      for (int i = 0; i < 6; ++i) in[14 + i] = so[i];
      B2D_NET_ROW(in, sl);
      const uint32_t cr[6] = {at(wr + 1, 4), at(wr + 2, 4), at(wr + 3, 4), at(wr + 4, 4), at(wr + 5, 4), at(wr + 6, 4)};
      sort6_lanes(cr, so);
#pragma unroll
      for (int i = 0; i < 6; ++i) in[14 + i] = so[i];
      B2D_NET_ROW(in, sr);
    }
    // level 3: private rows wr (top outputs) and wr + 7 (bottom outputs)
    uint32_t res[4];                                // (x, y), (x+1, y), (x, y+1), (x+1, y+1)
    auto finish = [&](const uint32_t (&row6)[6], uint32_t corner, const uint32_t (&surv)[8]) -> uint32_t {
      uint32_t in14[14], m[2];
#pragma unroll
      for (int i = 0; i < 8; ++i) in14[i] = surv[i];
#pragma unroll
      for (int i = 0; i < 6; ++i) in14[8 + i] = row6[i];
      B2D_NET_MID2(in14, m);                        // the two middle values of the 14
      return VMIN(VMAX(corner, m[0]), m[1]);        // ... and the 15th value between, below or above them
    };
    {
      uint32_t t6[6];
      sorted_row(wr, t6);
      res[0] = finish(t6, at(wr, -3), sl);
      res[1] = finish(t6, at(wr, 4), sr);
      sorted_row(wr + 7, nxt);                      // row y + 4: the next block's row y + 2
      res[2] = finish(nxt, at(wr + 7, -3), sl);
      res[3] = finish(nxt, at(wr + 7, 4), sr);
    }
    // stores: lane 0 -> bytes c (left) and c + 3 (right), lane 1 -> c + 6 and c + 9
#pragma unroll
    for (int br = 0; br < 2; ++br) {
      if (2 * k + br >= nrows) break;
      uint8_t* op = out + (size_t)(2 * k + br) * s.rowbytes;
      const uint32_t l = res[2 * br], r = res[2 * br + 1];
      if (v0) op[0] = (uint8_t)l;
      if (v1) op[3] = (uint8_t)r;
      if (v2) op[6] = (uint8_t)(l >> 16);
      if (v3) op[9] = (uint8_t)(r >> 16);
    }
  };
  const int nsteps = (nrows + 1) / 2;
  for (int k = 0; k < nsteps; k += 2) {
    step(k, true);
    if (k + 1 < nsteps) step(k + 1, false);
  }
}

// ---------------------------------------------------------------------------------------------------------
// The same scheme for cv2.medianBlur(u8, 5): core 4 x 4 (rows y-1 .. y+2, columns x-1 .. x+2) = two row pairs, so a step
// merges the pair it kept with the new one (16 -> middle 10), adds the sorted outer column (14 -> 6) and the sorted
// private row (10 -> two middle values + corner).  232 min / max per block = 58 per output byte (forgetful selection: 158).
// Tile row tr = image row y0 - 2 + tr; the state (one 8-list, one sorted row) stays in registers.
constexpr int S5_ROWS = S7_TILE_H + 5;

__global__ void __launch_bounds__(S7_THREADS, 5) k_median5_sep(ImageSet s) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * S7_TILE_WB, y0 = blockIdx.y * S7_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  load_tile(smem, s, img, c0, y0 - 2, S5_ROWS);
  __syncthreads();
  const int tx = tid & 31, ch = (tid >> 5) % 3, seg = tid / 96;
  const int c = c0 + 12 * tx + ch;
  if (12 * tx + c0 >= s.rowbytes) return;
  const int ys = y0 + seg * S7_SEG_H;
  if (ys >= s.h) return;
  const uint8_t* q0 = smem + (seg * S7_SEG_H) * S7_SROW + S7_APRON + 12 * tx + ch;
  auto at = [&](int r, int d) {                     // window row r = image row ys - 2 + r, column offset d pixels
    const uint8_t* q = q0 + r * S7_SROW + 3 * d;
    uint32_t v;
    asm("mad.lo.u32 %0, %1, 65536, %2;" : "=r"(v) : "r"((uint32_t)q[6]), "r"((uint32_t)q[0]));
    return v;
  };
  auto sorted_row = [&](int r, uint32_t (&o)[4]) {  // the 4 core columns d = -1 .. 2 of a row, sorted
    const uint32_t v[4] = {at(r, -1), at(r, 0), at(r, 1), at(r, 2)};
    B2D_NET_SORT4(v, o);
  };
  auto pair_list = [&](const uint32_t (&a)[4], const uint32_t (&b)[4], uint32_t (&o)[8]) {
    uint32_t in[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) { in[i] = a[i]; in[4 + i] = b[i]; }
    B2D_NET_PAIR5(in, o);
  };
  uint32_t Lp[8], nxt[4];                           // rows (y-1, y) merged; row y+1 sorted
  {
    uint32_t r1[4], r2[4];
    sorted_row(1, r1); sorted_row(2, r2); pair_list(r1, r2, Lp);
    sorted_row(3, nxt);
  }
  uint8_t* out = s.out + (uint64_t)img_local * s.elems + (uint64_t)ys * s.rowbytes + c;
  const int nrows = min(S7_SEG_H, s.h - ys);
  const bool v0 = c < s.rowbytes, v1 = c + 3 < s.rowbytes, v2 = c + 6 < s.rowbytes, v3 = c + 9 < s.rowbytes;
  const int nsteps = (nrows + 1) / 2;
  for (int k = 0; k < nsteps; ++k) {
    const int wr = 2 * k;                           // window rows wr .. wr + 5 = image rows y-2 .. y+3
    uint32_t Ln[8], core[10];
    {
      uint32_t r4[4];
      sorted_row(wr + 4, r4);                       // row y + 2
      pair_list(nxt, r4, Ln);                       // rows (y + 1, y + 2)
      uint32_t in[16];
#pragma unroll
      for (int i = 0; i < 8; ++i) { in[i] = Lp[i]; in[8 + i] = Ln[i]; }
      B2D_NET_CORE5(in, core);
#pragma unroll
      for (int i = 0; i < 8; ++i) Lp[i] = Ln[i];
    }
    uint32_t sl[6], sr[6];
    {
      uint32_t in[14], so[4];
#pragma unroll
      for (int i = 0; i < 10; ++i) in[i] = core[i];
      const uint32_t cl[4] = {at(wr + 1, -2), at(wr + 2, -2), at(wr + 3, -2), at(wr + 4, -2)};
      B2D_NET_SORT4(cl, so);
#pragma unroll
      for (int i = 0; i < 4; ++i) in[10 + i] = so[i];
      B2D_NET_ROW5(in, sl);
      const uint32_t cr[4] = {at(wr + 1, 3), at(wr + 2, 3), at(wr + 3, 3), at(wr + 4, 3)};
      B2D_NET_SORT4(cr, so);
#pragma unroll
      for (int i = 0; i < 4; ++i) in[10 + i] = so[i];
      B2D_NET_ROW5(in, sr);
    }
    auto finish = [&](const uint32_t (&row4)[4], uint32_t corner, const uint32_t (&surv)[6]) -> uint32_t {
      uint32_t in10[10], m[2];
#pragma unroll
      for (int i = 0; i < 6; ++i) in10[i] = surv[i];
#pragma unroll
      for (int i = 0; i < 4; ++i) in10[6 + i] = row4[i];
      B2D_NET_MID5(in10, m);
      return VMIN(VMAX(corner, m[0]), m[1]);
    };
    uint32_t res[4];
    {
      uint32_t t4[4];
      sorted_row(wr, t4);
      res[0] = finish(t4, at(wr, -2), sl);
      res[1] = finish(t4, at(wr, 3), sr);
      sorted_row(wr + 5, nxt);                      // row y + 3: the next block's row y + 1
      res[2] = finish(nxt, at(wr + 5, -2), sl);
      res[3] = finish(nxt, at(wr + 5, 3), sr);
    }
#pragma unroll
    for (int br = 0; br < 2; ++br) {
      if (2 * k + br >= nrows) break;
      uint8_t* op = out + (size_t)(2 * k + br) * s.rowbytes;
      const uint32_t l = res[2 * br], r = res[2 * br + 1];
      if (v0) op[0] = (uint8_t)l;
      if (v1) op[3] = (uint8_t)r;
      if (v2) op[6] = (uint8_t)(l >> 16);
      if (v3) op[9] = (uint8_t)(r >> 16);
    }
  }
}

cudaError_t launch_median5_sep(const ImageSet& s, cudaStream_t st) {
  const dim3 grid((s.rowbytes + S7_TILE_WB - 1) / S7_TILE_WB, (s.h + S7_TILE_H - 1) / S7_TILE_H, s.count);
  k_median5_sep<<<grid, S7_THREADS, (size_t)S5_ROWS * S7_SROW, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_median7_sep(const ImageSet& s, cudaStream_t st) {
  static bool attr_done[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!attr_done[dev & 63]) {
    const cudaError_t e = cudaFuncSetAttribute(k_median7_sep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S7_SMEM);
    if (e != cudaSuccess) return e;
    attr_done[dev & 63] = true;
  }
  const dim3 grid((s.rowbytes + S7_TILE_WB - 1) / S7_TILE_WB, (s.h + S7_TILE_H - 1) / S7_TILE_H, s.count);
  k_median7_sep<<<grid, S7_THREADS, S7_SMEM, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b2d
