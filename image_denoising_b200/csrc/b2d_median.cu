// b2d_median.cu - cv2.medianBlur(u8, k) for k = 5, 7 (BASELINE config 5 uses
// k = 7; the reference itself only calls k = 3, lib/roi_data_layer/minibatch.py:153,
// which is fused in b2d_fused3.cu).  Per channel median of the k x k window,
// border REPLICATE (OpenCV's medianBlur border for every k).
//
// Method: register "forgetful selection" (Perrot, Domas, Couturier 2014): keep
// ceil(n/2)+1 candidates, repeatedly drop their min and max (neither can be the
// median) and insert the next window element; the last three give the median.
// The network is data independent, so it runs on 2 x u16 lanes per register
// (even / odd bytes of a 4-byte output word; VIMNMX.U16x2 is native on sm_100a,
// u8x4 min/max is not).  INT-pipe bound: report against the FP32/INT roofline.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <cstdlib>

namespace b2d {

constexpr int MD_THREADS = 256;
constexpr int MD_TILE_WB = MD_THREADS * 4;
constexpr int MD_TILE_H = 16;
constexpr int MD_APRON = 16;                       // >= 3 * 3 bytes, multiple of 4
constexpr int MD_SROW = MD_TILE_WB + 2 * MD_APRON;

// compare-exchange of two u16x2 words: VIMNMX.U16x2 min + max, ALU pipe, 2 cycles each: the networks saturate that
// pipe (ncu: pipe_alu 93 %).  Tried: the maximum as HMNMX2 (`max.f16x2`; lanes 0x00vv are positive subnormal halves
// ordered like the integers, results identical) hoping for the FMA pipe - ncu attributes HMNMX2 to the ALU pipe as
// well and the kernel time does not move (10.10 ms either way), so the knob below stays off.
#ifndef B2D_CE_HMAX
#define B2D_CE_HMAX 0
#endif
__device__ __forceinline__ void ce(uint32_t& a, uint32_t& b) {
  const uint32_t lo = __vminu2(a, b);
#if B2D_CE_HMAX
  uint32_t hi;
  asm("max.f16x2 %0, %1, %2;" : "=r"(hi) : "r"(a), "r"(b));
  b = hi;
#else
  b = __vmaxu2(a, b);
#endif
  a = lo;
}

// min of a[0..M) -> a[0], max -> a[M-1]; all elements preserved.  Pair up low / high halves, then
// reduce the lows to their minimum and the highs to their maximum as TREES (same M/2 + M - 2
// compare-exchanges as a serial scan, dependency depth log2 instead of M/2: the networks are
// latency bound at the occupancy their register count allows).
template <int M, int CAP>
__device__ __forceinline__ void minmax(uint32_t (&a)[CAP]) {
  constexpr int half = (M + 1) / 2;
#pragma unroll
  for (int i = 0; i < M / 2; ++i) ce(a[i], a[i + half]);
  if (M & 1) ce(a[half - 1], a[M - 1]);
#pragma unroll
  for (int st = 1; st < half; st *= 2)
#pragma unroll
    for (int i = 0; i + st < half; i += 2 * st) ce(a[i], a[i + st]);                    // lows: a[0 .. half)
  constexpr int nh = M - half;                                                          // highs: a[half .. M)
#pragma unroll
  for (int st = 1; st < nh; st *= 2)
#pragma unroll
    for (int i = 0; i + st < nh; i += 2 * st) ce(a[M - 1 - i - st], a[M - 1 - i]);
}

template <int M, int CAP, int N>
struct Forget {
  // a[0..M) are candidates, `next` indexes the next unseen window element
  template <class Get>
  static __device__ __forceinline__ uint32_t run(uint32_t (&a)[CAP], Get get) {
    if constexpr (M == 3) {
      return __vmaxu2(__vminu2(a[0], a[1]), __vminu2(__vmaxu2(a[0], a[1]), a[2]));
    } else {
      minmax<M, CAP>(a);
      a[0] = get(N - (M - 3));          // elements CAP .. N-1 are inserted in order as M shrinks from CAP to 4
      return Forget<M - 1, CAP, N>::run(a, get);
    }
  }
};

template <int K>
__global__ void __launch_bounds__(MD_THREADS) k_median(ImageSet s) {
  constexpr int R = K / 2, N = K * K, CAP = (N + 1) / 2 + 1;
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * MD_TILE_WB, y0 = blockIdx.y * MD_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int rows = MD_TILE_H + 2 * R;

  // tile + apron, REPLICATE border resolved per byte (pixel clamp keeps the channel)
  for (int it = tid; it < rows * (MD_SROW / 4); it += MD_THREADS) {
    const int tr = it / (MD_SROW / 4), wc = it - tr * (MD_SROW / 4);
    const int y = min(max(y0 - R + tr, 0), s.h - 1);
    const uint8_t* row = img + (uint64_t)y * s.rowbytes;
    const int c = c0 - MD_APRON + 4 * wc;
    uint32_t w;
    if (c >= 0 && c + 4 <= s.rowbytes && ((reinterpret_cast<uintptr_t>(row + c) & 3u) == 0)) {
      w = __ldg(reinterpret_cast<const uint32_t*>(row + c));
    } else {
      uint32_t b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cc = c + k;
        int px = cc >= 0 ? cc / 3 : -((-cc + 2) / 3);
        const int ch = cc - 3 * px;
        px = min(max(px, 0), s.w - 1);
        b[k] = __ldg(row + px * 3 + ch);
      }
      w = pack4(b[0], b[1], b[2], b[3]);
    }
    *reinterpret_cast<uint32_t*>(smem + tr * MD_SROW + 4 * wc) = w;
  }
  __syncthreads();

  const int c = c0 + 4 * tid;
  if (c >= s.rowbytes) return;
  const uint8_t* sp = smem + MD_APRON + 4 * tid;
  const int nrows = min(MD_TILE_H, s.h - y0);
  for (int r = 0; r < nrows; ++r) {
    // window element e = (dy, dx): bytes at row r + dy, byte offset 3 * (dx - R)
    uint32_t win[N];
#pragma unroll
    for (int dy = 0; dy < K; ++dy) {
      const uint8_t* rp = sp + (r + dy) * MD_SROW;
      uint32_t al[2 * R + 2];                       // aligned words covering bytes [-4R, 4R + 4)
#pragma unroll
      for (int q = 0; q < 2 * R + 2; ++q) al[q] = *reinterpret_cast<const uint32_t*>(rp + 4 * (q - R));
#pragma unroll
      for (int dx = 0; dx < K; ++dx) {
        const int off = 3 * (dx - R) + 4 * R;       // byte offset from al[0], >= 0
        const int q = off / 4, sh = off % 4;
        win[dy * K + dx] = sh == 0 ? al[q] : __byte_perm(al[q], al[q + 1], sh == 1 ? 0x4321 : sh == 2 ? 0x5432 : 0x6543);
      }
    }
    uint32_t res[2];
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      auto get = [&](int e) { return half ? __byte_perm(win[e], 0u, 0x4341) : __byte_perm(win[e], 0u, 0x4240); };
      uint32_t a[CAP];
#pragma unroll
      for (int e = 0; e < CAP; ++e) a[e] = get(e);
      res[half] = Forget<CAP, CAP, N>::run(a, get);
    }
    const uint32_t o = __byte_perm(res[0], res[1], 0x6240);
    uint8_t* op = s.out + (uint64_t)img_local * s.elems + (uint64_t)(y0 + r) * s.rowbytes + c;
    if (c + 4 <= s.rowbytes && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
      *reinterpret_cast<uint32_t*>(op) = o;
    } else {
      for (int k = 0; k < 4 && c + k < s.rowbytes; ++k) op[k] = (uint8_t)(o >> (8 * k));
    }
  }
}

// ---------------------------------------------------------------------------
// Two horizontally adjacent outputs at a time (a "shared partial sort").  For 7 x 7 the windows of
// pixels x and x + 1 share 6 columns = 42 values.  A value of the shared set with >= 25 shared
// values above (below) it cannot be the median (rank 25 of 49) of either window, so forgetful
// selection first strips the 17 smallest and 17 largest shared values (start with 26 candidates,
// drop min and max, insert the next value, ...); the median of a window is then the median of
// the 8 survivors and its own private column (15 values).  ~262 compare-exchanges per output
// instead of 471 (5 x 5: 20 shared -> 6 survivors + 5 private, ~79 instead of 126).  A thread owns 4 pixels: lane 0 of a u16x2 word works on the pair
// (x, x + 1), lane 1 on (x + 2, x + 3) of the same channel, i.e. two bytes 6 apart.
constexpr int M7_THREADS = 128;
#ifndef B2D_MEDIAN_BLOCKS_MINB
#define B2D_MEDIAN_BLOCKS_MINB 6
#endif
constexpr int M7_TILE_WB = M7_THREADS * 12;        // 4 pixels per thread
constexpr int M7_TILE_H = 16;
constexpr int M7_APRON = 12;                       // >= 9 bytes to the left, 9 + 6 - ... to the right; multiple of 4
constexpr int M7_SROW = M7_TILE_WB + 2 * M7_APRON;

// strip min and max of a[0..M) while values remain to be inserted; ends with M = 2 CAP - N, then one more strip:
// survivors a[1 .. 2 CAP - N - 2]
template <int M, int CAP, int N>
struct Strip {
  template <class Get>
  static __device__ __forceinline__ void run(uint32_t (&a)[CAP], Get get) {
    minmax<M, CAP>(a);
    if constexpr (M > 2 * CAP - N) {
      a[0] = get(N - (M - (2 * CAP - N)));         // values CAP .. N-1 in order
      Strip<M - 1, CAP, N>::run(a, get);
    }
  }
};

template <int K>
__global__ void __launch_bounds__(M7_THREADS) k_median_pairs(ImageSet s) {
  constexpr int R = K / 2, N = K * K, NS = (K - 1) * K;      // window, shared values of a pair
  constexpr int CAP = (N + 1) / 2 + 1;                        // 26 (K = 7), 14 (K = 5)
  constexpr int KEEP = 2 * CAP - NS - 2;                      // 8, 6 survivors
  constexpr int CAP2 = KEEP + 1, N2 = KEEP + K;               // median of 15 / 11
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * M7_TILE_WB, y0 = blockIdx.y * M7_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int rows = M7_TILE_H + 2 * R;
  static_assert(3 * R <= M7_APRON && 3 * (R + 1) + 2 + 6 <= 11 + M7_APRON, "apron too small");

  for (int it = tid; it < rows * (M7_SROW / 4); it += M7_THREADS) {
    const int tr = it / (M7_SROW / 4), wc = it - tr * (M7_SROW / 4);
    const int y = min(max(y0 - R + tr, 0), s.h - 1);
    const uint8_t* row = img + (uint64_t)y * s.rowbytes;
    const int c = c0 - M7_APRON + 4 * wc;
    uint32_t w;
    if (c >= 0 && c + 4 <= s.rowbytes && ((reinterpret_cast<uintptr_t>(row + c) & 3u) == 0)) {
      w = __ldg(reinterpret_cast<const uint32_t*>(row + c));
    } else {                                        // REPLICATE: clamp the pixel, keep the channel
      uint32_t b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cc = c + k;
        int px = cc >= 0 ? cc / 3 : -((-cc + 2) / 3);
        const int ch = cc - 3 * px;
        px = min(max(px, 0), s.w - 1);
        b[k] = __ldg(row + px * 3 + ch);
      }
      w = pack4(b[0], b[1], b[2], b[3]);
    }
    *reinterpret_cast<uint32_t*>(smem + tr * M7_SROW + 4 * wc) = w;
  }
  __syncthreads();

  const int c = c0 + 12 * tid;
  if (c >= s.rowbytes) return;
  const uint8_t* sp = smem + M7_APRON + 12 * tid;
  const int nrows = min(M7_TILE_H, s.h - y0);
  for (int r = 0; r < nrows; ++r) {
    const uint8_t* rp = sp + r * M7_SROW;
    uint32_t L[3], Rt[3];                           // per channel: (px0, px2) and (px1, px3) medians
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      // value of window row dy, pixel column d (relative to px0 / px2), channel k as (lane0, lane1)
      auto at = [&](int dy, int d) {
        const uint8_t* q = rp + dy * M7_SROW + 3 * d + k;
        return __byte_perm((uint32_t)q[0], (uint32_t)q[6], 0x5410);
      };
      auto shared = [&](int e) { return at(e / (K - 1), e % (K - 1) - (R - 1)); };   // columns -(R-1) .. R
      uint32_t a[CAP];
#pragma unroll
      for (int e = 0; e < CAP; ++e) a[e] = shared(e);
      Strip<CAP, CAP, NS>::run(a, shared);          // survivors a[1 .. KEEP]
      uint32_t bl[CAP2], br[CAP2];
#pragma unroll
      for (int e = 0; e < KEEP; ++e) bl[e] = br[e] = a[1 + e];
      bl[KEEP] = at(0, -R);
      br[KEEP] = at(0, R + 1);
      L[k] = Forget<CAP2, CAP2, N2>::run(bl, [&](int e) { return at(e - KEEP, -R); });
      Rt[k] = Forget<CAP2, CAP2, N2>::run(br, [&](int e) { return at(e - KEEP, R + 1); });
    }
    // bytes: px0 = (L0.lo, L1.lo, L2.lo), px1 = (R*.lo), px2 = (L*.hi), px3 = (R*.hi)
    const uint32_t w0 = __byte_perm(__byte_perm(L[0], L[1], 0x0040), __byte_perm(L[2], Rt[0], 0x0040), 0x5410);
    const uint32_t w1 = __byte_perm(__byte_perm(Rt[1], Rt[2], 0x0040), __byte_perm(L[0], L[1], 0x0062), 0x5410);
    const uint32_t w2 = __byte_perm(__byte_perm(L[2], Rt[0], 0x0062), __byte_perm(Rt[1], Rt[2], 0x0062), 0x5410);
    uint8_t* op = s.out + (uint64_t)img_local * s.elems + (uint64_t)(y0 + r) * s.rowbytes + c;
    if (c + 12 <= s.rowbytes && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
      uint32_t* ow = reinterpret_cast<uint32_t*>(op);
      ow[0] = w0; ow[1] = w1; ow[2] = w2;
    } else {
      const uint32_t w[3] = {w0, w1, w2};
      for (int b = 0; b < 12 && c + b < s.rowbytes; ++b) op[b] = (uint8_t)(w[b >> 2] >> (8 * (b & 3)));
    }
  }
}

// ---------------------------------------------------------------------------
// 2 x 2 output blocks, two levels of sharing.  The four windows of pixels (x, y) .. (x+1, y+1)
// share (k-1)^2 values (7x7: 36).  Level 1 strips the 11 smallest / 11 largest of them (each has
// >= 25 shared values on one side: it is on that side of all four medians) -> 14 survivors, and
// every window is left with rank 14 of 27.  Level 2, per row of the block: the two windows share
// the survivors + 6 values of their outer row (20); strip 6 + 6 -> 8 survivors, rank 8 of 15.
// Level 3: median of survivors + the private column (15).  (324 + 2 x 100 + 4 x 50) / 4 = 181
// compare-exchanges per output (5x5: 16 -> 10, 14 -> 6, median of 11).  Lanes as in the pair kernel.
// one channel of one 2 x 2 block (x 2 lanes): q points at window row 0, pixel column 0 of the channel.
// Kept out of line so that the three channels do not interleave and triple the register pressure.
template <int K>
__device__ __noinline__ void median_block_channel(const uint8_t* q0, uint32_t (&out)[4]) {
  constexpr int R = K / 2, N = K * K, HALF = (N + 1) / 2;
  constexpr int C1 = (K - 1) * (K - 1), CAP1 = HALF + 1, KEEP1 = 2 * CAP1 - C1 - 2;       // 36, 26, 14
  constexpr int RANK1 = HALF - (C1 - HALF);                                                // 14
  constexpr int S2 = KEEP1 + (K - 1), CAP2 = RANK1 + 1, KEEP2 = 2 * CAP2 - S2 - 2;         // 20, 15, 8
  constexpr int CAP3 = KEEP2 + 1, N3 = KEEP2 + K;                                          // 9, 15
  static_assert(CAP2 == KEEP1 + 1 && 2 * (RANK1 - (S2 - RANK1)) - 1 == N3, "level sizes");
  auto at = [&](int dy, int d) {
    const uint8_t* q = q0 + dy * M7_SROW + 3 * d;
    return __byte_perm((uint32_t)q[0], (uint32_t)q[6], 0x5410);
  };
  auto common = [&](int e) { return at(1 + e / (K - 1), e % (K - 1) - (R - 1)); };       // rows 1 .. K-1, cols -(R-1) .. R
  uint32_t a[CAP1];
#pragma unroll
  for (int e = 0; e < CAP1; ++e) a[e] = common(e);
  Strip<CAP1, CAP1, C1>::run(a, common);            // survivors a[1 .. KEEP1]
#pragma unroll
  for (int br = 0; br < 2; ++br) {                  // top pair: outer row dy = 0; bottom pair: dy = K
    const int orow = br ? K : 0, first = br ? 1 : 0;              // private columns span dy = first .. first + K - 1
    auto outer = [&](int e) { return at(orow, e - KEEP1 - (R - 1)); };                     // e = KEEP1 .. S2-1
    uint32_t b[CAP2];
#pragma unroll
    for (int e = 0; e < KEEP1; ++e) b[e] = a[1 + e];
    b[KEEP1] = outer(KEEP1);
    Strip<CAP2, CAP2, S2>::run(b, outer);           // survivors b[1 .. KEEP2]
    uint32_t cl[CAP3], cr[CAP3];
#pragma unroll
    for (int e = 0; e < KEEP2; ++e) cl[e] = cr[e] = b[1 + e];
    cl[KEEP2] = at(first, -R);
    cr[KEEP2] = at(first, R + 1);
    out[2 * br] = Forget<CAP3, CAP3, N3>::run(cl, [&](int e) { return at(first + e - KEEP2, -R); });
    out[2 * br + 1] = Forget<CAP3, CAP3, N3>::run(cr, [&](int e) { return at(first + e - KEEP2, R + 1); });
  }
}

template <int K>
__global__ void __launch_bounds__(M7_THREADS, B2D_MEDIAN_BLOCKS_MINB) k_median_blocks(ImageSet s) {
  constexpr int R = K / 2;
  static_assert(3 * R <= M7_APRON && 3 * (R + 1) + 2 + 6 <= 11 + M7_APRON, "apron too small");
  extern __shared__ __align__(16) uint8_t smem[];
  const int tid = threadIdx.x;
  const int c0 = blockIdx.x * M7_TILE_WB, y0 = blockIdx.y * M7_TILE_H;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  const int rows = M7_TILE_H + 2 * R;

  for (int it = tid; it < rows * (M7_SROW / 4); it += M7_THREADS) {
    const int tr = it / (M7_SROW / 4), wc = it - tr * (M7_SROW / 4);
    const int y = min(max(y0 - R + tr, 0), s.h - 1);
    const uint8_t* row = img + (uint64_t)y * s.rowbytes;
    const int c = c0 - M7_APRON + 4 * wc;
    uint32_t w;
    if (c >= 0 && c + 4 <= s.rowbytes && ((reinterpret_cast<uintptr_t>(row + c) & 3u) == 0)) {
      w = __ldg(reinterpret_cast<const uint32_t*>(row + c));
    } else {                                        // REPLICATE: clamp the pixel, keep the channel
      uint32_t b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int cc = c + k;
        int px = cc >= 0 ? cc / 3 : -((-cc + 2) / 3);
        const int ch = cc - 3 * px;
        px = min(max(px, 0), s.w - 1);
        b[k] = __ldg(row + px * 3 + ch);
      }
      w = pack4(b[0], b[1], b[2], b[3]);
    }
    *reinterpret_cast<uint32_t*>(smem + tr * M7_SROW + 4 * wc) = w;
  }
  __syncthreads();

  const int c = c0 + 12 * tid;
  if (c >= s.rowbytes) return;
  const uint8_t* sp = smem + M7_APRON + 12 * tid;
  const int nrows = min(M7_TILE_H, s.h - y0);
  for (int r = 0; r < nrows; r += 2) {
    const uint8_t* rp = sp + r * M7_SROW;           // window rows dy = 0 .. K (tile rows r .. r + K)
    uint32_t res[2][2][3];                          // [block row][left / right][channel] as (lane0, lane1)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      uint32_t o4[4];
      median_block_channel<K>(rp + k, o4);
      res[0][0][k] = o4[0]; res[0][1][k] = o4[1]; res[1][0][k] = o4[2]; res[1][1][k] = o4[3];
    }
#pragma unroll
    for (int br = 0; br < 2; ++br) {
      if (r + br >= nrows) break;
      const uint32_t* L = res[br][0];
      const uint32_t* Rt = res[br][1];
      const uint32_t w0 = __byte_perm(__byte_perm(L[0], L[1], 0x0040), __byte_perm(L[2], Rt[0], 0x0040), 0x5410);
      const uint32_t w1 = __byte_perm(__byte_perm(Rt[1], Rt[2], 0x0040), __byte_perm(L[0], L[1], 0x0062), 0x5410);
      const uint32_t w2 = __byte_perm(__byte_perm(L[2], Rt[0], 0x0062), __byte_perm(Rt[1], Rt[2], 0x0062), 0x5410);
      uint8_t* op = s.out + (uint64_t)img_local * s.elems + (uint64_t)(y0 + r + br) * s.rowbytes + c;
      if (c + 12 <= s.rowbytes && (reinterpret_cast<uintptr_t>(op) & 3u) == 0) {
        uint32_t* ow = reinterpret_cast<uint32_t*>(op);
        ow[0] = w0; ow[1] = w1; ow[2] = w2;
      } else {
        const uint32_t w[3] = {w0, w1, w2};
        for (int b = 0; b < 12 && c + b < s.rowbytes; ++b) op[b] = (uint8_t)(w[b >> 2] >> (8 * (b & 3)));
      }
    }
  }
}

template <int K>
static cudaError_t launch_median_blocks(const ImageSet& s, cudaStream_t st) {
  const dim3 grid((s.rowbytes + M7_TILE_WB - 1) / M7_TILE_WB, (s.h + M7_TILE_H - 1) / M7_TILE_H, s.count);
  const size_t smem = (size_t)(M7_TILE_H + 2 * (K / 2)) * M7_SROW;
  k_median_blocks<K><<<grid, M7_THREADS, smem, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

template <int K>
static cudaError_t launch_median_pairs(const ImageSet& s, cudaStream_t st) {
  const dim3 grid((s.rowbytes + M7_TILE_WB - 1) / M7_TILE_WB, (s.h + M7_TILE_H - 1) / M7_TILE_H, s.count);
  const size_t smem = (size_t)(M7_TILE_H + 2 * (K / 2)) * M7_SROW;
  k_median_pairs<K><<<grid, M7_THREADS, smem, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

template <int K>
static cudaError_t launch_median_t(const ImageSet& s, cudaStream_t st) {
  const dim3 grid((s.rowbytes + MD_TILE_WB - 1) / MD_TILE_WB, (s.h + MD_TILE_H - 1) / MD_TILE_H, s.count);
  const size_t smem = (size_t)(MD_TILE_H + 2 * (K / 2)) * MD_SROW;
  k_median<K><<<grid, MD_THREADS, smem, st>>>(s);
  count_launch();
  return cudaGetLastError();
}

cudaError_t launch_median7_sep(const ImageSet& s, cudaStream_t st);     // b2d_median7.cu
cudaError_t launch_median5_sep(const ImageSet& s, cudaStream_t st);

cudaError_t launch_median_k(const ImageSet& s, int k, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  // B2D_MEDIAN_MODE = 1: one output at a time, 2: horizontal pairs, 3: 2 x 2 blocks by forgetful selection; default:
  // 2 x 2 blocks by merging sorted rows (b2d_median7.cu)  (measurement knob)
  static const int mode = [] { const char* e = knob("B2D_MEDIAN_MODE"); return e ? atoi(e) : 0; }();
  if (k == 7 && mode == 0) return launch_median7_sep(s, st);
  if (k == 5 && mode == 0) return launch_median5_sep(s, st);
  if (k == 5) return mode == 1 ? launch_median_t<5>(s, st) : mode == 2 ? launch_median_pairs<5>(s, st) : launch_median_blocks<5>(s, st);
  if (k == 7) return mode == 1 ? launch_median_t<7>(s, st) : mode == 2 ? launch_median_pairs<7>(s, st) : launch_median_blocks<7>(s, st);
  return cudaErrorInvalidValue;
}

}  // namespace b2d
