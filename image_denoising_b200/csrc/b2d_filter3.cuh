// b2d_filter3.cuh - the three 3x3 denoisers of the reference on packed bytes.
//
//   gaus_blur  cv2.GaussianBlur(u8,(3,3),0)  = (sum([1 2 1]x[1 2 1] p) + 8) >> 4, REFLECT_101
//   mean       cv2.blur(u8,(3,3))            = (S + 4) / 9  (== round-half-up(S/9)), REFLECT_101
//   median     cv2.medianBlur(u8,3)          = median of 9, REPLICATE
// (call sites lib/roi_data_layer/minibatch.py:119,136,153).
//
// The image is treated as an H x (3W) byte matrix: a per-channel 3x3 window on
// interleaved BGR is taps at byte offsets -3, 0, +3 and rows -1, 0, +1.  Each
// thread owns one 4-byte column word and slides down the tile: per row it
// forms a "row part" from the left/centre/right words (hpart) and combines
// the parts of three consecutive rows (vcomb).  Arithmetic is on 2 x u16 lanes
// (even bytes / odd bytes of the word) - sm_100a has native u16x2 min/max
// (VIMNMX.U16x2) but no native u8x4 min/max.
#pragma once
#include "b2d_common.cuh"

namespace b2d {

enum FilterKind { FK_NONE = 0, FK_GAUS3, FK_MEAN3, FK_MEDIAN3, FK_COUNT };
enum BorderKind { BORDER_REFLECT101 = 0, BORDER_REPLICATE = 1 };

__device__ __forceinline__ uint32_t lanes_even(uint32_t x) { return __byte_perm(x, 0u, 0x4240); }
__device__ __forceinline__ uint32_t lanes_odd(uint32_t x) { return __byte_perm(x, 0u, 0x4341); }
// low bytes of the four u16 lanes of (E, O) -> packed bytes [E0, O0, E1, O1]
__device__ __forceinline__ uint32_t lanes_merge(uint32_t E, uint32_t O) { return __byte_perm(E, O, 0x6240); }
// bytes b-3..b of the row from words (b-4..b-1), (b..b+3); bytes b+3..b+6 from (b..b+3), (b+4..b+7)
__device__ __forceinline__ uint32_t word_left3(uint32_t wl, uint32_t wc) { return __byte_perm(wl, wc, 0x4321); }
__device__ __forceinline__ uint32_t word_right3(uint32_t wc, uint32_t wr) { return __byte_perm(wc, wr, 0x6543); }

__device__ __forceinline__ uint32_t min2x16(uint32_t a, uint32_t b) { return __vminu2(a, b); }
__device__ __forceinline__ uint32_t max2x16(uint32_t a, uint32_t b) { return __vmaxu2(a, b); }
__device__ __forceinline__ uint32_t med3x16(uint32_t a, uint32_t b, uint32_t c) {
  return max2x16(min2x16(a, b), min2x16(max2x16(a, b), c));
}

template <int KIND>
struct Filter3;

template <>
struct Filter3<FK_GAUS3> {
  static constexpr int BORDER = BORDER_REFLECT101;
  struct Part { uint32_t e, o; };
  static __device__ __forceinline__ Part hpart(uint32_t L, uint32_t C, uint32_t R) {
    Part p;
    p.e = lanes_even(L) + 2u * lanes_even(C) + lanes_even(R);
    p.o = lanes_odd(L) + 2u * lanes_odd(C) + lanes_odd(R);
    return p;
  }
  static __device__ __forceinline__ uint32_t vcomb(const Part& t, const Part& m, const Part& b) {
    const uint32_t e = (t.e + 2u * m.e + b.e + 0x00080008u) >> 4;
    const uint32_t o = (t.o + 2u * m.o + b.o + 0x00080008u) >> 4;
    return lanes_merge(e, o);  // lane values <= 255 after the shift: low byte of each lane
  }
  static __device__ __forceinline__ uint32_t scalar(const uint32_t (&p)[9]) {
    return (p[0] + 2 * p[1] + p[2] + 2 * p[3] + 4 * p[4] + 2 * p[5] + p[6] + 2 * p[7] + p[8] + 8) >> 4;
  }
};

template <>
struct Filter3<FK_MEAN3> {
  static constexpr int BORDER = BORDER_REFLECT101;
  struct Part { uint32_t e, o; };
  static __device__ __forceinline__ Part hpart(uint32_t L, uint32_t C, uint32_t R) {
    Part p;
    p.e = lanes_even(L) + lanes_even(C) + lanes_even(R);
    p.o = lanes_odd(L) + lanes_odd(C) + lanes_odd(R);
    return p;
  }
  // (S + 4) / 9 for S <= 2295 as ((S + 4) * 7282) >> 16: the quotient is byte 2 of the product
  static __device__ __forceinline__ uint32_t div9(uint32_t s) { return s * 7282u + 4u * 7282u; }
  static __device__ __forceinline__ uint32_t vcomb(const Part& t, const Part& m, const Part& b) {
    const uint32_t se = t.e + m.e + b.e, so = t.o + m.o + b.o;
    const uint32_t q0 = div9(se & 0xffffu), q2 = div9(se >> 16);
    const uint32_t q1 = div9(so & 0xffffu), q3 = div9(so >> 16);
    const uint32_t lo = __byte_perm(q0, q1, 0x0062);  // [q0.b2, q1.b2, x, x]
    const uint32_t hi = __byte_perm(q2, q3, 0x0062);
    return __byte_perm(lo, hi, 0x5410);
  }
  static __device__ __forceinline__ uint32_t scalar(const uint32_t (&p)[9]) {
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < 9; ++i) s += p[i];
    return (2 * s + 9) / 18;
  }
};

template <>
struct Filter3<FK_MEDIAN3> {
  static constexpr int BORDER = BORDER_REPLICATE;
  struct Part { uint32_t lo_e, mid_e, hi_e, lo_o, mid_o, hi_o; };
  static __device__ __forceinline__ void sort3(uint32_t a, uint32_t b, uint32_t c, uint32_t& lo,
                                               uint32_t& mid, uint32_t& hi) {
    const uint32_t mn = min2x16(a, b), mx = max2x16(a, b);
    lo = min2x16(mn, c);
    hi = max2x16(mx, c);
    mid = max2x16(mn, min2x16(mx, c));
  }
  static __device__ __forceinline__ Part hpart(uint32_t L, uint32_t C, uint32_t R) {
    Part p;
    sort3(lanes_even(L), lanes_even(C), lanes_even(R), p.lo_e, p.mid_e, p.hi_e);
    sort3(lanes_odd(L), lanes_odd(C), lanes_odd(R), p.lo_o, p.mid_o, p.hi_o);
    return p;
  }
  // median of 9 = med3(max of the row minima, med3 of the row medians, min of the row maxima)
  static __device__ __forceinline__ uint32_t vcomb(const Part& t, const Part& m, const Part& b) {
    const uint32_t e = med3x16(max2x16(max2x16(t.lo_e, m.lo_e), b.lo_e),
                               med3x16(t.mid_e, m.mid_e, b.mid_e),
                               min2x16(min2x16(t.hi_e, m.hi_e), b.hi_e));
    const uint32_t o = med3x16(max2x16(max2x16(t.lo_o, m.lo_o), b.lo_o),
                               med3x16(t.mid_o, m.mid_o, b.mid_o),
                               min2x16(min2x16(t.hi_o, m.hi_o), b.hi_o));
    return lanes_merge(e, o);
  }
  static __device__ __forceinline__ uint32_t scalar(const uint32_t (&p)[9]) {
    uint32_t a[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) a[i] = p[i];
#pragma unroll
    for (int i = 0; i < 5; ++i)
#pragma unroll
      for (int j = i + 1; j < 9; ++j) {
        const uint32_t lo = min(a[i], a[j]), hi = max(a[i], a[j]);
        a[i] = lo;
        a[j] = hi;
      }
    return a[4];
  }
};

// border index maps (OpenCV borderInterpolate) for a 1-pixel apron
__device__ __forceinline__ int border_map(int i, int n, int kind) {
  if (kind == BORDER_REPLICATE) return i < 0 ? 0 : (i >= n ? n - 1 : i);
  if (n == 1) return 0;
  if (i < 0) i = -i;
  if (i >= n) i = 2 * n - 2 - i;
  return i < 0 ? 0 : (i >= n ? n - 1 : i);
}

}  // namespace b2d
