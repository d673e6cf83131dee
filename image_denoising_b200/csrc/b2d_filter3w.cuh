// b2d_filter3w.cuh - the three 3x3 denoisers on KW adjacent 4-byte words per thread.
//
//   gaus_blur  cv2.GaussianBlur(u8,(3,3),0)  = (sum([1 2 1]x[1 2 1] p) + 8) >> 4, REFLECT_101
//   mean       cv2.blur(u8,(3,3))            = (S + 4) / 9, REFLECT_101
//   median     cv2.medianBlur(u8,3)          = median of 9, REPLICATE
// (call sites lib/roi_data_layer/minibatch.py:119,136,153).
//
// The image is an H x (3W) byte matrix; the per-channel window of interleaved BGR
// is taps at byte offsets -3, 0, +3.  A word W_i = (b0 b1 b2 b3) is split once into
// two u16x2 registers E_i = (b0, b2), O_i = (b1, b3).  In that form the shifted taps
// need no byte extraction at all:
//     byte - 3 :  E_left  = O_{i-1}                  O_left  = (E_{i-1}.hi, E_i.lo)
//     byte + 3 :  E_right = (O_i.hi, O_{i+1}.lo)     O_right = E_{i+1}
// i.e. 2 PRMT per word for the taps + 2 (KW + 2) / KW PRMT per word for the split.
// All arithmetic is on the u16x2 lanes (sums <= 4080 never carry between lanes).
#pragma once
#include "b2d_filter3.cuh"

namespace b2d {

__device__ __forceinline__ uint32_t prmt_c(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
// (a & mask) | (b & ~mask) as ONE LOP3 (the compiler splits the two-constant form in two)
template <uint32_t MASK>
__device__ __forceinline__ uint32_t bitsel(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "n"(MASK));
  return d;
}

// row of KW + 2 words (index 0 = word to the left, KW + 1 = word to the right), split in lanes
template <int KW>
struct RowLanes {
  uint32_t e[KW + 2], o[KW + 2];
  __device__ __forceinline__ void split(const uint32_t (&w)[KW + 2]) {
#pragma unroll
    for (int i = 0; i < KW + 2; ++i) {
      e[i] = prmt_c(w[i], 0u, 0x4240);
      o[i] = prmt_c(w[i], 0u, 0x4341);
    }
  }
  // taps of word i (1-based inside the row): left / centre / right, for the E and O lanes
  __device__ __forceinline__ uint32_t el(int i) const { return o[i - 1]; }
  __device__ __forceinline__ uint32_t ol(int i) const { return prmt_c(e[i - 1], e[i], 0x5432); }
  __device__ __forceinline__ uint32_t er(int i) const { return prmt_c(o[i], o[i + 1], 0x5432); }
  __device__ __forceinline__ uint32_t orr(int i) const { return e[i + 1]; }
};

template <int KIND, int KW>
struct Filter3W;

// ---- Gaussian [1 2 1] x [1 2 1] / 16 ---------------------------------------
template <int KW>
struct Filter3W<FK_GAUS3, KW> {
  static constexpr int BORDER = BORDER_REFLECT101;
  struct Part { uint32_t e[KW], o[KW]; };
  static __device__ __forceinline__ void hpart(const RowLanes<KW>& r, Part& p) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      p.e[i] = r.el(i + 1) + 2u * r.e[i + 1] + r.er(i + 1);
      p.o[i] = r.ol(i + 1) + 2u * r.o[i + 1] + r.orr(i + 1);
    }
  }
  static __device__ __forceinline__ void vcomb(const Part& t, const Part& m, const Part& b, uint32_t (&out)[KW]) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      const uint32_t ve = t.e[i] + 2u * m.e[i] + b.e[i] + 0x00080008u;
      const uint32_t vo = t.o[i] + 2u * m.o[i] + b.o[i] + 0x00080008u;
      // result byte = bits 4..11 of each lane: E lanes to bytes 0 / 2, O lanes to bytes 1 / 3
      out[i] = bitsel<0x00ff00ffu>(ve >> 4, vo << 4);
    }
  }
};

// ---- box mean: (S + 4) / 9 --------------------------------------------------
template <int KW>
struct Filter3W<FK_MEAN3, KW> {
  static constexpr int BORDER = BORDER_REFLECT101;
  struct Part { uint32_t e[KW], o[KW]; };
  static __device__ __forceinline__ void hpart(const RowLanes<KW>& r, Part& p) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      p.e[i] = r.el(i + 1) + r.e[i + 1] + r.er(i + 1);
      p.o[i] = r.ol(i + 1) + r.o[i + 1] + r.orr(i + 1);
    }
  }
  // (S + 4) / 9 for S <= 2295 is floor((S + 4) * 7282 / 65536) (exact: 9 * 7282 = 2^16 + 2 and 2299 * 2 < 2^16; the
  // remainder term is at most 8 * 7282 + 2 * 255 = 58766).  The quotients are taken on the FMA pipe, straight from
  // the packed u16x2 sums, with 64-bit multiply-adds (IMAD.HI with a 64-bit addend) - no lane extraction on the ALU
  // pipe, which is the one that limits the fused kernels:
  //   high lane: hi32(S2 * 7282 + K), S2 = lo + hi * 2^16.  The low lane's product adds (lo + 4) * 7282 >> 16 <= 255 to
  //              the high lane's remainder: 58766 + 255 < 65536, the quotient is not disturbed;
  //   low lane : hi32((S2 << 16) * 7282 + K'), the shift drops the high lane.
  // Four clean byte values are then merged with multiply-adds by 2^8 / 2^16 (FMA pipe again).
  // Written in C so that ptxas picks IMAD.HI.U32 (one instruction per quotient); the quotient then passes an empty
  // asm so that the compiler cannot fuse ">> 32, * 256" into 64-bit funnel shifts + masks on the ALU pipe, and the
  // merges are explicit mad.lo (IMAD).
  static __device__ __forceinline__ uint32_t opaque(uint32_t v) {
    asm("" : "+r"(v));
    return v;
  }
  static __device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
  }
  static __device__ __forceinline__ uint32_t q_hi(uint32_t s2) {
    return opaque((uint32_t)(((uint64_t)s2 * 7282u + (uint64_t)(4u * 7282u) * 0x10001ull) >> 32));
  }
  static __device__ __forceinline__ uint32_t q_lo(uint32_t s2) {
    return opaque((uint32_t)(((uint64_t)mad_lo(s2, 65536u, 0u) * 7282u + ((uint64_t)(4u * 7282u) << 16)) >> 32));
  }
  static __device__ __forceinline__ void vcomb(const Part& t, const Part& m, const Part& b, uint32_t (&out)[KW]) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      const uint32_t se = t.e[i] + m.e[i] + b.e[i], so = t.o[i] + m.o[i] + b.o[i];
      const uint32_t lo = mad_lo(q_lo(so), 256u, q_lo(se));          // out bytes 1, 0
      const uint32_t hi = mad_lo(q_hi(so), 256u, q_hi(se));          // out bytes 3, 2
      out[i] = mad_lo(hi, 65536u, lo);
    }
  }
};

// ---- median of 9 -----------------------------------------------------------
// rows are sorted (lo, mid, hi) per column of three taps; median of 9 =
// med3(max of the lows, med3 of the mids, min of the highs).  VIMNMX3.U16x2 gives
// min3 / max3 in one instruction; the middle of three is a ^ b ^ c ^ min ^ max.
__device__ __forceinline__ uint32_t min3x16(uint32_t a, uint32_t b, uint32_t c) { return __vimin3_u16x2(a, b, c); }
__device__ __forceinline__ uint32_t max3x16(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_u16x2(a, b, c); }
// middle of three = their sum minus the smallest and the largest: byte values in u16 lanes never carry or borrow across
// a lane, and IADD3 issues at twice the rate of the LOP3 that the xor form (a ^ b ^ c ^ min ^ max) needs
__device__ __forceinline__ uint32_t mid3x16(uint32_t a, uint32_t b, uint32_t c) {
  return (a + b + c) - (min3x16(a, b, c) + max3x16(a, b, c));
}

template <int KW>
struct Filter3W<FK_MEDIAN3, KW> {
  static constexpr int BORDER = BORDER_REPLICATE;
  struct Part { uint32_t lo_e[KW], mid_e[KW], hi_e[KW], lo_o[KW], mid_o[KW], hi_o[KW]; };
  static __device__ __forceinline__ void sort3(uint32_t a, uint32_t b, uint32_t c, uint32_t& lo, uint32_t& mid,
                                               uint32_t& hi) {
    lo = min3x16(a, b, c);
    hi = max3x16(a, b, c);
    mid = (a + b + c) - (lo + hi);
  }
  static __device__ __forceinline__ void hpart(const RowLanes<KW>& r, Part& p) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      sort3(r.el(i + 1), r.e[i + 1], r.er(i + 1), p.lo_e[i], p.mid_e[i], p.hi_e[i]);
      sort3(r.ol(i + 1), r.o[i + 1], r.orr(i + 1), p.lo_o[i], p.mid_o[i], p.hi_o[i]);
    }
  }
  static __device__ __forceinline__ void vcomb(const Part& t, const Part& m, const Part& b, uint32_t (&out)[KW]) {
#pragma unroll
    for (int i = 0; i < KW; ++i) {
      const uint32_t e = mid3x16(max3x16(t.lo_e[i], m.lo_e[i], b.lo_e[i]), mid3x16(t.mid_e[i], m.mid_e[i], b.mid_e[i]),
                                 min3x16(t.hi_e[i], m.hi_e[i], b.hi_e[i]));
      const uint32_t o = mid3x16(max3x16(t.lo_o[i], m.lo_o[i], b.lo_o[i]), mid3x16(t.mid_o[i], m.mid_o[i], b.mid_o[i]),
                                 min3x16(t.hi_o[i], m.hi_o[i], b.hi_o[i]));
      out[i] = lanes_merge(e, o);
    }
  }
};

}  // namespace b2d
