// b2d_noise.cuh - per-element noise operators, fused into every consumer
// kernel's prologue.  One call handles one 8-element group (8 consecutive
// uint8 values of one image's flat H*W*3 index space) from ONE Philox4x32-10
// block: 16 random bits per element.
//
// Reference expressions (lib/roi_data_layer/minibatch.py, twins in
// lib/model/test.py) are cited per operator; ->u8 is the reference's
// (255 * x).astype(np.uint8), i.e. truncate toward zero, wrap modulo 256.
#pragma once
#include "b2d_common.cuh"

namespace b2d {

enum NoiseKind {   // device-side operator ids (not the ABI enum)
  NK_NONE = 0, NK_GAUSSIAN, NK_SPECKLE, NK_SAP, NK_POISSON, NK_UNIFORM,
  NK_RAYLEIGH, NK_GAMMA, NK_TABLE_ADD, NK_SHADER, NK_COUNT
};
enum NoiseMode { NM_GEN = 0, NM_INJECT = 1, NM_DUMP = 2 };

struct NoiseArgs {
  uint32_t k0, k1;          // Philox key (seed)
  uint64_t first_global;    // global index of batch image 0
  uint64_t elems;           // H*W*3
  float p0;                 // gaussian/speckle: sigma; uniform: high; rayleigh/gamma: scale
  float p1;
  double d0;                // inject mode: amount (sap)
  uint32_t u0;              // sap: 15-bit flip threshold; shader: 1 = keep BGR
  const double* f0;         // injected fields (batch base) or nullptr
  const double* f1;
  const uint8_t* tab;       // periodic table / brownian field
  uint64_t tab_stride;      // bytes between images in tab (0: shared by all images)
  const float* img_par;     // per-image parameter (poisson: vals), batch-local index
  float* dump0;             // NM_DUMP outputs (batch base), float32
  float* dump1;
};

struct GroupCtx {
  uint32_t img_local;       // index inside the batch
  uint32_t group;           // 8-element group inside the image
  const uint8_t* img;       // base of this image's input bytes
};

__device__ __forceinline__ uint4 group_random(const NoiseArgs& a, const GroupCtx& g, uint32_t sub) {
  const uint64_t gi = a.first_global + g.img_local;
  return philox4x32(g.group, sub, (uint32_t)gi, (uint32_t)(gi >> 32), a.k0, a.k1);
}
__device__ __forceinline__ uint32_t r16(const uint4& r, int i) {
  const uint32_t w = (i >> 1) == 0 ? r.x : (i >> 1) == 1 ? r.y : (i >> 1) == 2 ? r.z : r.w;
  return (i & 1) ? (w >> 16) : (w & 0xffffu);
}

template <int KIND, int MODE>
struct NoiseOp;

// ---- identity (retain_original, minibatch.py:1512-1516) -------------------
template <int MODE>
struct NoiseOp<NK_NONE, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs&, const GroupCtx&, uint32_t (&)[8], float* = nullptr) {}
};

// element index helpers for field access
__device__ __forceinline__ uint64_t field_index(const NoiseArgs& a, const GroupCtx& g, int i) {
  return (uint64_t)g.img_local * a.elems + (uint64_t)g.group * 8u + i;
}
__device__ __forceinline__ bool in_image(const NoiseArgs& a, const GroupCtx& g, int i) {
  return (uint64_t)g.group * 8u + i < a.elems;
}

// ---- gaussian: clip(f + N(0, sqrt(var)), 0, 1) ->u8  (minibatch.py:117-118)
// ---- speckle:  clip(f + f * N(0, sqrt(var)), 0, 1)   (minibatch.py:380-383)
// Two N(0, sigma^2) samples from one 32-bit Philox word (Box-Muller on two 16-bit
// uniforms).  Integer -> float goes through the 2^23 magic number (PRMT + FADD): I2F.U16
// and F2I are quarter-rate XU-pipe instructions on sm_100a (scripts/ubench/pipes.cu).
//   u1 = (ka + 0.5) / 65536, theta = 2 pi (kb + 0.5) / 65536
//   n  = sigma * sqrt(-2 ln u1) * (cos, sin)(theta);   sigma^2 is folded into the sqrt argument
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}
#define B2D_MAGIC 0x4B000000u   /* 2^23 as fp32 bits: 0x4B0000xx == 8388608 + xx */
__device__ __forceinline__ float magic_lo16(uint32_t w) { return __uint_as_float(prmt(w, B2D_MAGIC, 0x7410u)); }
__device__ __forceinline__ float magic_hi16(uint32_t w) { return __uint_as_float(prmt(w, B2D_MAGIC, 0x7432u)); }
template <int I>
__device__ __forceinline__ float magic_byte(uint32_t w) { return __uint_as_float(prmt(w, B2D_MAGIC, 0x7440u | I)); }

// Box-Muller in polar form: radius sigma * sqrt(-2 ln u1) and (cos, sin) of theta, from one 32-bit Philox word.
// theta comes from the magic float 2^23 + kb in ONE FFMA: its constant carries -2^23 * c1 (the product is exact inside the
// FFMA, so the cancellation costs nothing; the angle grid is shifted by < 3e-5 rad, a constant).
__device__ __forceinline__ void normal_polar(uint32_t r32, float sigma, float& rad, float& c, float& s) {
  const float a = magic_lo16(r32) - 8388607.5f;                       // ka + 0.5, exact
  const float s2 = sigma * sigma * 1.3862943611198906f;               // sigma^2 * 2 ln 2 (loop invariant)
  rad = fast_sqrt(fmaf(fast_lg2(a), -s2, 16.0f * s2));                // sigma * sqrt(-2 ln u1)
  constexpr float c1 = 6.283185307179586f / 65536.0f;
  constexpr float c0 = (float)(3.141592653589793 / 65536.0 - 8388608.0 * (double)c1);
  const float th = fmaf(magic_hi16(r32), c1, c0);                     // 2 pi (kb + 0.5) / 65536
  c = fast_cos(th);
  s = fast_sin(th);
}
__device__ __forceinline__ void normal_pair(uint32_t r32, float sigma, float& n0, float& n1) {
  float rad, c, s;
  normal_polar(r32, sigma, rad, c, s);
  n0 = rad * c;
  n1 = rad * s;
}

template <int MODE, bool SPECKLE>
struct NormalNoise {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (!in_image(a, g, i)) continue;
        const double n = a.f0[field_index(a, g, i)];
        const double f = (double)v[i] / 255.0;
        double o = SPECKLE ? __dadd_rn(f, __dmul_rn(f, n)) : __dadd_rn(f, n);
        o = fmin(fmax(o, 0.0), 1.0);
        if (xf) xf[i] = (float)__dmul_rn(255.0, o);
        v[i] = wrap_u8(__dmul_rn(255.0, o));
      }
      return;
    }
    const uint4 r = group_random(a, g, 0);
    const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      // gaussian: x = v + 255 n = cos * (255 rad) + v in ONE fma (the radius is drawn at sigma * 255), speckle: v + v n.
      // Exact value, one rounding toward zero - the packed operator below truncates the same exact value, so both
      // agree on the uint8 result.
      float rad, cs, sn;
      normal_polar(rw[j], SPECKLE ? a.p0 : a.p0 * 255.0f, rad, cs, sn);
      const float n0 = SPECKLE ? rad * cs : rad * cs * (1.0f / 255.0f), n1 = SPECKLE ? rad * sn : rad * sn * (1.0f / 255.0f);
      if (MODE == NM_DUMP) {
        if (in_image(a, g, 2 * j)) a.dump0[field_index(a, g, 2 * j)] = n0;
        if (in_image(a, g, 2 * j + 1)) a.dump0[field_index(a, g, 2 * j + 1)] = n1;
      }
      const float v0 = (float)v[2 * j], v1 = (float)v[2 * j + 1];
      const float x0 = SPECKLE ? __fmaf_rz(v0, n0, v0) : __fmaf_rz(cs, rad, v0);
      const float x1 = SPECKLE ? __fmaf_rz(v1, n1, v1) : __fmaf_rz(sn, rad, v1);
      const float c0 = fminf(fmaxf(x0, 0.0f), 255.0f), c1 = fminf(fmaxf(x1, 0.0f), 255.0f);
      if (xf) { xf[2 * j] = c0; xf[2 * j + 1] = c1; }
      v[2 * j] = (uint32_t)__float2int_rz(c0);
      v[2 * j + 1] = (uint32_t)__float2int_rz(c1);
    }
  }
};
template <int MODE> struct NoiseOp<NK_GAUSSIAN, MODE> : NormalNoise<MODE, false> {};
template <int MODE> struct NoiseOp<NK_SPECKLE, MODE> : NormalNoise<MODE, true> {};

// ---- s&p: flipped = u1 < amount; salted = u2 < 0.5   (minibatch.py:259)
// generator: the low 15 bits of a 16-bit draw decide the flip, bit 15 salt/pepper.
template <int MODE>
struct NoiseOp<NK_SAP, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (!in_image(a, g, i)) continue;
        const uint64_t e = field_index(a, g, i);
        if (a.f0[e] < a.d0) v[i] = (a.f1[e] < 0.5) ? 255u : 0u;
      }
      return;
    }
    const uint4 r = group_random(a, g, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t k = r16(r, i);
      const bool flip = (k & 0x7fffu) < a.u0;
      const bool salt = (k & 0x8000u) != 0u;
      if (MODE == NM_DUMP && in_image(a, g, i)) {
        a.dump0[field_index(a, g, i)] = ((float)(k & 0x7fffu) + 0.5f) * (1.0f / 32768.0f);
        a.dump1[field_index(a, g, i)] = salt ? 0.25f : 0.75f;
      }
      v[i] = flip ? (salt ? 255u : 0u) : v[i];
    }
  }
};

// ---- additive, NOT clipped, wrapping ->u8:
//   uniform  cv2.add(f, U(0, high))                 (minibatch.py:676-677,710)
//   rayleigh cv2.add(f, rayleigh.rvs(scale))        (minibatch.py:1356-1357,1387)
//   gamma    cv2.add(f, gamma.rvs(1.99, scale))     (minibatch.py:1216-1217,1247)
__device__ __forceinline__ uint32_t add_wrap_inject(uint32_t v, double s, float* xf) {
  const double f = (double)v / 255.0;
  const double x = __dmul_rn(255.0, __dadd_rn(f, s));
  if (xf) *xf = (float)x;
  return wrap_u8(x);
}

template <int MODE>
struct NoiseOp<NK_UNIFORM, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (in_image(a, g, i)) v[i] = add_wrap_inject(v[i], a.f0[field_index(a, g, i)], xf ? xf + i : nullptr);
      return;
    }
    const uint4 r = group_random(a, g, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float s = a.p0 * u16_to_unit(r16(r, i));
      if (MODE == NM_DUMP && in_image(a, g, i)) a.dump0[field_index(a, g, i)] = s;
      const float x = fmaf(s, 255.0f, (float)v[i]);
      if (xf) xf[i] = x;
      v[i] = wrap_u8(x);
    }
  }
};

template <int MODE>
struct NoiseOp<NK_RAYLEIGH, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (in_image(a, g, i)) v[i] = add_wrap_inject(v[i], a.f0[field_index(a, g, i)], xf ? xf + i : nullptr);
      return;
    }
    const uint4 r = group_random(a, g, 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      // inverse CDF: scale * sqrt(-2 ln u)
      const float s = a.p0 * fast_sqrt(-1.3862943611198906f * fast_lg2(u16_to_unit(r16(r, i))));
      if (MODE == NM_DUMP && in_image(a, g, i)) a.dump0[field_index(a, g, i)] = s;
      const float x = fmaf(s, 255.0f, (float)v[i]);
      if (xf) xf[i] = x;
      v[i] = wrap_u8(x);
    }
  }
};

// Test hook (NM_DUMP) of the table sampler: unfold K = R + 256 h with h drawn from the plan's fold
// table (substream 1), place the sample inside [K, K + 1) by inverting a linear density through the
// gamma pdf at both ends (substream 2; K = 0: cdf ~ x^1.99), and write (K + u) / 255.
static __device__ __noinline__ void gamma_dump8(const NoiseArgs& a, const GroupCtx& g, const uint32_t (&k)[8]) {
  const uint32_t folds = a.u0;
  const uint32_t* fold = reinterpret_cast<const uint32_t*>(a.tab + 512);
  const uint4 rh = group_random(a, g, 1), rj = group_random(a, g, 2);
  const float q1 = __expf(-1.0f / (255.0f * a.p0));
  for (int i = 0; i < 8; ++i) {
    if (!in_image(a, g, i)) continue;
    const uint32_t uh = r16(rh, i);
    uint32_t h = 0;
    while (h + 1 < folds && uh >= fold[k[i] * folds + h]) ++h;
    const float K = (float)(k[i] + 256u * h);
    const float U = ((float)r16(rj, i) + 0.5f) * (1.0f / 65536.0f);
    float u;
    if (K == 0.0f) {
      u = exp2f(log2f(U) * (1.0f / 1.99f));
    } else {
      const float q = powf((K + 1.0f) / K, 0.99f) * q1, dq = q - 1.0f;
      u = fabsf(dq) < 1e-3f ? U : (sqrtf(fmaf(q * q - 1.0f, U, 1.0f)) - 1.0f) / dq;
    }
    u = fminf(fmaxf(u, 0.004f), 0.996f);
    a.dump0[field_index(a, g, i)] = (K + u) / 255.0f;
  }
}

// Float view: Marsaglia-Tsang for shape a = 1.99 (d = a - 1/3, c = 1/sqrt(9 d)).  Each
// attempt t draws one Philox block per 4 elements (substream 2t + half):
// words 0,1 -> 4 normals, words 2,3 -> 4 acceptance uniforms.
template <int MODE>
struct NoiseOp<NK_GAMMA, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (in_image(a, g, i)) v[i] = add_wrap_inject(v[i], a.f0[field_index(a, g, i)], xf ? xf + i : nullptr);
      return;
    }
    if (xf == nullptr) {
      // uint8 consumers: out = (v + K) mod 256 with K = floor(255 * scale * G); R = K mod 256 comes from
      // the plan's alias table (b2d_gamma.cu), low byte of the 16 random bits = slot, high byte = u
      const uint16_t* tab = reinterpret_cast<const uint16_t*>(a.tab);
      const uint4 r = group_random(a, g, 0);
      uint32_t k[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint32_t rr = r16(r, i);
        const uint32_t e = __ldg(tab + (rr & 0xffu));
        k[i] = (rr < (e & 0xff00u) ? rr : e) & 0xffu;
      }
      if (MODE == NM_DUMP) gamma_dump8(a, g, k);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = (v[i] + k[i]) & 0xffu;
      return;
    }
    // float view (wavelet consumers): Marsaglia-Tsang samples
    const float d = 1.99f - 1.0f / 3.0f;
    const float c = 0.25897428f;  // 1 / sqrt(9 d)
    float smp[8];
    uint32_t pending = 0xffu;
#pragma unroll
    for (int i = 0; i < 8; ++i) smp[i] = d;
    for (uint32_t t = 0; t < 16u && pending; ++t) {
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        if (!((pending >> (4 * half)) & 0xfu)) continue;
        const uint4 r = group_random(a, g, 2u * t + half);
        float z[4];
        box_muller16(r.x & 0xffffu, r.x >> 16, z[0], z[1]);
        box_muller16(r.y & 0xffffu, r.y >> 16, z[2], z[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i = 4 * half + j;
          if (!((pending >> i) & 1u)) continue;
          const uint32_t ku = (j < 2 ? r.z : r.w) >> (16 * (j & 1)) & 0xffffu;
          const float x = z[j];
          float vv = fmaf(c, x, 1.0f);
          vv = vv * vv * vv;
          if (vv <= 0.0f) continue;
          const float lnu = 0.6931471805599453f * fast_lg2(u16_to_unit(ku));
          const float lnv = 0.6931471805599453f * fast_lg2(vv);
          if (lnu < 0.5f * x * x + d - d * vv + d * lnv) {
            smp[i] = d * vv;
            pending &= ~(1u << i);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float s = a.p0 * smp[i];
      if (MODE == NM_DUMP && in_image(a, g, i)) a.dump0[field_index(a, g, i)] = s;
      const float x = fmaf(s, 255.0f, (float)v[i]);
      if (xf) xf[i] = x;
      v[i] = wrap_u8(x);
    }
  }
};

// ---- poisson: vals = 2^ceil(log2(#unique)); clip(Poisson(f*vals)/vals, 0, 1)
// (skimage random_noise 'poisson', minibatch.py:211).  img_par[img] = vals.
// Sampler: table-free inversion for lambda < 12, PTRS (Hoermann 1993, the
// algorithm NumPy's legacy poisson uses for lambda >= 10) above; 24-bit
// uniforms, one Philox block per 2 elements per attempt.
__device__ __forceinline__ uint32_t poisson_to_u8(float k, float vals) {
  // 255 * min(k / vals, 1): vals is a power of two, so this is exact
  const float q = fminf(k, vals);
  return (uint32_t)__float2int_rz(255.0f * q / vals);
}

__device__ __forceinline__ float poisson_small(float lam, float u) {
  float p = __expf(-lam), F = p, k = 0.0f;
  while (u > F && k < 80.0f) {
    k += 1.0f;
    p *= lam / k;
    F += p;
  }
  return k;
}

// one PTRS attempt; returns k >= 0 on acceptance, -1 on rejection
__device__ __forceinline__ float poisson_ptrs_try(float lam, float U01, float V) {
  const float slam = sqrtf(lam), loglam = __logf(lam);
  const float b = 0.931f + 2.53f * slam;
  const float a = -0.059f + 0.02483f * b;
  const float invalpha = 1.1239f + 1.1328f / (b - 3.4f);
  const float vr = 0.9277f - 3.6224f / (b - 2.0f);
  const float U = U01 - 0.5f;
  const float us = 0.5f - fabsf(U);
  const float k = floorf((2.0f * a / us + b) * U + lam + 0.43f);
  if (us >= 0.07f && V <= vr) return k;
  if (k < 0.0f || (us < 0.013f && V > us)) return -1.0f;
  if (__logf(V) + __logf(invalpha) - __logf(a / (us * us) + b) <= -lam + k * loglam - lgammaf(k + 1.0f))
    return k;
  return -1.0f;
}

template <int MODE>
struct NoiseOp<NK_POISSON, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    const float vals = a.img_par[g.img_local];
    if (MODE == NM_INJECT) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (!in_image(a, g, i)) continue;
        const double k = a.f0[field_index(a, g, i)];
        const double o = fmin(fmax(k / (double)vals, 0.0), 1.0);
        v[i] = wrap_u8(__dmul_rn(255.0, o));
      }
      return;
    }
    float ks[8];
    uint32_t pending = 0xffu;
    for (uint32_t t = 0; t < 24u && pending; ++t) {
#pragma unroll
      for (int pr = 0; pr < 4; ++pr) {
        if (!((pending >> (2 * pr)) & 3u)) continue;
        const uint4 r = group_random(a, g, 4u * t + pr);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int i = 2 * pr + j;
          if (!((pending >> i) & 1u)) continue;
          const float lam = (float)v[i] * vals * (1.0f / 255.0f);
          const float u = u32_to_unit(j ? r.z : r.x), w = u32_to_unit(j ? r.w : r.y);
          float k;
          if (lam <= 0.0f) k = 0.0f;
          else if (lam < 12.0f) k = poisson_small(lam, u);
          else k = poisson_ptrs_try(lam, u, w);
          if (k >= 0.0f) {
            ks[i] = k;
            pending &= ~(1u << i);
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if ((pending >> i) & 1u) ks[i] = floorf((float)v[i] * vals * (1.0f / 255.0f));  // never in practice
      if (MODE == NM_DUMP && in_image(a, g, i)) a.dump0[field_index(a, g, i)] = ks[i];
      v[i] = poisson_to_u8(ks[i], vals);
    }
  }
};

// ---- periodic / brownian: saturating cv2.add(img, u8 field)
// (minibatch.py:828,1046).  The field comes from a table: shared by every
// image (periodic: depends on H, W, L only) or one per image (brownian).
template <int MODE>
struct NoiseOp<NK_TABLE_ADD, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
    const uint8_t* t = a.tab + (uint64_t)g.img_local * a.tab_stride + (uint64_t)g.group * 8u;
    if ((uint64_t)g.group * 8u + 8u <= a.elems && ((reinterpret_cast<uintptr_t>(t) & 7u) == 0)) {
      const uint2 w = *reinterpret_cast<const uint2*>(t);
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = min(v[i] + byte_of(i < 4 ? w.x : w.y, i & 3), 255u);
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (in_image(a, g, i)) v[i] = min(v[i] + (uint32_t)t[i], 255u);
    }
  }
};

// ---- shader: ImageEnhance.Brightness(PIL RGB image).enhance(3) = min(255, 3 v)
// on the RGB-ordered decode (minibatch.py:1505-1511): output channel c of a
// pixel is min(255, 3 * BGR[2 - c]) unless u0 = 1 (keep BGR).
template <int MODE>
struct NoiseOp<NK_SHADER, MODE> {
  static __device__ __forceinline__ void apply8(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8],
                                                float* xf = nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint64_t e = (uint64_t)g.group * 8u + i;
      uint32_t s = v[i];
      if (a.u0 == 0u && e < a.elems) {
        const uint32_t c = (uint32_t)(e % 3u);
        s = g.img[e + 2 - 2 * c];
      }
      v[i] = min(3u * s, 255u);
    }
  }
};


// float view of the noised values (0..255 scale, BEFORE the ->u8 cast) for the
// consumers that the reference feeds with the float image (denoise_wavelet on
// random_noise / cv2.add output, minibatch.py:91-94,677-679).  Operators whose
// reference result is already uint8 (s&p values, poisson_wavelet's explicit
// ->u8 at minibatch.py:212, periodic, brownian, shader) return the uint8 value.
template <int KIND, int MODE>
__device__ __forceinline__ void apply8_float(const NoiseArgs& a, const GroupCtx& g, uint32_t (&v)[8], float (&xf)[8]) {
  constexpr bool kFloat = KIND == NK_GAUSSIAN || KIND == NK_SPECKLE || KIND == NK_UNIFORM || KIND == NK_RAYLEIGH ||
                          KIND == NK_GAMMA;
  if (kFloat) {
#pragma unroll
    for (int i = 0; i < 8; ++i) xf[i] = (float)v[i];
    NoiseOp<KIND, MODE>::apply8(a, g, v, xf);
  } else {
    NoiseOp<KIND, MODE>::apply8(a, g, v);
#pragma unroll
    for (int i = 0; i < 8; ++i) xf[i] = (float)v[i];
  }
}


// ===========================================================================
// Packed operators: one 8-byte group held as two 32-bit words, never unpacked
// to one register per byte.  These are what the fused kernels and the
// noise-only kernel run in generator mode; they compute exactly the uint8
// values of the unpacked operators above (tests/test_gpu_filters3.py compares
// the two paths).  Cost model (scripts/ubench/pipes.cu, cycles per warp
// instruction per SM sub-partition): ALU-class 2, IMAD 2, IMAD.WIDE 4, FFMA 1,
// MUFU / F2I 8; one Philox4x32-10 block = 20 IMAD.WIDE + 20 LOP3.
// ===========================================================================
template <int KIND>
struct PackedNoise {                 // default: unpack -> per-element operator -> pack
  static constexpr bool kNative = false;
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    uint32_t v[8];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[i] = byte_of(w.x, i);
      v[4 + i] = byte_of(w.y, i);
    }
    NoiseOp<KIND, NM_GEN>::apply8(a, g, v);
    return make_uint2(pack4(v[0], v[1], v[2], v[3]), pack4(v[4], v[5], v[6], v[7]));
  }
};

template <>
struct PackedNoise<NK_NONE> {
  static constexpr bool kNative = true;
  static __device__ __forceinline__ uint2 apply(const NoiseArgs&, const GroupCtx&, uint2 w) { return w; }
};

// s&p on bytes: per 16-bit draw k, lane = 0x8000 + (k & 0x7fff) - T has bit 15 clear iff
// the element flips; PRMT's sign-replicate mode turns bits 15 of four lanes into a
// byte mask, one LOP3 selects {value, 255, 0}.  7 ALU instructions per 4 bytes.
template <>
struct PackedNoise<NK_SAP> {
  static constexpr bool kNative = true;
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    const uint4 r = group_random(a, g, 0);
    const uint32_t tt = a.u0 * 0x00010001u;
    const uint32_t y0 = (r.x | 0x80008000u) - tt, y1 = (r.y | 0x80008000u) - tt;
    const uint32_t y2 = (r.z | 0x80008000u) - tt, y3 = (r.w | 0x80008000u) - tt;
    const uint32_t keep0 = prmt(y0, y1, 0xFDB9u), keep1 = prmt(y2, y3, 0xFDB9u);   // 0xff: no flip
    const uint32_t salt0 = prmt(r.x, r.y, 0xFDB9u), salt1 = prmt(r.z, r.w, 0xFDB9u);
    return make_uint2((w.x & keep0) | (~keep0 & salt0), (w.y & keep1) | (~keep1 & salt1));
  }
};

// gaussian / speckle: x = v + 255 n (v + v n) evaluated exactly by one FFMA.RZ onto the
// magic number 2^23 + v, so the low 16 bits of the result are floor(x) as a signed
// 16-bit integer (x < 0 lands below 2^23 where the spacing is 0.5: still negative);
// VIMNMX.S16x2.RELU clamps two elements to [0, 255] in one instruction.
// Needs |x| < 16384: b2d_plan_create bounds var for these two noises.
template <bool SPECKLE>
struct PackedNormal {
  static constexpr bool kNative = true;
  // gaussian: trig = cos / sin, rad = 255 sigma sqrt(-2 ln u): x = trig * rad + v; speckle: trig * rad = n, x = v n + v
  template <int I>
  static __device__ __forceinline__ uint32_t elem(uint32_t w, float trig, float rad) {
    const float mv = magic_byte<I>(w);
    const float t = SPECKLE ? __fmaf_rz(mv - 8388608.0f, rad * trig, mv) : __fmaf_rz(trig, rad, mv);
    return __float_as_uint(t);
  }
  static __device__ __forceinline__ uint32_t clamp2(uint32_t t0, uint32_t t1) {
    return __vimin_s16x2_relu(prmt(t0, t1, 0x5410u), 0x00ff00ffu);
  }
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    const uint4 r = group_random(a, g, 0);
    const float sg = SPECKLE ? a.p0 : a.p0 * 255.0f;
    float r0, r1, r2, r3, n0, n1, n2, n3, n4, n5, n6, n7;               // n = (cos, sin) per Philox word
    normal_polar(r.x, sg, r0, n0, n1);
    normal_polar(r.y, sg, r1, n2, n3);
    normal_polar(r.z, sg, r2, n4, n5);
    normal_polar(r.w, sg, r3, n6, n7);
    const uint32_t l01 = clamp2(elem<0>(w.x, n0, r0), elem<1>(w.x, n1, r0));
    const uint32_t l23 = clamp2(elem<2>(w.x, n2, r1), elem<3>(w.x, n3, r1));
    const uint32_t l45 = clamp2(elem<0>(w.y, n4, r2), elem<1>(w.y, n5, r2));
    const uint32_t l67 = clamp2(elem<2>(w.y, n6, r3), elem<3>(w.y, n7, r3));
    return make_uint2(prmt(l01, l23, 0x6420u), prmt(l45, l67, 0x6420u));
  }
};
// shader: min(255, 3 v) of the RGB-ordered decode = swap bytes 0 and 2 of every pixel, then a
// saturating x3 on u16x2 lanes.  Output byte i of the group (channel c = (8 g + i) % 3) reads byte
// i + 2 - 2 c: the sources span [-2, +9] around the group, so the two neighbouring words are
// fetched (L1 hits) and three PRMTs with phase-dependent selectors do the swap.
template <>
struct PackedNoise<NK_SHADER> {
  static constexpr bool kNative = true;
  static __device__ __forceinline__ uint32_t times3_sat(uint32_t w) {
    const uint32_t e = __vminu2((w & 0x00ff00ffu) * 3u, 0x00ff00ffu);
    const uint32_t o = __vminu2(((w >> 8) & 0x00ff00ffu) * 3u, 0x00ff00ffu);
    return prmt(e, o, 0x6240u);
  }
  static __device__ __forceinline__ uint32_t word_at(const GroupCtx& g, const NoiseArgs& a, int64_t e) {
    if (e < 0) return 0u;
    const uint8_t* p = g.img + e;
    if ((uint64_t)e + 4u <= a.elems && (reinterpret_cast<uintptr_t>(p) & 3u) == 0)
      return __ldg(reinterpret_cast<const uint32_t*>(p));
    uint32_t w = 0u;
    for (int i = 0; i < 4; ++i)
      if ((uint64_t)e + i < a.elems) w |= (uint32_t)__ldg(p + i) << (8 * i);
    return w;
  }
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    if (a.u0 != 0u) return make_uint2(times3_sat(w.x), times3_sat(w.y));      // compat: keep BGR order
    const int64_t e0 = (int64_t)g.group * 8;
    const uint32_t prev = word_at(g, a, e0 - 4), next = word_at(g, a, e0 + 8);
    // per output byte: channel c -> source L (byte i - 2, c = 2), M (byte i, c = 1) or R (byte i + 2, c = 0)
    const uint32_t lx = prmt(prev, w.x, 0x5432u), rx = prmt(w.x, w.y, 0x5432u);
    const uint32_t ly = rx, ry = prmt(w.y, next, 0x5432u);
    // selector nibbles over {A = (L, M): 0-3 L, 4-7 M} then {(A', R): 0-3 A', 4-7 R}
    const uint32_t ph = (uint32_t)((g.group * 8u) % 3u);                      // channel of byte 0
    // channel pattern of bytes 0..3 for ph = 0, 1, 2: (0,1,2,0) (1,2,0,1) (2,0,1,2); bytes 4..7 start at ph + 1
    auto swap4 = [](uint32_t l, uint32_t m, uint32_t r, uint32_t phase) {
      // first PRMT: c = 2 -> L byte i, else M byte i;  second: c = 0 -> R byte i, else keep
      const uint32_t selA = phase == 0 ? 0x7254u : phase == 1 ? 0x7614u : 0x3650u;
      const uint32_t selB = phase == 0 ? 0x7214u : phase == 1 ? 0x3610u : 0x3250u;
      return prmt(prmt(l, m, selA), r, selB);
    };
    const uint32_t ox = swap4(lx, w.x, rx, ph), oy = swap4(ly, w.y, ry, (ph + 1u) % 3u);
    return make_uint2(times3_sat(ox), times3_sat(oy));
  }
};

// uniform / rayleigh: x = v + 255 s >= 0, not clipped, ->u8 keeps the low byte of trunc(x) (wrap).  One
// FFMA.RZ / FADD.RZ onto the magic number 2^23 + v leaves floor(x) in the mantissa: its low byte is the
// result.  k + 0.5 comes from the 16-bit draw by PRMT + one exact FADD, as in normal_pair.
template <bool RAYLEIGH>
struct PackedAdditive {
  static constexpr bool kNative = true;
  template <int I>
  static __device__ __forceinline__ uint32_t elem(uint32_t w, float kf, float c) {
    const float mv = magic_byte<I>(w);
    if (RAYLEIGH) {                                               // 255 * scale * sqrt(-2 ln((k + 0.5) / 65536))
      const float rad = fast_sqrt(fmaf(fast_lg2(kf), -c, 16.0f * c));
      return __float_as_uint(__fadd_rz(rad, mv));
    }
    return __float_as_uint(__fmaf_rz(kf, c, mv));                 // 255 * high * (k + 0.5) / 65536
  }
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    const uint4 r = group_random(a, g, 0);
    const float c = RAYLEIGH ? 1.3862943611198906f * (255.0f * a.p0) * (255.0f * a.p0) : a.p0 * (255.0f / 65536.0f);
    const uint32_t t0 = elem<0>(w.x, magic_lo16(r.x) - 8388607.5f, c), t1 = elem<1>(w.x, magic_hi16(r.x) - 8388607.5f, c);
    const uint32_t t2 = elem<2>(w.x, magic_lo16(r.y) - 8388607.5f, c), t3 = elem<3>(w.x, magic_hi16(r.y) - 8388607.5f, c);
    const uint32_t t4 = elem<0>(w.y, magic_lo16(r.z) - 8388607.5f, c), t5 = elem<1>(w.y, magic_hi16(r.z) - 8388607.5f, c);
    const uint32_t t6 = elem<2>(w.y, magic_lo16(r.w) - 8388607.5f, c), t7 = elem<3>(w.y, magic_hi16(r.w) - 8388607.5f, c);
    return make_uint2(prmt(prmt(t0, t1, 0x0040u), prmt(t2, t3, 0x0040u), 0x5410u),
                      prmt(prmt(t4, t5, 0x0040u), prmt(t6, t7, 0x0040u), 0x5410u));
  }
};
// gamma: R = K mod 256 of eight elements from the alias table, one wrapping byte add per word
template <>
struct PackedNoise<NK_GAMMA> {
  static constexpr bool kNative = true;
  static __device__ __forceinline__ uint2 apply(const NoiseArgs& a, const GroupCtx& g, uint2 w) {
    const uint16_t* tab = reinterpret_cast<const uint16_t*>(a.tab);
    const uint4 r = group_random(a, g, 0);
    const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t r0 = rw[j] & 0xffffu, r1 = rw[j] >> 16;
      const uint32_t e0 = __ldg(tab + (r0 & 0xffu)), e1 = __ldg(tab + (r1 & 0xffu));
      o[2 * j] = r0 < (e0 & 0xff00u) ? r0 : e0;                                  // low byte = R
      o[2 * j + 1] = r1 < (e1 & 0xff00u) ? r1 : e1;
    }
    const uint32_t kx = prmt(prmt(o[0], o[1], 0x0040u), prmt(o[2], o[3], 0x0040u), 0x5410u);
    const uint32_t ky = prmt(prmt(o[4], o[5], 0x0040u), prmt(o[6], o[7], 0x0040u), 0x5410u);
    return make_uint2(__vadd4(w.x, kx), __vadd4(w.y, ky));
  }
};

template <> struct PackedNoise<NK_UNIFORM> : PackedAdditive<false> {};
template <> struct PackedNoise<NK_RAYLEIGH> : PackedAdditive<true> {};

template <> struct PackedNoise<NK_GAUSSIAN> : PackedNormal<false> {};
template <> struct PackedNoise<NK_SPECKLE> : PackedNormal<true> {};

}  // namespace b2d
