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
  return philox4x32_10(g.group, sub, (uint32_t)gi, (uint32_t)(gi >> 32), a.k0, a.k1);
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
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float z0, z1;
      box_muller16(r16(r, 2 * j), r16(r, 2 * j + 1), z0, z1);
      const float n0 = a.p0 * z0, n1 = a.p0 * z1;
      if (MODE == NM_DUMP) {
        if (in_image(a, g, 2 * j)) a.dump0[field_index(a, g, 2 * j)] = n0;
        if (in_image(a, g, 2 * j + 1)) a.dump0[field_index(a, g, 2 * j + 1)] = n1;
      }
      const float v0 = (float)v[2 * j], v1 = (float)v[2 * j + 1];
      const float x0 = SPECKLE ? fmaf(v0, n0, v0) : fmaf(n0, 255.0f, v0);
      const float x1 = SPECKLE ? fmaf(v1, n1, v1) : fmaf(n1, 255.0f, v1);
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
// generator: 15 bits decide the flip, 1 bit salt/pepper.
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
      const bool flip = (k >> 1) < a.u0;
      const bool salt = (k & 1u) != 0u;
      if (MODE == NM_DUMP && in_image(a, g, i)) {
        a.dump0[field_index(a, g, i)] = ((float)(k >> 1) + 0.5f) * (1.0f / 32768.0f);
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

// Marsaglia-Tsang for shape a = 1.99 (d = a - 1/3, c = 1/sqrt(9 d)).  Each
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

}  // namespace b2d
