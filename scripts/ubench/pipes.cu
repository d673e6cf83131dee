// pipes.cu - per-instruction issue cost on B200 (sm_100a): cycles per warp-instruction
// per SM sub-partition at saturation.  Design input for the fused kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; run on the GPU box.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#define NCH 8
#define UNR 8
#define ITERS 512

#define KERNEL(NAME, INIT, ...)                                                      \
  __global__ void __launch_bounds__(256) k_##NAME(uint32_t* out, uint32_t s, long long* cyc) { \
    uint32_t a[NCH];                                                                 \
    uint32_t b = s * 2654435761u + threadIdx.x, c = s ^ 0x9E3779B9u;                 \
    _Pragma("unroll") for (int j = 0; j < NCH; ++j) a[j] = (threadIdx.x + j * 977u + s) INIT; \
    __syncthreads();                                                                 \
    const long long t0 = clock64();                                                  \
    for (int i = 0; i < ITERS; ++i) {                                                \
      _Pragma("unroll") for (int u = 0; u < UNR; ++u) {                              \
        _Pragma("unroll") for (int j = 0; j < NCH; ++j) { __VA_ARGS__; }                  \
      }                                                                              \
    }                                                                                \
    const long long t1 = clock64();                                                  \
    uint32_t r = b ^ c;                                                              \
    _Pragma("unroll") for (int j = 0; j < NCH; ++j) r ^= a[j];                       \
    out[blockIdx.x * 256 + threadIdx.x] = r;                                         \
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;                                 \
  }

KERNEL(imad, , asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c)))
KERNEL(imad_wide, , { unsigned long long w; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(w) : "r"(a[j]), "r"(b)); a[j] = (uint32_t)(w >> 32) ^ (uint32_t)w; })
KERNEL(imad_wide_only, , { unsigned long long w; asm volatile("mad.wide.u32 %0, %1, %2, %3;" : "=l"(w) : "r"(a[j]), "r"(b), "l"((unsigned long long)c)); a[j] = (uint32_t)(w >> 32); })
KERNEL(imad_hi, , asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c)))
KERNEL(lop3, , asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c)))
KERNEL(iadd3, , asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b)))
KERNEL(prmt, , asm volatile("prmt.b32 %0, %0, %1, 0x4321;" : "+r"(a[j]) : "r"(b)))
KERNEL(shf, , asm volatile("shf.l.wrap.b32 %0, %0, %1, 7;" : "+r"(a[j]) : "r"(b)))
KERNEL(vmin2, , a[j] = __vminu2(a[j], b) + 0u)
KERNEL(vmax2min2, , { uint32_t lo = __vminu2(a[j], b); b = __vmaxu2(a[j], b); a[j] = lo; })
KERNEL(vsad4, , a[j] = __vsadu4(a[j], b))
KERNEL(dp4a, , a[j] = __dp4a(a[j], 0x01010101u, b))
KERNEL(isetp_sel, , a[j] = (a[j] < b) ? c : a[j] + 1u)
KERNEL(popc, , a[j] = __popc(a[j]) + b)
KERNEL(ffma, | 0x3f800000u, { float f = __uint_as_float(a[j]); f = fmaf(f, 0.999f, 1e-3f); a[j] = __float_as_uint(f); })
KERNEL(ffma_reg, | 0x3f800000u, { float f = __uint_as_float(a[j]); f = fmaf(f, __uint_as_float(b | 0x3f000000u), __uint_as_float(c | 0x3f000000u)); a[j] = __float_as_uint(f); })
KERNEL(fadd_rz, | 0x3f800000u, { float f = __uint_as_float(a[j]); f = __fadd_rz(f, 1.0f); a[j] = __float_as_uint(f); })
KERNEL(hfma2, | 0x3c003c00u, asm volatile("fma.rn.f16x2 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(0x3bff3bffu), "r"(0x10001000u)))
KERNEL(mufu_lg2, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f) | 0x3f800000u; })
KERNEL(mufu_lg2_only, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(mufu_sin, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("sin.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(mufu_sqrt, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("sqrt.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(mufu_rsqrt, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("rsqrt.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(mufu_ex2, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(mufu_rcp, | 0x3f800000u, { float f = __uint_as_float(a[j]); asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); })
KERNEL(i2f, , { float f; asm volatile("cvt.rn.f32.u32 %0, %1;" : "=f"(f) : "r"(a[j])); a[j] = __float_as_uint(f); })
KERNEL(i2f_u16, , { float f; unsigned short h = (unsigned short)a[j]; asm volatile("cvt.rn.f32.u16 %0, %1;" : "=f"(f) : "h"(h)); a[j] = __float_as_uint(f); })
KERNEL(f2i, | 0x40000000u, { int q; asm volatile("cvt.rzi.s32.f32 %0, %1;" : "=r"(q) : "f"(__uint_as_float(a[j]))); a[j] = (uint32_t)q | 0x40000000u; })
KERNEL(f2i_only, | 0x40000000u, { int q; asm volatile("cvt.rzi.s32.f32 %0, %1;" : "=r"(q) : "f"(__uint_as_float(a[j]))); a[j] = (uint32_t)q; })
// one Philox4x32 round on (a[j], a[j^1]) pairs: 2 wide multiplies + 2 xor3
KERNEL(philox_round, , if (!(j & 1)) { unsigned long long p0, p1; asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p0) : "r"(a[j]), "r"(0xD2511F53u)); asm volatile("mul.wide.u32 %0, %1, %2;" : "=l"(p1) : "r"(a[j + 1]), "r"(0xCD9E8D57u)); a[j] = (uint32_t)(p1 >> 32) ^ (uint32_t)p0 ^ b; a[j + 1] = (uint32_t)(p0 >> 32) ^ (uint32_t)p1 ^ c; })
// alu + fma pipe mix: does the scheduler reach 1 warp-instr / clk / SMSP?
KERNEL(mix_prmt_imad, , if (j & 1) asm volatile("prmt.b32 %0, %0, %1, 0x4321;" : "+r"(a[j]) : "r"(b)); else asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c)))
KERNEL(mix_lop_ffma, | 0x3f800000u, if (j & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(a[j]) : "r"(0x007fffffu), "r"(0x3f800000u)); else { float f = __uint_as_float(a[j]); f = fmaf(f, 0.999f, 1e-3f); a[j] = __float_as_uint(f); })
KERNEL(mix_mufu_alu, | 0x3f800000u, if ((j & 7) == 0) { float f = __uint_as_float(a[j]); asm volatile("lg2.approx.ftz.f32 %0, %0;" : "+f"(f)); a[j] = __float_as_uint(f); } else asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(a[j]) : "r"(b), "r"(c)))

// packed fp32 pairs (sm_100 f32x2): one instruction, two lanes of a 64-bit register pair
#define PK(lo, hi) ({ unsigned long long _v; asm("mov.b64 %0, {%1,%2};" : "=l"(_v) : "r"(lo), "r"(hi)); _v; })
#define UPK(v, lo, hi) asm("mov.b64 {%0,%1}, %2;" : "=r"(lo), "=r"(hi) : "l"(v))
KERNEL(ffma2, | 0x3f800000u, if (!(j & 1)) { unsigned long long v = PK(a[j], a[j + 1]); const unsigned long long m = PK(0x3f7fbe77u, 0x3f7fbe77u), k = PK(0x3a83126fu, 0x3a83126fu); asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(m), "l"(k)); UPK(v, a[j], a[j + 1]); })
KERNEL(fadd2, | 0x3f800000u, if (!(j & 1)) { unsigned long long v = PK(a[j], a[j + 1]); const unsigned long long k = PK(0x3a83126fu, 0x3a83126fu); asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(k)); UPK(v, a[j], a[j + 1]); })
KERNEL(fmul2, | 0x3f800000u, if (!(j & 1)) { unsigned long long v = PK(a[j], a[j + 1]); const unsigned long long m = PK(0x3f7fbe77u, 0x3f7fbe77u); asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(m)); UPK(v, a[j], a[j + 1]); })
KERNEL(fmnmx, | 0x3f800000u, { float f = __uint_as_float(a[j]); f = fminf(f, __uint_as_float(b | 0x3f000000u)); a[j] = __float_as_uint(f) + 1u; })
KERNEL(fsel, | 0x3f800000u, { a[j] = (b & (1u << j)) ? a[j] + 1u : c; })
KERNEL(shfl, , a[j] = __shfl_xor_sync(0xffffffffu, a[j], 1))

// half2 min / max on the same 16-bit lanes as VIMNMX.U16x2 (bit patterns 0x00vv are positive subnormal halves, ordered
// like the integers): which pipe, and do the two co-issue?
KERNEL(hmnmx2, , { __half2 x = *reinterpret_cast<__half2*>(&a[j]); uint32_t bb = b & 0x00ff00ffu; __half2 y = *reinterpret_cast<__half2*>(&bb); x = __hmax2(x, y); a[j] = *reinterpret_cast<uint32_t*>(&x) + 1u; })
KERNEL(hmnmx2_only, | 0x3c003c00u, { __half2 x = *reinterpret_cast<__half2*>(&a[j]); uint32_t bb = b & 0x00ff00ffu; __half2 y = *reinterpret_cast<__half2*>(&bb); x = __hmax2(x, y); a[j] = *reinterpret_cast<uint32_t*>(&x); })
KERNEL(mix_vmin_hmax, , if (!(j & 1)) { uint32_t lo = __vminu2(a[j], a[j + 1]); __half2 x = *reinterpret_cast<__half2*>(&a[j]); __half2 y = *reinterpret_cast<__half2*>(&a[j + 1]); __half2 hi = __hmax2(x, y); a[j] = lo; a[j + 1] = *reinterpret_cast<uint32_t*>(&hi); })
KERNEL(mix_vmin_hmax_indep, , if (j & 1) { a[j] = __vminu2(a[j], b) + 0u; } else { __half2 x = *reinterpret_cast<__half2*>(&a[j]); uint32_t bb = c & 0x00ff00ffu; __half2 y = *reinterpret_cast<__half2*>(&bb); x = __hmax2(x, y); a[j] = *reinterpret_cast<uint32_t*>(&x); })
KERNEL(mix_vmin_vmax, , if (!(j & 1)) { uint32_t lo = __vminu2(a[j], a[j + 1]); uint32_t hi = __vmaxu2(a[j], a[j + 1]); a[j] = lo; a[j + 1] = hi; })

// shared-memory read throughput
template <int VEC>
__global__ void __launch_bounds__(256) k_lds(uint32_t* out, uint32_t s, long long* cyc) {
  __shared__ __align__(16) uint32_t sm[4096];
  for (int i = threadIdx.x; i < 4096; i += 256) sm[i] = i * s;
  __syncthreads();
  uint32_t acc = 0;
  uint32_t idx = (threadIdx.x * VEC) & 4095;
  const long long t0 = clock64();
  for (int i = 0; i < ITERS; ++i) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      if (VEC == 1) acc += sm[(idx + u * 256) & 4095];
      if (VEC == 2) { uint2 v = *reinterpret_cast<uint2*>(&sm[(idx + u * 512) & 4095]); acc += v.x ^ v.y; }
      if (VEC == 4) { uint4 v = *reinterpret_cast<uint4*>(&sm[(idx + u * 1024) & 4095]); acc += v.x ^ v.y ^ v.z ^ v.w; }
    }
    idx = (idx + acc) & 4095 & ~(VEC - 1);
  }
  const long long t1 = clock64();
  out[blockIdx.x * 256 + threadIdx.x] = acc;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

typedef void (*kern_t)(uint32_t*, uint32_t, long long*);
struct Entry { const char* name; kern_t fn; double ops_per_iter; };

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount, bps = 4, nb = nsm * bps;   // 4 x 256 threads = 8 warps per SMSP
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, sizeof(uint32_t) * nb * 256);
  cudaMalloc(&cyc, sizeof(long long) * nb);
  long long* h = (long long*)malloc(sizeof(long long) * nb);
  const double full = (double)UNR * NCH;
  Entry es[] = {
#define E(n, f) {#n, k_##n, f}
      E(imad, full), E(imad_wide, full), E(imad_wide_only, full), E(imad_hi, full), E(lop3, full), E(iadd3, full),
      E(prmt, full), E(shf, full), E(vmin2, full), E(vmax2min2, full), E(vsad4, full), E(dp4a, full),
      E(isetp_sel, full), E(popc, full), E(ffma, full), E(ffma_reg, full), E(fadd_rz, full), E(hfma2, full),
      E(mufu_lg2, full), E(mufu_lg2_only, full), E(mufu_sin, full), E(mufu_sqrt, full), E(mufu_rsqrt, full),
      E(mufu_ex2, full), E(mufu_rcp, full), E(i2f, full), E(i2f_u16, full), E(f2i, full), E(f2i_only, full),
      E(philox_round, full / 2), E(mix_prmt_imad, full), E(mix_lop_ffma, full), E(mix_mufu_alu, full),
      E(ffma2, full / 2), E(fadd2, full / 2), E(fmul2, full / 2), E(fmnmx, full), E(fsel, full), E(shfl, full),
      E(hmnmx2, full), E(hmnmx2_only, full), E(mix_vmin_hmax, full), E(mix_vmin_hmax_indep, full), E(mix_vmin_vmax, full),
      {"lds32", k_lds<1>, 16.0}, {"lds64", k_lds<2>, 16.0}, {"lds128", k_lds<4>, 16.0},
  };
  printf("device %s, %d SMs, %d MHz max\n", p.name, nsm, p.clockRate / 1000);
  printf("%-16s %12s %14s %12s\n", "op", "cycles", "cyc/step/SMSP", "ms");
  for (const Entry& e : es) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    e.fn<<<nb, 256>>>(out, 1u, cyc);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    e.fn<<<nb, 256>>>(out, 3u, cyc);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("%s: %s\n", e.name, cudaGetErrorString(err)); return 1; }
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaMemcpy(h, cyc, sizeof(long long) * nb, cudaMemcpyDeviceToHost);
    long long mx = 0;
    for (int i = 0; i < nb; ++i) mx = h[i] > mx ? h[i] : mx;
    // each SMSP hosts 8 warps (4 blocks x 8 warps / 4 SMSPs); "step" = one BODY execution per warp
    const double steps = (double)ITERS * e.ops_per_iter * 8.0;
    printf("%-16s %12lld %14.3f %12.4f\n", e.name, mx, (double)mx / steps, ms);
  }
  return 0;
}
