// b2d_api.cu - the C ABI (include/b200denoise.h): spec parsing, plans, dispatch.
#include "../../include/b200denoise.h"
#include "b2d_kernels.cuh"
#include "b2d_prepass.cuh"
#include "b2d_plan.h"

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

namespace b2d {

static thread_local char g_err[512] = "";
static thread_local uint64_t g_launches = 0;

void count_launch() { ++g_launches; }

int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

int sm_count() {
  static int sms[64] = {};
  const int dev = current_device() & 63;
  if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, current_device());
  return sms[dev] > 0 ? sms[dev] : 148;
}

const char* knob(const char* name) {
  static const bool enabled = [] { const char* e = getenv("B2D_DEBUG_KNOBS"); return e && e[0] == '1'; }();
  return enabled ? getenv(name) : nullptr;
}

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int fail_cuda(cudaError_t e, const char* what) {
  return fail(B2D_ERR_CUDA, "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
}

}  // namespace b2d

using namespace b2d;

#define CK(x)                                        \
  do {                                               \
    cudaError_t e__ = (x);                           \
    if (e__ != cudaSuccess) return fail_cuda(e__, #x); \
  } while (0)

// ---------------------------------------------------------------------------
// spec parsing: the reference's substring precedence
// ---------------------------------------------------------------------------
namespace {

struct LevelEntry { const char* tag; double value; };
struct NoiseEntry { const char* name; int id; LevelEntry levels[3]; int nlevels; };

// order of the top-level elif chain (minibatch.py:1575-1634) and, per noise,
// of the `'var<L>' in noise_type` tests inside its closure
const NoiseEntry kNoises[] = {
    {"gaussian", B2D_NOISE_GAUSSIAN, {{"var0.1", 0.1}, {"var1.0", 1.0}, {"var1.5", 1.5}}, 3},
    {"poisson", B2D_NOISE_POISSON, {}, 0},
    {"sap", B2D_NOISE_SAP, {{"var0.2", 0.2}, {"var0.4", 0.4}, {"var0.8", 0.8}}, 3},
    {"speckle", B2D_NOISE_SPECKLE, {{"var0.5", 0.5}, {"var1.0", 1.0}, {"var2.0", 2.0}}, 3},
    {"periodic", B2D_NOISE_PERIODIC, {{"var3.14", 3.141592653589793}, {"var100", 100.0}, {"varsize", B2D_LEVEL_SIZE}}, 3},
    {"brownian", B2D_NOISE_BROWNIAN, {{"var0.9", 0.9}, {"var0.09", 0.09}, {"var0.009", 0.009}}, 3},
    {"quant", B2D_NOISE_QUANT, {{"var3", 3}, {"var7", 7}, {"var10", 10}}, 3},
    {"uniform", B2D_NOISE_UNIFORM, {{"var0.2", 0.2}, {"var0.6", 0.6}, {"var1.2", 1.2}}, 3},
    {"gamma", B2D_NOISE_GAMMA, {{"var0.05", 0.05}, {"var0.1", 0.1}, {"var0.2", 0.2}}, 3},
    {"rayleigh", B2D_NOISE_RAYLEIGH, {{"var0.1", 0.1}, {"var0.2", 0.2}, {"var0.3", 0.3}}, 3},
    {"bloom", B2D_NOISE_BLOOM, {}, 0},
    {"shader", B2D_NOISE_SHADER, {}, 0},
};
struct DenoiseEntry { const char* name; int id; };
// order of the `<noise>_<denoise>` tests inside each closure (minibatch.py:89-184)
const DenoiseEntry kDenoisers[] = {
    {"wavelet", B2D_DENOISE_WAVELET}, {"gaus_blur", B2D_DENOISE_GAUS_BLUR}, {"mean", B2D_DENOISE_MEAN},
    {"median", B2D_DENOISE_MEDIAN},   {"bilateral", B2D_DENOISE_BILATERAL},
};
// order of the post-denoise pass on the top-level spec (minibatch.py:1636-1663),
// used only for noise-free specs such as `original_median`
const DenoiseEntry kPostDenoisers[] = {
    {"gaus_blur", B2D_DENOISE_GAUS_BLUR}, {"mean", B2D_DENOISE_MEAN},           {"median", B2D_DENOISE_MEDIAN},
    {"wavelet", B2D_DENOISE_WAVELET},     {"bilateral", B2D_DENOISE_BILATERAL},
};

void spec_defaults(b2d_spec* s) {
  memset(s, 0, sizeof(*s));
  s->ksize = 3;
  s->d = 9;
  s->sigma_color = 20.0f;
  s->sigma_space = 100.0f;
  s->wavelet = B2D_WAVELET_BIOR15;   // live files: wavelet='bior1.5' (minibatch.py:93)
  s->levels = 0;
}

}  // namespace

extern "C" int b2d_parse_spec(const char* spec, b2d_spec* out) {
  if (!spec || !out) return fail(B2D_ERR_INVALID, "b2d_parse_spec: null argument");
  spec_defaults(out);
  if (strstr(spec, "mix")) return fail(B2D_ERR_UNSUPPORTED, "'mix' specs are resolved by the host front-end");
  if (strstr(spec, "curvelet")) return fail(B2D_ERR_UNSUPPORTED, "curvelet denoising is out of scope (SURVEY 8(f) f3)");
  const NoiseEntry* ne = nullptr;
  for (const NoiseEntry& e : kNoises)
    if (strstr(spec, e.name)) { ne = &e; break; }
  if (!ne) {  // retain_original, optionally followed by the post-denoise pass
    out->noise = B2D_NOISE_NONE;
    for (const DenoiseEntry& d : kPostDenoisers)
      if (strstr(spec, d.name)) { out->denoise = d.id; break; }
    return B2D_OK;
  }
  out->noise = ne->id;
  if (ne->id == B2D_NOISE_BLOOM || ne->id == B2D_NOISE_SHADER) return B2D_OK;  // add_bloom()/add_shader() take no spec
  char key[64];
  for (const DenoiseEntry& d : kDenoisers) {
    snprintf(key, sizeof(key), "%s_%s", ne->name, d.name);
    if (strstr(spec, key)) { out->denoise = d.id; break; }
  }
  if (ne->nlevels) {
    bool found = false;
    for (int i = 0; i < ne->nlevels && !found; ++i)
      if (strstr(spec, ne->levels[i].tag)) { out->level = ne->levels[i].value; found = true; }
    if (!found)
      return fail(B2D_ERR_PARSE, "spec '%s': no known level for %s (the reference leaves `im` unbound)", spec, ne->name);
  }
  return B2D_OK;
}

// ---------------------------------------------------------------------------
// misc ABI
// ---------------------------------------------------------------------------
extern "C" uint64_t b2d_launch_count(void) { return g_launches; }
extern "C" void b2d_reset_launch_count(void) { g_launches = 0; }
extern "C" const char* b2d_last_error(void) { return g_err; }
extern "C" const char* b2d_version(void) { return "b200denoise 0.1 (sm_100a)"; }

extern "C" int b2d_host_register(void* ptr, size_t bytes) {
  if (!ptr) return fail(B2D_ERR_INVALID, "b2d_host_register: null pointer");
  CK(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
  return B2D_OK;
}
extern "C" int b2d_host_unregister(void* ptr) {
  if (!ptr) return fail(B2D_ERR_INVALID, "b2d_host_unregister: null pointer");
  CK(cudaHostUnregister(ptr));
  return B2D_OK;
}

// ---------------------------------------------------------------------------
// per-image checksum and squared error (8(e): the only cross-GPU payload)
// ---------------------------------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) k_stats(const uint8_t* ref, const uint8_t* out, uint64_t elems,
                                               unsigned long long* checksum, double* sse) {
  const int img = blockIdx.y;
  const uint8_t* o = out + (uint64_t)img * elems;
  const uint8_t* r = ref ? ref + (uint64_t)img * elems : nullptr;
  unsigned long long cs = 0;
  double se = 0.0;
  for (uint64_t e = (uint64_t)blockIdx.x * 256 + threadIdx.x; e < elems; e += (uint64_t)gridDim.x * 256) {
    const unsigned v = o[e];
    cs += (unsigned long long)v * (1u + (unsigned)(e % 251u));
    if (r) {
      const double d = (double)v - (double)r[e];
      se += d * d;
    }
  }
  for (int off = 16; off; off >>= 1) {
    cs += __shfl_down_sync(0xffffffffu, cs, off);
    se += __shfl_down_sync(0xffffffffu, se, off);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&checksum[img], cs);
    if (r) atomicAdd(&sse[img], se);
  }
}
}  // namespace

extern "C" int b2d_stats(b2d_plan* plan, const uint8_t* d_ref, const uint8_t* d_out, uint64_t* d_checksum,
                         double* d_sse, void* cuda_stream) {
  if (!plan || !d_out || !d_checksum) return fail(B2D_ERR_INVALID, "b2d_stats: null argument");
  if (d_ref && !d_sse) return fail(B2D_ERR_INVALID, "b2d_stats: d_sse required with d_ref");
  if (plan->n == 0) return B2D_OK;
  DeviceGuard guard(plan->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  CK(cudaMemsetAsync(d_checksum, 0, sizeof(uint64_t) * plan->n, st));
  if (d_sse) CK(cudaMemsetAsync(d_sse, 0, sizeof(double) * plan->n, st));
  unsigned bx = (unsigned)std::min<uint64_t>(64, (plan->elems + 256 * 16 - 1) / (256 * 16));
  if (bx < 1) bx = 1;
  for (int a = 0; a < plan->n; a += 65535) {          // the image index is grid.y
    const int cnt = std::min(65535, plan->n - a);
    const size_t off = (size_t)a * plan->elems;
    k_stats<<<dim3(bx, cnt), 256, 0, st>>>(d_ref ? d_ref + off : nullptr, d_out + off, plan->elems,
                                           reinterpret_cast<unsigned long long*>(d_checksum) + a, d_sse ? d_sse + a : nullptr);
  }
  count_launch();
  CK(cudaGetLastError());
  return B2D_OK;
}
