// b2d_run.cu - plan creation and the dispatch of one batch onto the kernels.
#include "b2d_kernels.cuh"
#include "b2d_prepass.cuh"
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>

using namespace b2d;

#define CK(x)                                          \
  do {                                                 \
    cudaError_t e__ = (x);                             \
    if (e__ != cudaSuccess) return fail_cuda(e__, #x); \
  } while (0)

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static int validate_spec(const b2d_spec& s, int h, int w) {
  if (s.noise < 0 || s.noise >= B2D_NOISE_COUNT) return fail(B2D_ERR_INVALID, "spec.noise %d out of range", s.noise);
  if (s.denoise < 0 || s.denoise >= B2D_DENOISE_COUNT)
    return fail(B2D_ERR_INVALID, "spec.denoise %d out of range", s.denoise);
  if (s.denoise == B2D_DENOISE_MEDIAN && s.ksize != 3 && s.ksize != 5 && s.ksize != 7)
    return fail(B2D_ERR_INVALID, "median ksize must be 3, 5 or 7 (got %d)", s.ksize);
  if (s.denoise == B2D_DENOISE_BILATERAL && (s.d < 1 || s.d > 15 || !(s.d & 1)))
    return fail(B2D_ERR_INVALID, "bilateral d must be odd in [1, 15] (got %d)", s.d);
  if (s.noise == B2D_NOISE_BLOOM && h / 100 - 2 < 1)
    return fail(B2D_ERR_INVALID, "bloom needs H >= 300 (randint(1, H//100-2), Automold.py:583)");
  if (s.noise == B2D_NOISE_QUANT && (s.level < 1 || s.level > 32))
    return fail(B2D_ERR_INVALID, "quant k must be in [1, 32]");
  if (s.noise == B2D_NOISE_GAMMA && !(s.level > 0.0 && s.level <= 100.0))
    return fail(B2D_ERR_INVALID, "gamma scale %g outside (0, 100]", s.level);
  if ((s.noise == B2D_NOISE_GAUSSIAN || s.noise == B2D_NOISE_SPECKLE) && !(s.level >= 0.0 && s.level <= 100.0))
    return fail(B2D_ERR_INVALID, "gaussian / speckle var must be in [0, 100] (packed generator range)");
  if (s.denoise == B2D_DENOISE_WAVELET) {
    if (s.wavelet != B2D_WAVELET_DB1 && s.wavelet != B2D_WAVELET_BIOR15)
      return fail(B2D_ERR_INVALID, "spec.wavelet %d: 0 (db1) or 1 (bior1.5)", s.wavelet);
    if (s.levels < 0 || s.levels > 8) return fail(B2D_ERR_INVALID, "spec.levels %d outside [0, 8] (0 = skimage's rule)", s.levels);
  }
  if (s.noise == B2D_NOISE_PERIODIC && s.level != B2D_LEVEL_SIZE && !(s.level > 0.0))
    return fail(B2D_ERR_INVALID, "periodic limit must be > 0 or B2D_LEVEL_SIZE");
  if (s.noise == B2D_NOISE_BROWNIAN && !(s.level > 0.0)) return fail(B2D_ERR_INVALID, "brownian dt must be > 0");
  if (s.noise == B2D_NOISE_SAP && !(s.level >= 0.0 && s.level <= 1.0))
    return fail(B2D_ERR_INVALID, "s&p amount must be in [0, 1]");
  if ((s.noise == B2D_NOISE_UNIFORM || s.noise == B2D_NOISE_RAYLEIGH) && !(s.level >= 0.0 && s.level <= 16.0))
    return fail(B2D_ERR_INVALID, "uniform / rayleigh level must be in [0, 16] (packed generator range)");
  (void)w;
  return B2D_OK;
}

extern "C" int b2d_plan_create(b2d_plan** out, int device, int n, int h, int w, const b2d_spec* specs,
                               int n_specs) {
  if (!out || !specs) return fail(B2D_ERR_INVALID, "b2d_plan_create: null argument");
  *out = nullptr;
  if (n < 0 || h <= 0 || w <= 0) return fail(B2D_ERR_INVALID, "b2d_plan_create: bad shape n=%d h=%d w=%d", n, h, w);
  if (n > (1 << 24)) return fail(B2D_ERR_INVALID, "b2d_plan_create: at most 2^24 images per plan");
  if (n_specs != 1 && n_specs != n) return fail(B2D_ERR_INVALID, "n_specs must be 1 or n (got %d, n=%d)", n_specs, n);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
    return fail(B2D_ERR_NO_DEVICE, "no CUDA device: libb200denoise has no CPU fallback");
  if (device < 0 || device >= ndev) return fail(B2D_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
  DeviceGuard guard(device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  for (int i = 0; i < n_specs; ++i) {
    const int rc = validate_spec(specs[i], h, w);
    if (rc) return rc;
  }
  b2d_plan* p = new (std::nothrow) b2d_plan();
  if (!p) return fail(B2D_ERR_INVALID, "out of host memory");
  p->device = device; p->n = n; p->h = h; p->w = w; p->rowbytes = 3 * w;
  p->elems = (uint64_t)h * w * 3;
  p->h_d_in = p->h_d_out = p->h_d_ws = nullptr;
  p->d_poisson = nullptr;
  p->h_ready = false;
  p->need_aux = false;
  p->wv_stats_dev = nullptr;
  p->wv_stats_first = p->wv_stats_count = 0;
  // group images by identical spec, keeping first-appearance order
  for (int i = 0; i < n; ++i) {
    const b2d_spec& s = specs[n_specs == 1 ? 0 : i];
    SpecGroup* g = nullptr;
    for (SpecGroup& c : p->groups)
      if (memcmp(&c.spec, &s, sizeof(s)) == 0) { g = &c; break; }
    if (!g) {
      p->groups.push_back(SpecGroup{s, {}, true, nullptr});
      g = &p->groups.back();
    }
    g->images.push_back(i);
  }
  for (SpecGroup& g : p->groups) {
    g.contiguous = g.images.back() - g.images.front() + 1 == (int)g.images.size();
    if (!g.contiguous) {
      if (cudaMalloc(&g.d_list, sizeof(int) * g.images.size()) != cudaSuccess ||
          cudaMemcpy(g.d_list, g.images.data(), sizeof(int) * g.images.size(), cudaMemcpyHostToDevice) != cudaSuccess) {
        b2d_plan_destroy(p);
        return fail(B2D_ERR_CUDA, "plan: cannot upload image list");
      }
    }
    const int nz = g.spec.noise;
    if (nz == B2D_NOISE_BROWNIAN) p->need_aux = true;
    if (nz == B2D_NOISE_POISSON && !p->d_poisson) {
      const cudaError_t e = upload_poisson_tables(&p->d_poisson);
      if (e != cudaSuccess) {
        b2d_plan_destroy(p);
        return fail_cuda(e, "poisson tables");
      }
    }
    if (nz == B2D_NOISE_GAMMA) {
      bool have = false;
      for (const GammaTable& t : p->gamma) have = have || t.level == g.spec.level;
      if (!have) {
        GammaTable t{g.spec.level, nullptr, gamma_folds(g.spec.level)};
        const cudaError_t e = upload_gamma_table(g.spec.level, &t.d_tab);
        if (e != cudaSuccess) {
          if (t.d_tab) cudaFree(t.d_tab);
          b2d_plan_destroy(p);
          return fail_cuda(e, "gamma table");
        }
        p->gamma.push_back(t);
      }
    }
    if (nz == B2D_NOISE_PERIODIC) {
      bool have = false;
      for (const PeriodicTable& t : p->periodic) have = have || t.level == g.spec.level;
      if (!have) {
        PeriodicTable t{g.spec.level, nullptr};
        const double lim = g.spec.level == B2D_LEVEL_SIZE ? (double)p->elems : g.spec.level;
        cudaError_t e = cudaMalloc(&t.d_tab, p->elems);
        if (e == cudaSuccess) e = launch_periodic_table(t.d_tab, p->elems, lim, 0);
        if (e == cudaSuccess) e = cudaStreamSynchronize(0);
        if (e != cudaSuccess) {
          if (t.d_tab) cudaFree(t.d_tab);
          b2d_plan_destroy(p);
          return fail_cuda(e, "periodic table");
        }
        p->periodic.push_back(t);
      }
    }
  }
  // workspace layout
  p->brown_chunks = brownian_chunks(p->elems);
  size_t off = 0;
  p->off_par = off;    off = align_up(off + sizeof(float) * (size_t)std::max(n, 1) * 4, 256);
  p->off_bitmap = off; off = align_up(off + sizeof(uint32_t) * 8 * (size_t)std::max(n, 1), 256);
  p->off_scan = off;   off = align_up(off + (p->need_aux ? sizeof(double) * (size_t)p->brown_chunks * n : 0), 256);
  p->off_stage = off;  off = align_up(off + (size_t)p->elems * n, 256);
  p->off_aux = off;    off = align_up(off + (p->need_aux ? (size_t)p->elems * n : 0), 256);
  p->off_pix = off;
  {
    size_t pix = 0;
    for (const SpecGroup& g : p->groups) {
      if (g.spec.noise == B2D_NOISE_QUANT) pix = std::max(pix, quant_table_bytes(n));
      if (g.spec.noise == B2D_NOISE_BLOOM) pix = std::max(pix, bloom_table_bytes(n, h));
    }
    off = align_up(off + pix, 256);
  }
  p->off_denoise = off;
  off += denoise_workspace_bytes(*p);
  p->ws_bytes = off;
  *out = p;
  return B2D_OK;
}

extern "C" int b2d_plan_workspace_bytes(const b2d_plan* plan, size_t* bytes) {
  if (!plan || !bytes) return fail(B2D_ERR_INVALID, "b2d_plan_workspace_bytes: null argument");
  *bytes = plan->ws_bytes;
  return B2D_OK;
}

extern "C" void b2d_plan_destroy(b2d_plan* p) {
  if (!p) return;
  DeviceGuard guard(p->device);
  for (SpecGroup& g : p->groups)
    if (g.d_list) cudaFree(g.d_list);
  for (PeriodicTable& t : p->periodic)
    if (t.d_tab) cudaFree(t.d_tab);
  for (GammaTable& t : p->gamma)
    if (t.d_tab) cudaFree(t.d_tab);
  if (p->d_poisson) cudaFree(p->d_poisson);
  if (p->h_d_in) cudaFree(p->h_d_in);
  if (p->h_d_out) cudaFree(p->h_d_out);
  if (p->h_d_ws) cudaFree(p->h_d_ws);
  if (p->h_ready)
    for (int i = 0; i < 3; ++i) cudaStreamDestroy(p->h_streams[i]);
  for (cudaEvent_t e : p->h_events) cudaEventDestroy(e);
  delete p;
}

