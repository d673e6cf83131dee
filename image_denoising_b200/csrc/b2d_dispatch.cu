// b2d_dispatch.cu - one batch -> kernel launches (b2d_run, b2d_run_host,
// b2d_noise_field).  Replaces the closure dispatch of lib/model/test.py:1612-1784.
#include "b2d_kernels.cuh"
#include "b2d_prepass.cuh"
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>

using namespace b2d;

#define CK(x)                                          \
  do {                                                 \
    cudaError_t e__ = (x);                             \
    if (e__ != cudaSuccess) return fail_cuda(e__, #x); \
  } while (0)

// ---------------------------------------------------------------------------
// dispatch
// ---------------------------------------------------------------------------
namespace {

struct RunCtx {
  b2d_plan* p;
  const uint8_t* in;     // batch base
  uint8_t* out;          // batch base
  uint8_t* ws;
  const b2d_fields* fields;
  uint64_t seed, first_global;
  cudaStream_t st;
  bool defer_denoise = false;   // run_range: every group stages its noised uint8 image, ONE denoise launch follows
};

ImageSet make_set(const b2d_plan* p, const uint8_t* in, uint8_t* out, const SpecGroup& g, int lo, int hi) {
  ImageSet s;
  s.in = in; s.out = out; s.h = p->h; s.w = p->w; s.rowbytes = p->rowbytes; s.elems = p->elems;
  s.count = hi - lo;
  s.first = g.images[lo];
  s.list = g.contiguous ? nullptr : g.d_list + lo;
  return s;
}

bool word_aligned(const b2d_plan* p, const void* a, const void* b) {
  return p->rowbytes % 4 == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 3u) == 0;
}

int filter3_kind(const b2d_spec& sp) {
  if (sp.denoise == B2D_DENOISE_GAUS_BLUR) return FK_GAUS3;
  if (sp.denoise == B2D_DENOISE_MEAN) return FK_MEAN3;
  if (sp.denoise == B2D_DENOISE_MEDIAN && sp.ksize == 3) return FK_MEDIAN3;
  return FK_NONE;
}

// denoise a uint8 image set that is already noised (src -> dst)
int denoise_u8(const RunCtx& c, const b2d_spec& sp, const ImageSet& s, bool noisy_input = true) {
  const int f3 = filter3_kind(sp);
  if (f3 != FK_NONE) {
    NoiseArgs none{};
    none.elems = c.p->elems;
    if (word_aligned(c.p, s.in, s.out)) CK(launch_fused3(NK_NONE, f3, s, none, c.st));
    else CK(launch_filter3_any(f3, s, c.st));
    return B2D_OK;
  }
  return run_denoise_pass(*c.p, sp, s, c.ws, c.st, noisy_input);
}

// Mixed batches whose groups all end in the same stand-alone uint8 denoiser (median 5 / 7, bilateral): the denoiser is
// launched once over the whole range instead of once per group (BASELINE config C5: 10 groups -> median 7 x 7; at
// 512 images per GPU the per-group launches are small and their tails add up).
bool same_denoiser(const b2d_spec& a, const b2d_spec& b) {
  return a.denoise == b.denoise && a.ksize == b.ksize && a.d == b.d && a.sigma_color == b.sigma_color &&
         a.sigma_space == b.sigma_space;
}
bool standalone_u8_denoiser(const b2d_spec& sp) {
  return (sp.denoise == B2D_DENOISE_MEDIAN && sp.ksize != 3) || sp.denoise == B2D_DENOISE_BILATERAL;
}

int run_group(const RunCtx& c, const SpecGroup& g, int lo, int hi) {
  b2d_plan* p = c.p;
  const b2d_spec& sp = g.spec;
  const bool inject = c.fields != nullptr && c.fields->f0 != nullptr;
  uint8_t* stage = c.ws + p->off_stage;
  ImageSet s = make_set(p, c.in, c.out, g, lo, hi);

  NoiseArgs na{};
  na.k0 = (uint32_t)c.seed; na.k1 = (uint32_t)(c.seed >> 32);
  na.first_global = c.first_global;
  na.elems = p->elems;
  if (inject) { na.f0 = c.fields->f0; na.f1 = c.fields->f1; }
  int kind = NK_NONE;
  switch (sp.noise) {
    case B2D_NOISE_NONE: kind = NK_NONE; break;
    case B2D_NOISE_GAUSSIAN: kind = NK_GAUSSIAN; na.p0 = (float)std::sqrt(sp.level); break;
    case B2D_NOISE_SPECKLE: kind = NK_SPECKLE; na.p0 = (float)std::sqrt(sp.level); break;
    case B2D_NOISE_SAP:
      kind = NK_SAP; na.d0 = sp.level;
      if ((sp.compat_flags & B2D_COMPAT_SAP_TRAIN06) && sp.denoise == B2D_DENOISE_NONE && sp.level == 0.8)
        na.d0 = 0.6;                                     // training closure, minibatch.py:367
      na.u0 = (uint32_t)std::max(0.0, std::min(32768.0, std::ceil(na.d0 * 32768.0 - 0.5)));
      if (inject && !na.f1) return fail(B2D_ERR_INVALID, "sap with injected fields needs f0 (u1) and f1 (u2)");
      break;
    case B2D_NOISE_UNIFORM: kind = NK_UNIFORM; na.p0 = (float)sp.level; break;
    case B2D_NOISE_RAYLEIGH: kind = NK_RAYLEIGH; na.p0 = (float)sp.level; break;
    case B2D_NOISE_GAMMA:
      kind = NK_GAMMA; na.p0 = (float)sp.level;
      for (const GammaTable& t : p->gamma)
        if (t.level == sp.level) { na.tab = t.d_tab; na.u0 = (uint32_t)t.folds; }
      break;
    case B2D_NOISE_SHADER: kind = NK_SHADER; na.u0 = (sp.compat_flags & B2D_COMPAT_SHADER_BGR) ? 1u : 0u; break;
    case B2D_NOISE_POISSON: {
      kind = NK_POISSON;
      float* par = reinterpret_cast<float*>(c.ws + p->off_par);
      CK(launch_poisson_vals(s, reinterpret_cast<uint32_t*>(c.ws + p->off_bitmap), par, c.st));
      na.img_par = par;
      if (inject) break;                               // injected counts: the per-element operator divides and clips
      // generator mode: alias-table sampler writes the corrupted uint8 image, the denoiser reads it
      uint8_t* dst = sp.denoise == B2D_DENOISE_NONE ? c.out : stage;
      ImageSet ps = s;
      ps.out = dst;
      CK(launch_poisson(ps, na, p->d_poisson, c.st));
      if (sp.denoise == B2D_DENOISE_NONE || c.defer_denoise) return B2D_OK;
      ImageSet ds = s;
      ds.in = stage;
      if (sp.denoise == B2D_DENOISE_WAVELET) {
        NoiseArgs none{};
        none.elems = p->elems;
        return run_wavelet(*p, sp, ds, NK_NONE, NM_GEN, none, c.ws, c.st);
      }
      return denoise_u8(c, sp, ds);
    }
    case B2D_NOISE_PERIODIC: {
      kind = NK_TABLE_ADD;
      for (const PeriodicTable& t : p->periodic)
        if (t.level == sp.level) na.tab = t.d_tab;
      na.tab_stride = 0;
      break;
    }
    case B2D_NOISE_BROWNIAN: {
      kind = NK_TABLE_ADD;
      BrownianArgs ba;
      ba.sqrt_dt = std::sqrt(sp.level);
      ba.chunk_sums = reinterpret_cast<double*>(c.ws + p->off_scan);
      ba.nchunks = p->brown_chunks;
      ba.field = c.ws + p->off_aux;
      ba.add_out = nullptr;
      // Consumers other than the fused 3x3 kernels and the wavelet (median 5 / 7, bilateral, none) read a corrupted uint8
      // image: the field kernel adds the image itself and writes that - no field in HBM, no separate table-add pass.
      const bool direct = sp.denoise != B2D_DENOISE_WAVELET &&
                          (filter3_kind(sp) == FK_NONE || !word_aligned(p, s.in, s.out));
      if (direct) {
        ba.add_out = sp.denoise == B2D_DENOISE_NONE ? c.out : stage;
        CK(launch_brownian_field(s, na, ba, inject, c.st));
        if (sp.denoise == B2D_DENOISE_NONE || c.defer_denoise) return B2D_OK;
        ImageSet ds = s;
        ds.in = stage;
        return denoise_u8(c, sp, ds);
      }
      CK(launch_brownian_field(s, na, ba, inject, c.st));
      na.tab = ba.field;
      na.tab_stride = p->elems;
      break;
    }
    case B2D_NOISE_QUANT:
    case B2D_NOISE_BLOOM: {
      // pixel-granular pre-pass writes the corrupted uint8 image; the denoiser then reads it
      uint8_t* dst = sp.denoise == B2D_DENOISE_NONE ? c.out : stage;
      ImageSet ps = s;
      ps.out = dst;
      uint8_t* tables = c.ws + p->off_pix;
      const int rc = sp.noise == B2D_NOISE_QUANT
                         ? run_quant(*p, sp, ps, tables, c.fields, c.seed, c.first_global, c.st)
                         : run_bloom(*p, ps, g, lo, hi, tables, c.fields, c.seed, c.first_global, c.st);
      if (rc) return rc;
      if (sp.denoise == B2D_DENOISE_NONE || c.defer_denoise) return B2D_OK;
      ImageSet ds = s;
      ds.in = stage;
      if (sp.denoise == B2D_DENOISE_WAVELET) {           // quant_wavelet: denoise_wavelet(uint8 palette image), minibatch.py:501-537
        NoiseArgs none{};
        none.elems = p->elems;
        return run_wavelet(*p, sp, ds, NK_NONE, NM_GEN, none, c.ws, c.st);
      }
      return denoise_u8(c, sp, ds);
    }
    default: return fail(B2D_ERR_INVALID, "unknown noise %d", sp.noise);
  }
  // brownian/periodic consume injected fields in their pre-pass, not in the operator
  const bool op_inject = inject && kind != NK_TABLE_ADD && kind != NK_NONE && kind != NK_SHADER;
  const int mode = op_inject ? NM_INJECT : NM_GEN;

  if (sp.denoise == B2D_DENOISE_NONE) {
    CK(launch_noise_only(kind, mode, s, na, c.st));
    return B2D_OK;
  }
  if (sp.denoise == B2D_DENOISE_WAVELET) return run_wavelet(*p, sp, s, kind, mode, na, c.ws, c.st);
  const int f3 = filter3_kind(sp);
  if (f3 != FK_NONE && !op_inject && word_aligned(p, s.in, s.out)) {
    CK(launch_fused3(kind, f3, s, na, c.st));          // the one-pass hot path
    return B2D_OK;
  }
  if (kind == NK_NONE && !c.defer_denoise) return denoise_u8(c, sp, s, false);    // `original_*`: nothing to stage, the denoiser reads the input
  // two-pass: noise -> stage (uint8) -> denoiser
  ImageSet ns = s;
  ns.out = stage;
  CK(launch_noise_only(kind, mode, ns, na, c.st));
  if (c.defer_denoise) return B2D_OK;
  ImageSet ds = s;
  ds.in = stage;
  return denoise_u8(c, sp, ds);
}

int run_range(const RunCtx& c0, int img0, int cnt) {
  RunCtx c = c0;
  // groups that have images in [img0, img0 + cnt)
  const SpecGroup* first = nullptr;
  int live = 0;
  bool same = true;
  for (const SpecGroup& g : c.p->groups) {
    const auto lo = std::lower_bound(g.images.begin(), g.images.end(), img0);
    if (lo == g.images.end() || *lo >= img0 + cnt) continue;
    ++live;
    if (!first) first = &g;
    else same = same && same_denoiser(first->spec, g.spec);
  }
  c.defer_denoise = live > 1 && same && standalone_u8_denoiser(first->spec);
  // the image index is a grid dimension (<= 65535) in every kernel: larger groups go out in several launches
  constexpr int kMaxLaunchImages = 65535;
  for (const SpecGroup& g : c.p->groups) {
    const int lo = (int)(std::lower_bound(g.images.begin(), g.images.end(), img0) - g.images.begin());
    const int hi = (int)(std::lower_bound(g.images.begin(), g.images.end(), img0 + cnt) - g.images.begin());
    for (int a = lo; a < hi; a += kMaxLaunchImages) {
      const int rc = run_group(c, g, a, std::min(hi, a + kMaxLaunchImages));
      if (rc) return rc;
    }
  }
  if (c.defer_denoise) {                              // every image of the range now sits noised in the staging set
    for (int a = 0; a < cnt; a += kMaxLaunchImages) {
      ImageSet s;
      s.in = c.ws + c.p->off_stage; s.out = c.out; s.h = c.p->h; s.w = c.p->w; s.rowbytes = c.p->rowbytes; s.elems = c.p->elems;
      s.count = std::min(kMaxLaunchImages, cnt - a); s.first = img0 + a; s.list = nullptr;
      const int rc = denoise_u8(c, first->spec, s);
      if (rc) return rc;
    }
  }
  return B2D_OK;
}

}  // namespace

extern "C" int b2d_run(b2d_plan* plan, const uint8_t* d_in, uint8_t* d_out, void* d_ws,
                       const b2d_fields* fields, uint64_t seed, uint64_t first_global_image_index,
                       void* cuda_stream) {
  if (!plan) return fail(B2D_ERR_INVALID, "b2d_run: null plan");
  if (plan->n == 0) return B2D_OK;
  if (!d_in || !d_out) return fail(B2D_ERR_INVALID, "b2d_run: null image buffer");
  if (!d_ws && plan->ws_bytes) return fail(B2D_ERR_INVALID, "b2d_run: workspace of %zu bytes required", plan->ws_bytes);
  DeviceGuard guard(plan->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  RunCtx c{plan, d_in, d_out, static_cast<uint8_t*>(d_ws), fields, seed, first_global_image_index,
           static_cast<cudaStream_t>(cuda_stream)};
  return run_range(c, 0, plan->n);
}

// ---------------------------------------------------------------------------
// host-buffer entry point: chunked H2D / kernels / D2H on three streams
// ---------------------------------------------------------------------------
extern "C" int b2d_run_host(b2d_plan* plan, const uint8_t* h_in, uint8_t* h_out, uint64_t seed,
                            uint64_t first_global_image_index) {
  if (!plan) return fail(B2D_ERR_INVALID, "b2d_run_host: null plan");
  if (plan->n == 0) return B2D_OK;
  if (!h_in || !h_out) return fail(B2D_ERR_INVALID, "b2d_run_host: null image buffer");
  DeviceGuard guard(plan->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  const size_t bytes = (size_t)plan->elems * plan->n;
  if (!plan->h_ready) {
    CK(cudaMalloc(&plan->h_d_in, bytes));
    CK(cudaMalloc(&plan->h_d_out, bytes));
    if (plan->ws_bytes) CK(cudaMalloc(&plan->h_d_ws, plan->ws_bytes));
    for (int i = 0; i < 3; ++i) CK(cudaStreamCreateWithFlags(&plan->h_streams[i], cudaStreamNonBlocking));
    plan->h_ready = true;
  }
  // chunks of >= 32 MiB (at least one image) so each copy runs at full PCIe rate (B2D_HOST_CHUNK_MB: diagnostic knob;
  // 32 is the optimum of a 4-128 MiB sweep)
  static const uint64_t chunk_bytes = [] {
    const char* e = knob("B2D_HOST_CHUNK_MB");
    return (uint64_t)(e && atoi(e) > 0 ? atoi(e) : 32) << 20;
  }();
  static const bool chunk_forced = knob("B2D_HOST_CHUNK_MB") != nullptr;
  int per = (int)std::max<uint64_t>(1, chunk_bytes / plan->elems);
  if (!chunk_forced) {
    // Chunking must not starve the launches: a mixed batch cuts every chunk into its spec groups (keep >= 4 images per
    // group and chunk), and the per-image pre-passes that are latency bound - k-means fit (quant), fp64 scan (brownian),
    // the flare tables (bloom) - cost the same for 2 images as for 200, so such plans run in at most 8 chunks.
    bool latency_bound = false;
    for (const SpecGroup& g : plan->groups)
      latency_bound = latency_bound || g.spec.noise == B2D_NOISE_QUANT || g.spec.noise == B2D_NOISE_BROWNIAN ||
                      g.spec.noise == B2D_NOISE_BLOOM;
    if (plan->groups.size() > 1) per = std::max(per, 4 * (int)plan->groups.size());
    if (latency_bound) per = std::max(per, (plan->n + 7) / 8);
  }
  per = std::min(per, plan->n);
  // (ramping the first and last chunks down to per/8 to shorten the one-way fill and drain phases measured no gain)
  std::vector<int> cut;
  for (int i = 0; i < plan->n; i += per) cut.push_back(i);
  cut.push_back(plan->n);
  const int nchunks = (int)cut.size() - 1;
  while ((int)plan->h_events.size() < 2 * nchunks) {
    cudaEvent_t e;
    CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    plan->h_events.push_back(e);
  }
  // Three streams, one per engine: every H2D copy is queued at once on the first (the device buffers hold the whole
  // batch: nothing is reused, the copies run ahead freely), chunk k computes on the second when its input has landed -
  // ONE compute stream, so the chunks use the plan's shared scratch (wavelet planes, tables, staging rows) one at a time -
  // and the third drains chunk k when its kernels are done.
  cudaStream_t s_in = plan->h_streams[0], s_run = plan->h_streams[1], s_out = plan->h_streams[2];
  int rc = B2D_OK;
  cudaError_t ce = cudaSuccess;
  for (int k = 0; k < nchunks && ce == cudaSuccess; ++k) {
    const size_t off = (size_t)cut[k] * plan->elems, len = (size_t)(cut[k + 1] - cut[k]) * plan->elems;
    ce = cudaMemcpyAsync(plan->h_d_in + off, h_in + off, len, cudaMemcpyHostToDevice, s_in);
    if (ce == cudaSuccess) ce = cudaEventRecord(plan->h_events[2 * k], s_in);
  }
  for (int k = 0; k < nchunks && rc == B2D_OK && ce == cudaSuccess; ++k) {
    const int img0 = cut[k], cnt = cut[k + 1] - cut[k];
    const size_t off = (size_t)img0 * plan->elems, len = (size_t)cnt * plan->elems;
    ce = cudaStreamWaitEvent(s_run, plan->h_events[2 * k], 0);
    if (ce != cudaSuccess) break;
    RunCtx c{plan, plan->h_d_in, plan->h_d_out, plan->h_d_ws, nullptr, seed, first_global_image_index, s_run};
    rc = run_range(c, img0, cnt);
    if (rc) break;
    ce = cudaEventRecord(plan->h_events[2 * k + 1], s_run);
    if (ce == cudaSuccess) ce = cudaStreamWaitEvent(s_out, plan->h_events[2 * k + 1], 0);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(h_out + off, plan->h_d_out + off, len, cudaMemcpyDeviceToHost, s_out);
  }
  // always drain all three streams: on an error the copies already queued still target the caller's memory
  for (int i = 0; i < 3; ++i) {
    const cudaError_t e = cudaStreamSynchronize(plan->h_streams[i]);
    if (ce == cudaSuccess) ce = e;
  }
  if (rc) return rc;
  if (ce != cudaSuccess) return fail_cuda(ce, "b2d_run_host");
  return B2D_OK;
}

// ---------------------------------------------------------------------------
// test hook: dump the generator's noise fields
// ---------------------------------------------------------------------------
extern "C" int b2d_noise_field(b2d_plan* plan, float* d_f0, float* d_f1, uint64_t seed,
                               uint64_t first_global_image_index, void* cuda_stream) {
  if (!plan || !d_f0) return fail(B2D_ERR_INVALID, "b2d_noise_field: null argument");
  DeviceGuard guard(plan->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  for (const SpecGroup& g : plan->groups) {
    const b2d_spec& sp = g.spec;
    NoiseArgs na{};
    na.k0 = (uint32_t)seed; na.k1 = (uint32_t)(seed >> 32);
    na.first_global = first_global_image_index;
    na.elems = plan->elems;
    na.dump0 = d_f0; na.dump1 = d_f1;
    int kind;
    switch (sp.noise) {
      case B2D_NOISE_GAUSSIAN: kind = NK_GAUSSIAN; na.p0 = (float)std::sqrt(sp.level); break;
      case B2D_NOISE_SPECKLE: kind = NK_SPECKLE; na.p0 = (float)std::sqrt(sp.level); break;
      case B2D_NOISE_SAP:
        kind = NK_SAP;
        na.u0 = (uint32_t)std::max(0.0, std::min(32768.0, std::ceil(sp.level * 32768.0 - 0.5)));
        if (!d_f1) return fail(B2D_ERR_INVALID, "b2d_noise_field: sap needs d_f1");
        break;
      case B2D_NOISE_UNIFORM: kind = NK_UNIFORM; na.p0 = (float)sp.level; break;
      case B2D_NOISE_RAYLEIGH: kind = NK_RAYLEIGH; na.p0 = (float)sp.level; break;
      case B2D_NOISE_GAMMA:
        kind = NK_GAMMA; na.p0 = (float)sp.level;
        for (const GammaTable& t : plan->gamma)
          if (t.level == sp.level) { na.tab = t.d_tab; na.u0 = (uint32_t)t.folds; }
        break;
      default: return fail(B2D_ERR_UNSUPPORTED, "b2d_noise_field: noise %d has no dumpable field", sp.noise);
    }
    // the operators never read pixels in dump mode except poisson (not dumpable): `in` may be any valid buffer
    ImageSet s = make_set(plan, reinterpret_cast<const uint8_t*>(d_f0), nullptr, g, 0, (int)g.images.size());
    CK(launch_noise_only(kind, NM_DUMP, s, na, st));
  }
  return B2D_OK;
}

// ---------------------------------------------------------------------------
// test hook: the 8-bit Lab conversions of the quant branch
// ---------------------------------------------------------------------------
extern "C" int b2d_lab_convert(const uint8_t* d_in, uint8_t* d_out, size_t n_pixels, int direction, void* cuda_stream) {
  if (!d_in || !d_out) return fail(B2D_ERR_INVALID, "b2d_lab_convert: null argument");
  if (direction != 0 && direction != 1) return fail(B2D_ERR_INVALID, "b2d_lab_convert: direction must be 0 or 1");
  if (n_pixels == 0) return B2D_OK;
  CK(launch_lab_convert(d_in, d_out, n_pixels, direction, static_cast<cudaStream_t>(cuda_stream)));
  return B2D_OK;
}
