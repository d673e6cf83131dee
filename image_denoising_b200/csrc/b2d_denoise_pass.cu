// b2d_denoise_pass.cu - dispatch of the stand-alone denoiser passes
// (median 5/7, bilateral) that read an already-noised uint8 image.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

namespace b2d {

cudaError_t launch_median_k(const ImageSet& s, int k, cudaStream_t st);
int run_bilateral(const b2d_spec& sp, const ImageSet& s, uint8_t* ws_tables, cudaStream_t st, bool noisy_input);

constexpr size_t kBilateralTableBytes = sizeof(float) * (768 + 15 * 15 + 31 + 10 * 768);   // lut, space, pad, lut2

size_t denoise_workspace_bytes(const b2d_plan& p) {
  size_t extra = 0;
  for (const SpecGroup& g : p.groups)
    if (g.spec.denoise == B2D_DENOISE_BILATERAL) extra = kBilateralTableBytes;
  return extra + wavelet_workspace_bytes(p);
}

int run_denoise_pass(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, uint8_t* ws, cudaStream_t st, bool noisy_input) {
  if (sp.denoise == B2D_DENOISE_MEDIAN) {
    const cudaError_t e = launch_median_k(s, sp.ksize, st);
    return e == cudaSuccess ? B2D_OK : fail_cuda(e, "median launch");
  }
  if (sp.denoise == B2D_DENOISE_BILATERAL) return run_bilateral(sp, s, ws + p.off_denoise, st, noisy_input);
  return fail(B2D_ERR_UNSUPPORTED, "denoiser %d cannot run as a uint8 pass", sp.denoise);
}

}  // namespace b2d
