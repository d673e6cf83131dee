// b2d_todo.cu - entry points whose kernels are not written yet fail loudly.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
namespace b2d {
size_t wavelet_workspace_bytes(const b2d_plan&) { return 0; }
int run_pixel_noise(const b2d_plan&, const b2d_spec& sp, const ImageSet&, uint8_t*, const b2d_fields*, uint64_t,
                    uint64_t, cudaStream_t) {
  return fail(B2D_ERR_UNSUPPORTED, "noise %d not implemented yet", sp.noise);
}
int run_wavelet(const b2d_plan&, const b2d_spec&, const ImageSet&, int, int, const NoiseArgs&, uint8_t*, cudaStream_t) {
  return fail(B2D_ERR_UNSUPPORTED, "wavelet denoiser not implemented yet");
}
}  // namespace b2d
