// b2d_todo.cu - entry points whose kernels are not written yet fail loudly.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"
namespace b2d {
int run_pixel_noise(const b2d_plan&, const b2d_spec& sp, const ImageSet&, uint8_t*, const b2d_fields*, uint64_t,
                    uint64_t, cudaStream_t) {
  return fail(B2D_ERR_UNSUPPORTED, "noise %d not implemented yet", sp.noise);
}
}  // namespace b2d
