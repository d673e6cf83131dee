// b2d_denoise.cuh - denoisers that run as their own pass over a uint8 image
// (input already noised): the any-width 3x3 fallback, median 5/7, bilateral,
// wavelet.  The aligned 3x3 filters live in b2d_fused3.cu.
#pragma once
#include "b2d_kernels.cuh"
#include "../../include/b200denoise.h"
#include "b2d_plan.h"

namespace b2d {

size_t denoise_workspace_bytes(const b2d_plan& p);
size_t wavelet_workspace_bytes(const b2d_plan& p);

// 3x3 filters for images whose row size is not a multiple of 4 bytes (W % 4 != 0)
cudaError_t launch_filter3_bytes(int filter, const ImageSet& s, cudaStream_t st);   // one thread per byte
// the same through re-aligning shared-memory tiles (b2d_filter3_any.cu): the path that runs
cudaError_t launch_filter3_any(int filter, const ImageSet& s, cudaStream_t st);

}  // namespace b2d

namespace b2d {
// median 5/7, bilateral (uint8 in -> uint8 out).  noisy_input: the image comes out of a noise pass (bilateral then clamps
// its colour-LUT index at the first all-zero entry: same bits, no shared-memory conflicts on the zero-weight taps)
int run_denoise_pass(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, uint8_t* ws, cudaStream_t st,
                     bool noisy_input = true);
// quant / bloom: pixel-granular corruption written as a uint8 image (s.out); `tables` = plan scratch
size_t quant_table_bytes(int n);
size_t bloom_table_bytes(int n, int h);
int run_quant(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, uint8_t* tables, const b2d_fields* fields,
              uint64_t seed, uint64_t first_global, cudaStream_t st);
int run_bloom(const b2d_plan& p, const ImageSet& s, const SpecGroup& g, int lo, int hi, uint8_t* tables,
              const b2d_fields* fields, uint64_t seed, uint64_t first_global, cudaStream_t st);
cudaError_t launch_lab_convert(const uint8_t* in, uint8_t* out, size_t n, int dir, cudaStream_t st);
// noise (operator `kind`, float where the reference keeps float) -> BayesShrink wavelet -> ->u8
int run_wavelet(const b2d_plan& p, const b2d_spec& sp, const ImageSet& s, int kind, int mode, const NoiseArgs& na,
                uint8_t* ws, cudaStream_t st);
// db1 / 3 levels / H, W % 8 == 0 / generator mode: register-resident two-pass path (b2d_haar.cu)
struct WvGeom;
bool haar_supported(const b2d_plan& p, const b2d_spec& sp, int mode);
int run_haar(const b2d_plan& p, const ImageSet& s, int kind, const NoiseArgs& na, const WvGeom& g_full, uint8_t* wbase,
             size_t avail, const float (&inv)[3][3], cudaStream_t st);
}  // namespace b2d
