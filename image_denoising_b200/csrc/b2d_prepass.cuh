// b2d_prepass.cuh - declarations for the per-image pre-passes.
#pragma once
#include "b2d_kernels.cuh"

namespace b2d {

struct BrownianArgs {
  double sqrt_dt;
  double* chunk_sums;   // [batch][nchunks] workspace
  uint32_t nchunks;
  uint8_t* field;       // [batch][elems] u8 noise field (batch-indexed)
  // non-null: the field kernel writes cv2.add(image, field) (saturating) here instead of the field - the corrupted
  // image in one step, for the consumers that are not the fused 3x3 kernels ([batch][elems], batch-indexed like s.in)
  uint8_t* add_out;
};

uint32_t brownian_chunks(uint64_t elems);
// generator-mode Poisson through per-`vals` alias tables (b2d_poisson.cu); img_par = vals per image
cudaError_t upload_poisson_tables(uint8_t** d_tables);
// 256 probabilities -> alias row (alias | threshold << 8), probabilities rounded to 2^-16
void alias_row_256(const double* P, uint16_t* row);
// generator-mode gamma: alias table over K mod 256 + fold table for the test hook (b2d_gamma.cu)
int gamma_folds(double scale);
cudaError_t upload_gamma_table(double scale, uint8_t** d_table);
cudaError_t launch_poisson(const ImageSet& s, const NoiseArgs& na, const uint8_t* tables, cudaStream_t st);
cudaError_t launch_poisson_vals(const ImageSet& s, uint32_t* bitmaps, float* vals, cudaStream_t st);
cudaError_t launch_periodic_table(uint8_t* tab, uint64_t n, double lim, cudaStream_t st);
cudaError_t launch_brownian_field(const ImageSet& s, const NoiseArgs& na, const BrownianArgs& ba, bool inject,
                                  cudaStream_t st);

}  // namespace b2d
