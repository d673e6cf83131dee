// b2d_kernels.cuh - launch-side declarations shared by the .cu files.
#pragma once
#include "b2d_noise.cuh"
#include "b2d_filter3.cuh"

namespace b2d {

// one launch = one group of images that share a spec
struct ImageSet {
  const uint8_t* in;     // batch base (device), image i at in + i * elems
  uint8_t* out;          // batch base (device)
  int h, w;
  int rowbytes;          // 3 * w
  uint64_t elems;        // h * w * 3
  int count;             // images in this launch
  int first;             // batch index of the first image when list == nullptr
  const int* list;       // optional device list of batch indices (mixed batches)
};

__device__ __forceinline__ int image_of(const ImageSet& s, int z) { return s.list ? s.list[z] : s.first + z; }

// thread-local launch counter (bench.py's gpu_launches)
void count_launch();
// diagnostic environment knob (the list is in include/b200denoise.h): getenv(name) when
// B2D_DEBUG_KNOBS=1 is set in the environment, nullptr otherwise - a production process cannot be
// steered onto a cross-check kernel by a stray variable
const char* knob(const char* name);
// current device index (cudaGetDevice) and its SM count (cached per device)
int current_device();
int sm_count();

// noise only (no denoiser), all modes
cudaError_t launch_noise_only(int kind, int mode, const ImageSet& s, const NoiseArgs& na, cudaStream_t st);
// fused noise -> ->u8 -> 3x3 filter (generator mode or NK_NONE)
cudaError_t launch_fused3(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st);

}  // namespace b2d
