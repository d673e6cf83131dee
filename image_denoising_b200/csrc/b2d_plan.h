// b2d_plan.h - host-side plan object behind the opaque b2d_plan handle.
#pragma once
#include "../../include/b200denoise.h"
#include <cuda_runtime.h>
#include <stdint.h>
#include <vector>

namespace b2d {

int fail(int code, const char* fmt, ...);
int fail_cuda(cudaError_t e, const char* what);

// The ABI entry points run on the plan's device and leave the caller's current device as they found it
// (under torch a changed current device silently redirects every later allocation).
struct DeviceGuard {
  int prev = -1;
  cudaError_t err = cudaSuccess;
  explicit DeviceGuard(int device) {
    err = cudaGetDevice(&prev);
    if (err == cudaSuccess && prev != device) err = cudaSetDevice(device);
    else if (err == cudaSuccess) prev = -1;          // nothing to restore
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  bool ok() const { return err == cudaSuccess; }
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;
};

struct SpecGroup {
  b2d_spec spec;
  std::vector<int> images;   // ascending batch indices
  bool contiguous;           // images == [first, first + count)
  int* d_list;               // device copy of `images` when !contiguous
};

struct PeriodicTable {
  double level;              // pi, 100 or B2D_LEVEL_SIZE
  uint8_t* d_tab;            // elems bytes
};

struct GammaTable {
  double level;              // scale
  uint8_t* d_tab;            // 512-byte alias table + fold table (b2d_gamma.cu)
  int folds;
};

}  // namespace b2d

struct b2d_plan {
  int device, n, h, w, rowbytes;
  uint64_t elems;
  std::vector<b2d::SpecGroup> groups;
  std::vector<b2d::PeriodicTable> periodic;
  std::vector<b2d::GammaTable> gamma;
  uint8_t* d_poisson;        // alias tables of the Poisson generator (b2d_poisson.cu), when a group needs them
  // workspace layout (byte offsets into the caller's d_ws)
  size_t ws_bytes, off_par, off_bitmap, off_stage, off_aux, off_scan, off_denoise, off_pix;
  uint32_t brown_chunks;
  bool need_aux;
  // test hook b2d_wavelet_stats: where the last wavelet chunk of the last run left its statistics
  mutable const void* wv_stats_dev;   // WvStats[3 * wv_stats_count] on the device
  mutable int wv_stats_first, wv_stats_count;   // group-relative index of the chunk's first image, images in it
  // b2d_run_host resources (lazily created)
  uint8_t *h_d_in, *h_d_out, *h_d_ws;
  cudaStream_t h_streams[3]; // copy-in, compute, copy-out
  std::vector<cudaEvent_t> h_events;   // per chunk: "input landed", "kernels done"
  bool h_ready;
};
