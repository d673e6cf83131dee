// b2d_wavelet.cuh - geometry / statistics records of the wavelet denoiser.
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace b2d {

constexpr int WV_MAXL = 8;      // levels
constexpr int WV_MAXF = 10;     // filter taps (bior1.5)
constexpr int WV_BINS = 2048;   // radix-select histogram bins per pass

struct WvGeom {
  int F, L;
  int h[WV_MAXL + 1], w[WV_MAXL + 1];           // band sizes per level; [0] = image
  float dec_lo[WV_MAXF], dec_hi[WV_MAXF], rec_lo[WV_MAXF], rec_hi[WV_MAXF];
  // float offsets inside one channel's workspace block
  size_t off_plane, off_tlo, off_thi, off_app[WV_MAXL + 1], off_det[WV_MAXL + 1][3];
  size_t chan_floats;
};

struct WvStats {                 // one per (image in chunk, channel)
  int mn_ord, mx_ord;            // ordered-int min / max of the channel
  double energy[WV_MAXL + 1][3]; // sum d^2 per level / band (0: 'da', 1: 'ad', 2: 'dd')
  unsigned int prefix[2];        // radix select state for ranks (n-1)/2 and n/2
  unsigned long long rank[2];
  unsigned long long count;      // non-zero |dd_1| coefficients
  float sigma;
  float thr[WV_MAXL + 1][3];
  float lo, hi;                  // clip range = channel min / max
};

void wavelet_geometry(WvGeom& g, int wavelet, int levels_req, int h, int w);

#ifdef __CUDACC__
// order-preserving float <-> int map for atomicMin / atomicMax
__device__ __forceinline__ int f2ord(float f) {
  const int i = __float_as_int(f);
  return i ^ ((i >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ord2f(int o) { return __int_as_float(o ^ ((o >> 31) & 0x7fffffff)); }

// pywt.threshold(d, thr, 'soft') = d * max(1 - thr / |d|, 0) = sign(d) * max(|d| - thr, 0): the
// second form needs no division (3 instructions) and differs from the first by rounding only
__device__ __forceinline__ float wv_soft(float d, float thr) { return copysignf(fmaxf(fabsf(d) - thr, 0.0f), d); }

// shared between the general path (b2d_wavelet.cu) and the Haar fast path (b2d_haar.cu)
void wavelet_stats_begin(WvStats* stats, unsigned int* hist, int planes, cudaStream_t st);
// median(|dd_1| != 0) -> sigma, then the BayesShrink thresholds; dd_1 of plane z at ws + z * g.chan_floats + g.off_det[1][2]
// first_pass = 1: the histogram of radix pass 0 (bits 30..20, hist[plane][0][*]) has been filled by the caller
void wavelet_sigma_thresholds(const WvGeom& g, const float* ws, WvStats* stats, unsigned int* hist, int planes,
                              cudaStream_t st, int first_pass = 0);
#endif

}  // namespace b2d
