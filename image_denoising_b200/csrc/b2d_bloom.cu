// b2d_bloom.cu - `bloom`: Automold.add_sun_flare(img, flare_center=(100,100), angle=-pi/4)
// (call at lib/roi_data_layer/minibatch.py:1487; tools/Automold.py:553-627).
//
// The reference draws 8 random filled circles on an `overlay` copy, each followed by a
// FULL-IMAGE cv2.addWeighted(overlay, a, output, 1-a, 0, output); then 40 concentric white
// circles (radius linspace(1,400,40), alpha = linspace(0,1,40)[39-i]^3) on a fresh overlay,
// each again followed by a full-image blend: 48 passes over the image.
// Here every pixel replays the 48 (inside?, alpha) steps in registers: one pass over HBM.
//   blend  : cv2 4.x addWeighted on uint8 = rint(fma(src1, (float)a, src2 * (float)b)), probed
//            bit-exact against cv2 4.13
//   circle : cv2.circle(..., thickness=-1) is the midpoint algorithm of cv::Circle: row cy+-dy gets
//            the span cx+-dx and row cy+-dx the span cx+-dy for every step (dx, dy); the host turns
//            a radius into a half-width-per-row table with that same loop.
#include "b2d_denoise.cuh"
#include "b2d_plan.h"

#include <cmath>
#include <vector>

namespace b2d {

constexpr int BLOOM_RANDOM = 8;       // no_of_flare_circles
constexpr int BLOOM_SOURCE = 40;      // src_radius // 10
constexpr int BLOOM_SRC_RADIUS = 400;

struct BloomCircle {
  float alpha, beta;      // (float)a, (float)(1 - a)
  int cx, cy, r;
  int hw_off;             // offset of the r + 1 half widths in the table
  float col[3];           // BGR order of the array (the reference passes src_color as is)
  int pad;
};

// half width of the filled circle per |row - cy| (cv::Circle, imgproc/src/drawing.cpp)
static void circle_halfwidths(int radius, std::vector<int>& hw) {
  hw.assign((size_t)radius + 1, -1);
  int err = 0, dx = radius, dy = 0, plus = 1, minus = (radius << 1) - 1;
  while (dx >= dy) {
    if (dx > hw[dy]) hw[dy] = dx;       // rows cy +- dy : cx - dx .. cx + dx
    if (dy > hw[dx]) hw[dx] = dy;       // rows cy +- dx : cx - dy .. cx + dy
    dy++;
    err += plus;
    plus += 2;
    const int mask = (err <= 0) - 1;
    err -= minus & mask;
    dx += mask;
    minus -= mask & 2;
  }
}

__device__ __forceinline__ float blend_u8(float ov, float out, float alpha, float beta) {
  // rint via the 1.5 * 2^23 magic number (round half to even, |x| <= 255)
  const float t = fmaf(ov, alpha, __fmul_rn(out, beta));
  return __fadd_rn(__fadd_rn(t, 12582912.0f), -12582912.0f);
}

__global__ void __launch_bounds__(256) k_bloom(ImageSet s, const BloomCircle* circles /* [batch][8] then [40] */,
                                               const int* hw, int n_batch) {
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  const BloomCircle* rc = circles + (size_t)img_local * BLOOM_RANDOM;
  const BloomCircle* sc = circles + (size_t)n_batch * BLOOM_RANDOM;
  const int npix = s.h * s.w;
  for (int p = blockIdx.x * 256 + threadIdx.x; p < npix; p += gridDim.x * 256) {
    const int y = p / s.w, x = p - y * s.w;
    float ov[3], o[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) ov[c] = o[c] = (float)img[(size_t)p * 3 + c];
    for (int k = 0; k < BLOOM_RANDOM; ++k) {             // add_sun_process, Automold.py:579-584
      const BloomCircle c = rc[k];
      const int ady = abs(y - c.cy), adx = abs(x - c.cx);
      if (ady <= c.r && adx <= hw[c.hw_off + ady]) { ov[0] = c.col[0]; ov[1] = c.col[1]; ov[2] = c.col[2]; }
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) o[ch] = blend_u8(ov[ch], o[ch], c.alpha, c.beta);
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) ov[ch] = o[ch];       // flare_source: overlay = image.copy()
    {
      const int ady = abs(y - sc[0].cy), adx = abs(x - sc[0].cx);
      bool inside = false;                               // the concentric circles are nested
      for (int k = 0; k < BLOOM_SOURCE; ++k) {           // flare_source, Automold.py:559-562
        const BloomCircle c = sc[k];
        if (!inside && ady <= c.r && adx <= hw[c.hw_off + ady]) {
          inside = true;
          ov[0] = c.col[0]; ov[1] = c.col[1]; ov[2] = c.col[2];
        }
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) o[ch] = blend_u8(ov[ch], o[ch], c.alpha, c.beta);
      }
    }
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) out[(size_t)p * 3 + ch] = (uint8_t)(int)o[ch];
  }
}

// the random draws of add_sun_process for image `gi` (Automold.py:579-582) from Philox
static void host_draws(uint64_t seed, uint64_t gi, int h, int w, double* d /* [8][7] */) {
  const double two_pi = 2.0 * M_PI;
  double ang = std::fmod(-M_PI / 4.0, two_pi);
  if (ang != 0.0 && ang < 0.0) ang += two_pi;            // Python's float %
  const double t = std::tan(ang);
  const int nline = (w + 9) / 10;                         // range(0, W, 10)
  const int rad_hi = h / 100 - 2;                         // randint(1, H//100 - 2), validated >= 1
  for (int k = 0; k < BLOOM_RANDOM; ++k) {
    const uint4 r0 = philox4x32((uint32_t)k, 0xB100u, (uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)seed,
                                   (uint32_t)(seed >> 32));
    const uint4 r1 = philox4x32((uint32_t)k, 0xB101u, (uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)seed,
                                   (uint32_t)(seed >> 32));
    const double u = ((double)r0.x + 0.5) / 4294967296.0;
    const int idx = (int)(((uint64_t)r0.y * (uint64_t)nline) >> 32);
    const int rad = 1 + (int)(((uint64_t)r0.z * (uint64_t)rad_hi) >> 32);
    const double rx = 10.0 * idx;
    const double ry = t * (rx - 100.0) + 100.0;           // add_sun_flare_line, Automold.py:569-571
    d[k * 7 + 0] = 0.05 + (0.2 - 0.05) * u;               // random.uniform(0.05, 0.2)
    d[k * 7 + 1] = (double)(int)rx;
    d[k * 7 + 2] = (double)(int)(2.0 * 100.0 - ry);
    d[k * 7 + 3] = (double)(rad * rad * rad);
    d[k * 7 + 4] = 205.0 + (double)(((uint64_t)r0.w * 51u) >> 32);   // randint(205, 255)
    d[k * 7 + 5] = 205.0 + (double)(((uint64_t)r1.x * 51u) >> 32);
    d[k * 7 + 6] = 205.0 + (double)(((uint64_t)r1.y * 51u) >> 32);
  }
}

size_t bloom_table_bytes(int n, int h) {
  const int rad_hi = std::max(h / 100 - 2, 1);
  const size_t rmax = (size_t)rad_hi * rad_hi * rad_hi;
  const size_t circles = ((size_t)n * BLOOM_RANDOM + BLOOM_SOURCE) * sizeof(BloomCircle);
  const size_t tab = ((size_t)n * BLOOM_RANDOM * (rmax + 1) + (size_t)BLOOM_SOURCE * (BLOOM_SRC_RADIUS + 1)) * sizeof(int);
  return circles + tab + 512;
}

// s.out receives the bloomed uint8 images.  tables: device scratch of bloom_table_bytes().
int run_bloom(const b2d_plan& p, const ImageSet& s, const SpecGroup& g, int lo, int hi, uint8_t* tables,
              const b2d_fields* fields, uint64_t seed, uint64_t first_global, cudaStream_t st) {
  const int n = p.n;
  std::vector<BloomCircle> circles((size_t)n * BLOOM_RANDOM + BLOOM_SOURCE);
  std::vector<int> table, hw;
  std::vector<double> draws((size_t)n * BLOOM_RANDOM * 7, 0.0);
  if (fields && fields->f0) {      // injected draws live on the device: fetch them (test path, synchronous)
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess)
      e = cudaMemcpy(draws.data(), fields->f0, draws.size() * sizeof(double), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return fail_cuda(e, "bloom draws");
  } else {
    for (int z = lo; z < hi; ++z) host_draws(seed, first_global + (uint64_t)g.images[z], p.h, p.w, &draws[(size_t)g.images[z] * 56]);
  }
  for (int z = lo; z < hi; ++z) {
    const int b = g.images[z];
    for (int k = 0; k < BLOOM_RANDOM; ++k) {
      const double* d = &draws[((size_t)b * BLOOM_RANDOM + k) * 7];
      BloomCircle& c = circles[(size_t)b * BLOOM_RANDOM + k];
      c.alpha = (float)d[0];
      c.beta = (float)(1.0 - d[0]);
      c.cx = (int)d[1]; c.cy = (int)d[2]; c.r = (int)d[3];
      if (c.r < 0 || c.r > (1 << 20)) return fail(B2D_ERR_INVALID, "bloom: circle radius %d out of range", c.r);
      c.col[0] = (float)d[4]; c.col[1] = (float)d[5]; c.col[2] = (float)d[6];
      c.pad = 0;
      circle_halfwidths(c.r, hw);
      c.hw_off = (int)table.size();
      table.insert(table.end(), hw.begin(), hw.end());
    }
  }
  for (int k = 0; k < BLOOM_SOURCE; ++k) {                 // flare_source(output, (100,100), 400, (255,255,255))
    BloomCircle& c = circles[(size_t)n * BLOOM_RANDOM + k];
    const double a_step = 1.0 / (BLOOM_SOURCE - 1), r_step = (BLOOM_SRC_RADIUS - 1.0) / (BLOOM_SOURCE - 1);
    const int j = BLOOM_SOURCE - k - 1;
    const double a = j == BLOOM_SOURCE - 1 ? 1.0 : j * a_step;                     // np.linspace(0, 1, 40)[j]
    const double rad = k == BLOOM_SOURCE - 1 ? (double)BLOOM_SRC_RADIUS : k * r_step + 1.0;  // np.linspace(1, 400, 40)[k]
    const double alp = a * a * a;
    c.alpha = (float)alp;
    c.beta = (float)(1.0 - alp);
    c.cx = 100; c.cy = 100; c.r = (int)rad;
    c.col[0] = c.col[1] = c.col[2] = 255.0f;
    c.pad = 0;
    circle_halfwidths(c.r, hw);
    c.hw_off = (int)table.size();
    table.insert(table.end(), hw.begin(), hw.end());
  }
  const size_t cbytes = circles.size() * sizeof(BloomCircle);
  if (cbytes + table.size() * sizeof(int) > bloom_table_bytes(n, p.h))
    return fail(B2D_ERR_INVALID, "bloom: circle tables exceed the workspace (radius larger than (H//100-2)^3?)");
  BloomCircle* d_c = reinterpret_cast<BloomCircle*>(tables);
  int* d_hw = reinterpret_cast<int*>(tables + cbytes);
  cudaError_t e = cudaMemcpyAsync(d_c, circles.data(), cbytes, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_hw, table.data(), table.size() * sizeof(int), cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);    // the host vectors die at return
  if (e != cudaSuccess) return fail_cuda(e, "bloom tables");
  const unsigned bx = (unsigned)std::min<size_t>(((size_t)p.h * p.w + 255) / 256, 148 * 8);
  k_bloom<<<dim3(bx, 1, s.count), 256, 0, st>>>(s, d_c, d_hw, n);
  count_launch();
  e = cudaGetLastError();
  return e == cudaSuccess ? B2D_OK : fail_cuda(e, "bloom launch");
}

}  // namespace b2d
