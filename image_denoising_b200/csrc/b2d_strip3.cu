// b2d_strip3.cu - noise -> ->u8 -> 3x3 denoiser, one WARP per column strip: no shared memory, no barrier.
//
// Replaces the same reference code as b2d_fused3w.cu (lib/roi_data_layer/minibatch.py:114-160:
// random_noise / np.random -> (255 x).astype(uint8) -> cv2.GaussianBlur / blur).
//
// Why a third schedule: the tile kernel (b2d_fused3w.cu) runs load -> generator -> barrier -> filter -> store as
// PHASES, and its CTAs move through them together - ncu shows the ALU pipe throttling in one phase while the warps
// wait on the long scoreboard or the barrier in the next (issue slots 54 % / 69 % used, no pipe above 62 %).  Here
// every warp is on its own and its instruction stream mixes all of it, row by row:
//   lane  : owns ONE 16-byte column (two 8-byte Philox groups) of a strip of 32 columns and marches down a segment
//           of rows; the raw bytes of the row ST_MR rows ahead are already in flight in registers;
//   noise : packed operators (b2d_noise.cuh) on the row's 16 bytes, in registers;
//   taps  : the word left / right of the lane's 16 bytes comes from the neighbouring lane by SHFL; lanes 0 and 31 are
//           the strip's apron (generated, never stored: 2 / 32 of the generator work), the image's border pixel is
//           built from the lane's own bytes;
//   filter: row part (b2d_filter3w.cuh) combined with the two previous parts kept in registers; 2 x STG.64.
// Only the first two rows of a segment are generated twice.  Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_kernels.cuh"
#include "b2d_filter3w.cuh"

#include <algorithm>

namespace b2d {

#ifndef B2D_STRIP_MINB
#define B2D_STRIP_MINB 8
#endif
constexpr int ST_WARPS = 4;                 // independent warps per CTA
constexpr int ST_OUT = 30;                  // columns a strip writes (lanes 1 .. 30)
constexpr int ST_MR = 3;                    // prefetch distance in rows; multiple of 3 so the part rotation closes

struct StripGeom {
  int upr;          // 16-byte columns per row
  int strips;       // strips per row
  int segs, seg_h;  // segments per image, rows per segment
  unsigned items;   // images * segs * strips
};

template <int KIND, int FILTER>
__global__ void __launch_bounds__(32 * ST_WARPS, B2D_STRIP_MINB) k_strip3(ImageSet s, NoiseArgs na, StripGeom g) {
  using F = Filter3W<FILTER, 4>;
  const unsigned item = blockIdx.x * ST_WARPS + (threadIdx.x >> 5);
  if (item >= g.items) return;                       // warp uniform; nothing below synchronises the CTA
  const int lane = threadIdx.x & 31;
  const unsigned per_img = (unsigned)(g.segs * g.strips);
  const int z = (int)(item / per_img), rem = (int)(item - (unsigned)z * per_img);
  const int seg = rem / g.strips, strip = rem - seg * g.strips;   // the warps of a CTA take neighbouring strips
  const int img_local = image_of(s, z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;

  const int u = strip * ST_OUT - 1 + lane;           // column of this lane
  const bool valid = u >= 0 && u < g.upr;
  const int c = valid ? 16 * u : 0;                  // idle lanes shadow column 0 (loads only)
  const bool v1 = c + 8 < s.rowbytes;                // second 8-byte group of the column is inside the row
  const bool first = u == 0, last = u == g.upr - 1;
  const bool writer = valid && lane >= 1 && lane <= ST_OUT;
  constexpr uint32_t kSelL = F::BORDER == BORDER_REPLICATE ? 0x2100 : 0x5430;   // [x, p0, p1, p2] of pixel 0 / 1
  constexpr uint32_t kSelR = F::BORDER == BORDER_REPLICATE ? 0x0765 : 0x0432;   // pixel W-1 / W-2 of the last 8 bytes
  const int top_src = F::BORDER == BORDER_REPLICATE ? 0 : 1;
  const int bot_src = F::BORDER == BORDER_REPLICATE ? s.h - 1 : s.h - 2;

  const int ya = seg * g.seg_h, yb = min(s.h, ya + g.seg_h);
  const int nv = yb - ya + 2;                        // virtual rows j = 0 .. nv-1  <->  image rows ya-1 .. yb
  uint8_t* optr = s.out + (uint64_t)img_local * s.elems + (uint64_t)ya * s.rowbytes + c;
  auto row_off = [&](int j) -> uint32_t {            // rows past the segment reload its last row
    j = min(j, nv - 1);
    int y = ya - 1 + j;
    y = y < 0 ? top_src : (y >= s.h ? bot_src : y);
    return (uint32_t)y * (uint32_t)s.rowbytes + (uint32_t)c;
  };
  uint2 ra[ST_MR], rb[ST_MR];
  auto fetch = [&](int k, int j) {
    const uint8_t* p = img + row_off(j);
    ra[k] = __ldg(reinterpret_cast<const uint2*>(p));
    rb[k] = __ldg(reinterpret_cast<const uint2*>(p + (v1 ? 8 : 0)));
  };
#pragma unroll
  for (int k = 0; k < ST_MR; ++k) fetch(k, k);

  typename F::Part p0, p1, p2;
  for (int j0 = 0; j0 < nv; j0 += ST_MR) {
#pragma unroll
    for (int k = 0; k < ST_MR; ++k) {
      const int j = j0 + k;
      if (j >= nv) break;                            // warp uniform
      const uint2 a = ra[k], b = rb[k];
      const uint32_t e = row_off(j);
      fetch(k, j + ST_MR);
      uint2 n0 = a, n1 = b;
      if (valid) {
        GroupCtx g0{(uint32_t)img_local, e >> 3, img};
        n0 = PackedNoise<KIND>::apply(na, g0, a);
        if (v1) {
          GroupCtx g1{(uint32_t)img_local, (e >> 3) + 1u, img};
          n1 = PackedNoise<KIND>::apply(na, g1, b);
        }
      }
      uint32_t w[6];
      w[1] = n0.x; w[2] = n0.y; w[3] = n1.x; w[4] = n1.y;
      w[0] = __shfl_up_sync(0xffffffffu, n1.y, 1);
      w[5] = __shfl_down_sync(0xffffffffu, n0.x, 1);
      if (first) w[0] = __byte_perm(n0.x, n0.y, kSelL);
      if (last) {
        if (v1) w[5] = __byte_perm(n1.x, n1.y, kSelR);
        else w[3] = __byte_perm(n0.x, n0.y, kSelR);
      }
      typename F::Part& pb = (k % 3 == 0) ? p0 : (k % 3 == 1) ? p1 : p2;
      typename F::Part& pt = (k % 3 == 0) ? p1 : (k % 3 == 1) ? p2 : p0;
      typename F::Part& pm = (k % 3 == 0) ? p2 : (k % 3 == 1) ? p0 : p1;
      RowLanes<4> rl;
      rl.split(w);
      F::hpart(rl, pb);
      if (j >= 2) {                                  // rows j-2, j-1, j -> output row ya + j - 2
        uint32_t o[4];
        F::vcomb(pt, pm, pb, o);
        if (writer) {
          *reinterpret_cast<uint2*>(optr) = make_uint2(o[0], o[1]);
          if (v1) *reinterpret_cast<uint2*>(optr + 8) = make_uint2(o[2], o[3]);
        }
        optr += s.rowbytes;
      }
    }
  }
}

template <int KIND, int FILTER>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  StripGeom g;
  g.upr = (s.rowbytes + 15) / 16;
  g.strips = (g.upr + ST_OUT - 1) / ST_OUT;
  static int occ_cache[64] = {};                                 // [device] -> CTAs per SM
  int dev = 0;
  cudaGetDevice(&dev);
  int& occ = occ_cache[dev & 63];
  if (!occ) {
    const cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_strip3<KIND, FILTER>, 32 * ST_WARPS, 0);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
  }
  const long slots = (long)sm_count() * occ * ST_WARPS;          // resident warps on the device
  // segments per image: short enough that the last wave of warps is a small part of the run, long enough that the
  // two regenerated rows per segment stay cheap
  static const char* seg_env = knob("B2D_STRIP_SEG");
  int best_segs = 1;
  double best = -1.0;
  for (int segs = 1; segs <= std::max(1, s.h / 16); ++segs) {
    const int seg_h = (s.h + segs - 1) / segs;
    if ((s.h + seg_h - 1) / seg_h != segs) continue;
    if (seg_env && atoi(seg_env) > 0) {
      if (seg_h <= atoi(seg_env)) { best_segs = segs; break; }
      continue;
    }
    const long items = (long)s.count * segs * g.strips;
    const double waves = (double)items / (double)slots;
    const double eff = waves / (waves + 0.5) * (double)seg_h / (double)(seg_h + 2);   // half a wave of tail on average
    if (eff > best + 1e-9) { best = eff; best_segs = segs; }
  }
  g.segs = best_segs;
  g.seg_h = (s.h + best_segs - 1) / best_segs;
  g.items = (unsigned)((long)s.count * g.segs * g.strips);
  const unsigned grid = (g.items + ST_WARPS - 1) / ST_WARPS;
  k_strip3<KIND, FILTER><<<grid, 32 * ST_WARPS, 0, st>>>(s, na, g);
  count_launch();
  return cudaGetLastError();
}

template <int KIND>
static cudaError_t launch_filter(int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (filter) {
    case FK_GAUS3: return launch_one<KIND, FK_GAUS3>(s, na, st);
    case FK_MEAN3: return launch_one<KIND, FK_MEAN3>(s, na, st);
  }
  return cudaErrorInvalidValue;
}

bool strip3_supported(int kind, int filter, const ImageSet& s) {
  return kind != NK_NONE && (filter == FK_GAUS3 || filter == FK_MEAN3) && s.rowbytes % 8 == 0 && s.rowbytes >= 24 &&
         s.h >= 2 && s.elems < (1ull << 31) && (uint64_t)s.count * s.h * ((s.rowbytes + 479) / 480) < (1ull << 31) &&
         ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 7u) == 0;
}

cudaError_t launch_strip3(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (kind) {
    case NK_GAUSSIAN: return launch_filter<NK_GAUSSIAN>(filter, s, na, st);
    case NK_SPECKLE: return launch_filter<NK_SPECKLE>(filter, s, na, st);
    case NK_SAP: return launch_filter<NK_SAP>(filter, s, na, st);
    case NK_POISSON: return launch_filter<NK_POISSON>(filter, s, na, st);
    case NK_UNIFORM: return launch_filter<NK_UNIFORM>(filter, s, na, st);
    case NK_RAYLEIGH: return launch_filter<NK_RAYLEIGH>(filter, s, na, st);
    case NK_GAMMA: return launch_filter<NK_GAMMA>(filter, s, na, st);
    case NK_TABLE_ADD: return launch_filter<NK_TABLE_ADD>(filter, s, na, st);
    case NK_SHADER: return launch_filter<NK_SHADER>(filter, s, na, st);
  }
  return cudaErrorInvalidValue;
}

}  // namespace b2d
