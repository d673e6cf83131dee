// b2d_march3.cu - noise -> ->u8 -> 3x3 denoiser, ROW-MARCHING persistent kernel: the hot path
// for images whose rows fit one CTA (3 W <= 4096 bytes, rows a multiple of 8 bytes).
//
// Replaces the same reference code as b2d_fused3w.cu (lib/roi_data_layer/minibatch.py:114-160:
// random_noise / np.random -> (255 x).astype(uint8) -> cv2.GaussianBlur / blur / medianBlur).
//
// A work item is a horizontal segment of one image (seg_h rows, all columns).  A thread owns ONE
// 16-byte column of the image for the whole segment and marches down it in chunks of MR rows:
//   prefetch : the raw bytes of the NEXT chunk are loaded into registers (2 x LDG.64 per row)
//              before the current chunk is processed - global latency hides behind compute;
//   noise    : packed operators (b2d_noise.cuh) on the current chunk's registers, STS.128 into
//              one half of a double buffer;
//   barrier  : ONE __syncthreads per chunk (double buffer, proof in DESIGN.md section 3);
//   filter   : the thread re-reads its 16 bytes + one word each side (the first / last column
//              builds the image's border pixel from its own bytes), forms the row part
//              (b2d_filter3w.cuh) and combines it with the two previous parts kept in
//              registers across chunks - no vertical apron inside a segment, no band overlap.
// Only the first two rows of a segment are generated twice (2 / seg_h of the generator work).
// Algorithmic HBM traffic: 3 B read + 3 B written per pixel.
#include "b2d_kernels.cuh"
#include "b2d_filter3w.cuh"

#include <algorithm>

namespace b2d {

#ifndef B2D_MR
#define B2D_MR 6
#endif
#ifndef B2D_MARCH_MINB
#define B2D_MARCH_MINB 3
#endif
constexpr int MR = B2D_MR;                  // rows per chunk; multiple of 3 so the part rotation closes
constexpr int MARCH_MAX_THREADS = 256;

struct MarchGeom {
  int upr;        // 16-byte units per row = threads that own a column
  int srow;       // shared row stride: 16 B apron + upr * 16 + 16 B apron
  int segs;       // segments per image
  int seg_h;      // rows per segment
  int items;      // images * segs
};

// A4: rows / image bases only 4-byte aligned (rows of 4 (mod 8) bytes, W = 500), no generator: the same
// march with 32-bit global accesses; the last column then owns 4, 8, 12 or 16 valid bytes.
template <int KIND, int FILTER, bool A4 = false>
__global__ void __launch_bounds__(MARCH_MAX_THREADS, FILTER == FK_MEDIAN3 ? 2 : B2D_MARCH_MINB) k_march3(ImageSet s, NoiseArgs na, MarchGeom g) {
  static_assert(!A4 || KIND == NK_NONE, "the 4-byte aligned variant has no generator");
  using F = Filter3W<FILTER, 4>;
  extern __shared__ __align__(16) uint8_t smem[];   // [2][MR][srow]
  const int tid = threadIdx.x;
  const bool active = tid < g.upr;
  const int c = active ? 16 * tid : 0;               // idle threads shadow column 0 (loads only)
  const bool v1 = c + 8 < s.rowbytes;                // second 8-byte group of the unit is inside the row
  const bool first = tid == 0, last = tid == g.upr - 1;
  const int nwv = min(4, (s.rowbytes - c) >> 2);     // A4: valid words of the unit (re-read the last one beyond)
  const int o1 = nwv > 1 ? 1 : 0, o2 = nwv > 2 ? 2 : nwv - 1, o3 = nwv - 1;
  uint8_t* const my = smem + 16 + 16 * tid;          // this thread's unit in row 0 of buffer 0
  // border rows of the image (cv2 REFLECT_101: -1 -> 1, h -> h-2; REPLICATE: -1 -> 0, h -> h-1)
  const int top_src = F::BORDER == BORDER_REPLICATE ? 0 : 1;
  const int bot_src = F::BORDER == BORDER_REPLICATE ? s.h - 1 : s.h - 2;
  constexpr uint32_t kSelL = F::BORDER == BORDER_REPLICATE ? 0x2100 : 0x5430;   // [x, p0, p1, p2] of pixel 0 / 1
  constexpr uint32_t kSelR = F::BORDER == BORDER_REPLICATE ? 0x0765 : 0x0432;   // pixel W-1 / W-2 of the last 8 bytes

  for (int item = blockIdx.x; item < g.items; item += gridDim.x) {
    const int z = item / g.segs, seg = item - z * g.segs;
    const int img_local = image_of(s, z);
    const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
    const int ya = seg * g.seg_h, yb = min(s.h, ya + g.seg_h);
    const int nv = yb - ya + 2;                      // virtual rows j = 0 .. nv-1  <->  image rows ya-1 .. yb
    const int nch = (nv + MR - 1) / MR;
    uint8_t* optr = s.out + (uint64_t)img_local * s.elems + (uint64_t)ya * s.rowbytes + c;
    // byte offset of virtual row j inside the image (rows past the segment reload its last row)
    auto row_off = [&](int j) -> uint32_t {
      j = min(j, nv - 1);
      int y = ya - 1 + j;
      y = y < 0 ? top_src : (y >= s.h ? bot_src : y);
      return (uint32_t)y * (uint32_t)s.rowbytes + (uint32_t)c;
    };

    uint2 ra[MR], rb[MR];
    auto fetch = [&](int ch) {
#pragma unroll
      for (int k = 0; k < MR; ++k) {
        const uint8_t* p = img + row_off(ch * MR + k);
        if (A4) {
          const uint32_t* pw = reinterpret_cast<const uint32_t*>(p);
          ra[k] = make_uint2(__ldg(pw), __ldg(pw + o1));
          rb[k] = make_uint2(__ldg(pw + o2), __ldg(pw + o3));
        } else {
          ra[k] = __ldg(reinterpret_cast<const uint2*>(p));
          rb[k] = __ldg(reinterpret_cast<const uint2*>(p + (v1 ? 8 : 0)));
        }
      }
    };
    fetch(0);
    typename F::Part p0, p1, p2;
    for (int ch = 0; ch < nch; ++ch) {
      uint8_t* const buf = my + (ch & 1) * MR * g.srow;
      const int jn = min(MR, nv - ch * MR);          // rows of this chunk (block uniform)
      // ---- noise: registers -> shared ------------------------------------------------
      if (active) {
#pragma unroll
        for (int k = 0; k < MR; ++k) {
          if (k >= jn) break;
          if (A4) {
            *reinterpret_cast<uint4*>(buf + k * g.srow) = make_uint4(ra[k].x, ra[k].y, rb[k].x, rb[k].y);
            continue;
          }
          const uint32_t e = row_off(ch * MR + k);
          GroupCtx g0{(uint32_t)img_local, e >> 3, img};
          const uint2 n0 = PackedNoise<KIND>::apply(na, g0, ra[k]);
          uint2 n1 = rb[k];
          if (v1) {
            GroupCtx g1{(uint32_t)img_local, (e >> 3) + 1u, img};
            n1 = PackedNoise<KIND>::apply(na, g1, rb[k]);
          }
          *reinterpret_cast<uint4*>(buf + k * g.srow) = make_uint4(n0.x, n0.y, n1.x, n1.y);
        }
      }
      if (ch + 1 < nch) fetch(ch + 1);               // in flight across the barrier and the filter phase
      __syncthreads();
      // ---- filter: shared -> registers -> global ----------------------------------------
      if (active) {
#pragma unroll
        for (int k = 0; k < MR; ++k) {
          if (k >= jn) break;
          typename F::Part& pb = (k % 3 == 0) ? p0 : (k % 3 == 1) ? p1 : p2;
          typename F::Part& pt = (k % 3 == 0) ? p1 : (k % 3 == 1) ? p2 : p0;
          typename F::Part& pm = (k % 3 == 0) ? p2 : (k % 3 == 1) ? p0 : p1;
          const uint8_t* rp = buf + k * g.srow;
          uint32_t w[6];
          w[0] = *reinterpret_cast<const uint32_t*>(rp - 4);
          const uint4 q = *reinterpret_cast<const uint4*>(rp);
          w[1] = q.x; w[2] = q.y; w[3] = q.z; w[4] = q.w;
          w[5] = *reinterpret_cast<const uint32_t*>(rp + 16);
          // image border pixels come from the thread's own bytes: pixel -1 for the first column,
          // pixel W for the last (which owns 16 or 8 valid bytes)
          if (first) w[0] = __byte_perm(q.x, q.y, kSelL);
          if (last) {
            if (A4) {                                // border pixel behind the nwv valid words
              if (nwv == 4) w[5] = __byte_perm(w[3], w[4], kSelR);
              else if (nwv == 3) w[4] = __byte_perm(w[2], w[3], kSelR);
              else if (nwv == 2) w[3] = __byte_perm(w[1], w[2], kSelR);
              else w[2] = __byte_perm(w[0], w[1], kSelR);
            } else if (v1) w[5] = __byte_perm(q.z, q.w, kSelR);
            else w[3] = __byte_perm(q.x, q.y, kSelR);
          }
          RowLanes<4> rl;
          rl.split(w);
          F::hpart(rl, pb);
          if (ch > 0 || k >= 2) {                    // rows j-2, j-1, j -> output row ya + j - 2
            uint32_t o[4];
            F::vcomb(pt, pm, pb, o);
            if (A4) {
#pragma unroll
              for (int i = 0; i < 4; ++i)
                if (i < nwv) reinterpret_cast<uint32_t*>(optr)[i] = o[i];
            } else {
              *reinterpret_cast<uint2*>(optr) = make_uint2(o[0], o[1]);
              if (v1) *reinterpret_cast<uint2*>(optr + 8) = make_uint2(o[2], o[3]);
            }
            optr += s.rowbytes;
          }
        }
      }
    }
    __syncthreads();                                 // the next item starts again in buffer 0
  }
}

struct MarchPlan { int grid, threads; size_t smem; MarchGeom g; };

template <int KIND, int FILTER, bool A4 = false>
static cudaError_t launch_one(const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  MarchPlan mp;
  mp.g.upr = (s.rowbytes + 15) / 16;
  mp.g.srow = mp.g.upr * 16 + 32;
  mp.threads = (mp.g.upr + 31) / 32 * 32;
  mp.smem = (size_t)2 * MR * mp.g.srow;
  static int occ_cache[64][MARCH_MAX_THREADS / 32 + 1] = {};     // [device][warps] -> CTAs per SM
  int dev = 0;
  cudaGetDevice(&dev);
  int& occ = occ_cache[dev & 63][mp.threads / 32];
  if (!occ) {
    cudaError_t e = cudaFuncSetAttribute(k_march3<KIND, FILTER, A4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         2 * MR * (MARCH_MAX_THREADS * 16 + 32));
    if (e == cudaSuccess)
      e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_march3<KIND, FILTER, A4>, mp.threads, mp.smem);
    if (e != cudaSuccess) return e;
    if (occ < 1) occ = 1;
  }
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int slots = sms * occ;
  // segments per image: fill whole waves of `slots` CTAs, pay 2 regenerated rows per segment
  int best_segs = 1;
  double best = -1.0;
  for (int segs = 1; segs <= std::max(1, s.h / 24); ++segs) {
    const int seg_h = (s.h + segs - 1) / segs;
    const int real_segs = (s.h + seg_h - 1) / seg_h;
    if (real_segs != segs) continue;
    const long items = (long)s.count * segs;
    const long waves = (items + slots - 1) / slots;
    const double eff = (double)items / (double)(waves * slots) * (double)seg_h / (double)(seg_h + 2);
    if (eff > best + 1e-9) { best = eff; best_segs = segs; }
  }
  mp.g.segs = best_segs;
  mp.g.seg_h = (s.h + best_segs - 1) / best_segs;
  mp.g.items = s.count * best_segs;
  mp.grid = std::min(mp.g.items, slots);
  k_march3<KIND, FILTER, A4><<<mp.grid, mp.threads, mp.smem, st>>>(s, na, mp.g);
  count_launch();
  return cudaGetLastError();
}

template <int KIND>
static cudaError_t launch_filter(int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (filter) {
    case FK_GAUS3: return launch_one<KIND, FK_GAUS3>(s, na, st);
    case FK_MEAN3: return launch_one<KIND, FK_MEAN3>(s, na, st);
    case FK_MEDIAN3: return launch_one<KIND, FK_MEDIAN3>(s, na, st);
  }
  return cudaErrorInvalidValue;
}

bool march3_supported(const ImageSet& s) {
  return s.rowbytes % 8 == 0 && s.rowbytes >= 24 && s.rowbytes <= MARCH_MAX_THREADS * 16 && s.h >= 2 &&
         s.elems < (1ull << 31) &&
         ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 7u) == 0;
}

// no generator, rows a multiple of 4 bytes only
bool march3_a4_supported(const ImageSet& s) {
  return s.rowbytes % 4 == 0 && s.rowbytes >= 24 && s.rowbytes <= MARCH_MAX_THREADS * 16 && s.h >= 2 &&
         s.elems < (1ull << 31) &&
         ((reinterpret_cast<uintptr_t>(s.in) | reinterpret_cast<uintptr_t>(s.out)) & 3u) == 0;
}
cudaError_t launch_march3_a4(int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (filter) {
    case FK_GAUS3: return launch_one<NK_NONE, FK_GAUS3, true>(s, na, st);
    case FK_MEAN3: return launch_one<NK_NONE, FK_MEAN3, true>(s, na, st);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_march3(int kind, int filter, const ImageSet& s, const NoiseArgs& na, cudaStream_t st) {
  switch (kind) {
    case NK_NONE: return launch_filter<NK_NONE>(filter, s, na, st);
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
