// b2d_poisson.cu - generator-mode Poisson corruption:
//     skimage.util.random_noise(image, mode='poisson')   (lib/roi_data_layer/minibatch.py:211)
//         vals = 2 ** ceil(log2(len(unique(image)))); out = clip(poisson(image * vals) / vals, 0, 1)
//     -> (255 * out).astype(uint8)                          (minibatch.py:212-213)
// For one image `vals` is fixed and the input takes 256 values, so the uint8 RESULT has one of
// 256 fixed discrete distributions: out = (255 * min(k, vals)) >> log2(vals), k ~ Poisson(v * vals / 255).
// Instead of running a rejection sampler per element (lgamma, logs, data-dependent loops: 29 ms for
// 512 x 600 x 1000) the plan holds, for each of the 9 possible `vals`, a Walker / Vose alias table
// over the output byte per input value: 256 rows x 256 slots x (8-bit threshold, 8-bit alias) =
// 128 KB, built on the host.  Each row's pmf (double precision) is first rounded to multiples of
// 2^-16 by largest remainders; Vose's construction in integer arithmetic then represents that
// rounded pmf EXACTLY with 8-bit thresholds (a slot holds 256 units of 2^-16).  The kernel keeps
// the table of its image in shared memory; one element = 16 random bits, one shared-memory gather.
#include "b2d_kernels.cuh"
#include "b2d_plan.h"
#include "b2d_noise.cuh"
#include "b2d_prepass.cuh"

#include <algorithm>

#include <cmath>
#include <vector>

namespace b2d {

constexpr int PO_ROWS = 256, PO_SLOTS = 256, PO_TABLES = 9;           // vals = 1, 2, ..., 256
constexpr size_t PO_TABLE_BYTES = (size_t)PO_ROWS * PO_SLOTS * 2;     // 131072
constexpr int PO_THREADS = 1024;

size_t poisson_tables_bytes() { return PO_TABLES * PO_TABLE_BYTES; }

// host: 256 probabilities -> integer weights (sum 65536, largest remainders) -> alias row.
// Vose on integers: every slot ends with exactly 256 units, thr of its own outcome + the rest of its alias.
void alias_row_256(const double* P, uint16_t* row) {
  double sum = 0.0;
  for (int i = 0; i < PO_SLOTS; ++i) sum += P[i];
  int w[PO_SLOTS], order[PO_SLOTS], units = 0;
  double frac[PO_SLOTS];
  for (int i = 0; i < PO_SLOTS; ++i) {
    const double x = P[i] / sum * 65536.0;
    w[i] = (int)std::floor(x);
    frac[i] = x - w[i];
    units += w[i];
    order[i] = i;
  }
  std::sort(order, order + PO_SLOTS, [&](int a, int b) { return frac[a] > frac[b] || (frac[a] == frac[b] && a < b); });
  for (int j = 0; units < 65536; ++j, ++units) ++w[order[j % PO_SLOTS]];
  int thr[PO_SLOTS], alias[PO_SLOTS], small[PO_SLOTS], large[PO_SLOTS], ns = 0, nl = 0;
  for (int i = 0; i < PO_SLOTS; ++i) {
    thr[i] = 256;
    alias[i] = i;
    if (w[i] < 256) small[ns++] = i; else large[nl++] = i;
  }
  while (ns && nl) {
    const int sm = small[--ns], lg = large[--nl];
    thr[sm] = w[sm];
    alias[sm] = lg;
    w[lg] -= 256 - w[sm];
    if (w[lg] < 256) small[ns++] = lg; else large[nl++] = lg;
  }
  for (int i = 0; i < PO_SLOTS; ++i) {
    // u (8 random bits) < thr keeps the slot; thr == 256 (always) is stored as alias = slot
    const int th = thr[i] >= 256 ? 255 : thr[i], al = thr[i] >= 256 ? i : alias[i];
    row[i] = (uint16_t)(al | (th << 8));
  }
}

// distribution of the output byte for input value v
static void build_row(int t, int v, uint16_t* row) {
  const int vals = 1 << t;
  const double lam = (double)v / 255.0 * (double)vals;                // img_as_float(image) * vals
  double P[PO_SLOTS] = {0.0};
  if (lam <= 0.0) {
    P[0] = 1.0;
  } else {
    double total = 0.0;
    const double ll = std::log(lam);
    for (int k = 0; k < vals; ++k) {
      const double p = std::exp(-lam + k * ll - std::lgamma((double)k + 1.0));
      P[(255 * k) >> t] += p;
      total += p;
    }
    P[255] += std::fmax(0.0, 1.0 - total);                            // k >= vals clips to 1.0 -> 255
  }
  alias_row_256(P, row);
}

void poisson_table_host(int t, uint16_t* tab) {
  for (int v = 0; v < PO_ROWS; ++v) build_row(t, v, tab + v * PO_SLOTS);
}

cudaError_t upload_poisson_tables(uint8_t** d_tables) {
  std::vector<uint8_t> blob(poisson_tables_bytes());
  for (int t = 0; t < PO_TABLES; ++t) {
    poisson_table_host(t, reinterpret_cast<uint16_t*>(blob.data() + t * PO_TABLE_BYTES));
  }
  cudaError_t e = cudaMalloc(d_tables, blob.size());
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*d_tables, blob.data(), blob.size(), cudaMemcpyHostToDevice);
}

// grid = (segments, images); one table per CTA in shared memory
__global__ void __launch_bounds__(PO_THREADS, 1) k_poisson(ImageSet s, NoiseArgs na, const uint8_t* tables, int seg_groups) {
  extern __shared__ __align__(16) uint8_t smem[];
  const uint16_t* tab = reinterpret_cast<const uint16_t*>(smem);
  const int img_local = image_of(s, blockIdx.y);
  const float vals = na.img_par[img_local];
  const int t = min(max(__float2int_rn(__log2f(fmaxf(vals, 1.0f))), 0), PO_TABLES - 1);
  {
    const uint4* src = reinterpret_cast<const uint4*>(tables + (size_t)t * PO_TABLE_BYTES);
    uint4* dst = reinterpret_cast<uint4*>(smem);
    for (int i = threadIdx.x; i < (int)(PO_TABLE_BYTES / 16); i += PO_THREADS) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  const uint32_t ngroups = (uint32_t)((s.elems + 7) / 8);
  const uint32_t g_lo = blockIdx.x * (uint32_t)seg_groups, g_hi = min(g_lo + (uint32_t)seg_groups, ngroups);
  const bool vec = ((reinterpret_cast<uintptr_t>(img) | reinterpret_cast<uintptr_t>(out)) & 7u) == 0;
  for (uint32_t grp = g_lo + threadIdx.x; grp < g_hi; grp += PO_THREADS) {
    const uint64_t e0 = (uint64_t)grp * 8u;
    const bool full = vec && e0 + 8u <= s.elems;
    uint2 w = make_uint2(0u, 0u);
    if (full) {
      w = __ldg(reinterpret_cast<const uint2*>(img + e0));
    } else {
      for (int i = 0; i < 8; ++i)
        if (e0 + i < s.elems) (i < 4 ? w.x : w.y) |= (uint32_t)img[e0 + i] << (8 * (i & 3));
    }
    GroupCtx gc{(uint32_t)img_local, grp, img};
    const uint4 ra = group_random(na, gc, 0);
    const uint32_t rw[4] = {ra.x, ra.y, ra.z, ra.w};
    // 16 random bits per element: low byte = slot, high byte = u.  One PRMT builds the table indices
    // (v << 8 | slot) of two elements; `r16 < (entry & 0xff00)` is u < threshold because slot < 256.
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t wj = j < 2 ? w.x : w.y;
      const uint32_t idx2 = __byte_perm(wj, rw[j], (j & 1) ? 0x3624u : 0x1604u);   // [slot0, v0, slot1, v1]
      const uint32_t r0 = rw[j] & 0xffffu, r1 = rw[j] >> 16;
      const uint32_t e0v = tab[idx2 & 0xffffu], e1v = tab[idx2 >> 16];
      o[2 * j] = r0 < (e0v & 0xff00u) ? r0 : e0v;                                // low byte = result
      o[2 * j + 1] = r1 < (e1v & 0xff00u) ? r1 : e1v;
    }
    const uint2 r = make_uint2(__byte_perm(__byte_perm(o[0], o[1], 0x0040), __byte_perm(o[2], o[3], 0x0040), 0x5410),
                               __byte_perm(__byte_perm(o[4], o[5], 0x0040), __byte_perm(o[6], o[7], 0x0040), 0x5410));
    if (full) {
      *reinterpret_cast<uint2*>(out + e0) = r;
    } else {
      for (int i = 0; i < 8; ++i)
        if (e0 + i < s.elems) out[e0 + i] = (uint8_t)((i < 4 ? r.x : r.y) >> (8 * (i & 3)));
    }
  }
}

// `na.img_par` must already hold vals per image (launch_poisson_vals)
cudaError_t launch_poisson(const ImageSet& s, const NoiseArgs& na, const uint8_t* tables, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  static bool attr_done[64] = {};               // per device: the attribute lives in the context
  const int dev = current_device() & 63;
  if (!attr_done[dev]) {
    cudaError_t e = cudaFuncSetAttribute(k_poisson, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PO_TABLE_BYTES);
    if (e != cudaSuccess) return e;
    attr_done[dev] = true;
  }
  const int sms = sm_count();
  const uint32_t ngroups = (uint32_t)((s.elems + 7) / 8);
  // one CTA per SM is resident (192 KB of shared memory): about two waves over the batch, and no
  // segment shorter than 16 passes of the CTA so the table load stays a small share
  int segs = std::max(1, (2 * sms + s.count - 1) / s.count);
  {                                                                   // fewest segments within 3 % of the best wave fill
    double best = 0.0;
    for (int c = segs; c < segs + 8; ++c) {
      const long total = (long)c * s.count;
      best = std::max(best, (double)total / (double)(((total + sms - 1) / sms) * sms));
    }
    for (int c = segs; c < segs + 8; ++c) {
      const long total = (long)c * s.count;
      if ((double)total / (double)(((total + sms - 1) / sms) * sms) >= best - 0.03) { segs = c; break; }
    }
  }
  segs = (int)std::min<uint32_t>((uint32_t)segs, std::max<uint32_t>(1u, ngroups / (16u * PO_THREADS)));
  const int seg_groups = (int)((ngroups + segs - 1) / segs);
  k_poisson<<<dim3((unsigned)segs, (unsigned)s.count, 1), PO_THREADS, PO_TABLE_BYTES, st>>>(s, na, tables, seg_groups);
  count_launch();
  return cudaGetLastError();
}

}  // namespace b2d

extern "C" int b2d_poisson_table(int log2_vals, uint16_t* h_table) {
  if (log2_vals < 0 || log2_vals >= b2d::PO_TABLES || !h_table)
    return b2d::fail(B2D_ERR_INVALID, "b2d_poisson_table: log2_vals must be in [0, 8] and h_table non-null");
  b2d::poisson_table_host(log2_vals, h_table);
  return B2D_OK;
}
