// b2d_gamma.cu - host tables of the generator-mode gamma corruption:
//     gamma_noise = scipy.stats.gamma.rvs(1.99, loc=0, scale=S, size=image.shape)   (minibatch.py:1216)
//     out = cv2.add(img_as_float(image), gamma_noise); (255 * out).astype(uint8)     (minibatch.py:1217,1247)
// The sum is not clipped and the cast wraps, so with K = floor(255 * S * G), G ~ Gamma(1.99, 1):
//     out = (v + K) mod 256.
// K has ONE fixed discrete distribution per plan (it depends on S only), so the uint8 consumers do
// not run a rejection sampler (Marsaglia-Tsang: normals, logs, data-dependent retries; 5.4 ms per
// 512 x 600 x 1000) but draw R = K mod 256 from a 256-slot Walker / Vose alias table (512 bytes,
// b2d_poisson.cu: alias_row_256, probabilities rounded to 2^-16 by largest remainders and represented
// exactly): one 16-bit random number and one cached gather per element.
// The test hook b2d_noise_field needs the unfolded K and a position inside [K, K + 1): a second
// table holds, per residue R, the cumulative distribution of the fold index h (K = R + 256 h).
#include "b2d_kernels.cuh"
#include "b2d_plan.h"
#include "b2d_prepass.cuh"

#include <algorithm>
#include <cmath>
#include <vector>

namespace b2d {

constexpr double kGammaShape = 1.99;

// regularised lower incomplete gamma function P(a, x): power series below a + 1, Lentz's continued
// fraction of Q above
static double gamma_p(double a, double x) {
  if (x <= 0.0) return 0.0;
  const double front = std::exp(-x + a * std::log(x) - std::lgamma(a));
  if (x < a + 1.0) {
    double term = 1.0 / a, sum = term;
    for (int n = 1; n < 10000; ++n) {
      term *= x / (a + n);
      sum += term;
      if (term < sum * 1e-17) break;
    }
    return std::min(1.0, sum * front);
  }
  const double tiny = 1e-300;
  double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, h = d;
  for (int i = 1; i < 10000; ++i) {
    const double an = -(double)i * ((double)i - a);
    b += 2.0;
    d = an * d + b;
    if (std::fabs(d) < tiny) d = tiny;
    c = b + an / c;
    if (std::fabs(c) < tiny) c = tiny;
    d = 1.0 / d;
    const double del = d * c;
    h *= del;
    if (std::fabs(del - 1.0) < 1e-16) break;
  }
  return std::max(0.0, 1.0 - front * h);
}

int gamma_folds(double scale) {
  // K beyond 255 * S * 40 has probability < 1e-15; at most 64 folds are tabulated (the rest lumps into the last)
  const double kmax = std::ceil(255.0 * scale * 40.0) + 1.0;
  return (int)std::min(64.0, std::max(1.0, std::ceil(kmax / 256.0)));
}

size_t gamma_table_bytes(double scale) { return 512 + (size_t)256 * gamma_folds(scale) * sizeof(uint32_t); }

// h_alias[256]: alias | threshold << 8 over R = K mod 256.  h_fold[256][folds]: 65536 * P(h <= j | R), rounded.
void gamma_table_host(double scale, uint16_t* h_alias, uint32_t* h_fold) {
  const int folds = gamma_folds(scale);
  const int kn = 256 * folds;
  std::vector<double> pk(kn);
  const double step = 1.0 / (255.0 * scale);
  double lo = 0.0;
  for (int k = 0; k < kn; ++k) {
    const double hi = k + 1 < kn ? gamma_p(kGammaShape, (k + 1) * step) : 1.0;
    pk[k] = std::max(0.0, hi - lo);
    lo = hi;
  }
  double P[256];
  for (int r = 0; r < 256; ++r) {
    P[r] = 0.0;
    for (int j = 0; j < folds; ++j) P[r] += pk[r + 256 * j];
  }
  alias_row_256(P, h_alias);
  if (!h_fold) return;
  for (int r = 0; r < 256; ++r) {
    double acc = 0.0;
    for (int j = 0; j < folds; ++j) {
      acc += pk[r + 256 * j];
      h_fold[r * folds + j] = P[r] > 0.0 ? (uint32_t)std::min(65536.0, std::floor(acc / P[r] * 65536.0 + 0.5)) : 65536u;
    }
    h_fold[r * folds + folds - 1] = 65536u;
  }
}

cudaError_t upload_gamma_table(double scale, uint8_t** d_table) {
  std::vector<uint8_t> blob(gamma_table_bytes(scale));
  gamma_table_host(scale, reinterpret_cast<uint16_t*>(blob.data()), reinterpret_cast<uint32_t*>(blob.data() + 512));
  cudaError_t e = cudaMalloc(d_table, blob.size());
  if (e != cudaSuccess) return e;
  return cudaMemcpy(*d_table, blob.data(), blob.size(), cudaMemcpyHostToDevice);
}

}  // namespace b2d

extern "C" int b2d_gamma_table(double scale, uint16_t* h_alias, uint32_t* h_fold, int* folds) {
  if (!(scale > 0.0) || !h_alias) return b2d::fail(B2D_ERR_INVALID, "b2d_gamma_table: scale must be > 0 and h_alias non-null");
  if (folds) *folds = b2d::gamma_folds(scale);
  b2d::gamma_table_host(scale, h_alias, h_fold);
  return B2D_OK;
}
