// b2d_filter3_bytes.cu - any-width variant of the 3x3 denoisers: one thread
// per output byte, neighbours fetched through the border map.  Used only when
// W % 4 != 0 (rows not word aligned); same arithmetic as Filter3<>::scalar.
#include "b2d_denoise.cuh"

namespace b2d {

template <int FILTER>
__global__ void __launch_bounds__(256) k_filter3_bytes(ImageSet s) {
  using F = Filter3<FILTER>;
  const int img_local = image_of(s, blockIdx.z);
  const uint8_t* img = s.in + (uint64_t)img_local * s.elems;
  uint8_t* out = s.out + (uint64_t)img_local * s.elems;
  for (uint64_t e = (uint64_t)blockIdx.x * 256 + threadIdx.x; e < s.elems; e += (uint64_t)gridDim.x * 256) {
    const int y = (int)(e / s.rowbytes);
    const int b = (int)(e - (uint64_t)y * s.rowbytes);
    const int x = b / 3, ch = b - 3 * x;
    uint32_t p[9];
#pragma unroll
    for (int dy = -1; dy <= 1; ++dy) {
      const int yy = border_map(y + dy, s.h, F::BORDER);
#pragma unroll
      for (int dx = -1; dx <= 1; ++dx) {
        const int xx = border_map(x + dx, s.w, F::BORDER);
        p[(dy + 1) * 3 + dx + 1] = __ldg(img + (uint64_t)yy * s.rowbytes + xx * 3 + ch);
      }
    }
    out[e] = (uint8_t)F::scalar(p);
  }
}

cudaError_t launch_filter3_bytes(int filter, const ImageSet& s, cudaStream_t st) {
  if (s.count <= 0) return cudaSuccess;
  uint32_t bx = (uint32_t)((s.elems + 255) / 256);
  if (bx > 148u * 16u) bx = 148u * 16u;
  const dim3 grid(bx, 1, s.count);
  switch (filter) {
    case FK_GAUS3: k_filter3_bytes<FK_GAUS3><<<grid, 256, 0, st>>>(s); break;
    case FK_MEAN3: k_filter3_bytes<FK_MEAN3><<<grid, 256, 0, st>>>(s); break;
    case FK_MEDIAN3: k_filter3_bytes<FK_MEDIAN3><<<grid, 256, 0, st>>>(s); break;
    default: return cudaErrorInvalidValue;
  }
  count_launch();
  return cudaGetLastError();
}

}  // namespace b2d
