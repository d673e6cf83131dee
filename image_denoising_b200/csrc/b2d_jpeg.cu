// b2d_jpeg.cu - SURVEY 8(f) f2: the step in front of the path, `cv2.imread(path)`
// (lib/model/test.py:191, lib/roi_data_layer/minibatch.py:85 and again inside most closures),
// for baseline JPEG - what VOC-style datasets and the reference's data/demo/*.jpg hold.
//
// cv2.imread = libjpeg-turbo with its defaults: JDCT_ISLOW, fancy (triangle) chroma
// upsampling, YCbCr -> BGR through 16-bit fixed-point tables.  Everything after the entropy
// decoder is integer arithmetic on independent 8x8 blocks / pixels and runs on the GPU,
// bit-exact with cv2.imdecode (oracle/jpeg.py restates it; tests/test_gpu_jpeg.py):
//   host, one thread per image : marker parsing, sequential Huffman decoding (ITU T.81 F.2.2)
//                                -> quantised coefficients, int16, natural order, pinned memory
//   k_jpeg_idct                : dequantise + jpeg_idct_islow (jidctint.c: CONST_BITS 13,
//                                PASS1_BITS 2), one thread per block, 64 values in registers
//   k_jpeg_color               : h2v2 / h2v1 fancy upsampling (jdsample.c) + ycc_rgb_convert
//                                (jdcolor.c) + BGR interleave, 4 pixels per thread
// The bit-serial Huffman stage stays on the host cores (it is a dependent chain per scan);
// images decode in parallel on `host_threads` threads.  Progressive / arithmetic / 12-bit /
// CMYK streams are refused with B2D_ERR_UNSUPPORTED - there is no fallback decoder.
#include "b2d_kernels.cuh"
#include "b2d_plan.h"

#include <atomic>
#include <chrono>
#include <algorithm>
#include <cstring>
#include <exception>
#include <new>
#include <string>
#include <thread>
#include <vector>

namespace b2d {

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct JpegImg {                 // per-image descriptor, read by both kernels
  int ncomp, hs, vs;             // luma sampling factors (chroma is 1 x 1)
  int bx[3], by[3];              // blocks per row / column of each component (padded to whole MCUs)
  int dw, dh;                    // real chroma samples: ceil(w / hs), ceil(h / vs)
  unsigned coef_off[3];          // int16 offset of the component's first block inside the image's coefficient area
  unsigned plane_off[3];         // byte offset of the component's sample plane
  uint16_t q[3][64];             // quantisation tables, natural order
};

// ------------------------------------------------------------------- host: parsing + Huffman
struct HuffTab {
  uint16_t fast[512];            // 9-bit look-ahead: (length << 8) | symbol, 0 = longer code
  int maxcode[18], valptr[17], mincode[17];
  uint8_t vals[256];
  int16_t fast_ac[512];          // AC tables: (value << 8) | (run << 4) | (code length + magnitude bits) when both fit in 9 bits
  bool present = false;
};

static bool build_huff(HuffTab& t, const uint8_t* counts, const uint8_t* symbols, int nsym) {
  memset(t.fast, 0, sizeof(t.fast));
  memcpy(t.vals, symbols, nsym);
  int code = 0, k = 0;
  for (int l = 1; l <= 16; ++l) {
    t.valptr[l] = k;
    t.mincode[l] = code;
    for (int i = 0; i < counts[l - 1]; ++i, ++k, ++code) {
      if (l <= 9) {
        const int lo = code << (9 - l), n = 1 << (9 - l);
        if (lo + n > 512) return false;
        for (int j = 0; j < n; ++j) t.fast[lo + j] = (uint16_t)((l << 8) | symbols[k]);
      }
    }
    t.maxcode[l] = counts[l - 1] ? code - 1 : -1;
    if (code > (1 << l)) return false;
    code <<= 1;
  }
  t.maxcode[17] = 0x7fffffff;
  t.present = true;
  // run / size / value of short AC codes in one lookup (the common case: small coefficients, short codes)
  for (int i = 0; i < 512; ++i) {
    t.fast_ac[i] = 0;
    const uint16_t f = t.fast[i];
    if (!f) continue;
    const int l = f >> 8, rs = f & 0xff, run = rs >> 4, mag = rs & 15;
    if (mag == 0 || l + mag > 9) continue;
    int v = ((i << l) & 511) >> (9 - mag);
    if (v < (1 << (mag - 1))) v += (int)((~0u) << mag) + 1;
    if (v >= -128 && v <= 127) t.fast_ac[i] = (int16_t)(v * 256 + run * 16 + (l + mag));
  }
  return k == nsym;
}

struct BitReader {
  const uint8_t* d;
  size_t pos, len;
  uint64_t acc = 0;
  int n = 0;
  bool marker = false;           // fill() ran into a marker: zeros are fed from here on (as libjpeg does)
  void fill() {
    if (!marker && pos + 8 <= len && n <= 56) {        // 8 bytes at once when none of them is 0xFF
      uint64_t w;
      memcpy(&w, d + pos, 8);
      w = __builtin_bswap64(w);
      const uint64_t x = ~w;
      if (!((x - 0x0101010101010101ull) & ~x & 0x8080808080808080ull)) {
        const int take = (64 - n) >> 3;
        const uint64_t keep = take == 8 ? ~0ull : ~(~0ull >> (8 * take));
        acc |= (w & keep) >> n;
        pos += take;
        n += 8 * take;
        return;
      }
    }
    while (n <= 56) {
      uint32_t b = 0;
      if (!marker && pos < len) {
        b = d[pos];
        if (b == 0xFF) {
          if (pos + 1 < len && d[pos + 1] == 0) pos += 2;
          else { marker = true; b = 0; }
        } else {
          ++pos;
        }
      } else {
        marker = true;
      }
      acc |= (uint64_t)b << (56 - n);
      n += 8;
    }
  }
  inline uint32_t peek(int k) { return (uint32_t)(acc >> (64 - k)); }
  inline void skip(int k) { acc <<= k; n -= k; }
  bool restart() {               // discard the padding bits, step over RSTn
    acc = 0; n = 0; marker = false;
    while (pos + 1 < len) {
      if (d[pos] == 0xFF && d[pos + 1] >= 0xD0 && d[pos + 1] <= 0xD7) { pos += 2; return true; }
      ++pos;
    }
    return false;
  }
};

static inline int decode_sym(BitReader& br, const HuffTab& t) {
  if (br.n < 16) br.fill();
  const uint32_t look = br.peek(16);
  const uint16_t f = t.fast[look >> 7];
  if (f) { br.skip(f >> 8); return f & 0xff; }
  for (int l = 10; l <= 16; ++l) {
    const int code = (int)(look >> (16 - l));
    if (code <= t.maxcode[l]) { br.skip(l); return t.vals[t.valptr[l] + code - t.mincode[l]]; }
  }
  return -1;
}
static inline int receive_extend(BitReader& br, int s) {
  if (br.n < s) br.fill();
  const int v = (int)br.peek(s);
  br.skip(s);
  return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
}

struct Parsed {
  int h = 0, w = 0, ncomp = 0, hs = 1, vs = 1, dri = 0;
  int tq[3] = {0, 0, 0}, td[3] = {0, 0, 0}, ta[3] = {0, 0, 0};
  uint16_t q[4][64];
  bool have_q[4] = {false, false, false, false};
  HuffTab dc[4], ac[4];
  size_t scan = 0;
};
static inline unsigned be16(const uint8_t* p) { return ((unsigned)p[0] << 8) | p[1]; }

// 0 = ok, B2D_ERR_* otherwise; msg filled on failure
static int parse_jpeg(const uint8_t* d, size_t len, Parsed& P, std::string& msg, bool header_only) {
  if (len < 4 || d[0] != 0xFF || d[1] != 0xD8) { msg = "not a JPEG stream"; return B2D_ERR_INVALID; }
  size_t i = 2;
  bool have_sof = false, jfif = false, adobe = false;
  int adobe_transform = -1, orientation = 1;
  int cid[3] = {0, 0, 0};
  while (i + 4 <= len) {
    if (d[i] != 0xFF) { msg = "marker expected"; return B2D_ERR_INVALID; }
    while (i + 1 < len && d[i + 1] == 0xFF) ++i;
    const int m = d[i + 1];
    if (m == 0xD8 || m == 0x01 || (m >= 0xD0 && m <= 0xD7)) { i += 2; continue; }
    const size_t L = be16(d + i + 2);
    if (L < 2 || i + 2 + L > len) { msg = "truncated segment"; return B2D_ERR_INVALID; }
    const uint8_t* seg = d + i + 4;
    const size_t sl = L - 2;
    if (m == 0xDB) {
      size_t k = 0;
      while (k < sl) {
        const int pq = seg[k] >> 4, t = seg[k] & 15;
        if (t > 3 || k + 1 + (pq ? 128 : 64) > sl) { msg = "bad DQT"; return B2D_ERR_INVALID; }
        for (int z = 0; z < 64; ++z) P.q[t][kZigzag[z]] = pq ? (uint16_t)be16(seg + k + 1 + 2 * z) : seg[k + 1 + z];
        P.have_q[t] = true;
        k += 1 + (pq ? 128 : 64);
      }
    } else if (m == 0xC4) {
      size_t k = 0;
      while (k + 17 <= sl) {
        const int tc = seg[k] >> 4, th = seg[k] & 15;
        int n = 0;
        for (int z = 0; z < 16; ++z) n += seg[k + 1 + z];
        if (tc > 1 || th > 3 || n > 256 || k + 17 + n > sl) { msg = "bad DHT"; return B2D_ERR_INVALID; }
        if (!build_huff(tc ? P.ac[th] : P.dc[th], seg + k + 1, seg + k + 17, n)) { msg = "inconsistent DHT"; return B2D_ERR_INVALID; }
        k += 17 + n;
      }
    } else if (m == 0xC0 || m == 0xC1) {
      if (sl < 6 || seg[0] != 8) { msg = "only 8-bit samples"; return B2D_ERR_UNSUPPORTED; }
      P.h = (int)be16(seg + 1); P.w = (int)be16(seg + 3); P.ncomp = seg[5];
      if (P.ncomp != 1 && P.ncomp != 3) { msg = "only 1- or 3-component (YCbCr) images"; return B2D_ERR_UNSUPPORTED; }
      if (sl < 6 + 3 * (size_t)P.ncomp || P.h <= 0 || P.w <= 0) { msg = "bad SOF"; return B2D_ERR_INVALID; }
      for (int c = 0; c < P.ncomp; ++c) {
        cid[c] = seg[6 + 3 * c];
        const int hh = seg[7 + 3 * c] >> 4, vv = seg[7 + 3 * c] & 15;
        P.tq[c] = seg[8 + 3 * c] & 3;
        if (c == 0) { P.hs = hh; P.vs = vv; }
        else if (hh != 1 || vv != 1) { msg = "chroma sampling factors other than 1x1"; return B2D_ERR_UNSUPPORTED; }
      }
      if (P.ncomp == 1) { P.hs = P.vs = 1; }
      if (!((P.hs == 1 && P.vs == 1) || (P.hs == 2 && P.vs == 2) || (P.hs == 2 && P.vs == 1))) {
        msg = "luma sampling must be 1x1, 2x1 or 2x2"; return B2D_ERR_UNSUPPORTED;
      }
      have_sof = true;
      if (header_only) return B2D_OK;
    } else if (m >= 0xC2 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
      msg = "only baseline sequential Huffman JPEG (SOF0 / SOF1) is decoded"; return B2D_ERR_UNSUPPORTED;
    } else if (m == 0xE0) {
      if (sl >= 5 && !memcmp(seg, "JFIF", 5)) jfif = true;
    } else if (m == 0xEE) {
      if (sl >= 12 && !memcmp(seg, "Adobe", 5)) { adobe = true; adobe_transform = seg[11]; }
    } else if (m == 0xE1) {
      // EXIF orientation (TIFF tag 0x0112 in IFD0): cv2.imread rotates / mirrors the decoded image for values 2..8
      if (sl >= 14 && !memcmp(seg, "Exif\0\0", 6)) {
        const uint8_t* t = seg + 6;
        const size_t tl = sl - 6;
        const bool le = t[0] == 'I' && t[1] == 'I', be = t[0] == 'M' && t[1] == 'M';
        auto r16 = [&](size_t o) -> unsigned { return le ? (t[o] | (t[o + 1] << 8)) : ((t[o] << 8) | t[o + 1]); };
        auto r32 = [&](size_t o) -> size_t { return le ? ((size_t)r16(o) | ((size_t)r16(o + 2) << 16)) : (((size_t)r16(o) << 16) | r16(o + 2)); };
        if ((le || be) && tl >= 8) {
          const size_t ifd = r32(4);
          if (ifd + 2 <= tl) {
            const unsigned cnt = r16(ifd);
            for (unsigned e = 0; e < cnt && ifd + 2 + 12 * (size_t)(e + 1) <= tl; ++e) {
              const size_t o = ifd + 2 + 12 * (size_t)e;
              if (r16(o) == 0x0112) orientation = (int)r16(o + 8);
            }
          }
        }
      }
    } else if (m == 0xDD) {
      if (sl < 2) { msg = "bad DRI"; return B2D_ERR_INVALID; }
      P.dri = (int)be16(seg);
    } else if (m == 0xDA) {
      if (!have_sof) { msg = "SOS before SOF"; return B2D_ERR_INVALID; }
      if (sl < 1 || seg[0] != P.ncomp || sl < 1 + 2 * (size_t)P.ncomp) { msg = "non-interleaved scans are not decoded"; return B2D_ERR_UNSUPPORTED; }
      for (int c = 0; c < P.ncomp; ++c) {
        if (seg[1 + 2 * c] != cid[c]) { msg = "scan component order differs from the frame"; return B2D_ERR_UNSUPPORTED; }
        P.td[c] = seg[2 + 2 * c] >> 4; P.ta[c] = seg[2 + 2 * c] & 15;
        if (P.td[c] > 3 || P.ta[c] > 3 || !P.dc[P.td[c]].present || !P.ac[P.ta[c]].present || !P.have_q[P.tq[c]]) {
          msg = "scan refers to a missing table"; return B2D_ERR_INVALID;
        }
      }
      if (orientation >= 2 && orientation <= 8) {
        msg = "EXIF orientation " + std::to_string(orientation) + ": cv2.imread would rotate / mirror this image";
        return B2D_ERR_UNSUPPORTED;
      }
      // libjpeg's colour-space guess for 3 components (jdapimin.c default_decompress_parms): RGB streams are not converted
      if (P.ncomp == 3) {
        bool rgb = false;
        if (jfif) rgb = false;
        else if (adobe) rgb = adobe_transform == 0;
        else rgb = cid[0] == 'R' && cid[1] == 'G' && cid[2] == 'B';
        if (rgb) { msg = "RGB-coded JPEG (Adobe transform 0 / 'RGB' component ids)"; return B2D_ERR_UNSUPPORTED; }
      }
      P.scan = i + 2 + L;
      return B2D_OK;
    }
    i += 2 + L;
  }
  msg = have_sof ? "no SOS marker" : "no SOF marker";
  return B2D_ERR_INVALID;
}

static void layout(const Parsed& P, JpegImg& J) {
  memset(&J, 0, sizeof(J));
  J.ncomp = P.ncomp; J.hs = P.hs; J.vs = P.vs;
  const int mcux = (P.w + 8 * P.hs - 1) / (8 * P.hs), mcuy = (P.h + 8 * P.vs - 1) / (8 * P.vs);
  unsigned co = 0, po = 0;
  for (int c = 0; c < P.ncomp; ++c) {
    J.bx[c] = mcux * (c == 0 ? P.hs : 1);
    J.by[c] = mcuy * (c == 0 ? P.vs : 1);
    J.coef_off[c] = co; co += (unsigned)J.bx[c] * J.by[c] * 64u;
    J.plane_off[c] = po; po += (unsigned)J.bx[c] * J.by[c] * 64u;
    memcpy(J.q[c], P.q[P.tq[c]], sizeof(J.q[c]));
  }
  J.dw = (P.w + P.hs - 1) / P.hs;
  J.dh = (P.h + P.vs - 1) / P.vs;
}

static int entropy_decode(const uint8_t* d, size_t len, const Parsed& P, const JpegImg& J, int16_t* coef, std::string& msg) {
  BitReader br{d, P.scan, len};
  const int mcux = J.bx[P.ncomp - 1] / (P.ncomp == 1 ? P.hs : 1), mcuy = J.by[P.ncomp - 1] / (P.ncomp == 1 ? P.vs : 1);
  int pred[3] = {0, 0, 0};
  int count = 0;
  for (int my = 0; my < mcuy; ++my)
    for (int mx = 0; mx < mcux; ++mx) {
      if (P.dri && count && count % P.dri == 0) {
        if (!br.restart()) { msg = "restart marker missing"; return B2D_ERR_INVALID; }
        pred[0] = pred[1] = pred[2] = 0;
      }
      ++count;
      for (int c = 0; c < P.ncomp; ++c) {
        const HuffTab& dc = P.dc[P.td[c]];
        const HuffTab& ac = P.ac[P.ta[c]];
        const int nh = c == 0 ? P.hs : 1, nv = c == 0 ? P.vs : 1;
        for (int by = 0; by < nv; ++by)
          for (int bx = 0; bx < nh; ++bx) {
            int16_t* blk = coef + J.coef_off[c] + ((size_t)(my * nv + by) * J.bx[c] + (mx * nh + bx)) * 64;
            memset(blk, 0, 128);
            const int t = decode_sym(br, dc);
            if (t < 0 || t > 15) { msg = "bad DC code"; return B2D_ERR_INVALID; }
            if (t) pred[c] += receive_extend(br, t);
            blk[0] = (int16_t)pred[c];
            for (int k = 1; k < 64;) {
              if (br.n < 16) br.fill();
              const int fa = ac.fast_ac[br.peek(9)];
              if (fa) {
                k += (fa >> 4) & 15;
                if (k > 63) { msg = "AC run past the block"; return B2D_ERR_INVALID; }
                br.skip(fa & 15);
                blk[kZigzag[k++]] = (int16_t)(fa >> 8);
                continue;
              }
              const int rs = decode_sym(br, ac);
              if (rs < 0) { msg = "bad AC code"; return B2D_ERR_INVALID; }
              const int r = rs >> 4, s = rs & 15;
              if (s == 0) {
                if (r != 15) break;
                k += 16;
                continue;
              }
              k += r;
              if (k > 63) { msg = "AC run past the block"; return B2D_ERR_INVALID; }
              blk[kZigzag[k]] = (int16_t)receive_extend(br, s);
              ++k;
            }
          }
      }
    }
  return B2D_OK;
}

// --------------------------------------------------------------------------- device: IDCT
#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172

// one 1-D pass of jpeg_idct_islow on v[0..7] (jidctint.c), result descaled by SHIFT
template <int SHIFT>
__device__ __forceinline__ void idct8(int (&v)[8]) {
  int z2 = v[2], z3 = v[6];
  int z1 = (z2 + z3) * FIX_0_541196100;
  int tmp2 = z1 + z3 * (-FIX_1_847759065);
  int tmp3 = z1 + z2 * FIX_0_765366865;
  z2 = v[0]; z3 = v[4];
  int tmp0 = (z2 + z3) << 13, tmp1 = (z2 - z3) << 13;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = v[7]; tmp1 = v[5]; tmp2 = v[3]; tmp3 = v[1];
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * FIX_1_175875602;
  tmp0 *= FIX_0_298631336; tmp1 *= FIX_2_053119869; tmp2 *= FIX_3_072711026; tmp3 *= FIX_1_501321110;
  z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447; z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  constexpr int R = 1 << (SHIFT - 1);
  v[0] = (tmp10 + tmp3 + R) >> SHIFT; v[7] = (tmp10 - tmp3 + R) >> SHIFT;
  v[1] = (tmp11 + tmp2 + R) >> SHIFT; v[6] = (tmp11 - tmp2 + R) >> SHIFT;
  v[2] = (tmp12 + tmp1 + R) >> SHIFT; v[5] = (tmp12 - tmp1 + R) >> SHIFT;
  v[3] = (tmp13 + tmp0 + R) >> SHIFT; v[4] = (tmp13 - tmp0 + R) >> SHIFT;
}

// grid (block chunks, component, image); one thread = one 8x8 block
__global__ void __launch_bounds__(128) k_jpeg_idct(const JpegImg* desc, const int16_t* coef, uint8_t* planes,
                                                   size_t coef_per_img, size_t plane_per_img) {
  __shared__ int s_q[64];
  const JpegImg& J = desc[blockIdx.z];
  const int c = blockIdx.y;
  if (c >= J.ncomp) return;
  if (threadIdx.x < 64) s_q[threadIdx.x] = J.q[c][threadIdx.x];
  __syncthreads();
  const int nb = J.bx[c] * J.by[c];
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  const uint4* src = reinterpret_cast<const uint4*>(coef + blockIdx.z * coef_per_img + J.coef_off[c] + (size_t)b * 64);
  int ws[8][8];
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    const uint4 t = __ldg(src + r);
    const uint32_t w4[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      ws[r][2 * k] = (int)(int16_t)(w4[k] & 0xffffu) * s_q[r * 8 + 2 * k];
      ws[r][2 * k + 1] = (int)(int16_t)(w4[k] >> 16) * s_q[r * 8 + 2 * k + 1];
    }
  }
  // pass 1: columns, scaled up by 2^PASS1_BITS
#pragma unroll
  for (int col = 0; col < 8; ++col) {
    int v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) v[r] = ws[r][col];
    idct8<13 - 2>(v);
#pragma unroll
    for (int r = 0; r < 8; ++r) ws[r][col] = v[r];
  }
  // pass 2: rows, + CENTERJSAMPLE, range limit
  const int byy = b / J.bx[c], bxx = b - byy * J.bx[c];
  const int stride = J.bx[c] * 8;
  uint8_t* dst = planes + blockIdx.z * plane_per_img + J.plane_off[c] + (size_t)(byy * 8) * stride + bxx * 8;
#pragma unroll
  for (int r = 0; r < 8; ++r) {
    idct8<13 + 2 + 3>(ws[r]);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      lo |= (uint32_t)min(max(ws[r][k] + 128, 0), 255) << (8 * k);
      hi |= (uint32_t)min(max(ws[r][4 + k] + 128, 0), 255) << (8 * k);
    }
    *reinterpret_cast<uint2*>(dst + (size_t)r * stride) = make_uint2(lo, hi);
  }
}

// --------------------------------------------------- device: upsampling + colour conversion
// chroma sample at full-resolution (x, y): jdsample.c h2v2_fancy_upsample / h2v1_fancy_upsample / fullsize.
// Fancy upsampling is selected only when downsampled_width > 2 (jinit_upsampler), else samples are replicated.
// The first / last column special cases of the C code equal the general formula with the
// neighbour index clamped; rows: the row above the first and below the last REAL row replicate it.
__device__ __forceinline__ int chroma_at(const uint8_t* p, int stride, int hs, int vs, int dw, int dh, int x, int y) {
  if (hs == 1) return p[(size_t)y * stride + x];
  if (dw <= 2) return p[(size_t)(vs == 2 ? y >> 1 : y) * stride + (x >> 1)];   // jinit_upsampler: box replication
  const int cx = x >> 1, ox = (x & 1) ? min(cx + 1, dw - 1) : max(cx - 1, 0);
  if (vs == 1) {
    const uint8_t* r = p + (size_t)y * stride;
    return (3 * r[cx] + r[ox] + ((x & 1) ? 2 : 1)) >> 2;
  }
  const int cy = y >> 1, oy = (y & 1) ? min(cy + 1, dh - 1) : max(cy - 1, 0);
  const uint8_t* r0 = p + (size_t)cy * stride;
  const uint8_t* r1 = p + (size_t)oy * stride;
  const int s0 = 3 * r0[cx] + r1[cx], s1 = 3 * r0[ox] + r1[ox];
  return (3 * s0 + s1 + ((x & 1) ? 7 : 8)) >> 4;
}

// grid (x chunks of 4 * 128 pixels, rows, images)
__global__ void __launch_bounds__(128) k_jpeg_color(const JpegImg* desc, const uint8_t* planes, size_t plane_per_img,
                                                    uint8_t* out, int h, int w) {
  const JpegImg& J = desc[blockIdx.z];
  const int y = blockIdx.y, x0 = (blockIdx.x * 128 + threadIdx.x) * 4;
  if (x0 >= w) return;
  const uint8_t* base = planes + blockIdx.z * plane_per_img;
  const uint8_t* py = base + J.plane_off[0] + (size_t)y * (J.bx[0] * 8);
  uint8_t* o = out + ((size_t)blockIdx.z * h + y) * (size_t)w * 3 + (size_t)x0 * 3;
  uint32_t bytes[12];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int x = min(x0 + k, w - 1);
    const int yy = py[x];
    int b = yy, g = yy, r = yy;
    if (J.ncomp == 3) {
      const int cb = chroma_at(base + J.plane_off[1], J.bx[1] * 8, J.hs, J.vs, J.dw, J.dh, x, y) - 128;
      const int cr = chroma_at(base + J.plane_off[2], J.bx[2] * 8, J.hs, J.vs, J.dw, J.dh, x, y) - 128;
      // jdcolor.c build_ycc_rgb_table: FIX(1.40200) = 91881, FIX(1.77200) = 116130, FIX(0.71414) = 46802, FIX(0.34414) = 22554
      r = yy + ((91881 * cr + 32768) >> 16);
      b = yy + ((116130 * cb + 32768) >> 16);
      g = yy + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
    }
    bytes[3 * k] = (uint32_t)min(max(b, 0), 255);
    bytes[3 * k + 1] = (uint32_t)min(max(g, 0), 255);
    bytes[3 * k + 2] = (uint32_t)min(max(r, 0), 255);
  }
  if (x0 + 4 <= w && (reinterpret_cast<uintptr_t>(o) & 3u) == 0) {
    uint32_t* ow = reinterpret_cast<uint32_t*>(o);
#pragma unroll
    for (int k = 0; k < 3; ++k)
      ow[k] = bytes[4 * k] | (bytes[4 * k + 1] << 8) | (bytes[4 * k + 2] << 16) | (bytes[4 * k + 3] << 24);
  } else {
    for (int k = 0; k < 12 && x0 * 3 + k < w * 3; ++k) o[k] = (uint8_t)bytes[k];
  }
}

}  // namespace b2d

using namespace b2d;

static const int kJpegChunk = 16, kJpegRing = 8;   // images per upload / launch, chunks in the staging ring (128 images)

struct b2d_jpeg {
  int device, max_n, cap, h, w, threads;     // cap = ring of staged images (pinned + device scratch), a multiple of the chunk size
  size_t coef_per_img, plane_per_img;      // int16 / bytes
  int16_t *h_coef, *d_coef;
  JpegImg *h_desc, *d_desc;
  uint8_t* d_planes;
  cudaEvent_t chunk_ev[16];                 // upload of ring chunk k finished -> its pinned slots are free again
  cudaEvent_t uploaded;                      // last upload of the previous call
  bool pending;
};

#define CKJ(x)                                          \
  do {                                                  \
    cudaError_t e__ = (x);                              \
    if (e__ != cudaSuccess) return fail_cuda(e__, #x);  \
  } while (0)

extern "C" {

int b2d_jpeg_info(const uint8_t* data, size_t len, int* h, int* w, int* components, int* hsamp, int* vsamp) {
  if (!data) return fail(B2D_ERR_INVALID, "b2d_jpeg_info: null data");
  Parsed* P = new (std::nothrow) Parsed();
  if (!P) return fail(B2D_ERR_INVALID, "out of host memory");
  std::string msg;
  const int rc = parse_jpeg(data, len, *P, msg, true);
  if (rc == B2D_OK) {
    if (h) *h = P->h;
    if (w) *w = P->w;
    if (components) *components = P->ncomp;
    if (hsamp) *hsamp = P->hs;
    if (vsamp) *vsamp = P->vs;
  }
  delete P;
  return rc == B2D_OK ? B2D_OK : fail(rc, "b2d_jpeg_info: %s", msg.c_str());
}

int b2d_jpeg_entropy(const uint8_t* data, size_t len, int16_t* h_coef, size_t capacity, int* blocks_x, int* blocks_y) {
  if (!data || !h_coef || !blocks_x || !blocks_y) return fail(B2D_ERR_INVALID, "b2d_jpeg_entropy: null argument");
  Parsed* P = new (std::nothrow) Parsed();
  if (!P) return fail(B2D_ERR_INVALID, "out of host memory");
  std::string msg;
  int rc = parse_jpeg(data, len, *P, msg, false);
  JpegImg J;
  if (rc == B2D_OK) {
    layout(*P, J);
    size_t need = 0;
    for (int c = 0; c < 3; ++c) {
      blocks_x[c] = c < J.ncomp ? J.bx[c] : 0;
      blocks_y[c] = c < J.ncomp ? J.by[c] : 0;
      need += (size_t)blocks_x[c] * blocks_y[c] * 64;
    }
    if (need > capacity) { rc = B2D_ERR_INVALID; msg = "coefficient buffer too small"; }
    else rc = entropy_decode(data, len, *P, J, h_coef, msg);
  }
  delete P;
  return rc == B2D_OK ? B2D_OK : fail(rc, "b2d_jpeg_entropy: %s", msg.c_str());
}

int b2d_jpeg_create(b2d_jpeg** out, int device, int max_n, int h, int w, int host_threads) {
  if (!out) return fail(B2D_ERR_INVALID, "b2d_jpeg_create: null argument");
  *out = nullptr;
  if (max_n <= 0 || h <= 0 || w <= 0 || h > 65535 || w > 65535 || (size_t)h * w > (1u << 28)) return fail(B2D_ERR_INVALID, "b2d_jpeg_create: bad shape n=%d h=%d w=%d", max_n, h, w);
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(B2D_ERR_NO_DEVICE, "no CUDA device: the JPEG stage has no CPU fallback");
  }
  if (device < 0 || device >= ndev) return fail(B2D_ERR_INVALID, "device %d out of range (%d devices)", device, ndev);
  DeviceGuard guard(device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  b2d_jpeg* D = new (std::nothrow) b2d_jpeg();
  if (!D) return fail(B2D_ERR_INVALID, "out of host memory");
  memset(D, 0, sizeof(*D));
  D->device = device; D->max_n = max_n; D->h = h; D->w = w;
  D->cap = std::min((max_n + kJpegChunk - 1) / kJpegChunk, kJpegRing) * kJpegChunk;
  D->threads = host_threads > 0 ? host_threads : (int)std::max(1u, std::thread::hardware_concurrency());
  const size_t hp = ((size_t)h + 15) / 16 * 16, wp = ((size_t)w + 15) / 16 * 16;   // whole MCUs for every sampling
  D->coef_per_img = 3 * hp * wp;
  D->plane_per_img = 3 * hp * wp;
  cudaError_t e = cudaMallocHost(&D->h_coef, D->coef_per_img * D->cap * sizeof(int16_t));
  if (e == cudaSuccess) e = cudaMallocHost(&D->h_desc, sizeof(JpegImg) * D->cap);
  if (e == cudaSuccess) e = cudaMalloc(&D->d_coef, D->coef_per_img * D->cap * sizeof(int16_t));
  if (e == cudaSuccess) e = cudaMalloc(&D->d_desc, sizeof(JpegImg) * D->cap);
  if (e == cudaSuccess) e = cudaMalloc(&D->d_planes, D->plane_per_img * D->cap);
  if (e == cudaSuccess) e = cudaEventCreateWithFlags(&D->uploaded, cudaEventDisableTiming);
  for (int k = 0; k < kJpegRing && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&D->chunk_ev[k], cudaEventDisableTiming);
  if (e != cudaSuccess) {
    b2d_jpeg_destroy(D);
    return fail_cuda(e, "b2d_jpeg_create allocations");
  }
  *out = D;
  return B2D_OK;
}

void b2d_jpeg_destroy(b2d_jpeg* D) {
  if (!D) return;
  DeviceGuard guard(D->device);
  if (D->pending) cudaEventSynchronize(D->uploaded);
  if (D->h_coef) cudaFreeHost(D->h_coef);
  if (D->h_desc) cudaFreeHost(D->h_desc);
  if (D->d_coef) cudaFree(D->d_coef);
  if (D->d_desc) cudaFree(D->d_desc);
  if (D->d_planes) cudaFree(D->d_planes);
  if (D->uploaded) cudaEventDestroy(D->uploaded);
  for (int k = 0; k < kJpegRing; ++k)
    if (D->chunk_ev[k]) cudaEventDestroy(D->chunk_ev[k]);
  delete D;
}

static int jpeg_decode_impl(b2d_jpeg* D, int n, const uint8_t* const* data, const size_t* len, uint8_t* d_out, void* cuda_stream);

int b2d_jpeg_decode(b2d_jpeg* D, int n, const uint8_t* const* data, const size_t* len, uint8_t* d_out, void* cuda_stream) {
  if (!D || !data || !len || (!d_out && n > 0)) return fail(B2D_ERR_INVALID, "b2d_jpeg_decode: null argument");
  if (n < 0 || n > D->max_n) return fail(B2D_ERR_INVALID, "b2d_jpeg_decode: n=%d outside [0, %d]", n, D->max_n);
  try {                                                        // std::thread / std::vector may throw: nothing crosses the ABI
    return jpeg_decode_impl(D, n, data, len, d_out, cuda_stream);
  } catch (const std::exception& e) {
    return fail(B2D_ERR_INVALID, "b2d_jpeg_decode: %s", e.what());
  } catch (...) {
    return fail(B2D_ERR_INVALID, "b2d_jpeg_decode: unknown host error");
  }
}

static int jpeg_decode_impl(b2d_jpeg* D, int n, const uint8_t* const* data, const size_t* len, uint8_t* d_out, void* cuda_stream) {
  if (n == 0) return B2D_OK;
  DeviceGuard guard(D->device);
  if (!guard.ok()) return fail_cuda(guard.err, "cudaSetDevice");
  if (D->pending) { CKJ(cudaEventSynchronize(D->uploaded)); D->pending = false; }   // the whole pinned ring is free again
  // Host threads decode images in order into a ring of `cap` staging slots (image i -> slot i % cap); the calling
  // thread uploads and launches chunk after chunk as soon as a chunk's images are done and releases a chunk's slots
  // when its upload event has fired, so copies and kernels run under the Huffman decoding of the following chunks
  // and the pinned memory stays bounded for any batch size.
  constexpr int CH = kJpegChunk;
  const int cap = D->cap, ring = cap / CH;
  std::vector<int> rc(n, B2D_OK);
  std::vector<std::string> msg(n);
  std::vector<size_t> used(n, 0);
  std::vector<std::atomic<int>> done(n);
  for (auto& f : done) f.store(0, std::memory_order_relaxed);
  std::atomic<int> next(0), free_upto(0);                      // images < free_upto have left the staging ring
  std::atomic<bool> abort(false);
  auto decode_one = [&](int i, Parsed* P) {
    *P = Parsed();
    if (!data[i]) { rc[i] = B2D_ERR_INVALID; msg[i] = "null stream"; return; }
    rc[i] = parse_jpeg(data[i], len[i], *P, msg[i], false);
    if (rc[i] != B2D_OK) return;
    if (P->h != D->h || P->w != D->w) {
      rc[i] = B2D_ERR_INVALID;
      msg[i] = "image is " + std::to_string(P->h) + " x " + std::to_string(P->w) + ", decoder was created for " +
               std::to_string(D->h) + " x " + std::to_string(D->w);
      return;
    }
    JpegImg& J = D->h_desc[i % cap];
    layout(*P, J);
    used[i] = (size_t)J.coef_off[J.ncomp - 1] + (size_t)J.bx[J.ncomp - 1] * J.by[J.ncomp - 1] * 64;
    rc[i] = entropy_decode(data[i], len[i], *P, J, D->h_coef + (size_t)(i % cap) * D->coef_per_img, msg[i]);
  };
  auto work = [&]() {
    Parsed* P = new (std::nothrow) Parsed();
    for (int i = next.fetch_add(1); i < n; i = next.fetch_add(1)) {
      while (i >= free_upto.load(std::memory_order_acquire) + cap && !abort.load(std::memory_order_relaxed))
        std::this_thread::sleep_for(std::chrono::microseconds(20));      // slot still being uploaded
      if (!P) { rc[i] = B2D_ERR_INVALID; msg[i] = "out of host memory"; }
      else if (!abort.load(std::memory_order_relaxed)) decode_one(i, P);
      done[i].store(1, std::memory_order_release);
    }
    delete P;
  };
  const int nt = std::min(D->threads, n);
  std::vector<std::thread> pool;
  if (nt > 1) {
    try {
      for (int t = 0; t < nt; ++t) pool.emplace_back(work);
    } catch (...) {                                            // thread limit: stop the ones that started, report
      abort.store(true);
      for (auto& t : pool) t.join();
      return fail(B2D_ERR_INVALID, "b2d_jpeg_decode: cannot start %d host threads", nt);
    }
  }
  cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
  const size_t hp = ((size_t)D->h + 15) / 16 * 16, wp = ((size_t)D->w + 15) / 16 * 16;
  const unsigned max_blocks = (unsigned)(hp * wp / 64);
  Parsed* mine = nt > 1 ? nullptr : new (std::nothrow) Parsed();
  int status = B2D_OK, bad = -1, enqueued = 0, released = 0;   // chunks
  cudaError_t ce = cudaSuccess;
  auto release = [&](bool block) {                             // publish the slots of chunks whose upload has finished
    while (released < enqueued) {
      const cudaError_t q = block ? cudaEventSynchronize(D->chunk_ev[released % ring]) : cudaEventQuery(D->chunk_ev[released % ring]);
      if (q == cudaErrorNotReady) return;
      if (q != cudaSuccess) { ce = q; return; }
      ++released;
      free_upto.store(std::min(n, released * CH), std::memory_order_release);
      block = false;
    }
  };
  for (int first = 0; first < n && status == B2D_OK && ce == cudaSuccess; first += CH) {
    const int cnt = std::min(CH, n - first);
    size_t width = 0;
    for (int i = first; i < first + cnt && ce == cudaSuccess; ++i) {
      if (nt > 1) {
        while (!done[i].load(std::memory_order_acquire) && ce == cudaSuccess) {
          release(false);
          std::this_thread::sleep_for(std::chrono::microseconds(30));   // not a spin: every core is decoding
        }
      } else if (mine) {
        while (i >= free_upto.load(std::memory_order_relaxed) + cap && ce == cudaSuccess) release(true);
        decode_one(i, mine);
      } else {
        rc[i] = B2D_ERR_INVALID; msg[i] = "out of host memory";
      }
      if (rc[i] != B2D_OK && status == B2D_OK) { status = rc[i]; bad = i; }
      width = std::max(width, used[i]);
    }
    if (status != B2D_OK || ce != cudaSuccess) break;
    const int slot = first % cap;
    const size_t pitch = D->coef_per_img * sizeof(int16_t);
    ce = cudaMemcpyAsync(D->d_desc + slot, D->h_desc + slot, sizeof(JpegImg) * cnt, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess)
      ce = cudaMemcpy2DAsync(D->d_coef + (size_t)slot * D->coef_per_img, pitch, D->h_coef + (size_t)slot * D->coef_per_img, pitch,
                             width * sizeof(int16_t), cnt, cudaMemcpyHostToDevice, st);
    if (ce == cudaSuccess) ce = cudaEventRecord(D->chunk_ev[enqueued % ring], st);
    if (ce != cudaSuccess) break;
    ++enqueued;
    k_jpeg_idct<<<dim3((max_blocks + 127) / 128, 3, cnt), 128, 0, st>>>(D->d_desc + slot, D->d_coef + (size_t)slot * D->coef_per_img,
                                                                       D->d_planes + (size_t)slot * D->plane_per_img,
                                                                       D->coef_per_img, D->plane_per_img);
    k_jpeg_color<<<dim3((D->w + 511) / 512, D->h, cnt), 128, 0, st>>>(D->d_desc + slot, D->d_planes + (size_t)slot * D->plane_per_img,
                                                                     D->plane_per_img, d_out + (size_t)first * D->h * D->w * 3, D->h, D->w);
    count_launch(); count_launch();
    ce = cudaGetLastError();
  }
  if (status != B2D_OK || ce != cudaSuccess) abort.store(true);
  for (auto& t : pool) t.join();
  delete mine;
  if (enqueued) {                                              // the ring is busy until the last upload of this call
    cudaEventRecord(D->uploaded, st);
    D->pending = true;
  }
  if (status != B2D_OK) return fail(status, "b2d_jpeg_decode: image %d: %s", bad, msg[bad].c_str());
  if (ce != cudaSuccess) return fail_cuda(ce, "b2d_jpeg_decode upload / kernels");
  return B2D_OK;
}

}  // extern "C"
