/* b200denoise.h - C ABI of libb200denoise.so
 *
 * B200 (sm_100a) implementation of the image-corruption + denoising
 * preprocessing path of mahesh-kashyap-ml/image-denoising.  The reference has
 * no FFI for this path: it is a set of Python closures
 *   add_<noise>_noise(spec)   lib/roi_data_layer/minibatch.py:87-1516
 *                             lib/model/test.py:193-1607
 * selected by substring tests on the `--noise` spec string
 *   (minibatch.py:1518-1673, test.py:1612-1831).
 * Each entry point below names the reference expression it replaces.  Plain
 * pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Conventions
 *   - every function returns 0 on success or a negative b2d_status; nothing
 *     throws or aborts across the ABI; b2d_last_error() gives the message of
 *     the last failure on the calling thread;
 *   - the caller owns all device buffers (input, output, workspace, fields);
 *     the library allocates only inside b2d_plan_create / b2d_run_host and
 *     frees in b2d_plan_destroy;
 *   - b2d_run is asynchronous on the caller's CUDA stream;
 *   - images are N x H x W x 3 uint8, interleaved BGR, C-contiguous (the
 *     layout cv2.imread gives the reference, minibatch.py:85);
 *   - randomness is Philox4x32-10 keyed by (seed) with counter
 *     (element group, substream, global image index): results do not depend on
 *     batch split, launch geometry or GPU count.  There is no global RNG.
 */
#ifndef B200DENOISE_H
#define B200DENOISE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  B2D_OK = 0,
  B2D_ERR_INVALID = -1,     /* bad argument                                  */
  B2D_ERR_UNSUPPORTED = -2, /* valid request this build cannot run           */
  B2D_ERR_CUDA = -3,        /* CUDA runtime error (message has the detail)   */
  B2D_ERR_PARSE = -4,       /* spec string not understood / unknown level    */
  B2D_ERR_NO_DEVICE = -5    /* no CUDA device: there is NO CPU fallback      */
} b2d_status;

/* order = the reference's top-level elif chain, minibatch.py:1575-1634 */
typedef enum {
  B2D_NOISE_NONE = 0,   /* retain_original, minibatch.py:1512-1516 */
  B2D_NOISE_GAUSSIAN,   /* add_gaussian_noise  minibatch.py:87-201   */
  B2D_NOISE_POISSON,    /* add_poisson_noise   minibatch.py:203-251  */
  B2D_NOISE_SAP,        /* add_sap_noise       minibatch.py:253-372  */
  B2D_NOISE_SPECKLE,    /* add_speckle_noise   minibatch.py:374-490  */
  B2D_NOISE_PERIODIC,   /* add_periodic_noise  minibatch.py:1034-1206 */
  B2D_NOISE_BROWNIAN,   /* add_brownian_noise  minibatch.py:809-1032 */
  B2D_NOISE_QUANT,      /* add_quant_noise     minibatch.py:492-667  */
  B2D_NOISE_UNIFORM,    /* add_uniform_noise   minibatch.py:669-807  */
  B2D_NOISE_GAMMA,      /* add_gamma_noise     minibatch.py:1208-1347 */
  B2D_NOISE_RAYLEIGH,   /* add_rayleigh_noise  minibatch.py:1349-1484 */
  B2D_NOISE_BLOOM,      /* add_bloom -> Automold.add_sun_flare, minibatch.py:1485-1489 */
  B2D_NOISE_SHADER,     /* add_shader (PIL Brightness x3), minibatch.py:1505-1511 */
  B2D_NOISE_COUNT
} b2d_noise;

typedef enum {
  B2D_DENOISE_NONE = 0,
  B2D_DENOISE_GAUS_BLUR, /* cv2.GaussianBlur(u8,(3,3),0)      minibatch.py:119 */
  B2D_DENOISE_MEAN,      /* cv2.blur(u8,(3,3))                minibatch.py:136 */
  B2D_DENOISE_MEDIAN,    /* cv2.medianBlur(u8,k) k=3 (5,7)    minibatch.py:153 */
  B2D_DENOISE_BILATERAL, /* cv2.bilateralFilter(u8,9,20,100,BORDER_CONSTANT) minibatch.py:172 */
  B2D_DENOISE_WAVELET,   /* skimage denoise_wavelet BayesShrink soft YCbCr  minibatch.py:92-96 */
  B2D_DENOISE_COUNT
} b2d_denoise;

enum { B2D_WAVELET_DB1 = 0, B2D_WAVELET_BIOR15 = 1 };

/* compat_flags: reference quirks that are NOT part of the canonical semantics, opt-in */
enum {
  B2D_COMPAT_SHADER_BGR = 1,  /* keep BGR order for `shader` (the reference returns RGB, quirk Q5)              */
  B2D_COMPAT_SAP_TRAIN06 = 2, /* noise-only `sap_var0.8` draws with amount 0.6, as the TRAINING closure does
                                 (lib/roi_data_layer/minibatch.py:367; test.py:434 uses 0.8)                    */
  B2D_COMPAT_GAUSSIAN_RANDOM_LEVEL = 16 /* quirk Q3 (minibatch.py:1575-1580, test.py:1678-1682): a spec that
                                 contains `gaussian` ignores its own level and denoiser and draws one of
                                 gaussian_var{0.1,1.0,1.5} with Python's random.choice.  HOST front-end flag
                                 (image_denoising_b200.spec.resolve): the choice uses the caller's `random`
                                 state, which does not exist below this ABI; b2d_plan_create ignores the bit. */
};

/* Diagnostic environment variables.  They select cross-check kernels inside b2d_run / b2d_run_host and
 * are honoured ONLY when B2D_DEBUG_KNOBS=1 is also set (read once per process); otherwise ignored.
 *   B2D_MARCH=0|1        3x3 filters: force the tile kernel (0) or the row-marching kernel (1)
 *   B2D_BROWN_MARCH=0|1  brownian: chunk-sum / scan / field kernels (0) or one CTA per image (1) whatever the batch size
 *   B2D_A4=0             rows of 4 (mod 8) bytes: group-wise kernel instead of the 32-bit tile / marching kernels
 *   B2D_FILTER_BYTES=1   W % 4 != 0: one-thread-per-byte 3x3 filter instead of the re-aligning tile kernel
 *   B2D_MEDIAN_MODE=1|2|3  median 5 / 7: one-output (1), pair (2) or 2x2-block (3) selection-network kernels instead
 *                        of the default (sorted rows merged, b2d_median7.cu)
 *   B2D_BILATERAL=1      bilateral d = 9: the 4 x 1-output kernel of round 1 instead of the 4 x 2-output one
 *   B2D_HAAR=0           db1 / 3 levels: general multi-pass wavelet path instead of the register-resident one
 *   B2D_WV_FUSED=0       general wavelet path: one launch per axis instead of the fused 2-D level kernels
 *   B2D_WV_GENERIC=1     general wavelet path: per-output reference kernels
 *   B2D_HOST_CHUNK_MB=n  b2d_run_host: pipeline chunk size in MiB (default 32)
 * The Python loader additionally reads B2D_LIB (path of an alternative build of this library; not gated). */

/* B2D_LEVEL_SIZE: periodic `varsize` (L = H*W*3, minibatch.py:1066) */
#define B2D_LEVEL_SIZE (-1.0)

typedef struct b2d_spec {
  int32_t noise;        /* b2d_noise                                             */
  int32_t denoise;      /* b2d_denoise                                           */
  double level;         /* var / amount / k / high / dt / L / scale (README.md:90-100) */
  int32_t ksize;        /* median window 3|5|7 (gaus_blur/mean are fixed 3)      */
  int32_t d;            /* bilateral diameter (9)                                */
  float sigma_color;    /* bilateral (20)                                        */
  float sigma_space;    /* bilateral (100)                                       */
  int32_t wavelet;      /* B2D_WAVELET_*                                         */
  int32_t levels;       /* wavelet levels, 0 = skimage's automatic choice        */
  int32_t compat_flags;
  int32_t reserved;
} b2d_spec;

/* Optional injected noise fields (DEVICE pointers, float64, one value per
 * uint8 element, N*H*W*3 each unless noted) so that the CUDA path and the CPU
 * oracle consume identical randomness.  Meaning by noise:
 *   gaussian/speckle f0 = n (already scaled by sqrt(var))  rayleigh/gamma/uniform f0 = samples
 *   sap       f0 = u1, f1 = u2                             poisson  f0 = k (integer draws)
 *   brownian  f0 = z, N*(H*W*3-1) standard normals         quant    f0 = centres, N*k*3 (Lab)
 *   bloom     f0 = N*8*7 doubles (alpha,cx,cy,radius,c0,c1,c2 per circle)            */
typedef struct b2d_fields {
  const double* f0;
  const double* f1;
} b2d_fields;

typedef struct b2d_plan b2d_plan; /* opaque */

/* Parse a reference `--noise` spec string (tools/test_net.py:47-49,
 * tools/trainval_net.py:52-54) with the substring precedence of the closures
 * (minibatch.py:89-199 and twins).  `mix`/`curvelet` specs are host-side
 * policy (random.choice over a list) and are rejected here with
 * B2D_ERR_UNSUPPORTED: the Python front-end resolves them to a concrete spec. */
int b2d_parse_spec(const char* spec, b2d_spec* out);

/* A plan fixes (device, N, H, W) and one spec per image (n_specs == N) or one
 * spec for the whole batch (n_specs == 1).  Replaces the per-image closure
 * definition + dispatch of test.py:189-1784.  N <= 2^24 (groups of more than 65535 images go out in
 * several launches internally: the image index is a grid dimension). */
int b2d_plan_create(b2d_plan** out, int device, int n, int h, int w,
                    const b2d_spec* specs, int n_specs);
int b2d_plan_workspace_bytes(const b2d_plan* plan, size_t* bytes);
void b2d_plan_destroy(b2d_plan* plan);

/* corrupt + denoise the whole batch:  d_out[i] = closure(spec_i)(d_in[i]).
 * d_ws: workspace of b2d_plan_workspace_bytes bytes (may be NULL when 0).
 * Image i of the batch uses Philox counter `first_global_image_index + i`. */
int b2d_run(b2d_plan* plan, const uint8_t* d_in, uint8_t* d_out, void* d_ws,
            const b2d_fields* fields_or_null, uint64_t seed,
            uint64_t first_global_image_index, void* cuda_stream);

/* Same call with HOST buffers (what `im = add_<noise>_noise(spec)` is for a
 * NumPy caller): chunked H2D copy, kernels and D2H copy pipelined on internal
 * streams with internal device buffers; returns after h_out is complete.
 * Pinned host memory (b2d_host_register) is needed for the copies to overlap. */
int b2d_run_host(b2d_plan* plan, const uint8_t* h_in, uint8_t* h_out,
                 uint64_t seed, uint64_t first_global_image_index);
int b2d_host_register(void* ptr, size_t bytes);
int b2d_host_unregister(void* ptr);

/* Test hook: write the noise field(s) the generator produces for this plan's
 * noise (float32, N*H*W*3 each; f1 only for sap) in the form the oracle's
 * `fields=` argument consumes, WITHOUT touching the image. */
int b2d_noise_field(b2d_plan* plan, float* d_f0, float* d_f1, uint64_t seed,
                    uint64_t first_global_image_index, void* cuda_stream);

/* Test hook: the per-channel statistics the wavelet denoiser derived for one image during the LAST b2d_run of this
 * plan (skimage _wavelet_threshold / _sigma_est_dwt, reached from lib/roi_data_layer/minibatch.py:92-94).
 * image_in_group = position of the image inside its spec group (= batch index when the plan has one spec); it must
 * lie in the last workspace chunk the run processed (small test batches are one chunk).  h_out[12] = for each of the
 * three "YCbCr" channels {sigma, min, max, count}: sigma = median(|dd_1| != 0) / 0.6745 in the channel's OWN units
 * (divide by max - min for skimage's normalised value), min / max = the channel range used for the normalisation and
 * the inner clip, count = how many dd_1 coefficients were non-zero.  Synchronises cuda_stream.  The parity tests use it
 * to hand the oracle the same exact-zero census (tests/test_gpu_wavelet.py). */
int b2d_wavelet_stats(const b2d_plan* plan, int image_in_group, float* h_out, void* cuda_stream);

/* Test hook: the uint8 colour conversions of the `quant` branch alone, on n_pixels interleaved
 * 3-byte pixels (device pointers): direction 0 = cv2.cvtColor(COLOR_BGR2LAB), 1 = COLOR_LAB2BGR
 * (lib/roi_data_layer/minibatch.py:496,507). */
int b2d_lab_convert(const uint8_t* d_in, uint8_t* d_out, size_t n_pixels, int direction, void* cuda_stream);

/* Test hook (host only, no GPU needed): the alias table of the generator-mode Poisson corruption
 * (skimage random_noise 'poisson', lib/roi_data_layer/minibatch.py:211-213) for vals = 2^log2_vals,
 * 0 <= log2_vals <= 8.  h_table[v * 256 + slot] = alias | threshold << 8: input byte v, 8 random
 * bits pick the slot, 8 more are compared with the threshold (u < threshold keeps the slot's own
 * value, otherwise the alias is the output byte). */
int b2d_poisson_table(int log2_vals, uint16_t* h_table);

/* Test hook (host only, no GPU needed): the tables of the generator-mode gamma corruption
 * (scipy.stats.gamma.rvs(1.99, scale) added with cv2.add and cast with wrap-around,
 * lib/roi_data_layer/minibatch.py:1216-1217,1247; twin lib/model/test.py).  With
 * K = floor(255 * scale * G) the uint8 result is (v + K) mod 256; h_alias[256] is the alias table
 * (alias | threshold << 8, as above) over R = K mod 256.  If h_fold is non-null it receives
 * h_fold[R * folds + j] = round(65536 * P(K <= R + 256 j | K mod 256 = R)), which b2d_noise_field uses
 * to unfold K; *folds (at most 64) is written when `folds` is non-null, so h_fold needs 256 * 64 entries. */
int b2d_gamma_table(double scale, uint16_t* h_alias, uint32_t* h_fold, int* folds);

/* ---- next stage: the network input blob, on the device ------------------------
 * b2d_blob_shape: geometry of prep_im_for_blob (lib/utils/blob.py:37-45) for an h x w image:
 *   im_scale = target_size / min(h, w), capped so that round(im_scale * max(h, w)) <= max_size;
 *   out_h, out_w = cv2.resize's dsize = (cvRound(h * im_scale), cvRound(w * im_scale)).  Host only.
 * b2d_blob_prep: d_out[i] (blob_h x blob_w x 3 float32, zero padded like im_list_to_blob,
 *   blob.py:17-30) = cv2.resize(d_in[i].astype(float32) - pixel_means, fx = fy = im_scale,
 *   INTER_LINEAR) (blob.py:35,43-44); d_flip[i] != 0 mirrors image i horizontally first
 *   (lib/roi_data_layer/minibatch.py:1676-1677).  d_in: n x h x w x 3 uint8; d_flip: n bytes on the
 *   device or NULL; pixel_means: 3 host doubles (cfg.PIXEL_MEANS, lib/model/config.py:252).
 *   flags: B2D_BLOB_F32_COORDS = interpolation fractions from the float32 source coordinate, bit-exact with
 *   OpenCV's own C++ resize (cv2.setUseOptimized(False)); 0 = double-precision coordinate, which is what the
 *   IPP-backed cv2 wheels return (they differ from the former by up to 4e-5 relative on downscales). */
#define B2D_BLOB_F32_COORDS 1
int b2d_blob_shape(int h, int w, int target_size, int max_size, double* im_scale, int* out_h, int* out_w);
int b2d_blob_prep(const uint8_t* d_in, int n, int h, int w, float* d_out, int blob_h, int blob_w,
                  double im_scale, const double* pixel_means, const uint8_t* d_flip, int flags, void* cuda_stream);

/* SURVEY 8(f) f2 - the stage in front of the path: `cv2.imread(path)` (lib/model/test.py:191,
 *   lib/roi_data_layer/minibatch.py:85, and again inside most closures: minibatch.py:206,256,377,...) for baseline
 *   JPEG streams.  cv2.imread = libjpeg-turbo defaults (JDCT_ISLOW, fancy upsampling, YCbCr -> BGR); the result is
 *   bit-exact with cv2.imdecode(IMREAD_COLOR).  Huffman decoding runs on `host_threads` host threads (0 = all cores),
 *   dequantisation + IDCT + chroma upsampling + colour conversion on the GPU.
 *   b2d_jpeg_info   : frame header of one stream (any pointer may be NULL).
 *   b2d_jpeg_create : decoder for batches of up to max_n images of exactly h x w (owns pinned staging + device scratch).
 *   b2d_jpeg_decode : data[i] / len[i] = the i-th file's bytes (host); d_out = n x h x w x 3 uint8 BGR on the device,
 *                     written asynchronously on cuda_stream.  Sampling 4:4:4, 4:2:2 (h2v1), 4:2:0 and greyscale (returned
 *                     as 3 equal channels, like IMREAD_COLOR), restart intervals; progressive / arithmetic / 12-bit /
 *                     CMYK streams fail with B2D_ERR_UNSUPPORTED - there is no fallback decoder. */
typedef struct b2d_jpeg b2d_jpeg;
int b2d_jpeg_info(const uint8_t* data, size_t len, int* h, int* w, int* components, int* hsamp, int* vsamp);
int b2d_jpeg_create(b2d_jpeg** out, int device, int max_n, int h, int w, int host_threads);
int b2d_jpeg_decode(b2d_jpeg* dec, int n, const uint8_t* const* data, const size_t* len, uint8_t* d_out, void* cuda_stream);
void b2d_jpeg_destroy(b2d_jpeg* dec);
/* host-only hook (no GPU needed): the quantised coefficients the Huffman stage hands to the GPU - component after
 * component, blocks in raster order over whole MCUs, 64 int16 per block in natural (row-major) order - for the CPU tests
 * that compare it with oracle/jpeg.py::entropy_decode.  blocks_x / blocks_y: 3 ints each. */
int b2d_jpeg_entropy(const uint8_t* data, size_t len, int16_t* h_coef, size_t capacity, int* blocks_x, int* blocks_y);

/* SURVEY 8(f) f4 - the reference's only native code: greedy NMS of score-sorted boxes, `_nms` of lib/nms/gpu_nms.hpp:1
 *   (lib/nms/nms_kernel.cu:91-144; bound by lib/nms/gpu_nms.pyx:16-31, dispatched by lib/model/nms_wrapper.py:15-23).
 *   b2d_nms has `_nms`'s exact signature and meaning: boxes_host = boxes_num x boxes_dim float32 rows
 *   [x1, y1, x2, y2, score, ...] ALREADY sorted by descending score (gpu_nms.pyx sorts before the call); keep_out receives
 *   the kept row positions in ascending order, *num_out their count; box j is dropped when a kept i < j has
 *   IoU > nms_overlap_thresh ("+1" pixel convention, float32, the kernel's operation order).  Returns a status instead of
 *   aborting (the reference's CUDA_CHECK exits the process).  The mask stays on the device and the sweep runs there.
 *   b2d_nms_device: the same on device buffers (d_ws of b2d_nms_workspace_bytes, asynchronous on cuda_stream). */
int b2d_nms(int* keep_out, int* num_out, const float* boxes_host, int boxes_num, int boxes_dim, float nms_overlap_thresh,
            int device_id);
int b2d_nms_workspace_bytes(int boxes_num, size_t* bytes);
int b2d_nms_device(const float* d_boxes, int boxes_num, int boxes_dim, float nms_overlap_thresh, int* d_keep, int* d_num,
                   void* d_ws, void* cuda_stream);

/* b2d_nms_device without any allocation and without a caller workspace.  NO COUNTERPART IN THE REFERENCE: added
 *   for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 *   serving loops, harnesses that forbid it).  Same arguments and result as b2d_nms_device minus d_ws: the suppression
 *   mask lives in a static device array inside the library, 2 MiB (B2D_NMS_STATIC_MAX_BOXES boxes x 64 words), held
 *   by every GPU that loads the library until the process exits.  Runs on the CURRENT device, asynchronous on
 *   cuda_stream; calls neither cudaMalloc / cudaFree nor any allocating library routine (its first call on a device
 *   resolves the array's address with cudaGetSymbolAddress - make that call before starting a capture).
 *   boxes_num > B2D_NMS_STATIC_MAX_BOXES fails with B2D_ERR_INVALID and a message naming the limit.
 *   Every call on a device shares the array and leaves its bits there; each call rewrites all the words it reads, so
 *   results never depend on earlier calls - but calls that OVERLAP on the GPU (two streams, no ordering) race. */
#define B2D_NMS_STATIC_MAX_BOXES 4096
int b2d_nms_device_static(const float* d_boxes, int boxes_num, int boxes_dim, float nms_overlap_thresh, int* d_keep,
                          int* d_num, void* cuda_stream);

/* Per-image result statistics gathered after the run (8(e): the only data
 * that ever crosses GPUs): checksum = sum of bytes * (1 + index mod 251)
 * mod 2^64 of d_out[i]; sse = sum (d_out - d_ref)^2 (for PSNR vs clean). */
int b2d_stats(b2d_plan* plan, const uint8_t* d_ref, const uint8_t* d_out,
              uint64_t* d_checksum, double* d_sse, void* cuda_stream);

/* number of kernel launches issued by this library on the calling thread
 * since the last b2d_reset_launch_count() (bench.py's gpu_launches) */
uint64_t b2d_launch_count(void);
void b2d_reset_launch_count(void);

const char* b2d_last_error(void);
const char* b2d_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200DENOISE_H */
