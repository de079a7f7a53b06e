/*
 * b200img.h -- C ABI of libb200img.so: image-rs/image's data-parallel pixel
 * hot path (resampler, Gaussian / box-approximated blur, pixel-format
 * conversions, BGR(A) swizzle) as hand-written sm_100a CUDA kernels.
 *
 * The reference has no FFI or operator-plugin interface for imageops (its only
 * plugin API is decoder hooks, src/hooks.rs:77-152), so the drop-in boundary
 * is this library: plain pointers and sizes, no torch / C++ types.  Each entry
 * point cites the reference function it replaces (file:line relative to the
 * image-rs/image repository).  INTEGRATION.md shows the Rust `extern "C"`
 * block a maintainer adds on the reference side.
 *
 * Conventions
 *   - Images are packed row-major, interleaved channels, no row padding -- the
 *     packed case of the reference's #[repr(C)] SampleLayout
 *     (src/images/flat.rs:97-118; src/images/buffer.rs:550-557).
 *   - dtype: B200IMG_U8 / U16 / F32 subpixels; channels 1..4 = Luma, LumaA,
 *     Rgb, Rgba.
 *   - Every function returns 0 on success or a negative b200img_status and
 *     never unwinds; b200img_last_error() gives the thread's last message.
 *     Where the reference panics (asserts in sample.rs:1134-1158, 1211-1214,
 *     `unimplemented!` for > 4 channels) the call returns B200IMG_ERR_BAD_ARG;
 *     where it returns ImageError::Limits(DimensionError)
 *     (src/imageops/filter_1d.rs:638-647) it returns B200IMG_ERR_DIMENSION.
 *   - `*_dev` entry points take DEVICE pointers valid on the current CUDA
 *     device and enqueue on `stream` (a cudaStream_t passed as void*, NULL =
 *     the legacy default stream) without synchronising.  The entry points
 *     without the suffix take HOST pointers, copy in, run, copy out, and
 *     return after the result is in `dst`.
 *   - There is no CPU fallback: without a usable GPU every compute entry point
 *     fails with B200IMG_ERR_CUDA.
 *   - Thread-safe.  Scratch memory and weight tables are cached per device.
 */
#ifndef B200IMG_H
#define B200IMG_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    B200IMG_OK = 0,
    B200IMG_ERR_BAD_ARG = -1,     /* the reference would panic / not compile */
    B200IMG_ERR_DIMENSION = -2,   /* ImageError::Limits(DimensionError) */
    B200IMG_ERR_OOM = -3,         /* device or pinned-host allocation failed */
    B200IMG_ERR_CUDA = -4,        /* no device, launch failure, ... */
    B200IMG_ERR_UNSUPPORTED = -5
} b200img_status;

typedef enum { B200IMG_U8 = 0, B200IMG_U16 = 1, B200IMG_F32 = 2 } b200img_dtype;

/* FilterType, src/imageops/sample.rs:98-115 */
typedef enum {
    B200IMG_NEAREST = 0, B200IMG_TRIANGLE = 1, B200IMG_CATMULLROM = 2,
    B200IMG_GAUSSIAN = 3, B200IMG_LANCZOS3 = 4
} b200img_filter;

/* Arithmetic flags.  FAST lets f32 multiply-adds fuse (FFMA); results of the
 * f32-accumulating filters are then within +-1 LSB (u8/u16) or 1e-5 relative
 * (f32) of the reference.  STRICT keeps every multiply and add separately
 * rounded, in the reference's order: bit-exact with a default x86-64 build of
 * the reference given the same libm.  Integer paths (conversions, swizzle,
 * Nearest, u8 box blur, Q0.15 Gaussian) are bit-exact under either flag.
 * One more difference under FAST, for f32 images only: the resampler's grouped kernel walks the
 * union of four neighbouring windows with zero weights outside each sample's own window, so a
 * NON-FINITE source sample (0 * inf = NaN) can also turn up to three destination samples next to
 * the ones whose windows contain it non-finite.  The Gaussian blur's FAST kernels pad their
 * kernels with zero taps in the same way (up to 7 samples beyond the window).  STRICT treats such
 * samples exactly as the reference does. */
#define B200IMG_FP_FAST 0
#define B200IMG_FP_STRICT 1
/* multiply_accumulate of the reference fuses when it is built with +fma or for
 * aarch64 (src/math/utils.rs:7-47).  With STRICT, add this bit to reproduce
 * such a build in the Gaussian engine and the Luma<->RGB fallback. */
#define B200IMG_FP_FUSED_MAC 2
/* Resampler only: run the reference's own structure (vertical kernel, f32
 * intermediate in device scratch, horizontal kernel) instead of the fused
 * tile kernel.  Chosen automatically when a downscale is so extreme that a
 * tile's source footprint does not fit shared memory. */
#define B200IMG_RESIZE_UNFUSED 256
/* Resampler only: use the tile kernel even where the column-streaming kernel (downscales of
 * u8 / u16 images) applies.  Same results; exists so that both kernels can be tested. */
#define B200IMG_RESIZE_TILED 512
/* Resampler only: do not use the fixed-window kernels -- downscales by an integer ratio
 * (2x, 3x, 4x) and exact 2x upscales of four-channel images, where every window has the same
 * shape, the weights are one or two vectors and border windows are renormalised after the sum -- but
 * the table-driven kernels that serve every other size.  Both are within the FAST contract (+-1 LSB /
 * 1e-5 relative of the reference); they may differ from each other by rounding.  For tests. */
#define B200IMG_RESIZE_GENERAL 8192
/* thumbnail only: evaluate every output pixel on its own (the kernel that serves upscales and
 * unaligned rows) even where the column-sum kernel applies.  Same results; for tests. */
#define B200IMG_THUMBNAIL_PER_PIXEL 1024

/* Gaussian blur only: run the row pass and the column pass as two kernels with the intermediate
 * image in device scratch, even where the single fused kernel (intermediate in shared memory)
 * applies.  Same results; exists so that both can be tested. */
#define B200IMG_GAUSS_TWO_KERNELS 2048
/* ... and the opposite: use the fused kernel wherever it applies, also on jobs where the library
 * would pick the kernel pair for speed (short kernels, small images). */
#define B200IMG_GAUSS_FUSED 4096

const char *b200img_last_error(void);
int b200img_version(void);
int b200img_device_count(int *count);

/* Pinned host memory for the host-pointer entry points (optional; any host
 * pointer works, pinned ones copy at full PCIe rate). */
int b200img_pinned_alloc(size_t bytes, void **out);
int b200img_pinned_free(void *p);
/* Drop this thread's current device's cached scratch / tables. */
int b200img_release_caches(void);

/* ---- host-side tables (what the reference computes per call on the CPU) -- */

/* Filter{kernel,support} table in resize(), src/imageops/sample.rs:993-1014;
 * kernel functions 127-205. */
float b200img_filter_support(int filter);
float b200img_filter_kernel(int filter, float x);
/* `max_ks`, src/imageops/sample.rs:242. */
int b200img_resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len);
/* Weight generation of vertical_sample (axis 0, f32 centre;
 * src/imageops/sample.rs:514-544) and horizontal_sample (axis 1, f64 centre;
 * 237-307).  left/count have dst_len entries; weights is dst_len x stride,
 * zero-padded, normalised by division exactly as the reference does. */
int b200img_build_weights(int filter, uint32_t src_len, uint32_t dst_len, int axis,
                          int32_t *left, int32_t *count, float *weights, int stride);
/* resize_dimensions, src/math/utils.rs:56-84 (DynamicImage::resize /
 * resize_to_fill / thumbnail call it before resizing). */
int b200img_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, int fill,
                              uint32_t *out_w, uint32_t *out_h);

/* GaussianBlurParameters, src/imageops/sample.rs:1081-1295 */
typedef struct {
    uint32_t x_kernel_size; float x_sigma;
    uint32_t y_kernel_size; float y_sigma;
} b200img_gauss_params;
int b200img_gauss_params_from_radius(float radius, b200img_gauss_params *out);       /* 1132-1142 */
int b200img_gauss_params_from_kernel_size(float ks, b200img_gauss_params *out);      /* 1148-1166 */
int b200img_gauss_params_from_anisotropic(float xks, float yks, b200img_gauss_params *out); /* 1172-1204 */
int b200img_gauss_params_from_sigma(float sigma, b200img_gauss_params *out);         /* 1210-1249 */
/* get_gaussian_kernel_1d, src/imageops/sample.rs:1057-1077 */
int b200img_gaussian_kernel_1d(uint32_t width, float sigma, float *out);
/* boxes_for_gauss(sigma, 3), src/imageops/fast_blur.rs:201-226 */
int b200img_boxes_for_gauss(float sigma, uint32_t out_radii[3]);

/* ---- resampler ----------------------------------------------------------
 * imageops::resize, src/imageops/sample.rs:970-1019 (vertical_sample 494-562
 * then horizontal_sample 217-340, store rule src/primitive_sealed.rs:107-147).
 * acc_init is the vertical accumulator's start value: 1.0f reproduces the
 * reference literally (`Rgba([1.0; 4])`, sample.rs:547), 0.0f is the
 * mathematically intended filter.  An empty source gives a zero-filled
 * destination (977-984); equal dimensions give a copy (987-991). */
int b200img_resize(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                   void *dst, uint32_t nw, uint32_t nh, int filter, float acc_init, int flags);
int b200img_resize_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                       void *d_dst, uint32_t nw, uint32_t nh, int filter, float acc_init,
                       int flags, void *stream);
/* Device-resident batch variant: n_images equally-shaped images,
 * src_stride / dst_stride bytes apart (0 = tightly packed).  One launch. */
int b200img_resize_batch_dev(const void *d_src, size_t src_stride, uint32_t w, uint32_t h,
                             int channels, int dtype, void *d_dst, size_t dst_stride,
                             uint32_t nw, uint32_t nh, uint32_t n_images, int filter,
                             float acc_init, int flags, void *stream);
/* Host-resident batch: n_images equally-shaped, tightly packed images in HOST memory (pinned
 * memory recommended).  Uploads, kernels and downloads of consecutive images overlap on three
 * internal streams; returns when every result is in `dst`. */
int b200img_resize_batch(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                         uint32_t nw, uint32_t nh, uint32_t n_images, int filter, float acc_init, int flags);
/* One row band of a larger resize (multi-GPU single-image split): computes
 * destination rows [dst_row0, dst_row0 + dst_rows) of the nw x nh result.
 * d_src_band holds source rows [src_row0, src_row0 + src_rows) of the w x h
 * source, which must cover every row those destination rows tap (query with
 * b200img_resize_band_src_rows).  d_dst_band receives dst_rows x nw pixels. */
int b200img_resize_band_src_rows(uint32_t h, uint32_t nh, int filter, uint32_t dst_row0,
                                 uint32_t dst_rows, uint32_t *src_row0, uint32_t *src_rows);
int b200img_resize_band_dev(const void *d_src_band, uint32_t src_row0, uint32_t src_rows,
                            uint32_t w, uint32_t h, int channels, int dtype, void *d_dst_band,
                            uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh,
                            int filter, float acc_init, int flags, void *stream);

/* ---- Gaussian blur -------------------------------------------------------
 * path GENERIC = imageops::blur / blur_advanced (src/imageops/sample.rs:
 * 1034-1055, gaussian_blur_indirect 1492-1604): f32 engine for every dtype,
 * one rounding at the end.  path DYN = DynamicImage::blur / blur_advanced
 * (src/images/dynimage.rs:1226-1245, gaussian_blur_dyn_image sample.rs:
 * 1297-1490): Q0.15 fixed point with a u32 accumulator for u8 and u16
 * (bit-exact), f32 engine for f32.  Engine: filter_2d_separable,
 * src/imageops/filter_1d.rs:615-761 (row 323-390, column 171-316, clamp-to-
 * edge 60-130, store 132-164, Q15 transform 392-420, zero-end strip 423-437). */
#define B200IMG_BLUR_GENERIC 0
#define B200IMG_BLUR_DYN 1
int b200img_gaussian_blur(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                          void *dst, const b200img_gauss_params *params, int path, int flags);
int b200img_gaussian_blur_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                              void *d_dst, const b200img_gauss_params *params, int path,
                              int flags, void *stream);
/* host-resident batch, pipelined like b200img_resize_batch */
int b200img_gaussian_blur_batch(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                                uint32_t n_images, const b200img_gauss_params *params, int path, int flags);
int b200img_gaussian_blur_batch_dev(const void *d_src, uint32_t w, uint32_t h, int channels,
                                    int dtype, void *d_dst, uint32_t n_images,
                                    const b200img_gauss_params *params, int path, int flags,
                                    void *stream);

/* ---- box-approximated blur ----------------------------------------------
 * imageops::fast_blur / DynamicImage::fast_blur, src/imageops/fast_blur.rs:
 * 111-191 (horizontal pass 303-413, vertical pass 416-473, accumulators
 * 36-94).  u8: exact integer; u16: exact while kernel*65535 < 2^24. */
int b200img_fast_blur(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                      void *dst, float sigma);
int b200img_fast_blur_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                          void *d_dst, float sigma, void *stream);
/* host-resident batch, pipelined like b200img_resize_batch */
int b200img_fast_blur_batch(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                            uint32_t n_images, float sigma);
int b200img_fast_blur_batch_dev(const void *d_src, uint32_t w, uint32_t h, int channels,
                                int dtype, void *d_dst, uint32_t n_images, float sigma,
                                void *stream);

/* No counterpart in the reference.  Same result as b200img_fast_blur_dev for one image, but
 * this entry point allocates nothing: the blur's transient image lives in a fixed 32 MiB
 * array inside the library (held by every GPU that loads it, until the process exits)
 * instead of a cudaMalloc'd pool.  It is for callers that cannot let a library allocate --
 * a serving process whose own pool owns all device memory, or a latency-critical loop that
 * cannot take cudaMalloc's device-wide synchronisation.  Limits and rules:
 *   - w * h * channels * sizeof(subpixel) must be <= 33,554,432 bytes (3840x2160 Rgba8
 *     fits); larger images are refused with B200IMG_ERR_UNSUPPORTED.
 *   - The first call on each device looks up the array's device address; make that first
 *     call outside any stream capture.
 *   - All calls on a device share the one array and it keeps whatever the last call left in
 *     it (every call overwrites what it reads, nothing needs clearing).  Two calls whose GPU
 *     work can overlap -- different streams, or different host threads -- race on it: run
 *     them on one stream or order them yourself. */
int b200img_fast_blur_noalloc_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                                  void *d_dst, float sigma, void *stream);

/* ---- one image over several GPUs: row bands with ghost rows ------------------
 * The path has no reduction across shards, so nothing here is a collective and NCCL is not
 * involved: a band reads a few source rows beyond its own (the rows its vertical taps touch,
 * src/imageops/sample.rs:494-562; taps/2 rows for the Gaussian, whose column pass clamps at the
 * IMAGE's edges only, src/imageops/filter_1d.rs:102-130; r1+r2+r3 rows for the three composed
 * boxes of fast_blur, src/imageops/fast_blur.rs:303-473).  Those "ghost rows" are copied once
 * from the bands that own them (cudaMemcpyPeerAsync: NVLink where peer access exists), then every
 * device runs the ordinary kernels on its band.  Bands are never split in x: the horizontal
 * passes need whole rows (sample.rs:217-340).  Results are those of the whole-image call
 * (bit-identical where that is bit-exact).
 *
 * A b200img_band describes one device's share.  d_src is a buffer on `device` holding source rows
 * [src_row0, src_row0 + src_rows); of these the device OWNS [own_row0, own_row0 + own_rows), which
 * must be valid on entry (ordered on the band's stream), and the call fills the others.  d_dst
 * receives destination rows [dst_row0, dst_row0 + dst_rows).  The plan functions fill the six row
 * fields for an even split over n_bands (the caller sets device, d_src, d_dst afterwards).
 *
 * `streams`: NULL = internal per-device streams, the call returns when every band is done; or an
 * array of n_bands cudaStream_t (one per band, on that band's device): work is enqueued and the
 * call returns without waiting.  One host thread drives all devices; the calling thread's current
 * device is restored. */
typedef struct {
    int device;
    void *d_src;
    uint32_t src_row0, src_rows;
    uint32_t own_row0, own_rows;
    void *d_dst;
    uint32_t dst_row0, dst_rows;
} b200img_band;
int b200img_plan_resize_bands(uint32_t h, uint32_t nh, int filter, int n_bands, b200img_band *bands);
int b200img_plan_gaussian_blur_bands(uint32_t h, const b200img_gauss_params *params, int n_bands, b200img_band *bands);
int b200img_plan_fast_blur_bands(uint32_t h, float sigma, int n_bands, b200img_band *bands);
/* only the ghost-row copies (row_bytes = bytes of one source row) */
int b200img_exchange_ghost_rows(const b200img_band *bands, int n_bands, uint32_t h, size_t row_bytes,
                                void *const *streams);
/* imageops::resize of one w x h image into nw x nh, band by band */
int b200img_resize_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int channels, int dtype,
                         uint32_t nw, uint32_t nh, int filter, float acc_init, int flags, void *const *streams);
/* imageops::blur_advanced / DynamicImage::blur_advanced of one w x h image, band by band */
int b200img_gaussian_blur_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int channels,
                                int dtype, const b200img_gauss_params *params, int path, int flags,
                                void *const *streams);
/* imageops::fast_blur of one w x h image, band by band */
int b200img_fast_blur_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int channels, int dtype,
                            float sigma, void *const *streams);
/* One band on the current device, ghost rows already in place (what the calls above run per
 * device; also the building block of a one-process-per-GPU split).  d_src_band holds source rows
 * [src_row0, src_row0 + src_rows), which must include every row the destination rows read
 * (b200img_*_band_src_rows). */
int b200img_gaussian_blur_band_src_rows(uint32_t h, const b200img_gauss_params *params, uint32_t dst_row0,
                                        uint32_t dst_rows, uint32_t *src_row0, uint32_t *src_rows);
int b200img_gaussian_blur_band_dev(const void *d_src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w,
                                   uint32_t h, int channels, int dtype, void *d_dst_band, uint32_t dst_row0,
                                   uint32_t dst_rows, const b200img_gauss_params *params, int path, int flags,
                                   void *stream);
int b200img_fast_blur_band_src_rows(uint32_t h, float sigma, uint32_t dst_row0, uint32_t dst_rows,
                                    uint32_t *src_row0, uint32_t *src_rows);
int b200img_fast_blur_band_dev(const void *d_src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h,
                               int channels, int dtype, void *d_dst_band, uint32_t dst_row0, uint32_t dst_rows,
                               float sigma, void *stream);

/* ---- thumbnail (the first caller next to the core path) --------------------
 * imageops::thumbnail, src/imageops/sample.rs:564-886 (DynamicImage::thumbnail_exact,
 * src/images/dynimage.rs:1175).  Integer images are bit-exact, including the reference's
 * literal treatment of windows above 2^16 pixels (256x256 sub-block averages divided by
 * (w/256)*(h/256)); f32 images sum in f64 and agree to f32 rounding.  An integer image whose
 * window exceeds 2^16 pixels while being under 256 wide or high makes the reference divide
 * by zero: B200IMG_ERR_BAD_ARG here. */
int b200img_thumbnail(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                      uint32_t nw, uint32_t nh, int flags);
int b200img_thumbnail_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                          void *d_dst, uint32_t nw, uint32_t nh, int flags, void *stream);
/* n_images equally-shaped, tightly packed device images; one launch */
int b200img_thumbnail_batch_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                                void *d_dst, uint32_t nw, uint32_t nh, uint32_t n_images, int flags,
                                void *stream);

/* ---- 3x3 filter and unsharp mask (the callers next to the core path) -------
 * imageops::filter3x3, src/imageops/sample.rs:897-956: f32 accumulation in tap order, times
 * 1/sum(kernel) (1 if the sum is 0), clamp to [0, max], truncating cast; the one-pixel border
 * keeps the zero fill.  imageops::unsharpen, sample.rs:1617-1654: blur_advanced(new_from_sigma)
 * then the integer rule |c - blurred| > threshold ? clamp(2c - blurred, 0, max) : c.
 * `flags` are the Gaussian engine's (STRICT = bit-exact). */
int b200img_filter3x3(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                      const float kernel[9]);
int b200img_filter3x3_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                          const float kernel[9], void *stream);
int b200img_unsharpen(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                      float sigma, int threshold, int flags);
int b200img_unsharpen_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                          float sigma, int threshold, int flags, void *stream);

/* ---- either side of DynamicImage::resize (src/imageops/resize.rs) -----------
 * has_constant_alpha: resize.rs:155-218, restated exactly (integer alphas XOR-ed with the first
 * pixel's and summed; f32 alphas compared with ==, so a NaN alpha is never constant; no alpha
 * channel or an empty image = constant).  *result = 1 / 0.  The _dev form synchronises `stream`.
 * premultiply / unpremultiply (in place; LumaA and Rgba only, others untouched): the reference
 * calls the unvendored pic-scale-safe crate here -- PARITY UNPINNED; this library rounds to
 * nearest: (c*a + max/2)/max, and min((c*max + a/2)/a, max) with a = 0 keeping c; f32: c*a, c/a.
 * convert_resize: the decode hand-off -- one upload of the decoded buffer, DynamicImage::to_*
 * (B200IMG_CAST) on the device, imageops::resize, one download. */
int b200img_has_constant_alpha(const void *src, uint32_t w, uint32_t h, int channels, int dtype, int *result);
int b200img_has_constant_alpha_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                                   int *result, void *stream);
int b200img_premultiply_alpha(void *buf, uint32_t w, uint32_t h, int channels, int dtype);
int b200img_premultiply_alpha_dev(void *d_buf, uint32_t w, uint32_t h, int channels, int dtype, void *stream);
int b200img_unpremultiply_alpha(void *buf, uint32_t w, uint32_t h, int channels, int dtype);
int b200img_unpremultiply_alpha_dev(void *d_buf, uint32_t w, uint32_t h, int channels, int dtype, void *stream);
int b200img_convert_resize(const void *src, int src_dtype, int src_channels, uint32_t w, uint32_t h,
                           void *dst, int dst_dtype, int dst_channels, uint32_t nw, uint32_t nh,
                           int filter, float acc_init, int flags);

/* ---- elementwise colour operators (src/imageops/colorops.rs) ---------------
 * All bit-exact.  `dst` may be `src`: that is the reference's *_in_place form.
 *   invert      colorops.rs:70-83 (Invert impls, src/color.rs:973-1015): max - c, alpha kept
 *   brighten    colorops.rs:161-186 (brighten_in_place 194-221): clamp(c + value, 0, max) in i32 on
 *               the colour channels; f32 subpixels go through the reference's i32 cast too
 *   contrast    colorops.rs:92-118 (contrast_in_place 126-152): f32, every channel, truncating cast
 *   huerotate   colorops.rs:232-290 (huerotate_in_place 295-349): f64 matrix on the first three
 *               channels, clamp to [0, 255] whatever the subpixel type; < 3 channels: copy
 *   grayscale / grayscale_alpha  colorops.rs:14-66: pixel.to_luma() in the source subpixel type
 *               (dst has 1 channel, or 2 with keep_alpha) */
int b200img_invert(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst);
int b200img_invert_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst, void *stream);
int b200img_brighten(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, int32_t value);
int b200img_brighten_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                         int32_t value, void *stream);
int b200img_contrast(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, float contrast);
int b200img_contrast_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                         float contrast, void *stream);
int b200img_huerotate(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, int32_t degrees);
int b200img_huerotate_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                          int32_t degrees, void *stream);
int b200img_grayscale(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, int keep_alpha);
int b200img_grayscale_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype, void *d_dst,
                          int keep_alpha, void *stream);

/* ---- pixel-format conversions --------------------------------------------
 * mode CAST    = DynamicImage::to_* / into_* (src/images/dynimage.rs:320-568
 *                -> CicpRgb::cast_pixels src/metadata/cicp.rs:706-1033;
 *                Luma<->RGB through the f32 fallback 782-887 with `coefs`,
 *                NULL = [0.2126, 0.7152, 0.0722], src/images/buffer.rs:1743).
 * mode CONVERT = ImageBuffer::convert (src/images/buffer.rs:1643-1654 ->
 *                FromColor src/color.rs:625-850, integer sRGB luma
 *                src/primitive_sealed.rs:187-231).
 * Per-channel rule: FromPrimitive, src/color.rs:540-622.  Bit-exact. */
#define B200IMG_CAST 0
#define B200IMG_CONVERT 1
int b200img_convert(const void *src, int src_dtype, int src_channels, void *dst, int dst_dtype,
                    int dst_channels, size_t npixels, int mode, const float coefs[3]);
int b200img_convert_dev(const void *d_src, int src_dtype, int src_channels, void *d_dst,
                        int dst_dtype, int dst_channels, size_t npixels, int mode,
                        const float coefs[3], void *stream);

/* BgraSwizzle, src/primitive_sealed.rs:34-67; callers ImageBuffer::
 * {from_raw_bgr, into_raw_bgr, from_raw_bgra, into_raw_bgra},
 * src/images/buffer.rs:1544-1582.  In place; swaps subpixel 0 and 2 of every
 * full 3- or 4-subpixel chunk; a trailing partial chunk is left untouched. */
int b200img_swizzle_02(void *buf, int dtype, int channels, size_t nsubpixels);
int b200img_swizzle_02_dev(void *d_buf, int dtype, int channels, size_t nsubpixels, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200IMG_H */
