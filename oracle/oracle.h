/*
 * oracle.h -- CPU restatement of image-rs/image's pixel hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This library is the checker the CUDA path is
 * compared against; nothing under image_b200/ may link, load or call it.  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs use it.
 *
 * The reference is Rust and there is no Rust toolchain in this image, so the
 * reference itself cannot be compiled here (no oracle/_ref).  Every function
 * below restates one reference function literally (same operation order, same
 * f32/f64 types, no FP contraction: build with -ffp-contract=off) and cites
 * the file:line it follows, relative to the reference repository root.  The
 * oracle is pinned by the reference's own known-answer tests where they exist
 * (conversions, the u8 box-blur division, constant-image invariants, the
 * resize invariants) -- see tests/test_oracle_kat.py.  For the resampler and
 * the Gaussian blur the reference holds no value tests: parity for those is
 * "restated, invariants pinned, values unpinned" (DESIGN.md says so too).
 *
 * Conventions: images are packed row-major, interleaved channels, no row
 * padding (src/images/buffer.rs:550-557).  dtype: 0 = u8, 1 = u16, 2 = f32.
 * channels: 1 = Luma, 2 = LumaA, 3 = Rgb, 4 = Rgba.
 */
#ifndef B200IMG_ORACLE_H
#define B200IMG_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_U8 = 0, ORC_U16 = 1, ORC_F32 = 2 };
enum { ORC_NEAREST = 0, ORC_TRIANGLE = 1, ORC_CATMULLROM = 2, ORC_GAUSSIAN = 3, ORC_LANCZOS3 = 4 };
enum { ORC_CAST = 0 /* DynamicImage::to_* */, ORC_CONVERT = 1 /* ImageBuffer::convert */ };

/* multiply_accumulate (src/math/utils.rs:7-47): 0 = acc + a*b (default x86-64
 * build), 1 = fused (aarch64 / +fma builds).  Process-global; default 0. */
void orc_set_fused_mac(int fused);

/* ---- resampler (src/imageops/sample.rs) -------------------------------- */
float orc_filter_support(int filter);
float orc_filter_kernel(int filter, float x);
/* max taps of any output sample: sample.rs:242 */
int orc_resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len);
/* axis 0 = vertical_sample's rule (f32 centre), 1 = horizontal_sample's (f64
 * centre).  left[dst_len], count[dst_len], weights[dst_len * stride]. */
int orc_resize_weights(int filter, uint32_t src_len, uint32_t dst_len, int axis,
                       int32_t *left, int32_t *count, float *weights, int stride);
/* imageops::resize; acc_init = 1.0f is the literal reference. */
int orc_resize(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
               void *dst, uint32_t nw, uint32_t nh, int filter, float acc_init);
/* same, rows of the destination split over `threads` host threads (harness-
 * level parallelism for the timed CPU baseline; the reference is 1 thread). */
int orc_resize_mt(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                  void *dst, uint32_t nw, uint32_t nh, int filter, float acc_init,
                  int threads);

/* ---- Gaussian blur (sample.rs:1034-1604, filter_1d.rs) ------------------ */
typedef struct {
    uint32_t x_kernel_size; float x_sigma;
    uint32_t y_kernel_size; float y_sigma;
} orc_gauss_params;
int orc_gauss_params_from_radius(float radius, orc_gauss_params *out);
int orc_gauss_params_from_kernel_size(float ks, orc_gauss_params *out);
int orc_gauss_params_from_anisotropic(float xks, float yks, orc_gauss_params *out);
int orc_gauss_params_from_sigma(float sigma, orc_gauss_params *out);
void orc_gaussian_kernel_1d(uint32_t width, float sigma, float *out);
/* imageops::blur / blur_advanced: always the f32 engine, one rounding at the end */
int orc_gaussian_blur_generic(const void *src, uint32_t w, uint32_t h, int channels,
                              int dtype, void *dst, const orc_gauss_params *p);
/* DynamicImage::blur / blur_advanced: Q0.15 for u8 and u16, f32 engine for f32 */
int orc_gaussian_blur_dyn(const void *src, uint32_t w, uint32_t h, int channels,
                          int dtype, void *dst, const orc_gauss_params *p);
/* The same two calls through filters restated with the reference's own 4- / 16- / 32-wide chunk
 * structure (filter_1d.rs:171-316, 340-370) so that g++ vectorises them as rustc does: what the
 * TIMED CPU baseline runs.  Bit-identical to the scalar checker above. */
int orc_gaussian_blur_generic_chunked(const void *src, uint32_t w, uint32_t h, int channels,
                                      int dtype, void *dst, const orc_gauss_params *p);
int orc_gaussian_blur_dyn_chunked(const void *src, uint32_t w, uint32_t h, int channels,
                                  int dtype, void *dst, const orc_gauss_params *p);
/* Q15 transform + strip of zero ends; returns scanned length (filter_1d.rs:396-437) */
int orc_q15_kernel(const float *k, int n, uint32_t *out);

/* ---- fast_blur (src/imageops/fast_blur.rs) ------------------------------ */
void orc_boxes_for_gauss(float sigma, size_t out[3]);
int orc_fast_blur(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                  void *dst, float sigma);
uint8_t orc_u8_box_store(uint32_t acc, uint32_t kernel_size);

/* ---- conversions and swizzle (color.rs, cicp.rs, primitive_sealed.rs) --- */
int orc_convert(const void *src, int src_dtype, int src_channels, void *dst,
                int dst_dtype, int dst_channels, size_t npixels, int mode,
                const float coefs[3]);
int orc_swizzle_02(void *buf, int dtype, int channels, size_t nsubpixels);

/* ---- "next" rows of SURVEY 8(f): filter3x3 (sample.rs:897-956), unsharpen (1617-1654) ---- */
int orc_filter3x3(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                  const float kernel[9]);
int orc_unsharpen(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                  float sigma, int32_t threshold);

/* imageops::thumbnail (sample.rs:564-886); -2 = the reference's integer division by zero */
int orc_thumbnail(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                  uint32_t nw, uint32_t nh);

/* colorops.rs: op 0 invert (70-83), 1 brighten(ivalue) (161-186), 2 contrast(fvalue) (92-118),
 * 3 huerotate(ivalue degrees) (232-290).  dst may alias src (the *_in_place forms). */
int orc_colorop(int op, const void *src, void *dst, size_t npixels, int channels, int dtype,
                int32_t ivalue, float fvalue);
void orc_huerotate_matrix(int32_t value, double m[9]);

/* has_constant_alpha (src/imageops/resize.rs:155-218): 1 / 0 */
int orc_has_constant_alpha(const void *src, uint32_t w, uint32_t h, int channels, int dtype);
/* NOT reference arithmetic (pic-scale-safe is unvendored): the product's own documented
 * premultiply (un = 0) / unpremultiply (un = 1) rule, for a consistency check only */
int orc_alpha_mul(void *buf, size_t npixels, int channels, int dtype, int un);

/* ---- host helper (src/math/utils.rs:56-84) ------------------------------ */
void orc_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, int fill,
                           uint32_t *ow, uint32_t *oh);

#ifdef __cplusplus
}
#endif
#endif
