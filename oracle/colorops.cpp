/*
 * colorops.cpp -- oracle restatement of the elementwise colour operators, SURVEY.md section 8(f)
 * rank 3: invert, brighten, contrast, huerotate (src/imageops/colorops.rs).  grayscale /
 * grayscale_alpha are `pixel.to_luma()` = Luma::from_color (src/color.rs:436-446), i.e. orc_convert
 * in ORC_CONVERT mode to 1 or 2 channels of the same subpixel type.
 * TEST INFRASTRUCTURE ONLY; see oracle.h.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <algorithm>
#include <cmath>

namespace {

template <typename T> struct Max;
template <> struct Max<uint8_t> { static constexpr int i = 255; };
template <> struct Max<uint16_t> { static constexpr int i = 65535; };
template <> struct Max<float> { static constexpr int i = 1; };  // NumCast::from(1.0f32) as i32

// alpha channel of a pixel layout (map_with_alpha, Invert: src/color.rs:973-1015): LumaA -> 1, Rgba -> 3
inline int alpha_of(int C) { return C == 2 ? 1 : (C == 4 ? 3 : -1); }

// Invert for Luma / LumaA / Rgb / Rgba, src/color.rs:973-1015: max - c on every colour channel
template <typename T>
void invert_t(const T *src, T *dst, size_t npix, int C) {
    const int a = alpha_of(C);
    const T mx = (T)Max<T>::i;
    for (size_t p = 0; p < npix; ++p)
        for (int c = 0; c < C; ++c) dst[p * C + c] = c == a ? src[p * C + c] : (T)(mx - src[p * C + c]);
}

// brighten, colorops.rs:161-186: i32 arithmetic on the colour channels, alpha kept
template <typename T> int32_t to_i32(T v) { return (int32_t)v; }
template <> int32_t to_i32<float>(float v) { return rs::sat_cast<int32_t>(v); }  // NumCast truncates (NaN: the reference panics)
template <typename T>
void brighten_t(const T *src, T *dst, size_t npix, int C, int32_t value) {
    const int a = alpha_of(C);
    const int32_t mx = Max<T>::i;
    for (size_t p = 0; p < npix; ++p)
        for (int c = 0; c < C; ++c) {
            const T b = src[p * C + c];
            if (c == a) { dst[p * C + c] = b; continue; }
            const int32_t s = (int32_t)((uint32_t)to_i32<T>(b) + (uint32_t)value);  // release build: wrapping add
            dst[p * C + c] = (T)rs::clamp<int32_t>(s, 0, mx);
        }
}

// contrast, colorops.rs:92-118: every channel (alpha included) through f32
template <typename T> T from_f32(float e) { return rs::sat_cast<T>(e); }
template <> float from_f32<float>(float e) { return e; }
template <typename T>
void contrast_t(const T *src, T *dst, size_t n, float contrast) {
    const float max = (float)Max<T>::i;
    const float percent = rs::powi((100.0f + contrast) / 100.0f, 2);
    for (size_t i = 0; i < n; ++i) {
        const float c = (float)src[i];
        const float d = ((c / max - 0.5f) * percent + 0.5f) * max;
        dst[i] = from_f32<T>(rs::clamp(d, 0.0f, max));
    }
}

// huerotate, colorops.rs:232-290: f64 matrix on the first three channels, clamp to [0, 255] for
// every subpixel type (sic), NumCast back
template <typename T> T from_f64(double e) { return (T)e; }  // in [0, 255]: truncation
template <> float from_f64<float>(double e) { return (float)e; }
template <typename T>
void huerotate_t(const T *src, T *dst, size_t npix, int C, const double m[9]) {
    for (size_t p = 0; p < npix; ++p) {
        const T *s = src + p * C;
        T *d = dst + p * C;
        if (C < 3) { for (int c = 0; c < C; ++c) d[c] = s[c]; continue; }
        const double r = (double)s[0], g = (double)s[1], b = (double)s[2];
        const double nr = m[0] * r + m[1] * g + m[2] * b;
        const double ng = m[3] * r + m[4] * g + m[5] * b;
        const double nb = m[6] * r + m[7] * g + m[8] * b;
        if (C == 4) d[3] = s[3];
        d[0] = from_f64<T>(rs::clamp(nr, 0.0, 255.0));
        d[1] = from_f64<T>(rs::clamp(ng, 0.0, 255.0));
        d[2] = from_f64<T>(rs::clamp(nb, 0.0, 255.0));
    }
}

}  // namespace

/* colorops.rs:246-262 */
extern "C" void orc_huerotate_matrix(int32_t value, double m[9]) {
    const double angle = (double)value;
    const double rad = angle * (3.14159265358979323846264338327950288 / 180.0);  // f64::to_radians
    const double cosv = cos(rad), sinv = sin(rad);
    m[0] = 0.213 + cosv * 0.787 - sinv * 0.213;
    m[1] = 0.715 - cosv * 0.715 - sinv * 0.715;
    m[2] = 0.072 - cosv * 0.072 + sinv * 0.928;
    m[3] = 0.213 - cosv * 0.213 + sinv * 0.143;
    m[4] = 0.715 + cosv * 0.285 + sinv * 0.140;
    m[5] = 0.072 - cosv * 0.072 - sinv * 0.283;
    m[6] = 0.213 - cosv * 0.213 - sinv * 0.787;
    m[7] = 0.715 - cosv * 0.715 + sinv * 0.715;
    m[8] = 0.072 + cosv * 0.928 + sinv * 0.072;
}

/* op: 0 invert, 1 brighten(ivalue), 2 contrast(fvalue), 3 huerotate(ivalue degrees).  dst may alias src. */
extern "C" int orc_colorop(int op, const void *src, void *dst, size_t npixels, int channels, int dtype, int32_t ivalue,
                           float fvalue) {
    if (channels < 1 || channels > 4) return -1;
    double m[9];
    if (op == 3) orc_huerotate_matrix(ivalue, m);
#define DISPATCH(T)                                                                                    \
    switch (op) {                                                                                      \
        case 0: invert_t((const T *)src, (T *)dst, npixels, channels); return 0;                       \
        case 1: brighten_t((const T *)src, (T *)dst, npixels, channels, ivalue); return 0;             \
        case 2: contrast_t((const T *)src, (T *)dst, npixels * channels, fvalue); return 0;            \
        case 3: huerotate_t((const T *)src, (T *)dst, npixels, channels, m); return 0;                 \
    }                                                                                                  \
    return -1;
    switch (dtype) {
        case ORC_U8: DISPATCH(uint8_t)
        case ORC_U16: DISPATCH(uint16_t)
        case ORC_F32: DISPATCH(float)
    }
#undef DISPATCH
    return -1;
}

/* ---- either side of DynamicImage::resize (src/imageops/resize.rs) -------------------------- */
namespace {
// has_constant_alpha_integer, resize.rs:173-202: XOR with the first alpha, summed per row
template <typename T>
int constant_alpha_int(const T *src, uint32_t w, uint32_t h, int C) {
    if ((size_t)w * h == 0) return 1;
    const T first = src[C - 1];
    uint64_t sum_of_diffs = 0;
    for (uint32_t y = 0; y < h; ++y) {
        const T *row = src + (size_t)y * w * C;
        for (uint32_t x = 0; x < w; ++x) sum_of_diffs += (uint64_t)(T)(row[x * C + C - 1] ^ first);
        if (sum_of_diffs != 0) return 0;
    }
    return 1;
}
// has_constant_alpha_f32, resize.rs:205-218
int constant_alpha_f32(const float *src, uint32_t w, uint32_t h, int C) {
    const size_t n = (size_t)w * h;
    if (n == 0) return 1;
    const float first = src[C - 1];
    for (size_t i = 0; i < n; ++i)
        if (!(src[i * C + C - 1] == first)) return 0;
    return 1;
}
// The premultiply arithmetic of the reference lives in pic-scale-safe (unvendored): this is the
// product's documented own rule, restated -- a consistency check, not reference parity.
template <typename T>
void alpha_mul_int(T *buf, size_t npix, int C, uint32_t mx, int un) {
    for (size_t i = 0; i < npix; ++i) {
        const uint32_t a = buf[i * C + C - 1];
        for (int c = 0; c < C - 1; ++c) {
            const uint32_t v = buf[i * C + c];
            if (!un) buf[i * C + c] = (T)((v * a + mx / 2) / mx);
            else if (a != 0) buf[i * C + c] = (T)std::min<uint32_t>((v * mx + a / 2) / a, mx);
        }
    }
}
}  // namespace

extern "C" int orc_has_constant_alpha(const void *src, uint32_t w, uint32_t h, int channels, int dtype) {
    if (channels != 2 && channels != 4) return 1;  /* resize.rs:155-171: Luma / Rgb variants */
    switch (dtype) {
        case ORC_U8: return constant_alpha_int((const uint8_t *)src, w, h, channels);
        case ORC_U16: return constant_alpha_int((const uint16_t *)src, w, h, channels);
        case ORC_F32: return constant_alpha_f32((const float *)src, w, h, channels);
    }
    return -1;
}

extern "C" int orc_alpha_mul(void *buf, size_t npixels, int channels, int dtype, int un) {
    if (channels != 2 && channels != 4) return 0;
    if (dtype == ORC_U8) alpha_mul_int((uint8_t *)buf, npixels, channels, 255u, un);
    else if (dtype == ORC_U16) alpha_mul_int((uint16_t *)buf, npixels, channels, 65535u, un);
    else {
        float *f = (float *)buf;
        for (size_t i = 0; i < npixels; ++i) {
            const float a = f[i * channels + channels - 1];
            for (int c = 0; c < channels - 1; ++c) {
                if (!un) f[i * channels + c] = f[i * channels + c] * a;
                else if (a != 0.0f) f[i * channels + c] = f[i * channels + c] / a;
            }
        }
    }
    return 0;
}
