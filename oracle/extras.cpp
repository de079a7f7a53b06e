/*
 * extras.cpp -- oracle restatement of the "next" rows of SURVEY.md section 8(f):
 * filter3x3 (src/imageops/sample.rs:897-956) and unsharpen (sample.rs:1617-1654).
 * TEST INFRASTRUCTURE ONLY; see oracle.h.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

namespace {

template <typename T> float default_max_f();
template <> float default_max_f<uint8_t>() { return 255.0f; }
template <> float default_max_f<uint16_t>() { return 65535.0f; }
template <> float default_max_f<float>() { return 1.0f; }

// NumCast::from(f32) -> S for a value already clamped into range: truncation toward zero
template <typename T> T numcast_from_f32(float v) { return rs::sat_cast<T>(v); }
template <> float numcast_from_f32<float>(float v) { return v; }

// sample.rs:897-956: border pixels keep the zero fill of buffer_like()
template <typename T>
void filter3x3_t(const T *src, uint32_t w, uint32_t h, int C, T *dst, const float k[9]) {
    memset(dst, 0, (size_t)w * h * C * sizeof(T));
    if (w < 3 || h < 3) return;
    const float maxv = default_max_f<T>();
    float sum = 0.0f;
    for (int i = 0; i < 9; ++i) sum += k[i];
    const float inverse_sum = sum == 0.0f ? 1.0f : 1.0f / sum;
    static const int taps[9][2] = {{-1, -1}, {0, -1}, {1, -1}, {-1, 0}, {0, 0}, {1, 0}, {-1, 1}, {0, 1}, {1, 1}};
    for (uint32_t y = 1; y < h - 1; ++y)
        for (uint32_t x = 1; x < w - 1; ++x) {
            float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            for (int i = 0; i < 9; ++i) {
                const T *p = src + ((size_t)(y + taps[i][1]) * w + (x + taps[i][0])) * C;
                for (int c = 0; c < C; ++c) t[c] += (float)p[c] * k[i];
            }
            T *d = dst + ((size_t)y * w + x) * C;
            for (int c = 0; c < C; ++c) d[c] = numcast_from_f32<T>(rs::clamp(t[c] * inverse_sum, 0.0f, maxv));
        }
}

// NumCast::from(S) -> i32: integers exact; f32 truncates toward zero
template <typename T> int32_t to_i32(T v) { return (int32_t)v; }
template <> int32_t to_i32<float>(float v) { return rs::sat_cast<int32_t>(v); }

// sample.rs:1617-1654 (the blur is blur_advanced(new_from_sigma(sigma)): f32 engine)
template <typename T>
int unsharpen_t(const T *src, uint32_t w, uint32_t h, int C, T *dst, float sigma, int32_t threshold) {
    orc_gauss_params p;
    if (orc_gauss_params_from_sigma(sigma, &p)) return -1;
    int dtype = sizeof(T) == 1 ? ORC_U8 : (sizeof(T) == 2 ? ORC_U16 : ORC_F32);
    int rc = orc_gaussian_blur_generic(src, w, h, C, dtype, dst, &p);
    if (rc) return rc;
    const int32_t maxv = (int32_t)default_max_f<T>();
    const size_t n = (size_t)w * h * C;
    for (size_t i = 0; i < n; ++i) {
        const int32_t ic = to_i32<T>(src[i]), id = to_i32<T>(dst[i]);
        const int32_t diff = ic - id;
        if ((diff < 0 ? -diff : diff) > threshold) dst[i] = (T)rs::clamp<int32_t>(ic + diff, 0, maxv);
        else dst[i] = src[i];
    }
    return 0;
}

// ---- thumbnail (sample.rs:564-886) ---------------------------------------------------
// Enlargeable::Larger (src/traits.rs:94-131): u8, u16 -> u32; f32 -> f64
template <typename T> struct Larger { typedef uint32_t type; };
template <> struct Larger<float> { typedef double type; };

// Enlargeable::clamp_from, src/traits.rs:79-87
inline uint8_t clamp_from(uint32_t n, uint8_t) { return n > 255u ? 255 : (uint8_t)n; }
inline uint16_t clamp_from(uint32_t n, uint16_t) { return n > 65535u ? 65535 : (uint16_t)n; }
inline float clamp_from(double n, float) {
    const double mx = (double)std::numeric_limits<float>::max();
    if (n > mx) return std::numeric_limits<float>::max();
    if (n < -mx) return -std::numeric_limits<float>::max();
    return (float)n;
}

template <typename T> struct Thumb {
    typedef typename Larger<T>::type L;
    const T *src; uint32_t w, h; int C;
    const T *px(uint32_t x, uint32_t y) const { return src + ((size_t)y * w + x) * C; }

    // thumbnail_sample_block, sample.rs:680-742.  Returns false where the reference divides an
    // integer by zero (a block of more than 2^16 pixels that is under 256 wide or high).
    bool block(uint32_t left, uint32_t right, uint32_t bottom, uint32_t top, T *out) const {
        L sum[4] = {0, 0, 0, 0};
        const uint64_t pixel_count = (uint64_t)(right - left) * (uint64_t)(top - bottom);
        uint64_t sample_count;
        if (pixel_count > 256 * 256) {
            const uint32_t STEP = 256;
            for (uint64_t y = bottom; y < top; y += STEP)
                for (uint64_t x = left; x < right; x += STEP) {
                    T k[4];
                    const uint32_t xe = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(x + STEP, 0xFFFFFFFFull), right);
                    const uint32_t ye = (uint32_t)std::min<uint64_t>(std::min<uint64_t>(y + STEP, 0xFFFFFFFFull), top);
                    if (!block((uint32_t)x, xe, (uint32_t)y, ye, k)) return false;
                    for (int c = 0; c < C; ++c) sum[c] += (L)k[c];
                }
            // the literal count of the reference: truncating quotients, not the number of sub-blocks
            sample_count = ((uint64_t)(right - left) / STEP) * ((uint64_t)(top - bottom) / STEP);
        } else {
            for (uint32_t y = bottom; y < top; ++y)
                for (uint32_t x = left; x < right; ++x) {
                    const T *k = px(x, y);
                    for (int c = 0; c < C; ++c) sum[c] += (L)k[c];
                }
            sample_count = pixel_count;
        }
        const L n = (L)sample_count;
        if (std::numeric_limits<L>::is_integer) {
            if (n == 0) return false;
            const L round = n / 2;
            for (int c = 0; c < C; ++c) out[c] = clamp_from((L)((sum[c] + round) / n), T());
        } else {
            for (int c = 0; c < C; ++c) out[c] = clamp_from((L)(sum[c] / n), T());
        }
        return true;
    }

    // thumbnail_sample_fraction_horizontal, sample.rs:745-789
    void frac_h(uint32_t left, float fract, uint32_t bottom, uint32_t top, T *out) const {
        L sl[4] = {0, 0, 0, 0}, sr[4] = {0, 0, 0, 0};
        for (uint32_t y = bottom; y < top; ++y) {
            const T *a = px(left, y), *b = px(left + 1, y);
            for (int c = 0; c < C; ++c) { sl[c] += (L)a[c]; sr[c] += (L)b[c]; }
        }
        const float fact_right = fract / (float)(top - bottom);
        const float fact_left = (1.0f - fract) / (float)(top - bottom);
        for (int c = 0; c < C; ++c) out[c] = rs::nearest_from<T>(fact_left * (float)sl[c] + fact_right * (float)sr[c]);
    }

    // thumbnail_sample_fraction_vertical, sample.rs:792-836
    void frac_v(uint32_t left, uint32_t right, uint32_t bottom, float fract, T *out) const {
        L sb[4] = {0, 0, 0, 0}, st[4] = {0, 0, 0, 0};
        for (uint32_t x = left; x < right; ++x) {
            const T *a = px(x, bottom), *b = px(x, bottom + 1);
            for (int c = 0; c < C; ++c) { sb[c] += (L)a[c]; st[c] += (L)b[c]; }
        }
        const float fact_top = fract / (float)(right - left);
        const float fact_bot = (1.0f - fract) / (float)(right - left);
        for (int c = 0; c < C; ++c) out[c] = rs::nearest_from<T>(fact_bot * (float)sb[c] + fact_top * (float)st[c]);
    }

    // thumbnail_sample_fraction_both, sample.rs:839-886
    void frac_both(uint32_t left, float frac_v_, uint32_t bottom, float frac_h_, T *out) const {
        const T *bl = px(left, bottom), *tl = px(left, bottom + 1), *br = px(left + 1, bottom), *tr = px(left + 1, bottom + 1);
        const float fact_tr = frac_v_ * frac_h_;
        const float fact_tl = frac_v_ * (1.0f - frac_h_);
        const float fact_br = (1.0f - frac_v_) * frac_h_;
        const float fact_bl = (1.0f - frac_v_) * (1.0f - frac_h_);
        for (int c = 0; c < C; ++c)
            out[c] = rs::nearest_from<T>(fact_br * (float)br[c] + fact_tr * (float)tr[c] + fact_bl * (float)bl[c] + fact_tl * (float)tl[c]);
    }
};

inline float fract(float v) { return v - truncf(v); }  // f32::fract

// thumbnail, sample.rs:595-677
template <typename T>
int thumbnail_t(const T *src, uint32_t width, uint32_t height, int C, T *dst, uint32_t nw, uint32_t nh) {
    if (width == 0 || height == 0) {  // zero-filled output
        memset(dst, 0, (size_t)nw * nh * C * sizeof(T));
        return 0;
    }
    Thumb<T> im = {src, width, height, C};
    const float x_ratio = (float)width / (float)nw;
    const float y_ratio = (float)height / (float)nh;
    for (uint32_t outy = 0; outy < nh; ++outy) {
        const float bottomf = (float)outy * y_ratio;
        const float topf = bottomf + y_ratio;
        const uint32_t bottom = rs::clamp<uint32_t>(rs::sat_cast<uint32_t>(ceilf(bottomf)), 0, height - 1);
        const uint32_t top = rs::clamp<uint32_t>(rs::sat_cast<uint32_t>(ceilf(topf)), bottom, height);
        for (uint32_t outx = 0; outx < nw; ++outx) {
            const float leftf = (float)outx * x_ratio;
            const float rightf = leftf + x_ratio;
            const uint32_t left = rs::clamp<uint32_t>(rs::sat_cast<uint32_t>(ceilf(leftf)), 0, width - 1);
            const uint32_t right = rs::clamp<uint32_t>(rs::sat_cast<uint32_t>(ceilf(rightf)), left, width);
            T *out = dst + ((size_t)outy * nw + outx) * C;
            if (bottom != top && left != right) {
                if (!im.block(left, right, bottom, top, out)) return -2;
            } else if (bottom != top) {
                im.frac_h(right - 1, (fract(leftf) + fract(rightf)) / 2.0f, bottom, top, out);
            } else if (left != right) {
                im.frac_v(left, right, top - 1, (fract(topf) + fract(bottomf)) / 2.0f, out);
            } else {
                // the reference's names are crossed here; the values are what matters
                const float from_rows = (fract(topf) + fract(bottomf)) / 2.0f;
                const float from_cols = (fract(leftf) + fract(rightf)) / 2.0f;
                im.frac_both(right - 1, from_rows, top - 1, from_cols, out);
            }
        }
    }
    return 0;
}

}  // namespace

extern "C" int orc_filter3x3(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, const float kernel[9]) {
    if (channels < 1 || channels > 4) return -1;
    switch (dtype) {
        case ORC_U8: filter3x3_t((const uint8_t *)src, w, h, channels, (uint8_t *)dst, kernel); return 0;
        case ORC_U16: filter3x3_t((const uint16_t *)src, w, h, channels, (uint16_t *)dst, kernel); return 0;
        case ORC_F32: filter3x3_t((const float *)src, w, h, channels, (float *)dst, kernel); return 0;
    }
    return -1;
}

extern "C" int orc_unsharpen(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, float sigma,
                             int32_t threshold) {
    if (channels < 1 || channels > 4) return -1;
    if (w == 0 || h == 0) return 0;
    switch (dtype) {
        case ORC_U8: return unsharpen_t((const uint8_t *)src, w, h, channels, (uint8_t *)dst, sigma, threshold);
        case ORC_U16: return unsharpen_t((const uint16_t *)src, w, h, channels, (uint16_t *)dst, sigma, threshold);
        case ORC_F32: return unsharpen_t((const float *)src, w, h, channels, (float *)dst, sigma, threshold);
    }
    return -1;
}

/* returns -2 where the reference panics with an integer division by zero (see Thumb::block) */
extern "C" int orc_thumbnail(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst, uint32_t nw,
                             uint32_t nh) {
    if (channels < 1 || channels > 4) return -1;
    switch (dtype) {
        case ORC_U8: return thumbnail_t((const uint8_t *)src, w, h, channels, (uint8_t *)dst, nw, nh);
        case ORC_U16: return thumbnail_t((const uint16_t *)src, w, h, channels, (uint16_t *)dst, nw, nh);
        case ORC_F32: return thumbnail_t((const float *)src, w, h, channels, (float *)dst, nw, nh);
    }
    return -1;
}
