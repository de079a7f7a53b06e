/*
 * extras.cpp -- oracle restatement of the "next" rows of SURVEY.md section 8(f):
 * filter3x3 (src/imageops/sample.rs:897-956) and unsharpen (sample.rs:1617-1654).
 * TEST INFRASTRUCTURE ONLY; see oracle.h.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <cstring>
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
