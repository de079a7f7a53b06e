/*
 * convert.cpp -- oracle restatement of the pixel-format conversion loops and
 * the BGR(A) swizzle (TEST INFRASTRUCTURE ONLY; see oracle.h).
 *
 * mode ORC_CAST    = DynamicImage::to_* / into_*: src/images/dynimage.rs:320-568
 *                    -> ImageBuffer::cast_in_color_space src/images/buffer.rs:1731-1748
 *                    -> CicpRgb::cast_pixels src/metadata/cicp.rs:706-1033
 *                       (same-family layouts: per-channel FromPrimitive;
 *                        Luma<->RGB: f32 fallback 782-887 with 620-699, 1064-1108)
 * mode ORC_CONVERT = ImageBuffer::convert src/images/buffer.rs:1643-1654
 *                    -> FromColor impls src/color.rs:625-850, RgbToLuma
 *                       src/primitive_sealed.rs:187-231, Enlargeable::clamp_from
 *                       src/traits.rs:79-92
 * Per-channel map FromPrimitive: src/color.rs:540-622.
 * Swizzle: src/primitive_sealed.rs:34-67.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <cstring>

namespace {

// ---- FromPrimitive, color.rs:540-622 ----
template <typename D, typename S> struct FromPrim {
    static D f(S v) { return (D)v; }  // identity (same type only)
};
template <> struct FromPrim<uint8_t, float> {
    static uint8_t f(float v) { float x = fminf(v, 1.0f); return rs::sat_cast<uint8_t>(x * 255.0f + 0.5f); }
};
template <> struct FromPrim<uint16_t, float> {
    static uint16_t f(float v) { float x = fminf(v, 1.0f); return rs::sat_cast<uint16_t>(x * 65535.0f + 0.5f); }
};
template <> struct FromPrim<uint8_t, uint16_t> {
    static uint8_t f(uint16_t c) { uint32_t x = c; return (uint8_t)((x + 128u) / 257u); }
};
template <> struct FromPrim<float, uint16_t> {
    static float f(uint16_t v) { return (float)v / 65535.0f; }
};
template <> struct FromPrim<float, uint8_t> {
    static float f(uint8_t v) { return (float)v / 255.0f; }
};
template <> struct FromPrim<uint16_t, uint8_t> {
    static uint16_t f(uint8_t c) { uint64_t x = c; return (uint16_t)((x << 8) | x); }
};

// DEFAULT_MAX_VALUE, traits.rs:54-72
template <typename T> T default_max();
template <> uint8_t default_max<uint8_t>() { return 255; }
template <> uint16_t default_max<uint16_t>() { return 65535; }
template <> float default_max<float>() { return 1.0f; }

// ---- ColorComponentForCicp, cicp.rs:1064-1108 ----
template <typename T> struct Cicp;
template <> struct Cicp<uint8_t> {
    static float expand(uint8_t v) { const float R = 1.0f / 255.0f; return (float)v * R; }
    static uint8_t clamp(float v) { return rs::sat_cast<uint8_t>(v * 255.0f + 0.5f); }
};
template <> struct Cicp<uint16_t> {
    static float expand(uint16_t v) { const float R = 1.0f / 65535.0f; return (float)v * R; }
    static uint16_t clamp(float v) { return rs::sat_cast<uint16_t>(v * 65535.0f + 0.5f); }
};
template <> struct Cicp<float> {
    static float expand(float v) { return v; }
    static float clamp(float v) { return v; }
};

// ---- RgbToLuma, primitive_sealed.rs:187-231 ----
uint8_t rgb_to_luma(uint8_t r, uint8_t g, uint8_t b) {
    const uint32_t W_R = (uint32_t)((((uint64_t)1 << 24) * 2126 + 9999) / 10000);
    const uint32_t W_G = (uint32_t)((((uint64_t)1 << 24) * 7152 + 9999) / 10000);
    const uint32_t W_B = (uint32_t)((((uint64_t)1 << 24) * 722 + 9999) / 10000);
    return (uint8_t)(((uint32_t)r * W_R + (uint32_t)g * W_G + (uint32_t)b * W_B + 0x800000u) >> 24);
}
uint16_t rgb_to_luma(uint16_t r, uint16_t g, uint16_t b) {
    // default impl with Larger = u32; clamp_from (traits.rs:79-92)
    uint32_t l = 2126u * (uint32_t)r + 7152u * (uint32_t)g + 722u * (uint32_t)b;
    uint32_t q = l / 10000u;
    return q > 65535u ? (uint16_t)65535u : (uint16_t)q;
}
float rgb_to_luma(float r, float g, float b) {
    const float SR = 2126.0f / 10000.0f, SG = 7152.0f / 10000.0f, SB = 722.0f / 10000.0f;
    return SR * r + SG * g + SB * b;
}

inline bool is_luma(int ch) { return ch <= 2; }
inline bool has_alpha(int ch) { return ch == 2 || ch == 4; }

// One pixel through CicpRgb::cast_pixels.
template <typename S, typename D>
void cast_pixel(const S *s, int sc, D *d, int dc, const float coefs[3]) {
    if (is_luma(sc) == is_luma(dc)) {
        // cast_pixels_from_subpixels, cicp.rs:926-1033
        int colour = is_luma(sc) ? 1 : 3;
        for (int c = 0; c < colour; ++c) d[c] = FromPrim<D, S>::f(s[c]);
        if (has_alpha(dc)) d[colour] = has_alpha(sc) ? FromPrim<D, S>::f(s[colour]) : default_max<D>();
        return;
    }
    // cast_pixels_by_fallback, cicp.rs:782-887
    float rgba[4];
    if (is_luma(sc)) {  // expand_luma_rgb / expand_luma_rgba, 620-639
        float l = Cicp<S>::expand(s[0]);
        rgba[0] = rgba[1] = rgba[2] = l;
        rgba[3] = has_alpha(sc) ? Cicp<S>::expand(s[1]) : 1.0f;
    } else {  // expand_rgb / expand_rgba, 641-651
        for (int c = 0; c < 3; ++c) rgba[c] = Cicp<S>::expand(s[c]);
        rgba[3] = has_alpha(sc) ? Cicp<S>::expand(s[3]) : 1.0f;
    }
    if (is_luma(dc)) {  // clamp_rgb_luma / clamp_rgba_luma, 667-699
        float luma = 0.0f;
        for (int c = 0; c < 3; ++c) luma = orc_mac(luma, rgba[c], coefs[c]);
        d[0] = Cicp<D>::clamp(luma);
        if (dc == 2) d[1] = Cicp<D>::clamp(rgba[3]);
    } else {  // clamp_rgb / clamp_rgba, 653-665
        for (int c = 0; c < dc; ++c) d[c] = Cicp<D>::clamp(rgba[c]);
    }
}

// One pixel through FromColor, color.rs:652-850.
template <typename S, typename D>
void convert_pixel(const S *s, int sc, D *d, int dc) {
    if (is_luma(sc) && is_luma(dc)) {
        d[0] = FromPrim<D, S>::f(s[0]);
        if (dc == 2) d[1] = sc == 2 ? FromPrim<D, S>::f(s[1]) : default_max<D>();
    } else if (!is_luma(sc) && is_luma(dc)) {
        d[0] = FromPrim<D, S>::f(rgb_to_luma(s[0], s[1], s[2]));
        if (dc == 2) d[1] = sc == 4 ? FromPrim<D, S>::f(s[3]) : default_max<D>();
    } else if (is_luma(sc) && !is_luma(dc)) {
        D g = FromPrim<D, S>::f(s[0]);
        d[0] = d[1] = d[2] = g;
        if (dc == 4) d[3] = sc == 2 ? FromPrim<D, S>::f(s[1]) : default_max<D>();
    } else {
        for (int c = 0; c < 3; ++c) d[c] = FromPrim<D, S>::f(s[c]);
        if (dc == 4) d[3] = sc == 4 ? FromPrim<D, S>::f(s[3]) : default_max<D>();
    }
}

template <typename S, typename D>
int convert_t(const S *src, int sc, D *dst, int dc, size_t n, int mode, const float coefs[3]) {
    for (size_t i = 0; i < n; ++i) {
        if (mode == ORC_CAST) cast_pixel(src + i * sc, sc, dst + i * dc, dc, coefs);
        else convert_pixel(src + i * sc, sc, dst + i * dc, dc);
    }
    return 0;
}

template <typename S>
int convert_s(const S *src, int sc, void *dst, int dd, int dc, size_t n, int mode, const float coefs[3]) {
    switch (dd) {
        case ORC_U8: return convert_t(src, sc, (uint8_t *)dst, dc, n, mode, coefs);
        case ORC_U16: return convert_t(src, sc, (uint16_t *)dst, dc, n, mode, coefs);
        case ORC_F32: return convert_t(src, sc, (float *)dst, dc, n, mode, coefs);
    }
    return -1;
}

// BgraSwizzle, primitive_sealed.rs:34-56: swap element 0 and 2 of every full chunk
template <typename T>
void swizzle_t(T *p, int ch, size_t n) {
    for (size_t i = 0; i + ch <= n; i += ch) {
        T t = p[i];
        p[i] = p[i + 2];
        p[i + 2] = t;
    }
}

}  // namespace

extern "C" int orc_convert(const void *src, int src_dtype, int src_channels, void *dst,
                           int dst_dtype, int dst_channels, size_t npixels, int mode,
                           const float coefs[3]) {
    static const float srgb[3] = {0.2126f, 0.7152f, 0.0722f};  // buffer.rs:1743 fallback
    if (!coefs) coefs = srgb;
    if (src_channels < 1 || src_channels > 4 || dst_channels < 1 || dst_channels > 4) return -1;
    switch (src_dtype) {
        case ORC_U8: return convert_s((const uint8_t *)src, src_channels, dst, dst_dtype, dst_channels, npixels, mode, coefs);
        case ORC_U16: return convert_s((const uint16_t *)src, src_channels, dst, dst_dtype, dst_channels, npixels, mode, coefs);
        case ORC_F32: return convert_s((const float *)src, src_channels, dst, dst_dtype, dst_channels, npixels, mode, coefs);
    }
    return -1;
}

extern "C" int orc_swizzle_02(void *buf, int dtype, int channels, size_t nsubpixels) {
    if (channels != 3 && channels != 4) return -1;
    switch (dtype) {
        case ORC_U8: swizzle_t((uint8_t *)buf, channels, nsubpixels); return 0;
        case ORC_U16: swizzle_t((uint16_t *)buf, channels, nsubpixels); return 0;
        case ORC_F32: swizzle_t((float *)buf, channels, nsubpixels); return 0;
    }
    return -1;
}
