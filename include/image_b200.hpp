// image_b200.hpp -- header-only C++ mirror of the reference's interface for the pixel hot path,
// over the C ABI in b200img.h.  Same names, argument meaning and error behaviour as the Rust
// originals (cited per function); the reference's panics become image_b200::Panic, its
// ImageError::Limits becomes image_b200::LimitError.  Link with -lb200img.  No CPU fallback.
#pragma once

#include <cstdint>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "b200img.h"

namespace image_b200 {

struct Panic : std::runtime_error { using std::runtime_error::runtime_error; };       // assert!/unwrap/unimplemented! in the reference
struct LimitError : std::runtime_error { using std::runtime_error::runtime_error; };  // ImageError::Limits(DimensionError)
struct DeviceError : std::runtime_error { using std::runtime_error::runtime_error; }; // no GPU / CUDA failure / out of memory

inline void check(int rc) {
    if (rc == B200IMG_OK) return;
    const std::string msg = b200img_last_error();
    if (rc == B200IMG_ERR_BAD_ARG) throw Panic(msg);
    if (rc == B200IMG_ERR_DIMENSION) throw LimitError(msg);
    throw DeviceError(msg);
}

template <typename S> struct SubpixelCode;
template <> struct SubpixelCode<uint8_t> { static constexpr int value = B200IMG_U8; };
template <> struct SubpixelCode<uint16_t> { static constexpr int value = B200IMG_U16; };
template <> struct SubpixelCode<float> { static constexpr int value = B200IMG_F32; };

// src/imageops/sample.rs:98-115
enum class FilterType { Nearest = 0, Triangle = 1, CatmullRom = 2, Gaussian = 3, Lanczos3 = 4 };

// ImageBuffer<P, Vec<S>> with P = `Channels` subpixels of type S: packed, row-major (src/images/buffer.rs:550-557)
template <typename S, int Channels>
struct ImageBuffer {
    static_assert(Channels >= 1 && Channels <= 4, "1..4 channels");
    uint32_t width = 0, height = 0;
    std::vector<S> data;
    ImageBuffer() = default;
    ImageBuffer(uint32_t w, uint32_t h) : width(w), height(h), data((size_t)w * h * Channels) {}  // zero-filled, as ImageBuffer::new
    S *pixel(uint32_t x, uint32_t y) { return &data[((size_t)y * width + x) * Channels]; }
    const S *pixel(uint32_t x, uint32_t y) const { return &data[((size_t)y * width + x) * Channels]; }

    // ImageBuffer::convert::<To>()  src/images/buffer.rs:1643-1654
    template <typename S2, int C2>
    ImageBuffer<S2, C2> convert() const {
        ImageBuffer<S2, C2> out(width, height);
        check(b200img_convert(data.data(), SubpixelCode<S>::value, Channels, out.data.data(), SubpixelCode<S2>::value, C2,
                              (size_t)width * height, B200IMG_CONVERT, nullptr));
        return out;
    }
    // DynamicImage::to_*  src/images/dynimage.rs:320-424 (coefs = derived_luminance(); nullptr = sRGB fallback)
    template <typename S2, int C2>
    ImageBuffer<S2, C2> cast(const float *coefs = nullptr) const {
        ImageBuffer<S2, C2> out(width, height);
        check(b200img_convert(data.data(), SubpixelCode<S>::value, Channels, out.data.data(), SubpixelCode<S2>::value, C2,
                              (size_t)width * height, B200IMG_CAST, coefs));
        return out;
    }
    // ImageBuffer::from_raw_bgr(a) / into_raw_bgr(a) core  src/images/buffer.rs:1544-1582
    void swizzle_02() {
        static_assert(Channels == 3 || Channels == 4, "bgr(a) swizzle needs 3 or 4 channels");
        check(b200img_swizzle_02(data.data(), SubpixelCode<S>::value, Channels, data.size()));
    }
};

// GaussianBlurParameters  src/imageops/sample.rs:1081-1295
struct GaussianBlurParameters {
    b200img_gauss_params raw{};
    static GaussianBlurParameters new_from_radius(float r) { GaussianBlurParameters p; check(b200img_gauss_params_from_radius(r, &p.raw)); return p; }
    static GaussianBlurParameters new_from_kernel_size(float k) { GaussianBlurParameters p; check(b200img_gauss_params_from_kernel_size(k, &p.raw)); return p; }
    static GaussianBlurParameters new_anisotropic_kernel_size(float x, float y) { GaussianBlurParameters p; check(b200img_gauss_params_from_anisotropic(x, y, &p.raw)); return p; }
    static GaussianBlurParameters new_from_sigma(float s) { GaussianBlurParameters p; check(b200img_gauss_params_from_sigma(s, &p.raw)); return p; }
};

namespace imageops {

// imageops::resize  src/imageops/sample.rs:970.  acc_init = 1.0f is the literal reference (sample.rs:547).
template <typename S, int C>
ImageBuffer<S, C> resize(const ImageBuffer<S, C> &image, uint32_t nwidth, uint32_t nheight, FilterType filter,
                         float acc_init = 1.0f, int flags = B200IMG_FP_FAST) {
    ImageBuffer<S, C> out(nwidth, nheight);
    check(b200img_resize(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(), nwidth,
                         nheight, (int)filter, acc_init, flags));
    return out;
}

// imageops::blur_advanced  src/imageops/sample.rs:1050
template <typename S, int C>
ImageBuffer<S, C> blur_advanced(const ImageBuffer<S, C> &image, const GaussianBlurParameters &p, int flags = B200IMG_FP_FAST) {
    ImageBuffer<S, C> out(image.width, image.height);
    check(b200img_gaussian_blur(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(),
                                &p.raw, B200IMG_BLUR_GENERIC, flags));
    return out;
}
// imageops::blur(image, radius)  src/imageops/sample.rs:1034 -- radius, not sigma, at this revision
template <typename S, int C>
ImageBuffer<S, C> blur(const ImageBuffer<S, C> &image, float radius) { return blur_advanced(image, GaussianBlurParameters::new_from_radius(radius)); }

// DynamicImage::blur_advanced arithmetic (Q0.15 for u8/u16)  src/imageops/sample.rs:1297
template <typename S, int C>
ImageBuffer<S, C> blur_dyn(const ImageBuffer<S, C> &image, const GaussianBlurParameters &p) {
    ImageBuffer<S, C> out(image.width, image.height);
    check(b200img_gaussian_blur(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(),
                                &p.raw, B200IMG_BLUR_DYN, B200IMG_FP_FAST));
    return out;
}

// imageops::fast_blur  src/imageops/fast_blur.rs:111
template <typename S, int C>
ImageBuffer<S, C> fast_blur(const ImageBuffer<S, C> &image, float sigma) {
    if (image.width == 0 || image.height == 0) return image;  // clone
    ImageBuffer<S, C> out(image.width, image.height);
    check(b200img_fast_blur(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(), sigma));
    return out;
}

// imageops::thumbnail  src/imageops/sample.rs:595
template <typename S, int C>
ImageBuffer<S, C> thumbnail(const ImageBuffer<S, C> &image, uint32_t new_width, uint32_t new_height) {
    ImageBuffer<S, C> out(new_width, new_height);
    check(b200img_thumbnail(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(),
                            new_width, new_height, 0));
    return out;
}

// imageops::filter3x3  src/imageops/sample.rs:897
template <typename S, int C>
ImageBuffer<S, C> filter3x3(const ImageBuffer<S, C> &image, const float (&kernel)[9]) {
    ImageBuffer<S, C> out(image.width, image.height);
    check(b200img_filter3x3(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(), kernel));
    return out;
}

// imageops::unsharpen  src/imageops/sample.rs:1617
template <typename S, int C>
ImageBuffer<S, C> unsharpen(const ImageBuffer<S, C> &image, float sigma, int32_t threshold, int flags = B200IMG_FP_FAST) {
    ImageBuffer<S, C> out(image.width, image.height);
    check(b200img_unsharpen(image.data.data(), image.width, image.height, C, SubpixelCode<S>::value, out.data.data(), sigma,
                            threshold, flags));
    return out;
}

}  // namespace imageops
}  // namespace image_b200
