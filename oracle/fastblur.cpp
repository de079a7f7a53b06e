/*
 * fastblur.cpp -- oracle restatement of imageops::fast_blur
 * (TEST INFRASTRUCTURE ONLY; see oracle.h).
 *
 * Follows src/imageops/fast_blur.rs: BlurAccumulator impls 36-94, fast_blur
 * 111-191, boxes_for_gauss 201-217, ceil_to_odd 220-226, the accumulator
 * choice 229-300, box_blur_horizontal_pass 303-413, box_blur_vertical_pass
 * 416-473; WithBlurAcc src/primitive_sealed.rs:168-184.
 *
 * The horizontal pass's leading / interior / trailing phases (345-411) all
 * apply the same recurrence -- store, then `sum += p[min(x+r+1, W-1)]`,
 * `sum -= p[max(x-r, 0)]` -- so they are restated as one loop; the order of
 * the add and the subtract is kept because it matters for f32 sums.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <cstring>
#include <vector>

namespace {

// ---- accumulators, fast_blur.rs:36-94 ----
struct AccU32 {
    typedef uint32_t acc_t;
    struct Weight { uint32_t reciprocal, rounding_bias; };
    static acc_t from(uint8_t v) { return v; }
    static acc_t scale(acc_t a, size_t count) { return a * (uint32_t)count; }
    static Weight weight(size_t ks) { return {UINT32_MAX / (uint32_t)ks + 1u, (uint32_t)ks / 2u}; }
    static uint8_t store(acc_t a, Weight w) {
        uint32_t n = a + w.rounding_bias;
        return (uint8_t)(((uint64_t)n * (uint64_t)w.reciprocal) >> 32);
    }
};

template <typename T>
struct AccF32 {
    typedef float acc_t;
    typedef float Weight;
    static acc_t from(T v) { return (float)v; }
    static acc_t scale(acc_t a, size_t count) { return a * (float)count; }
    static Weight weight(size_t ks) { return 1.0f / (float)ks; }
    static T store(acc_t a, Weight w) { return rs::clamp_nearest_from<T>(a * w); }
};

// box_blur_horizontal_pass, fast_blur.rs:303-413
template <typename P, typename A>
void horizontal_pass(const P *src, P *dst, size_t width, size_t height, int CN, size_t radius) {
    const size_t kernel_size = radius * 2 + 1;
    const size_t edge_count = kernel_size / 2 + 1;
    const size_t half_kernel = kernel_size / 2;
    const size_t width_bound = width - 1;
    const typename A::Weight weight = A::weight(kernel_size);
    const size_t stride = width * CN;
    for (size_t y = 0; y < height; ++y) {
        const P *s = src + y * stride;
        P *d = dst + y * stride;
        typename A::acc_t sums[4];
        for (int c = 0; c < CN; ++c) sums[c] = A::scale(A::from(s[c]), edge_count);
        for (size_t x = 1; x <= half_kernel; ++x) {
            size_t px = (x < width_bound ? x : width_bound) * CN;
            for (int c = 0; c < CN; ++c) sums[c] += A::from(s[px + c]);
        }
        for (size_t x = 0; x < width; ++x) {
            size_t nx = x + half_kernel + 1;
            size_t next = (nx < width_bound ? nx : width_bound) * CN;
            size_t previous = (x > half_kernel ? x - half_kernel : 0) * CN;
            for (int c = 0; c < CN; ++c) d[x * CN + c] = A::store(sums[c], weight);
            for (int c = 0; c < CN; ++c) {
                sums[c] += A::from(s[next + c]);
                sums[c] -= A::from(s[previous + c]);
            }
        }
    }
}

// box_blur_vertical_pass, fast_blur.rs:416-473
template <typename P, typename A>
void vertical_pass(const P *src, P *dst, size_t width, size_t height, int n, size_t radius) {
    const size_t kernel_size = radius * 2 + 1;
    const size_t edge_count = kernel_size / 2 + 1;
    const size_t half_kernel = kernel_size / 2;
    const size_t height_bound = height - 1;
    const typename A::Weight weight = A::weight(kernel_size);
    const size_t buf_size = width * n;
    std::vector<typename A::acc_t> buffer(buf_size);
    for (size_t x = 0; x < buf_size; ++x) {
        typename A::acc_t w = A::scale(A::from(src[x]), edge_count);
        for (size_t y = 1; y <= half_kernel; ++y) {
            size_t shift = (y < height_bound ? y : height_bound) * buf_size;
            w += A::from(src[shift + x]);
        }
        buffer[x] = w;
    }
    for (size_t y = 0; y < height; ++y) {
        size_t ny = y + half_kernel + 1;
        const P *next_row = src + (ny < height_bound ? ny : height_bound) * buf_size;
        const P *prev_row = src + (y > half_kernel ? y - half_kernel : 0) * buf_size;
        P *d = dst + y * buf_size;
        for (size_t x = 0; x < buf_size; ++x) {
            typename A::acc_t acc = buffer[x];
            d[x] = A::store(acc, weight);
            buffer[x] = acc + A::from(next_row[x]) - A::from(prev_row[x]);
        }
    }
}

template <typename P, typename Native>
int fast_blur_t(const P *src, uint32_t w, uint32_t h, int C, P *dst, float sigma,
                size_t native_max_ks) {
    const size_t n = (size_t)w * h * C;
    if (w == 0 || h == 0) return 0;  // clone of an empty buffer, 117-119
    size_t boxes[3];
    orc_boxes_for_gauss(sigma, boxes);
    std::vector<P> transient(n);
    const P *cur = src;
    for (int pass = 0; pass < 3; ++pass) {
        size_t r = boxes[pass];
        // accumulator choice, 229-247 / 283-300
        if (r * 2 + 1 <= native_max_ks) {
            horizontal_pass<P, Native>(cur, transient.data(), w, h, C, r);
            vertical_pass<P, Native>(transient.data(), dst, w, h, C, r);
        } else {
            horizontal_pass<P, AccF32<P>>(cur, transient.data(), w, h, C, r);
            vertical_pass<P, AccF32<P>>(transient.data(), dst, w, h, C, r);
        }
        cur = dst;
    }
    return 0;
}

}  // namespace

extern "C" uint8_t orc_u8_box_store(uint32_t acc, uint32_t kernel_size) {
    return AccU32::store(acc, AccU32::weight(kernel_size));
}

extern "C" void orc_boxes_for_gauss(float sigma, size_t out[3]) {
    // fast_blur.rs:201-226, n = 3
    const float n = 3.0f;
    float w_ideal = sqrtf((12.0f * rs::powi(sigma, 2) / n) + 1.0f);
    float w_l = floorf(w_ideal);
    if (fmodf(w_l, 2.0f) == 0.0f) w_l -= 1.0f;
    float w_u = w_l + 2.0f;
    float m_ideal = 0.25f * n * (w_l + 3.0f) - 3.0f * rs::powi(sigma, 2) * (1.0f / (w_l + 1.0f));
    size_t m = rs::sat_cast<uint64_t>(roundf(m_ideal));
    for (size_t i = 0; i < 3; ++i) {
        size_t wi = i < m ? rs::sat_cast<uint64_t>(w_l) : rs::sat_cast<uint64_t>(w_u);
        size_t x = (wi > 0 ? wi - 1 : 0) / 2;  // saturating_sub(1) / 2
        out[i] = (x % 2 == 0) ? x + 1 : x;     // ceil_to_odd
    }
}

extern "C" int orc_fast_blur(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                             void *dst, float sigma) {
    if (channels < 1 || channels > 4) return -1;
    switch (dtype) {
        case ORC_U8:
            return fast_blur_t<uint8_t, AccU32>((const uint8_t *)src, w, h, channels, (uint8_t *)dst,
                                                sigma, (size_t)1 << 24);
        case ORC_U16:
            return fast_blur_t<uint16_t, AccF32<uint16_t>>((const uint16_t *)src, w, h, channels,
                                                           (uint16_t *)dst, sigma, SIZE_MAX);
        case ORC_F32:
            return fast_blur_t<float, AccF32<float>>((const float *)src, w, h, channels, (float *)dst,
                                                     sigma, SIZE_MAX);
    }
    return -1;
}
