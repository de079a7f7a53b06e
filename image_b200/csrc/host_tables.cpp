// host_tables.cpp -- host-side table builders of libb200img (see host_tables.h).
// Compiled with -ffp-contract=off: rustc does not fuse `a * b + c`, and the
// tables must match the reference's to the last bit.
//
// References: src/imageops/sample.rs 127-205 (kernel functions), 237-307 and
// 514-544 (tap ranges and weights), 1057-1077 (1-D Gaussian), 1132-1294
// (GaussianBlurParameters); src/imageops/filter_1d.rs 392-437 (Q0.15 transform,
// zero-end strip); src/imageops/fast_blur.rs 201-226 (box radii);
// src/math/utils.rs 56-84 (resize_dimensions).
#include "host_tables.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

namespace b200 {
namespace {

const float kPi = 3.14159265358979323846264338327950288f;

// Rust float -> integer `as`: truncating, saturating, NaN -> 0.
template <typename I, typename F>
I rust_as(F v) {
    if (std::isnan(v)) return 0;
    if (v <= (F)std::numeric_limits<I>::min()) return std::numeric_limits<I>::min();
    if (v >= std::ldexp((F)1, std::numeric_limits<I>::digits)) return std::numeric_limits<I>::max();
    return (I)v;
}

inline float sq(float a) { return a * a; }          // powi(2)
inline float cube(float a) { return a * (a * a); }  // powi(3): square-and-multiply

float sinc(float t) {
    const float a = t * kPi;
    return t == 0.0f ? 1.0f : sinf(a) / a;
}

float catmull_rom(float x) {
    // bc_cubic_spline with b = 0, c = 0.5; the coefficient expressions are
    // written out so that they fold exactly as the reference's do.
    const float b = 0.0f, c = 0.5f;
    const float a = fabsf(x);
    float k;
    if (a < 1.0f)
        k = (12.0f - 9.0f * b - 6.0f * c) * cube(a) + (-18.0f + 12.0f * b + 6.0f * c) * sq(a) + (6.0f - 2.0f * b);
    else if (a < 2.0f)
        k = (-b - 6.0f * c) * cube(a) + (6.0f * b + 30.0f * c) * sq(a) + (-12.0f * b - 48.0f * c) * a + (8.0f * b + 24.0f * c);
    else
        k = 0.0f;
    return k / 6.0f;
}

}  // namespace

float filter_support(int filter) {
    static const float s[5] = {0.0f, 1.0f, 2.0f, 3.0f, 3.0f};
    return (filter >= 0 && filter < 5) ? s[filter] : -1.0f;
}

float filter_kernel(int filter, float x) {
    switch (filter) {
        case B200IMG_NEAREST: return 1.0f;
        case B200IMG_TRIANGLE: return fabsf(x) < 1.0f ? 1.0f - fabsf(x) : 0.0f;
        case B200IMG_CATMULLROM: return catmull_rom(x);
        case B200IMG_GAUSSIAN: {
            const float r = 0.5f;
            return (1.0f / (sqrtf(2.0f * kPi) * r)) * expf(-sq(x) / (2.0f * sq(r)));
        }
        case B200IMG_LANCZOS3: return fabsf(x) < 3.0f ? sinc(x) * sinc(x / 3.0f) : 0.0f;
    }
    return 0.0f;
}

int resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len) {
    const float ratio = (float)src_len / (float)dst_len;
    const float sratio = ratio < 1.0f ? 1.0f : ratio;
    return (int)rust_as<uint32_t>(ceilf(2.0f * (filter_support(filter) * sratio))) + 2;
}

// See AxisTable::fixed_R / up2.
static void classify_axis(AxisTable &t, uint32_t src_len, uint32_t dst_len) {
    t.fixed_R = 0; t.fixed_L0 = 0; t.up2 = false;
    const int len = (int)src_len, NT = t.max_count;
    if (dst_len >= 2 && src_len % dst_len == 0 && src_len / dst_len >= 2 && src_len / dst_len <= 64 && NT >= 2) {
        const int R = (int)(src_len / dst_len);
        const uint32_t mid = dst_len / 2;
        bool ok = t.count[mid] == NT;
        const int L0 = t.left[mid] - R * (int)mid;
        const float *wm = &t.w[(size_t)mid * t.stride];
        for (int k = 0; ok && k < NT / 2; ++k) ok = std::memcmp(&wm[k], &wm[NT - 1 - k], 4) == 0;
        for (uint32_t o = 0; ok && o < dst_len; ++o) {
            const int first = R * (int)o + L0, l = std::max(first, 0), r = std::min(first + NT, len);
            ok = t.left[o] == l && t.count[o] == r - l &&
                 (r - l != NT || std::memcmp(&t.w[(size_t)o * t.stride], wm, sizeof(float) * NT) == 0);
        }
        if (ok) { t.fixed_R = R; t.fixed_L0 = L0; }
    }
    if (dst_len == 2 * (uint64_t)src_len && (NT == 3 || NT == 5 || NT == 7) && dst_len >= 32) {
        const int S = (NT - 1) / 2;
        const uint32_t mid = (dst_len / 2) & ~1u;
        bool ok = t.count[mid] == NT && t.count[mid + 1] == NT;
        const float *w0 = &t.w[(size_t)mid * t.stride], *w1 = &t.w[(size_t)(mid + 1) * t.stride];
        for (uint32_t o = 0; ok && o < dst_len; ++o) {
            const int first = (int)(o >> 1) - S, l = std::max(first, 0), r = std::min(first + NT, len);
            ok = t.left[o] == l && t.count[o] == r - l &&
                 (r - l != NT || std::memcmp(&t.w[(size_t)o * t.stride], (o & 1) ? w1 : w0, sizeof(float) * NT) == 0);
        }
        t.up2 = ok;
    }
}

void build_axis_table(int filter, uint32_t src_len, uint32_t dst_len, int axis, AxisTable &t) {
    const float ratio = (float)src_len / (float)dst_len;
    const float sratio = ratio < 1.0f ? 1.0f : ratio;
    const float ss = filter_support(filter) * sratio;

    t.left.assign(dst_len, 0);
    t.count.assign(dst_len, 0);
    // pass 1: ranges
    std::vector<float> centre(dst_len);
    int max_count = 1;
    for (uint32_t o = 0; o < dst_len; ++o) {
        int64_t left, right;
        if (axis == 0) {
            const float c = ((float)o + 0.5f) * ratio;
            left = std::clamp<int64_t>(rust_as<int64_t>(floorf(c - ss)), 0, (int64_t)src_len - 1);
            right = std::clamp<int64_t>(rust_as<int64_t>(ceilf(c + ss)), left + 1, (int64_t)src_len);
            centre[o] = c - 0.5f;
        } else {
            const double c = ((double)o + 0.5) * (double)ratio;
            left = std::min<int64_t>(rust_as<uint32_t>(c - (double)ss), (int64_t)src_len - 1);
            right = std::clamp<int64_t>(rust_as<int64_t>(std::ceil(c + (double)ss)), left + 1, (int64_t)src_len);
            centre[o] = (float)c - 0.5f;
        }
        t.left[o] = (int32_t)left;
        t.count[o] = (int32_t)(right - left);
        max_count = std::max(max_count, t.count[o]);
    }
    t.max_count = max_count;
    t.stride = max_count;
    t.w.assign((size_t)dst_len * t.stride, 0.0f);
    // pass 2: weights, summed in tap order, normalised by division
    for (uint32_t o = 0; o < dst_len; ++o) {
        float *w = &t.w[(size_t)o * t.stride];
        float sum = 0.0f;
        for (int i = 0; i < t.count[o]; ++i) {
            w[i] = filter_kernel(filter, ((float)(t.left[o] + i) - centre[o]) / sratio);
            sum += w[i];
        }
        for (int i = 0; i < t.count[o]; ++i) w[i] /= sum;
    }
    classify_axis(t, src_len, dst_len);
}

// ---- Gaussian parameters ------------------------------------------------------
namespace {
uint32_t round_to_nearest_odd(float x) {
    const uint32_t n = rust_as<uint32_t>(roundf(x));
    if (n & 1u) return n;
    const uint32_t lower = n - 1, upper = n + 1;
    return fabsf(x - (float)lower) <= fabsf(x - (float)upper) ? lower : upper;
}
float sigma_from_kernel_size(float ks) {
    const float safe = ks <= 1.0f ? 0.8f : ks;
    return 0.3f * ((safe - 1.0f) * 0.5f - 1.0f) + 0.8f;
}
bool normal(float v) { return std::fpclassify(v) == FP_NORMAL; }
}  // namespace

int gauss_params_from_kernel_size(float ks, b200img_gauss_params *out) {
    if (!(ks > 0.0f) || !normal(ks)) return B200IMG_ERR_BAD_ARG;
    const uint32_t k = round_to_nearest_odd(ks);
    if ((k & 1u) == 0) return B200IMG_ERR_BAD_ARG;
    const float s = sigma_from_kernel_size(ks);
    *out = {k, s, k, s};
    return B200IMG_OK;
}

int gauss_params_from_radius(float radius, b200img_gauss_params *out) {
    if (!(radius >= 0.0f)) return B200IMG_ERR_BAD_ARG;
    if (radius != 0.0f && !normal(radius)) return B200IMG_ERR_BAD_ARG;
    return gauss_params_from_kernel_size(radius * 2.0f + 1.0f, out);
}

int gauss_params_from_anisotropic(float xks, float yks, b200img_gauss_params *out) {
    // the reference checks is_normal() on the y size only (sample.rs:1176-1191)
    if (!(xks > 0.0f) || !(yks > 0.0f) || !normal(yks)) return B200IMG_ERR_BAD_ARG;
    const uint32_t xk = round_to_nearest_odd(xks), yk = round_to_nearest_odd(yks);
    if ((xk & 1u) == 0 || (yk & 1u) == 0) return B200IMG_ERR_BAD_ARG;
    *out = {xk, sigma_from_kernel_size(xks), yk, sigma_from_kernel_size(yks)};
    return B200IMG_OK;
}

int gauss_params_from_sigma(float sigma, b200img_gauss_params *out) {
    if (!(sigma >= 0.0f) || !std::isfinite(sigma)) return B200IMG_ERR_BAD_ARG;
    if (sigma < 0.16986436f) {
        *out = {1, 1.0f, 1, 1.0f};
        return B200IMG_OK;
    }
    const uint32_t k = rust_as<uint32_t>(fmaxf(6.66666f * sigma - 2.3333333f, 3.0f)) | 1u;
    *out = {k, sigma, k, sigma};
    return B200IMG_OK;
}

std::vector<float> gaussian_kernel_1d(uint32_t width, float sigma) {
    std::vector<float> k(width);
    const float scale = 1.0f / (sqrtf(2.0f * kPi) * sigma);
    const float mean = (float)(width / 2);
    float sum = 0.0f;
    for (uint32_t x = 0; x < width; ++x) {
        const float d = ((float)x - mean) / sigma;
        k[x] = expf(-0.5f * (d * d)) * scale;  // powf(d, 2.0) == d * d
        sum += k[x];
    }
    if (sum != 0.0f) {
        const float inv = 1.0f / sum;
        for (float &v : k) v = v * inv;
    }
    return k;
}

std::vector<uint32_t> q15_scanned_kernel(const std::vector<float> &k) {
    std::vector<uint32_t> q(k.size());
    for (size_t i = 0; i < k.size(); ++i) q[i] = rust_as<uint32_t>(fmaxf(fminf(k[i] * 32768.0f, 32767.0f), 0.0f));
    size_t a = 0, b = q.size();
    while (b - a > 2 && q[a] == 0 && q[b - 1] == 0) { ++a; --b; }
    return std::vector<uint32_t>(q.begin() + a, q.begin() + b);
}

std::vector<float> f32_scanned_kernel(const std::vector<float> &k) {
    size_t a = 0, b = k.size();
    while (b - a > 2 && k[a] == 0.0f && k[b - 1] == 0.0f) { ++a; --b; }
    return std::vector<float>(k.begin() + a, k.begin() + b);
}

void boxes_for_gauss(float sigma, uint32_t out[3]) {
    const float n = 3.0f;
    const float w_ideal = sqrtf(12.0f * sq(sigma) / n + 1.0f);
    float w_l = floorf(w_ideal);
    if (fmodf(w_l, 2.0f) == 0.0f) w_l -= 1.0f;
    const float w_u = w_l + 2.0f;
    const float m_ideal = 0.25f * n * (w_l + 3.0f) - 3.0f * sq(sigma) * (1.0f / (w_l + 1.0f));
    const uint64_t m = rust_as<uint64_t>(roundf(m_ideal));
    for (uint64_t i = 0; i < 3; ++i) {
        const uint64_t w = i < m ? rust_as<uint64_t>(w_l) : rust_as<uint64_t>(w_u);
        const uint64_t half = (w ? w - 1 : 0) / 2;
        const uint64_t r = (half % 2 == 0) ? half + 1 : half;  // ceil_to_odd
        out[i] = (uint32_t)std::min<uint64_t>(r, UINT32_MAX);
    }
}

void resize_dimensions(uint32_t width, uint32_t height, uint32_t nwidth, uint32_t nheight, bool fill,
                       uint32_t *ow, uint32_t *oh) {
    const double wr = (double)nwidth / (double)width, hr = (double)nheight / (double)height;
    const double ratio = fill ? std::fmax(wr, hr) : std::fmin(wr, hr);
    const uint64_t nw = std::max<uint64_t>(rust_as<uint64_t>(std::round((double)width * ratio)), 1);
    const uint64_t nh = std::max<uint64_t>(rust_as<uint64_t>(std::round((double)height * ratio)), 1);
    if (nw > UINT32_MAX) {
        const double r = (double)UINT32_MAX / (double)width;
        *ow = UINT32_MAX;
        *oh = std::max<uint32_t>(rust_as<uint32_t>(std::round((double)height * r)), 1);
    } else if (nh > UINT32_MAX) {
        const double r = (double)UINT32_MAX / (double)height;
        *ow = std::max<uint32_t>(rust_as<uint32_t>(std::round((double)width * r)), 1);
        *oh = UINT32_MAX;
    } else {
        *ow = (uint32_t)nw;
        *oh = (uint32_t)nh;
    }
}

}  // namespace b200
