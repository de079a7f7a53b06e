/*
 * resample.cpp -- oracle restatement of imageops::resize
 * (TEST INFRASTRUCTURE ONLY; see oracle.h).
 *
 * Follows src/imageops/sample.rs: kernels 127-205, horizontal_sample 217-340,
 * vertical_sample 494-562, resize 970-1019; store rule
 * src/primitive_sealed.rs:107-147; clamp src/utils/mod.rs:96-107;
 * resize_dimensions src/math/utils.rs:56-84.
 */
#include "oracle.h"
#include "rust_sem.h"

#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

int g_orc_fused_mac = 0;
extern "C" void orc_set_fused_mac(int fused) { g_orc_fused_mac = fused ? 1 : 0; }

namespace {

const float PI_F32 = 3.14159265358979323846264338327950288f;  // f32::consts::PI

// sample.rs:127-135
float sinc(float t) {
    float a = t * PI_F32;
    if (t == 0.0f) return 1.0f;
    return sinf(a) / a;
}

// sample.rs:138-144
float lanczos(float x, float t) {
    if (fabsf(x) < t) return sinc(x) * sinc(x / t);
    return 0.0f;
}

// sample.rs:148-165
float bc_cubic_spline(float x, float b, float c) {
    float a = fabsf(x);
    float k;
    if (a < 1.0f) {
        k = (12.0f - 9.0f * b - 6.0f * c) * rs::powi(a, 3) +
            (-18.0f + 12.0f * b + 6.0f * c) * rs::powi(a, 2) + (6.0f - 2.0f * b);
    } else if (a < 2.0f) {
        k = (-b - 6.0f * c) * rs::powi(a, 3) + (6.0f * b + 30.0f * c) * rs::powi(a, 2) +
            (-12.0f * b - 48.0f * c) * a + (8.0f * b + 24.0f * c);
    } else {
        k = 0.0f;
    }
    return k / 6.0f;
}

// sample.rs:169-171
float gaussian(float x, float r) {
    return (1.0f / (sqrtf(2.0f * PI_F32) * r)) * expf(-rs::powi(x, 2) / (2.0f * rs::powi(r, 2)));
}

// sample.rs:192-198
float triangle_kernel(float x) {
    if (fabsf(x) < 1.0f) return 1.0f - fabsf(x);
    return 0.0f;
}

}  // namespace

// Filter table in resize(), sample.rs:993-1014
extern "C" float orc_filter_support(int filter) {
    switch (filter) {
        case ORC_NEAREST: return 0.0f;
        case ORC_TRIANGLE: return 1.0f;
        case ORC_CATMULLROM: return 2.0f;
        case ORC_GAUSSIAN: return 3.0f;
        case ORC_LANCZOS3: return 3.0f;
    }
    return -1.0f;
}

extern "C" float orc_filter_kernel(int filter, float x) {
    switch (filter) {
        case ORC_NEAREST: return 1.0f;                              // box_kernel, 203-205
        case ORC_TRIANGLE: return triangle_kernel(x);               // 192-198
        case ORC_CATMULLROM: return bc_cubic_spline(x, 0.0f, 0.5f);  // 186-188
        case ORC_GAUSSIAN: return gaussian(x, 0.5f);                // 180-182
        case ORC_LANCZOS3: return lanczos(x, 3.0f);                 // 174-176
    }
    return 0.0f;
}

namespace {

struct AxisScale {
    float ratio, sratio, src_support;
};

// sample.rs:237-239 == 514-516
AxisScale axis_scale(int filter, uint32_t src_len, uint32_t dst_len) {
    AxisScale s;
    s.ratio = (float)src_len / (float)dst_len;
    s.sratio = s.ratio < 1.0f ? 1.0f : s.ratio;
    s.src_support = orc_filter_support(filter) * s.sratio;
    return s;
}

// One output sample's taps.  Vertical rule: sample.rs:521-544.  Horizontal
// rule: sample.rs:279-306.  Returns the tap count; weights normalised by division.
int taps_for(int filter, const AxisScale &s, uint32_t src_len, uint32_t o, int axis,
             uint32_t *left_out, std::vector<float> &ws) {
    uint32_t left, right;
    float centre;
    if (axis == 0) {
        float inputy = ((float)o + 0.5f) * s.ratio;
        int64_t l = rs::sat_cast<int64_t>(floorf(inputy - s.src_support));
        left = (uint32_t)rs::clamp<int64_t>(l, 0, (int64_t)src_len - 1);
        int64_t r = rs::sat_cast<int64_t>(ceilf(inputy + s.src_support));
        right = (uint32_t)rs::clamp<int64_t>(r, (int64_t)left + 1, (int64_t)src_len);
        centre = inputy - 0.5f;
    } else {
        double inputx = ((double)o + 0.5) * (double)s.ratio;
        left = rs::clamp<uint32_t>(rs::sat_cast<uint32_t>(inputx - (double)s.src_support), 0,
                                   src_len - 1);
        int64_t r = rs::sat_cast<int64_t>(ceil(inputx + (double)s.src_support));
        right = (uint32_t)rs::clamp<int64_t>(r, (int64_t)left + 1, (int64_t)src_len);
        centre = (float)inputx - 0.5f;
    }
    ws.clear();
    float sum = 0.0f;
    for (uint32_t i = left; i < right; ++i) {
        float w = orc_filter_kernel(filter, ((float)i - centre) / s.sratio);
        ws.push_back(w);
        sum += w;
    }
    for (float &w : ws) w /= sum;
    *left_out = left;
    return (int)ws.size();
}

template <typename T>
void resize_rows(const T *src, uint32_t w, uint32_t h, int C, T *dst, uint32_t nw, uint32_t nh,
                 int filter, float acc_init, uint32_t oy0, uint32_t oy1) {
    const AxisScale vs = axis_scale(filter, h, nh);
    const AxisScale hs = axis_scale(filter, w, nw);

    // horizontal tables once (the reference batches them, 244-307: no numeric effect)
    std::vector<uint32_t> hleft(nw);
    std::vector<size_t> hstart(nw + 1, 0);
    std::vector<float> hws;
    {
        std::vector<float> ws;
        for (uint32_t ox = 0; ox < nw; ++ox) {
            taps_for(filter, hs, w, ox, 1, &hleft[ox], ws);
            hws.insert(hws.end(), ws.begin(), ws.end());
            hstart[ox + 1] = hws.size();
        }
    }

    std::vector<float> tmp((size_t)w * 4);  // one Rgba32F row of vertical_sample's output
    std::vector<float> ws;
    for (uint32_t oy = oy0; oy < oy1; ++oy) {
        uint32_t left;
        taps_for(filter, vs, h, oy, 0, &left, ws);
        // vertical_sample inner loop, 546-558: pix = [acc_init; 4]; only the
        // first C lanes accumulate (zip with p.channels()).
        for (uint32_t x = 0; x < w; ++x) {
            float pix[4] = {acc_init, acc_init, acc_init, acc_init};
            for (size_t i = 0; i < ws.size(); ++i) {
                const T *p = src + ((size_t)(left + i) * w + x) * C;
                for (int c = 0; c < C; ++c) pix[c] += (float)p[c] * ws[i];
            }
            memcpy(&tmp[(size_t)x * 4], pix, sizeof pix);
        }
        // horizontal_sample inner loop, 310-334
        for (uint32_t ox = 0; ox < nw; ++ox) {
            float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
            const float *wsx = &hws[hstart[ox]];
            const size_t n = hstart[ox + 1] - hstart[ox];
            for (size_t i = 0; i < n; ++i) {
                const float *s4 = &tmp[(size_t)(hleft[ox] + i) * 4];
                for (int c = 0; c < 4; ++c) t[c] += s4[c] * wsx[i];
            }
            T *d = dst + ((size_t)oy * nw + ox) * C;
            for (int c = 0; c < C; ++c) d[c] = rs::clamp_nearest_from<T>(t[c]);
        }
    }
}

template <typename T>
int resize_t(const T *src, uint32_t w, uint32_t h, int C, T *dst, uint32_t nw, uint32_t nh,
             int filter, float acc_init, int threads) {
    // resize(), 977-991
    if (w == 0 || h == 0) {
        memset(dst, 0, (size_t)nw * nh * C * sizeof(T));
        return 0;
    }
    if (nw == w && nh == h) {
        memcpy(dst, src, (size_t)w * h * C * sizeof(T));
        return 0;
    }
    if (nw == 0 || nh == 0) return 0;  // empty output; nothing to store
    if (threads <= 1) {
        resize_rows(src, w, h, C, dst, nw, nh, filter, acc_init, 0, nh);
        return 0;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        uint32_t a = (uint32_t)((uint64_t)nh * t / threads);
        uint32_t b = (uint32_t)((uint64_t)nh * (t + 1) / threads);
        if (a < b)
            pool.emplace_back([=] { resize_rows(src, w, h, C, dst, nw, nh, filter, acc_init, a, b); });
    }
    for (auto &th : pool) th.join();
    return 0;
}

}  // namespace

extern "C" int orc_resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len) {
    // sample.rs:242: (2.0 * src_support).ceil() as usize + 2
    AxisScale s = axis_scale(filter, src_len, dst_len);
    return (int)rs::sat_cast<uint32_t>(ceilf(2.0f * s.src_support)) + 2;
}

extern "C" int orc_resize_weights(int filter, uint32_t src_len, uint32_t dst_len, int axis,
                                  int32_t *left, int32_t *count, float *weights, int stride) {
    if (orc_filter_support(filter) < 0.0f || src_len == 0) return -1;
    AxisScale s = axis_scale(filter, src_len, dst_len);
    std::vector<float> ws;
    for (uint32_t o = 0; o < dst_len; ++o) {
        uint32_t l;
        int n = taps_for(filter, s, src_len, o, axis, &l, ws);
        if (n > stride) return -2;
        left[o] = (int32_t)l;
        count[o] = n;
        for (int i = 0; i < stride; ++i) weights[(size_t)o * stride + i] = i < n ? ws[i] : 0.0f;
    }
    return 0;
}

extern "C" int orc_resize_mt(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                             void *dst, uint32_t nw, uint32_t nh, int filter, float acc_init,
                             int threads) {
    if (channels < 1 || channels > 4 || orc_filter_support(filter) < 0.0f) return -1;
    switch (dtype) {
        case ORC_U8:
            return resize_t((const uint8_t *)src, w, h, channels, (uint8_t *)dst, nw, nh, filter,
                            acc_init, threads);
        case ORC_U16:
            return resize_t((const uint16_t *)src, w, h, channels, (uint16_t *)dst, nw, nh, filter,
                            acc_init, threads);
        case ORC_F32:
            return resize_t((const float *)src, w, h, channels, (float *)dst, nw, nh, filter,
                            acc_init, threads);
    }
    return -1;
}

extern "C" int orc_resize(const void *src, uint32_t w, uint32_t h, int channels, int dtype,
                          void *dst, uint32_t nw, uint32_t nh, int filter, float acc_init) {
    return orc_resize_mt(src, w, h, channels, dtype, dst, nw, nh, filter, acc_init, 1);
}

// src/math/utils.rs:56-84
extern "C" void orc_resize_dimensions(uint32_t width, uint32_t height, uint32_t nwidth,
                                      uint32_t nheight, int fill, uint32_t *ow, uint32_t *oh) {
    double wratio = (double)nwidth / (double)width;
    double hratio = (double)nheight / (double)height;
    double ratio = fill ? fmax(wratio, hratio) : fmin(wratio, hratio);
    uint64_t nw = std::max<uint64_t>(rs::sat_cast<uint64_t>(round((double)width * ratio)), 1);
    uint64_t nh = std::max<uint64_t>(rs::sat_cast<uint64_t>(round((double)height * ratio)), 1);
    if (nw > (uint64_t)UINT32_MAX) {
        double r = (double)UINT32_MAX / (double)width;
        *ow = UINT32_MAX;
        *oh = std::max<uint32_t>(rs::sat_cast<uint32_t>(round((double)height * r)), 1);
    } else if (nh > (uint64_t)UINT32_MAX) {
        double r = (double)UINT32_MAX / (double)height;
        *ow = std::max<uint32_t>(rs::sat_cast<uint32_t>(round((double)width * r)), 1);
        *oh = UINT32_MAX;
    } else {
        *ow = (uint32_t)nw;
        *oh = (uint32_t)nh;
    }
}
