// boxblur.cu -- imageops::fast_blur: three box blurs (Kovesi), each a horizontal
// then a vertical pass that stores back to the pixel type.
//
// Reference: src/imageops/fast_blur.rs -- fast_blur 111-191, boxes_for_gauss
// 201-226, box_blur_horizontal_pass 303-413, box_blur_vertical_pass 416-473,
// accumulators 36-94 (u8: u32 sums and an exact multiply-shift division; other
// types: f32 sums times 1/ks, then clamp_nearest_from).
//
// The reference slides one running sum along each row / column.  A window sum
// of integers is the same number in any order, so for u8 (u32 sums) and for
// u16 while ks * 65535 < 2^24 (f32 sums of integers below 2^24 are exact) the
// result here is bit-identical.  For f32 pixels the running sum's rounding
// history is not reproduced; results agree to f32 round-off.
#include "common.cuh"
#include "host_tables.h"

namespace b200 {
namespace {

constexpr int THREADS = 256;

template <typename T> struct Box;
template <> struct Box<uint8_t> {
    typedef uint32_t acc_t;
    static __device__ __forceinline__ uint32_t lift(uint8_t v) { return v; }
    // to_store, fast_blur.rs:66-70: ((acc + ks/2) * ceil(2^32/ks)) >> 32
    static __device__ __forceinline__ uint8_t store(uint32_t acc, uint32_t ks, uint32_t recip, float) {
        return (uint8_t)__umulhi(acc + ks / 2u, recip);
    }
};
template <> struct Box<uint16_t> {
    typedef float acc_t;
    static __device__ __forceinline__ float lift(uint16_t v) { return (float)v; }
    static __device__ __forceinline__ uint16_t store(float acc, uint32_t, uint32_t, float inv) {
        return Px<uint16_t>::clamp_nearest(__fmul_rn(acc, inv));
    }
};
template <> struct Box<float> {
    typedef float acc_t;
    static __device__ __forceinline__ float lift(float v) { return v; }
    static __device__ __forceinline__ float store(float acc, uint32_t, uint32_t, float inv) {
        return Px<float>::clamp_nearest(__fmul_rn(acc, inv));
    }
};

struct BoxGeom {
    uint32_t w, h, C;
    uint32_t r, ks, recip;  // radius, 2r+1, ceil(2^32 / ks)
    float inv;              // 1.0f / ks
};

// window sum along x with clamp-to-edge: sum_{d=-r..r} p[clamp(x+d)]
template <typename T, typename B>
__global__ void __launch_bounds__(THREADS)
box_h_kernel(const T *src, T *dst, BoxGeom g, size_t total) {
    const size_t row_sub = (size_t)g.w * g.C;
    const int wmax = (int)g.w - 1, r = (int)g.r;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;
        const uint32_t s = (uint32_t)(idx - row * row_sub);
        const int x = (int)(s / g.C), c = (int)(s - (uint32_t)x * g.C);
        const T *p = src + row * row_sub + c;
        typename B::acc_t acc = 0;
        // replicated edges contribute count * edge value; the interior is a plain sum
        const int lo = x - r, hi = x + r;
        const int nlo = lo < 0 ? -lo : 0;           // taps clamped to column 0
        const int nhi = hi > wmax ? hi - wmax : 0;  // taps clamped to column w-1
        const int a = max(lo, 0), b = min(hi, wmax);
        for (int xx = a; xx <= b; ++xx) acc += B::lift(p[(size_t)xx * g.C]);
        if (nlo) acc += B::lift(p[0]) * (typename B::acc_t)nlo;
        if (nhi) acc += B::lift(p[(size_t)wmax * g.C]) * (typename B::acc_t)nhi;
        dst[idx] = B::store(acc, g.ks, g.recip, g.inv);
    }
}

template <typename T, typename B>
__global__ void __launch_bounds__(THREADS)
box_v_kernel(const T *src, T *dst, BoxGeom g, size_t total) {
    const size_t row_sub = (size_t)g.w * g.C;
    const int hmax = (int)g.h - 1, r = (int)g.r;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;
        const size_t s = idx - row * row_sub;
        const size_t img = row / g.h;
        const int y = (int)(row - img * g.h);
        const T *p = src + img * (size_t)g.h * row_sub + s;
        typename B::acc_t acc = 0;
        const int lo = y - r, hi = y + r;
        const int nlo = lo < 0 ? -lo : 0;
        const int nhi = hi > hmax ? hi - hmax : 0;
        const int a = max(lo, 0), b = min(hi, hmax);
        for (int yy = a; yy <= b; ++yy) acc += B::lift(p[(size_t)yy * row_sub]);
        if (nlo) acc += B::lift(p[0]) * (typename B::acc_t)nlo;
        if (nhi) acc += B::lift(p[(size_t)hmax * row_sub]) * (typename B::acc_t)nhi;
        dst[idx] = B::store(acc, g.ks, g.recip, g.inv);
    }
}

template <typename T>
int run(const void *src, void *dst, void *tmp, uint32_t w, uint32_t h, int C, uint32_t n_images, const uint32_t radii[3],
        int sm_count, cudaStream_t st) {
    const size_t total = (size_t)w * h * C * n_images;
    const unsigned grid = (unsigned)std::min<size_t>((total + THREADS - 1) / THREADS, (size_t)sm_count * 32);
    const T *cur = (const T *)src;
    for (int pass = 0; pass < 3; ++pass) {
        BoxGeom g;
        g.w = w; g.h = h; g.C = (uint32_t)C;
        g.r = radii[pass];
        g.ks = 2 * g.r + 1;
        g.recip = UINT32_MAX / g.ks + 1u;
        g.inv = 1.0f / (float)g.ks;
        box_h_kernel<T, Box<T>><<<grid, THREADS, 0, st>>>(cur, (T *)tmp, g, total);
        box_v_kernel<T, Box<T>><<<grid, THREADS, 0, st>>>((const T *)tmp, (T *)dst, g, total);
        cur = (const T *)dst;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace

int fast_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     float sigma, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    uint32_t radii[3];
    boxes_for_gauss(sigma, radii);
    for (int i = 0; i < 3; ++i)
        if (radii[i] >= (1u << 23))  // beyond here the reference's u8 path switches accumulators (fast_blur.rs:238-247)
            return fail(B200IMG_ERR_UNSUPPORTED, "fast_blur: box radius %u is beyond the supported range", radii[i]);
    size_t bytes;
    if (!checked_mul((size_t)w * h, (size_t)C, dtype_size(dtype), n_images, &bytes))
        return fail(B200IMG_ERR_DIMENSION, "fast_blur: size overflows");
    void *tmp;
    int rc = scratch(SCRATCH_A, bytes, &tmp);
    if (rc) return rc;
    switch (dtype) {
        case B200IMG_U8: return run<uint8_t>(src, dst, tmp, w, h, C, n_images, radii, ctx->sm_count, st);
        case B200IMG_U16: return run<uint16_t>(src, dst, tmp, w, h, C, n_images, radii, ctx->sm_count, st);
        default: return run<float>(src, dst, tmp, w, h, C, n_images, radii, ctx->sm_count, st);
    }
}

}  // namespace b200
