// colorops.cu -- the elementwise colour operators of src/imageops/colorops.rs (SURVEY.md section 8f
// rank 3): invert (70-83), brighten (161-186), contrast (92-118), huerotate (232-290), each also
// in its *_in_place form (dst == src).  grayscale / grayscale_alpha are pixel.to_luma() =
// Luma::from_color (src/color.rs:436-446) and go through the conversion kernel (convert.cu).
//
// One streaming pass, HBM-bound at 2 * W*H*C*sz bytes.  A thread owns C consecutive 16-byte
// vectors = 16/sz whole pixels, so the channel of every element is a compile-time constant and
// the alpha / first-three-channels rules cost nothing at run time.  Unaligned buffers and the
// last few pixels take a per-pixel path.
#include "common.cuh"

namespace b200 {
namespace {

enum { OP_INVERT = 0, OP_BRIGHTEN = 1, OP_CONTRAST = 2, OP_HUEROTATE = 3 };

struct ColorParams {
    int32_t value;      // brighten
    float percent;      // contrast: ((100 + contrast) / 100)^2
    double m[9];        // huerotate matrix (built on the host in f64, colorops.rs:246-262)
};

template <typename T> struct CMax;
template <> struct CMax<uint8_t> { static constexpr int i = 255; };
template <> struct CMax<uint16_t> { static constexpr int i = 65535; };
template <> struct CMax<float> { static constexpr int i = 1; };  // NumCast::from(1.0f32) -> i32

template <typename T> __device__ __forceinline__ int32_t to_i32(T v) { return (int32_t)v; }
template <> __device__ __forceinline__ int32_t to_i32<float>(float v) { return __float2int_rz(v); }  // saturating, NaN -> 0

template <typename T> __device__ __forceinline__ T from_f64_trunc(double e) { return (T)__double2uint_rz(e); }
template <> __device__ __forceinline__ float from_f64_trunc<float>(double e) { return __double2float_rn(e); }

// contrast's f32 round trip without the quarter-rate conversion pipe: integers enter through the
// 2^23 mantissa trick, c / max is the FMA-corrected quotient convert.cu validates exhaustively
// against IEEE division for every u8 / u16 value, and the truncating cast of a value in
// [0, max] is one round-toward-zero add of 2^23.
template <typename T> __device__ __forceinline__ float int_to_f32(T v) { return __uint_as_float(0x4B000000u | (uint32_t)v) - 8388608.0f; }
template <> __device__ __forceinline__ float int_to_f32<float>(float v) { return v; }
template <typename T> __device__ __forceinline__ float div_max(float x);
template <> __device__ __forceinline__ float div_max<float>(float x) { return x; }  // x / 1.0
__device__ __forceinline__ float div_exact(float x, float d, float r) {
    const float q = __fmul_rn(x, r);
    return __fmaf_rn(__fmaf_rn(-q, d, x), r, q);
}
template <> __device__ __forceinline__ float div_max<uint8_t>(float x) { return div_exact(x, 255.0f, 1.0f / 255.0f); }
template <> __device__ __forceinline__ float div_max<uint16_t>(float x) { return div_exact(x, 65535.0f, 1.0f / 65535.0f); }
template <typename T> __device__ __forceinline__ T trunc_store(float e) { return (T)(__float_as_uint(__fadd_rz(e, 8388608.0f)) & 0xFFFFu); }
template <> __device__ __forceinline__ float trunc_store<float>(float e) { return e; }

__device__ __forceinline__ double clamp255(double v) { return v < 0.0 ? 0.0 : (v > 255.0 ? 255.0 : v); }

// one pixel, in registers
template <typename T, int C, int OP>
__device__ __forceinline__ void color_pixel(T (&px)[C], const ColorParams &p) {
    constexpr int ALPHA = C == 2 ? 1 : (C == 4 ? 3 : -1);
    if (OP == OP_INVERT) {
#pragma unroll
        for (int c = 0; c < C; ++c)
            if (c != ALPHA) px[c] = (T)((T)CMax<T>::i - px[c]);
    } else if (OP == OP_BRIGHTEN) {
#pragma unroll
        for (int c = 0; c < C; ++c)
            if (c != ALPHA) {
                const int32_t s = (int32_t)((uint32_t)to_i32<T>(px[c]) + (uint32_t)p.value);  // wrapping, as a release build
                px[c] = (T)min(max(s, 0), CMax<T>::i);
            }
    } else if (OP == OP_CONTRAST) {
        const float mx = (float)CMax<T>::i;
#pragma unroll
        for (int c = 0; c < C; ++c) {
            const float v = int_to_f32<T>(px[c]);
            // ((c / max - 0.5) * percent + 0.5) * max, every operation rounded on its own
            float d = __fmul_rn(__fadd_rn(__fmul_rn(__fsub_rn(div_max<T>(v), 0.5f), p.percent), 0.5f), mx);
            d = d < 0.0f ? 0.0f : (d > mx ? mx : d);   // NaN (percent = inf on a mid-grey) falls through, as in utils::clamp
            px[c] = trunc_store<T>(d);
        }
    } else if (C >= 3) {  // OP_HUEROTATE; fewer than three channels: copy
        const double r = (double)px[0], g = (double)px[1], b = (double)px[2];
        const double nr = __dadd_rn(__dadd_rn(__dmul_rn(p.m[0], r), __dmul_rn(p.m[1], g)), __dmul_rn(p.m[2], b));
        const double ng = __dadd_rn(__dadd_rn(__dmul_rn(p.m[3], r), __dmul_rn(p.m[4], g)), __dmul_rn(p.m[5], b));
        const double nb = __dadd_rn(__dadd_rn(__dmul_rn(p.m[6], r), __dmul_rn(p.m[7], g)), __dmul_rn(p.m[8], b));
        px[0] = from_f64_trunc<T>(clamp255(nr));   // max = 255 for every subpixel type (colorops.rs:276)
        px[1] = from_f64_trunc<T>(clamp255(ng));
        px[2] = from_f64_trunc<T>(clamp255(nb));
    }
}

// vector body: thread = C uint4 = 16/sizeof(T) pixels
template <typename T, int C, int OP>
__global__ void __launch_bounds__(256)
colorop_vec_kernel(const uint4 *src, uint4 *dst, size_t units, const __grid_constant__ ColorParams p) {
    constexpr int VEC = 16 / sizeof(T);  // pixels per unit
    for (size_t u = (size_t)blockIdx.x * 256 + threadIdx.x; u < units; u += (size_t)gridDim.x * 256) {
        union { uint4 v[C]; T e[VEC * C]; } buf;
#pragma unroll
        for (int k = 0; k < C; ++k) buf.v[k] = src[u * C + k];   // plain loads: dst may be src
#pragma unroll
        for (int q = 0; q < VEC; ++q) {
            T px[C];
#pragma unroll
            for (int c = 0; c < C; ++c) px[c] = buf.e[q * C + c];
            color_pixel<T, C, OP>(px, p);
#pragma unroll
            for (int c = 0; c < C; ++c) buf.e[q * C + c] = px[c];
        }
#pragma unroll
        for (int k = 0; k < C; ++k) dst[u * C + k] = buf.v[k];
    }
}

// per-pixel path: pixels [first, npix)
template <typename T, int C, int OP>
__global__ void __launch_bounds__(256)
colorop_pixel_kernel(const T *src, T *dst, size_t first, size_t npix, const __grid_constant__ ColorParams p) {
    for (size_t i = first + (size_t)blockIdx.x * 256 + threadIdx.x; i < npix; i += (size_t)gridDim.x * 256) {
        T px[C];
#pragma unroll
        for (int c = 0; c < C; ++c) px[c] = src[i * C + c];
        color_pixel<T, C, OP>(px, p);
#pragma unroll
        for (int c = 0; c < C; ++c) dst[i * C + c] = px[c];
    }
}

template <typename T, int C, int OP>
int launch_op(const T *src, T *dst, size_t npix, const ColorParams &p, int sm_count, cudaStream_t st) {
    constexpr int VEC = 16 / sizeof(T);
    const bool aligned = (((uintptr_t)src | (uintptr_t)dst) & 15) == 0;
    const size_t units = aligned ? npix / VEC : 0;
    if (units) {
        const unsigned g = (unsigned)std::min<size_t>((units + 255) / 256, (size_t)sm_count * 16);
        colorop_vec_kernel<T, C, OP><<<g, 256, 0, st>>>((const uint4 *)src, (uint4 *)dst, units, p);
    }
    const size_t first = units * VEC;
    if (first < npix) {
        const unsigned g = (unsigned)std::min<size_t>((npix - first + 255) / 256, (size_t)sm_count * 16);
        colorop_pixel_kernel<T, C, OP><<<g, 256, 0, st>>>(src, dst, first, npix, p);
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename T, int C>
int launch_c(int op, const T *src, T *dst, size_t npix, const ColorParams &p, int sm, cudaStream_t st) {
    switch (op) {
        case OP_INVERT: return launch_op<T, C, OP_INVERT>(src, dst, npix, p, sm, st);
        case OP_BRIGHTEN: return launch_op<T, C, OP_BRIGHTEN>(src, dst, npix, p, sm, st);
        case OP_CONTRAST: return launch_op<T, C, OP_CONTRAST>(src, dst, npix, p, sm, st);
        default: return launch_op<T, C, OP_HUEROTATE>(src, dst, npix, p, sm, st);
    }
}

template <typename T>
int launch_t(int op, int C, const T *src, T *dst, size_t npix, const ColorParams &p, int sm, cudaStream_t st) {
    switch (C) {
        case 1: return launch_c<T, 1>(op, src, dst, npix, p, sm, st);
        case 2: return launch_c<T, 2>(op, src, dst, npix, p, sm, st);
        case 3: return launch_c<T, 3>(op, src, dst, npix, p, sm, st);
        default: return launch_c<T, 4>(op, src, dst, npix, p, sm, st);
    }
}

}  // namespace

// colorops.rs:246-262, f64 on the host (glibc cos / sin, as the reference's std does on Linux)
void huerotate_matrix(int32_t value, double m[9]) {
    const double rad = (double)value * (3.14159265358979323846264338327950288 / 180.0);
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

// op: 0 invert, 1 brighten(ivalue), 2 contrast(fvalue), 3 huerotate(ivalue degrees); dst may equal src
int colorop_launch(int op, const void *src, void *dst, size_t npixels, int C, int dtype, int32_t ivalue, float fvalue,
                   cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    ColorParams p = {};
    p.value = ivalue;
    {
        volatile float base = (100.0f + fvalue) / 100.0f;  // f32, then powi(2) = base * base
        volatile float sq = base * base;
        p.percent = sq;
    }
    if (op == OP_HUEROTATE) huerotate_matrix(ivalue, p.m);
    switch (dtype) {
        case B200IMG_U8: return launch_t<uint8_t>(op, C, (const uint8_t *)src, (uint8_t *)dst, npixels, p, ctx->sm_count, st);
        case B200IMG_U16: return launch_t<uint16_t>(op, C, (const uint16_t *)src, (uint16_t *)dst, npixels, p, ctx->sm_count, st);
        default: return launch_t<float>(op, C, (const float *)src, (float *)dst, npixels, p, ctx->sm_count, st);
    }
}

}  // namespace b200
