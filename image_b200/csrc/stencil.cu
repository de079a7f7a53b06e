// stencil.cu -- the first "next" rows after the core path:
//   filter3x3  src/imageops/sample.rs:897-956   9-tap stencil, f32 accumulate in tap order,
//              t * (1 / sum(kernel)), clamp to [0, max], truncating cast; border pixels stay 0.
//   unsharpen  src/imageops/sample.rs:1617-1654 blur_advanced(new_from_sigma(sigma)) followed by
//              the integer rule  diff = c - blurred; |diff| > threshold ? clamp(c + diff, 0, max) : c.
// Both are one streaming pass (HBM-bound); unsharpen reuses the Gaussian engine of gauss*.cu.
#include "common.cuh"

namespace b200 {

int gaussian_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                         const b200img_gauss_params *p, int path, int flags, cudaStream_t st);

namespace {

constexpr int THREADS = 256;

template <typename T> struct Lim;
template <> struct Lim<uint8_t> { static constexpr float fmax_ = 255.0f; static constexpr int imax = 255; };
template <> struct Lim<uint16_t> { static constexpr float fmax_ = 65535.0f; static constexpr int imax = 65535; };
template <> struct Lim<float> { static constexpr float fmax_ = 1.0f; static constexpr int imax = 1; };

struct K9 { float k[9]; float inv; };

// thread = one subpixel of the output; the 9 taps of a subpixel are C elements apart in x
template <typename T>
__global__ void __launch_bounds__(THREADS)
filter3x3_kernel(const T *__restrict__ src, T *__restrict__ dst, uint32_t w, uint32_t h, uint32_t C, size_t total, K9 k) {
    const size_t row_sub = (size_t)w * C;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;
        const uint32_t s = (uint32_t)(idx - row * row_sub);
        const uint32_t y = (uint32_t)(row % h), x = s / C;
        T out = (T)0;
        if (x >= 1 && x + 1 < w && y >= 1 && y + 1 < h) {
            const T *p = src + idx;
            float t = 0.0f;
#pragma unroll
            for (int j = 0; j < 3; ++j) {
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const float v = (float)p[(ptrdiff_t)(j - 1) * (ptrdiff_t)row_sub + (ptrdiff_t)(i - 1) * (ptrdiff_t)C];
                    t = __fadd_rn(t, __fmul_rn(v, k.k[j * 3 + i]));
                }
            }
            t = __fmul_rn(t, k.inv);
            // clamp(ti, 0, max) of src/utils/mod.rs:96-107, then a truncating NumCast
            t = t < 0.0f ? 0.0f : (t > Lim<T>::fmax_ ? Lim<T>::fmax_ : t);
            if (sizeof(T) == 4) out = (T)t;
            else out = (T)__float2uint_rz(t);
        }
        dst[idx] = out;
    }
}

// in place on `blurred`
template <typename T>
__global__ void __launch_bounds__(THREADS)
unsharpen_kernel(const T *__restrict__ src, T *__restrict__ blurred, size_t total, int threshold) {
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const T c = src[idx], d = blurred[idx];
        int ic, id;
        if (sizeof(T) == 4) { ic = __float2int_rz((float)c); id = __float2int_rz((float)d); }  // NumCast f32 -> i32 truncates
        else { ic = (int)c; id = (int)d; }
        const int diff = ic - id;
        blurred[idx] = abs(diff) > threshold ? (T)min(max(ic + diff, 0), Lim<T>::imax) : c;
    }
}

unsigned grid_for(size_t total, int sm_count) {
    return (unsigned)std::min<size_t>((total + THREADS - 1) / THREADS, (size_t)sm_count * 32);
}

}  // namespace

int filter3x3_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     const float kernel[9], cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    K9 k;
    float sum = 0.0f;
    for (int i = 0; i < 9; ++i) { k.k[i] = kernel[i]; sum += kernel[i]; }
    k.inv = sum == 0.0f ? 1.0f : 1.0f / sum;
    size_t total;
    if (!checked_mul(w, h, (size_t)C, n_images, &total)) return fail(B200IMG_ERR_DIMENSION, "filter3x3: size overflows");
    const unsigned g = grid_for(total, ctx->sm_count);
    switch (dtype) {
        case B200IMG_U8: filter3x3_kernel<uint8_t><<<g, THREADS, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, w, h, C, total, k); break;
        case B200IMG_U16: filter3x3_kernel<uint16_t><<<g, THREADS, 0, st>>>((const uint16_t *)src, (uint16_t *)dst, w, h, C, total, k); break;
        default: filter3x3_kernel<float><<<g, THREADS, 0, st>>>((const float *)src, (float *)dst, w, h, C, total, k); break;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

int unsharpen_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images, float sigma,
                     int threshold, int flags, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    b200img_gauss_params p;
    int rc = b200img_gauss_params_from_sigma(sigma, &p);
    if (rc) return rc;
    if ((rc = gaussian_blur_launch(src, w, h, C, dtype, dst, n_images, &p, B200IMG_BLUR_GENERIC, flags, st))) return rc;
    const size_t total = (size_t)w * h * C * n_images;
    const unsigned g = grid_for(total, ctx->sm_count);
    switch (dtype) {
        case B200IMG_U8: unsharpen_kernel<uint8_t><<<g, THREADS, 0, st>>>((const uint8_t *)src, (uint8_t *)dst, total, threshold); break;
        case B200IMG_U16: unsharpen_kernel<uint16_t><<<g, THREADS, 0, st>>>((const uint16_t *)src, (uint16_t *)dst, total, threshold); break;
        default: unsharpen_kernel<float><<<g, THREADS, 0, st>>>((const float *)src, (float *)dst, total, threshold); break;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace b200
