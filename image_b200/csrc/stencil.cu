// stencil.cu -- the first "next" rows after the core path:
//   filter3x3  src/imageops/sample.rs:897-956   9-tap stencil, f32 accumulate in tap order,
//              t * (1 / sum(kernel)), clamp to [0, max], truncating cast; border pixels stay 0.
//   unsharpen  src/imageops/sample.rs:1617-1654 blur_advanced(new_from_sigma(sigma)) followed by
//              the integer rule  diff = c - blurred; |diff| > threshold ? clamp(c + diff, 0, max) : c.
// Both are one streaming pass (HBM-bound); unsharpen reuses the Gaussian engine of gauss*.cu.
#include "common.cuh"
#include "packed.cuh"

namespace b200 {

int gaussian_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                         const b200img_gauss_params *p, int path, int flags, cudaStream_t st, bool image_top = true);

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

// ---- word-wise filter3x3 ------------------------------------------------------------------
// The subpixel kernel above spends its time in byte loads and I2F (9 of each per output).  Here a
// thread owns one 32-bit word of the output row (4 u8 / 2 u16 / 1 f32 subpixels) and walks V rows
// down the image with a rolling three-row window of already converted floats, so every source word
// is loaded and unpacked once per thread (PRMT + FADD against 2^23 instead of I2F) and the 18
// separately rounded FMUL/FADD per subpixel dominate.  Needs rows and base 4-byte aligned.
template <typename T> __device__ __forceinline__ float elem_of(uint32_t w, int q);
template <> __device__ __forceinline__ float elem_of<uint8_t>(uint32_t w, int q) {
    return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540 + q)) - 8388608.0f;
}
template <> __device__ __forceinline__ float elem_of<uint16_t>(uint32_t w, int q) {
    return __uint_as_float(__byte_perm(w, 0x4B000000u, q ? 0x7532 : 0x7510)) - 8388608.0f;
}
template <> __device__ __forceinline__ float elem_of<float>(uint32_t w, int) { return __uint_as_float(w); }

// f[k][d]: subpixel k of the thread's word, tap column d (0 = one pixel left, 1 = centre, 2 = right)
template <typename T, int C>
__device__ __forceinline__ void load_taps(const uint32_t *__restrict__ row, int i, int nwr, float (&f)[4 / sizeof(T)][3]) {
    constexpr int EPW = 4 / sizeof(T);
    if (sizeof(T) == 4) {
#pragma unroll
        for (int d = 0; d < 3; ++d) f[0][d] = __uint_as_float(__ldg(row + min(max(i + (d - 1) * C, 0), nwr - 1)));
    } else {
        constexpr int LW = (C + EPW - 1) / EPW;  // words to fetch on either side
        uint32_t wv[2 * LW + 1];
#pragma unroll
        for (int j = 0; j <= 2 * LW; ++j) wv[j] = __ldg(row + min(max(i + j - LW, 0), nwr - 1));
#pragma unroll
        for (int k = 0; k < EPW; ++k)
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                constexpr int base = LW * EPW - C;
                const int q = base + k + d * C;
                f[k][d] = elem_of<T>(wv[q / EPW], q % EPW);
            }
    }
}

template <typename T, int C, int V>
__global__ void __launch_bounds__(THREADS)
filter3x3_words_kernel(const T *__restrict__ src, T *__restrict__ dst, uint32_t w, uint32_t h, K9 k) {
    constexpr int EPW = 4 / sizeof(T);
    const int nwr = (int)((size_t)w * C / EPW);
    const int i = blockIdx.x * THREADS + threadIdx.x;
    if (i >= nwr) return;
    const size_t img_words = (size_t)nwr * h;
    const uint32_t *s = reinterpret_cast<const uint32_t *>(src) + blockIdx.z * img_words;
    uint32_t *d = reinterpret_cast<uint32_t *>(dst) + blockIdx.z * img_words;
    const uint32_t y0 = blockIdx.y * V;

    bool xin[EPW];
#pragma unroll
    for (int e = 0; e < EPW; ++e) {
        const uint32_t x = (uint32_t)(i * EPW + e) / C;
        xin[e] = x >= 1 && x + 1 < w;
    }
    float a[EPW][3], b[EPW][3], c[EPW][3];
    load_taps<T, C>(s + (size_t)(y0 ? y0 - 1 : 0) * nwr, i, nwr, a);
    load_taps<T, C>(s + (size_t)y0 * nwr, i, nwr, b);
#pragma unroll
    for (int r = 0; r < V; ++r) {
        const uint32_t y = y0 + r;
        if (y >= h) break;
        load_taps<T, C>(s + (size_t)min(y + 1, h - 1) * nwr, i, nwr, c);
        const bool yin = y >= 1 && y + 1 < h;
        uint32_t out = 0;
#pragma unroll
        for (int e = 0; e < EPW; ++e) {
            float t = 0.0f;  // tap order of sample.rs:907-917, every product and sum rounded on its own
#pragma unroll
            for (int dd = 0; dd < 3; ++dd) t = __fadd_rn(t, __fmul_rn(a[e][dd], k.k[dd]));
#pragma unroll
            for (int dd = 0; dd < 3; ++dd) t = __fadd_rn(t, __fmul_rn(b[e][dd], k.k[3 + dd]));
#pragma unroll
            for (int dd = 0; dd < 3; ++dd) t = __fadd_rn(t, __fmul_rn(c[e][dd], k.k[6 + dd]));
            t = __fmul_rn(t, k.inv);
            t = t < 0.0f ? 0.0f : (t > Lim<T>::fmax_ ? Lim<T>::fmax_ : t);
            const uint32_t bits = sizeof(T) == 4 ? __float_as_uint(t) : __float2uint_rz(t);
            if (xin[e] && yin) out |= bits << (8 * sizeof(T) * e % 32);
        }
        d[(size_t)y * nwr + i] = out;
#pragma unroll
        for (int e = 0; e < EPW; ++e)
#pragma unroll
            for (int dd = 0; dd < 3; ++dd) { a[e][dd] = b[e][dd]; b[e][dd] = c[e][dd]; }
    }
}

template <typename T, int V>
void launch_words(const T *src, T *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, const K9 &k, cudaStream_t st) {
    constexpr int EPW = 4 / sizeof(T);
    const unsigned nwr = (unsigned)((size_t)w * C / EPW);
    dim3 grid((nwr + THREADS - 1) / THREADS, (h + V - 1) / V, n_images);
    switch (C) {
        case 1: filter3x3_words_kernel<T, 1, V><<<grid, THREADS, 0, st>>>(src, dst, w, h, k); break;
        case 2: filter3x3_words_kernel<T, 2, V><<<grid, THREADS, 0, st>>>(src, dst, w, h, k); break;
        case 3: filter3x3_words_kernel<T, 3, V><<<grid, THREADS, 0, st>>>(src, dst, w, h, k); break;
        default: filter3x3_words_kernel<T, 4, V><<<grid, THREADS, 0, st>>>(src, dst, w, h, k); break;
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
    const size_t esz = dtype == B200IMG_U8 ? 1 : (dtype == B200IMG_U16 ? 2 : 4);
    const bool words = ((size_t)w * C * esz) % 4 == 0 && ((uintptr_t)src | (uintptr_t)dst) % 4 == 0 && w >= 3 && h >= 3 &&
                       (size_t)w * C < (1u << 30) && (h + 7) / 8 <= 65535 && n_images <= 65535;
    if (words) {
        switch (dtype) {
            case B200IMG_U8: launch_words<uint8_t, 8>((const uint8_t *)src, (uint8_t *)dst, w, h, C, n_images, k, st); break;
            case B200IMG_U16: launch_words<uint16_t, 8>((const uint16_t *)src, (uint16_t *)dst, w, h, C, n_images, k, st); break;
            default: launch_words<float, 8>((const float *)src, (float *)dst, w, h, C, n_images, k, st); break;
        }
        B200_CUDA(cudaGetLastError());
        return B200IMG_OK;
    }
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
