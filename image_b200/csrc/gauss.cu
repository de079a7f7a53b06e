// gauss.cu -- separable symmetric convolution engine behind the Gaussian blur.
//
// Restates filter_2d_separable (src/imageops/filter_1d.rs:615-761): a row pass
// over every row into an intermediate (f32 for the f32 engine, the *pixel type*
// -- quantised -- for the Q0.15 engine), then a column pass, both clamp-to-edge
// (make_arena_row 60-87, make_columns_arenas 102-130).  Per output subpixel
// (filter_symmetric_row 323-390, filter_symmetric_column 171-316):
//     acc = X(s0) * k[h];  for i in 0..h: acc = mac(acc, X(s[-(h-i)]) + X(s[+(h-i)]), k[i])
// and the store rules of 132-164.  The 1-tap quirks of the reference are kept:
// a 1-tap row kernel reads the pixel to the left (pad = max(len/2, 1) but the
// centre index is len/2), and in the ring-queue variant (< 55 column taps) a
// 1-tap column kernel leaves destination row 0 at its initial zero.
//
// Front ends: imageops::blur / blur_advanced (src/imageops/sample.rs:1034-1055,
// 1492-1604: f32 engine for every pixel type, `nearest_from` at the end) and
// DynamicImage::blur / blur_advanced (sample.rs:1297-1490: Q0.15 for u8/u16).
#include "common.cuh"
#include "host_tables.h"

#include <cstring>

namespace b200 {

// gauss_tiled.cu: register-tiled passes for reorderable sums (FAST f32 engine, Q0.15 engine)
constexpr int GT_MAX_TAPS = 384;
template <typename T, typename Mid, typename Acc>
int gauss_tiled_run(const T *src, T *dst, Mid *mid, uint32_t w, uint32_t h, int C, uint32_t n_images,
                    const uint32_t *rk, int rlen, const uint32_t *ck, int clen, int row_shift, int zero_first_row,
                    cudaStream_t st);

// gauss_fused.cu: both passes in one kernel, intermediate in shared memory (1 = does not apply)
template <typename T>
int gauss_fused_run(const T *src, T *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t src_row0,
                    uint32_t src_rows_buf, uint32_t dst_row0, uint32_t dst_rows, const uint32_t *rk, int rlen,
                    const uint32_t *ck, int clen, bool q15, bool force, const DeviceCtx *ctx, cudaStream_t st);

namespace {

constexpr int THREADS = 256;
constexpr int MAX_TAPS = 1 << 16;

// arithmetic of one engine
template <typename Acc, bool STRICT, bool FUSED> struct Eng;
template <bool STRICT, bool FUSED> struct Eng<float, STRICT, FUSED> {
    static __device__ __forceinline__ float k(uint32_t bits) { return __uint_as_float(bits); }
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float mac(float acc, float a, float b) {
        if (!STRICT || FUSED) return __fmaf_rn(a, b, acc);
        return __fadd_rn(acc, __fmul_rn(a, b));
    }
};
template <bool STRICT, bool FUSED> struct Eng<uint32_t, STRICT, FUSED> {
    static __device__ __forceinline__ uint32_t k(uint32_t bits) { return bits; }
    static __device__ __forceinline__ uint32_t mul(uint32_t a, uint32_t b) { return a * b; }
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) { return a + b; }
    static __device__ __forceinline__ uint32_t mac(uint32_t acc, uint32_t a, uint32_t b) { return acc + a * b; }
};

// ToStorage, src/imageops/filter_1d.rs:132-164, plus the generic path's final
// NearestFrom::nearest_from (src/imageops/sample.rs:1596-1601)
template <typename Out, typename Acc> struct Store;
template <> struct Store<uint8_t, uint32_t> { static __device__ __forceinline__ uint8_t f(uint32_t a) { return (uint8_t)min((a + 16384u) >> 15, 255u); } };
template <> struct Store<uint16_t, uint32_t> { static __device__ __forceinline__ uint16_t f(uint32_t a) { return (uint16_t)min((a + 16384u) >> 15, 65535u); } };
template <> struct Store<float, float> { static __device__ __forceinline__ float f(float a) { return a; } };
template <> struct Store<uint8_t, float> { static __device__ __forceinline__ uint8_t f(float a) { return Px<uint8_t>::nearest(a); } };
template <> struct Store<uint16_t, float> { static __device__ __forceinline__ uint16_t f(float a) { return Px<uint16_t>::nearest(a); } };

struct BlurGeom {
    uint32_t w, h, C;
    int row_len, col_len;  // scanned kernel lengths (odd)
    int row_shift;         // 1 for a 1-tap row kernel (reads the pixel to the left), else 0
    int zero_first_row;    // ring-queue + 1-tap column kernel: row 0 stays zero
};

// Row pass: out[y][x][c] over one image; thread = one subpixel.
template <typename In, typename Mid, typename Acc, bool STRICT, bool FUSED>
__global__ void __launch_bounds__(THREADS)
row_kernel(const In *src, Mid *mid, BlurGeom g, size_t n_sub_total, const uint32_t *__restrict__ c_row_k) {
    using E = Eng<Acc, STRICT, FUSED>;
    const size_t row_sub = (size_t)g.w * g.C;
    const int half = g.row_len / 2;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < n_sub_total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;  // row index across the whole batch
        const uint32_t s = (uint32_t)(idx - row * row_sub);
        const int x = (int)(s / g.C), c = (int)(s - (uint32_t)x * g.C);
        const In *r = src + row * row_sub + c;
        const int xc = x - g.row_shift;
        const int wmax = (int)g.w - 1;
        Acc acc = E::mul((Acc)r[(size_t)min(max(xc, 0), wmax) * g.C], E::k(c_row_k[half]));
        for (int i = 0; i < half; ++i) {
            const int d = half - i;
            const Acc a = (Acc)r[(size_t)min(max(xc - d, 0), wmax) * g.C];
            const Acc b = (Acc)r[(size_t)min(max(xc + d, 0), wmax) * g.C];
            acc = E::mac(acc, E::add(a, b), E::k(c_row_k[i]));
        }
        mid[idx] = Store<Mid, Acc>::f(acc);
    }
}

// Column pass: thread = one subpixel of the output.
template <typename Mid, typename Out, typename Acc, bool STRICT, bool FUSED>
__global__ void __launch_bounds__(THREADS)
col_kernel(const Mid *mid, Out *dst, BlurGeom g, size_t n_sub_total, const uint32_t *__restrict__ c_col_k) {
    using E = Eng<Acc, STRICT, FUSED>;
    const size_t row_sub = (size_t)g.w * g.C;
    const int half = g.col_len / 2;
    const int hmax = (int)g.h - 1;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < n_sub_total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;
        const size_t s = idx - row * row_sub;
        const size_t img = row / g.h;
        const int y = (int)(row - img * g.h);
        if (g.zero_first_row && y == 0) { dst[idx] = (Out)0; continue; }
        const Mid *col = mid + img * (size_t)g.h * row_sub + s;
        Acc acc = E::mul((Acc)col[(size_t)y * row_sub], E::k(c_col_k[half]));
        for (int i = 0; i < half; ++i) {
            const int d = half - i;
            const Acc a = (Acc)col[(size_t)max(y - d, 0) * row_sub];
            const Acc b = (Acc)col[(size_t)min(y + d, hmax) * row_sub];
            acc = E::mac(acc, E::add(a, b), E::k(c_col_k[i]));
        }
        dst[idx] = Store<Out, Acc>::f(acc);
    }
}

template <typename T, typename Mid, typename Acc>
int run(const void *src, void *dst, void *mid, const uint32_t *d_rk, const uint32_t *d_ck, const BlurGeom &g, size_t total,
        int flags, int sm_count, cudaStream_t st) {
    const unsigned grid = (unsigned)std::min<size_t>((total + THREADS - 1) / THREADS, (size_t)sm_count * 32);
    const bool strict = (flags & B200IMG_FP_STRICT) != 0, fused = (flags & B200IMG_FP_FUSED_MAC) != 0;
#define GO(S, F)                                                                                               \
    row_kernel<T, Mid, Acc, S, F><<<grid, THREADS, 0, st>>>((const T *)src, (Mid *)mid, g, total, d_rk);        \
    col_kernel<Mid, T, Acc, S, F><<<grid, THREADS, 0, st>>>((const Mid *)mid, (T *)dst, g, total, d_ck);
    if (!strict) { GO(false, false) }
    else if (fused) { GO(true, true) }
    else { GO(true, false) }
#undef GO
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace

// Rows [dst_row0, dst_row0 + dst_rows) of the blur of a w x h image, from a buffer holding its source rows
// [src_row0, src_row0 + src_rows) (which must include the rows those read), into a buffer whose row 0 is row
// dst_row0: the fused kernel's row window (no scratch, no copy).  Returns 1 when the fused kernel does not cover
// the case (the caller blurs a haloed band into scratch instead).
int gaussian_blur_rows_launch(const void *src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h, int C, int dtype,
                              void *dst_band, uint32_t dst_row0, uint32_t dst_rows, const b200img_gauss_params *p, int path, int flags,
                              cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    const bool q15 = path == B200IMG_BLUR_DYN && dtype != B200IMG_F32;
    const bool strict = (flags & B200IMG_FP_STRICT) != 0;
    if (!((q15 && dtype == B200IMG_U8) || (!q15 && !strict)) || (flags & B200IMG_GAUSS_TWO_KERNELS)) return 1;
    const std::vector<float> kx = gaussian_kernel_1d(p->x_kernel_size, p->x_sigma);
    const std::vector<float> ky = gaussian_kernel_1d(p->y_kernel_size, p->y_sigma);
    std::vector<uint32_t> rk, ck;
    if (q15) {
        rk = q15_scanned_kernel(kx);
        ck = q15_scanned_kernel(ky);
    } else {
        for (float v : f32_scanned_kernel(kx)) { uint32_t b; memcpy(&b, &v, 4); rk.push_back(b); }
        for (float v : f32_scanned_kernel(ky)) { uint32_t b; memcpy(&b, &v, 4); ck.push_back(b); }
    }
    const int rl = (int)rk.size(), cl = (int)ck.size();
    switch (dtype) {
        case B200IMG_U8: return gauss_fused_run<uint8_t>((const uint8_t *)src_band, (uint8_t *)dst_band, w, h, C, 1, src_row0, src_rows, dst_row0, dst_rows, rk.data(), rl, ck.data(), cl, q15, true, ctx, st);
        case B200IMG_U16: return gauss_fused_run<uint16_t>((const uint16_t *)src_band, (uint16_t *)dst_band, w, h, C, 1, src_row0, src_rows, dst_row0, dst_rows, rk.data(), rl, ck.data(), cl, q15, true, ctx, st);
        default: return gauss_fused_run<float>((const float *)src_band, (float *)dst_band, w, h, C, 1, src_row0, src_rows, dst_row0, dst_rows, rk.data(), rl, ck.data(), cl, q15, true, ctx, st);
    }
}

// image_top = false: row 0 of `src` is not the image's first row (a row band of a larger image), so the
// 1-tap ring-queue quirk, which is about the image's row 0, does not apply to it.
int gaussian_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                         const b200img_gauss_params *p, int path, int flags, cudaStream_t st, bool image_top) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    const std::vector<float> kx = gaussian_kernel_1d(p->x_kernel_size, p->x_sigma);
    const std::vector<float> ky = gaussian_kernel_1d(p->y_kernel_size, p->y_sigma);
    const bool q15 = path == B200IMG_BLUR_DYN && dtype != B200IMG_F32;

    std::vector<uint32_t> rk, ck;
    if (q15) {
        rk = q15_scanned_kernel(kx);
        ck = q15_scanned_kernel(ky);
    } else {
        for (float v : f32_scanned_kernel(kx)) { uint32_t b; memcpy(&b, &v, 4); rk.push_back(b); }
        for (float v : f32_scanned_kernel(ky)) { uint32_t b; memcpy(&b, &v, 4); ck.push_back(b); }
    }
    if (rk.size() > MAX_TAPS || ck.size() > MAX_TAPS)
        return fail(B200IMG_ERR_UNSUPPORTED, "gaussian_blur: kernels longer than %d taps are not supported", MAX_TAPS);

    BlurGeom g;
    g.w = w; g.h = h; g.C = (uint32_t)C;
    g.row_len = (int)rk.size(); g.col_len = (int)ck.size();
    g.row_shift = rk.size() == 1 ? 1 : 0;
    // ring-queue variant is chosen on the unscanned column length (filter_1d.rs:673)
    g.zero_first_row = (image_top && ky.size() < 55 && ck.size() == 1) ? 1 : 0;

    size_t total;
    if (!checked_mul(w, h, (size_t)C, n_images, &total)) return fail(B200IMG_ERR_DIMENSION, "gaussian_blur: size overflows");

    const bool strict = (flags & B200IMG_FP_STRICT) != 0;
    // One kernel for both passes (gauss_fused.cu) wherever reordering the sum is allowed and the job is one it
    // wins on (long kernels, enough chunks per SM; B200IMG_GAUSS_FUSED skips that test); it declines (returns 1)
    // what it does not cover.
    if (((q15 && dtype == B200IMG_U8) || (!q15 && !strict)) && image_top && !(flags & B200IMG_GAUSS_TWO_KERNELS)) {
        int fr;
        const bool force = (flags & B200IMG_GAUSS_FUSED) != 0;
        const int rl = (int)rk.size(), cl = (int)ck.size();
        if (dtype == B200IMG_U8) fr = gauss_fused_run<uint8_t>((const uint8_t *)src, (uint8_t *)dst, w, h, C, n_images, 0, h, 0, h, rk.data(), rl, ck.data(), cl, q15, force, ctx, st);
        else if (dtype == B200IMG_U16) fr = gauss_fused_run<uint16_t>((const uint16_t *)src, (uint16_t *)dst, w, h, C, n_images, 0, h, 0, h, rk.data(), rl, ck.data(), cl, q15, force, ctx, st);
        else fr = gauss_fused_run<float>((const float *)src, (float *)dst, w, h, C, n_images, 0, h, 0, h, rk.data(), rl, ck.data(), cl, q15, force, ctx, st);
        if (fr != 1) return fr;
    }
    // the kernel pairs below keep the intermediate image in device scratch
    const size_t mid_elem = q15 ? dtype_size(dtype) : 4;
    void *mid;
    int rc = scratch(SCRATCH_A, total * mid_elem, &mid);
    if (rc) return rc;
    if ((q15 || !strict) && rk.size() <= (size_t)GT_MAX_TAPS && ck.size() <= (size_t)GT_MAX_TAPS) {
        const int rl = (int)rk.size(), cl = (int)ck.size();
#define TILED(T, Mid, Acc)                                                                                                  \
    {                                                                                                                      \
        const int tr = gauss_tiled_run<T, Mid, Acc>((const T *)src, (T *)dst, (Mid *)mid, w, h, C, n_images, rk.data(), rl, \
                                                    ck.data(), cl, g.row_shift, g.zero_first_row, st);                     \
        if (tr != 1) return tr; /* 1 = tile too large for shared memory: direct kernels below */                           \
        break;                                                                                                             \
    }
        do {
            if (q15) {
                if (dtype == B200IMG_U8) TILED(uint8_t, uint8_t, uint32_t)
                TILED(uint16_t, uint16_t, uint32_t)
            }
            if (dtype == B200IMG_U8) TILED(uint8_t, float, float)
            if (dtype == B200IMG_U16) TILED(uint16_t, float, float)
            TILED(float, float, float)
        } while (0);
#undef TILED
    }

    // weights travel in this stream's table scratch, ordered on the stream: back-to-back calls
    // with different kernels are safe (pageable source: staged before the call returns)
    void *tab;
    if ((rc = scratch(SCRATCH_TABLES, (rk.size() + ck.size()) * 4, &tab))) return rc;
    uint32_t *d_rk = (uint32_t *)tab, *d_ck = d_rk + rk.size();
    B200_CUDA(cudaMemcpyAsync(d_rk, rk.data(), rk.size() * 4, cudaMemcpyHostToDevice, st));
    B200_CUDA(cudaMemcpyAsync(d_ck, ck.data(), ck.size() * 4, cudaMemcpyHostToDevice, st));

    if (q15) {
        if (dtype == B200IMG_U8) return run<uint8_t, uint8_t, uint32_t>(src, dst, mid, d_rk, d_ck, g, total, flags, ctx->sm_count, st);
        return run<uint16_t, uint16_t, uint32_t>(src, dst, mid, d_rk, d_ck, g, total, flags, ctx->sm_count, st);
    }
    switch (dtype) {
        case B200IMG_U8: return run<uint8_t, float, float>(src, dst, mid, d_rk, d_ck, g, total, flags, ctx->sm_count, st);
        case B200IMG_U16: return run<uint16_t, float, float>(src, dst, mid, d_rk, d_ck, g, total, flags, ctx->sm_count, st);
        default: return run<float, float, float>(src, dst, mid, d_rk, d_ck, g, total, flags, ctx->sm_count, st);
    }
}

}  // namespace b200
