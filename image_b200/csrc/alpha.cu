// alpha.cu -- the steps the reference takes either side of a DynamicImage::resize
// (src/imageops/resize.rs:13-93), SURVEY.md section 8f rank 4:
//   has_constant_alpha  resize.rs:155-218, in-tree and restated exactly: integer alphas are
//                       XOR-ed with the first pixel's and summed (non-zero sum <=> some alpha
//                       differs); f32 alphas are compared with `==` (so a NaN first alpha is never
//                       "constant").  Images without an alpha channel are constant.
//   premultiply / unpremultiply   the reference calls pic-scale-safe (premultiply_rgba8 & co.),
//                       which is not vendored and not version-locked here: PARITY UNPINNED.  The
//                       arithmetic below is this library's own round-to-nearest rule,
//                       (c*a + max/2) / max and (c*max + a/2) / a clamped to max (a = 0 keeps c);
//                       f32: c*a and c/a (a = 0 keeps c).
// One streaming pass each; the scan reads 16-byte vectors and tests the alpha lanes with masks.
#include "common.cuh"

namespace b200 {
namespace {

template <typename T> struct AMax;
template <> struct AMax<uint8_t> { static constexpr uint32_t v = 255; };
template <> struct AMax<uint16_t> { static constexpr uint32_t v = 65535; };

// ---- constant-alpha scan ------------------------------------------------------------------
template <typename T, int C>
__global__ void __launch_bounds__(256)
alpha_scan_kernel(const T *__restrict__ src, size_t npix, int *__restrict__ differs) {
    const T first = src[C - 1];
    bool bad = false;
    const bool aligned = ((uintptr_t)src & 15) == 0;
    constexpr int VEC = 16 / (sizeof(T) * C);  // pixels per 16-byte vector (>= 1 for every T, C in {2, 4})
    const size_t nvec = aligned ? npix / VEC : 0;
    const uint4 *v4 = reinterpret_cast<const uint4 *>(src);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (size_t)gridDim.x * 256) {
        const uint4 q = ldg_stream(v4 + i);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
        if (sizeof(T) == 4) {
            const float f = __uint_as_float(*reinterpret_cast<const uint32_t *>(&first));
#pragma unroll
            for (int k = C - 1; k < 4; k += C) bad |= !(__uint_as_float(w[k]) == f);
        } else if (sizeof(T) == 2) {
            const uint32_t a = *reinterpret_cast<const uint16_t *>(&first);
            // C = 2: alpha is the high half of every word; C = 4: of words 1 and 3
#pragma unroll
            for (int k = (C == 2 ? 0 : 1); k < 4; k += (C == 2 ? 1 : 2)) bad |= (w[k] >> 16) != a;
        } else {
            const uint32_t a = *reinterpret_cast<const uint8_t *>(&first);
            const uint32_t mask = C == 2 ? 0xFF00FF00u : 0xFF000000u, pat = C == 2 ? (a << 24 | a << 8) : (a << 24);
#pragma unroll
            for (int k = 0; k < 4; ++k) bad |= ((w[k] ^ pat) & mask) != 0;
        }
    }
    // the pixels after the last whole vector (or everything, when the buffer is not 16-byte aligned)
    for (size_t i = nvec * VEC + (size_t)blockIdx.x * 256 + threadIdx.x; i < npix; i += (size_t)gridDim.x * 256)
        bad |= !(src[i * C + C - 1] == first);
    if (__any_sync(0xFFFFFFFFu, bad) && (threadIdx.x & 31) == 0) atomicOr(differs, 1);
}

// ---- premultiply / unpremultiply (own arithmetic, see the header) ---------------------------
template <typename T, bool UN> __device__ __forceinline__ T alpha_mul(T c, T a) {
    constexpr uint32_t mx = AMax<T>::v;
    if (!UN) return (T)(((uint32_t)c * a + mx / 2) / mx);
    if (a == 0) return c;
    return (T)min(((uint32_t)c * mx + a / 2) / a, mx);
}
template <> __device__ __forceinline__ float alpha_mul<float, false>(float c, float a) { return __fmul_rn(c, a); }
template <> __device__ __forceinline__ float alpha_mul<float, true>(float c, float a) { return a == 0.0f ? c : __fdiv_rn(c, a); }

template <typename T, int C, bool UN>
__global__ void __launch_bounds__(256)
alpha_mul_kernel(T *buf, size_t npix) {
    constexpr int VEC = 16 / (sizeof(T) * C);
    const bool aligned = ((uintptr_t)buf & 15) == 0;
    const size_t nvec = aligned ? npix / VEC : 0;
    uint4 *v4 = reinterpret_cast<uint4 *>(buf);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (size_t)gridDim.x * 256) {
        union { uint4 v; T e[VEC * C]; } u;
        u.v = v4[i];
#pragma unroll
        for (int p = 0; p < VEC; ++p) {
            const T a = u.e[p * C + C - 1];
#pragma unroll
            for (int c = 0; c < C - 1; ++c) u.e[p * C + c] = alpha_mul<T, UN>(u.e[p * C + c], a);
        }
        v4[i] = u.v;
    }
    for (size_t i = nvec * VEC + (size_t)blockIdx.x * 256 + threadIdx.x; i < npix; i += (size_t)gridDim.x * 256) {
        const T a = buf[i * C + C - 1];
#pragma unroll
        for (int c = 0; c < C - 1; ++c) buf[i * C + c] = alpha_mul<T, UN>(buf[i * C + c], a);
    }
}

unsigned grid_of(size_t n, int sm) { return (unsigned)std::max<size_t>(1, std::min<size_t>((n + 255) / 256, (size_t)sm * 16)); }

template <typename T>
void scan_t(const T *src, size_t npix, int C, int *flag, int sm, cudaStream_t st) {
    constexpr int VECB = 16 / sizeof(T);
    if (C == 2) alpha_scan_kernel<T, 2><<<grid_of(npix / (VECB / 2), sm), 256, 0, st>>>(src, npix, flag);
    else alpha_scan_kernel<T, 4><<<grid_of(npix / (VECB / 4), sm), 256, 0, st>>>(src, npix, flag);
}

template <typename T>
void mul_t(T *buf, size_t npix, int C, bool un, int sm, cudaStream_t st) {
    constexpr int VECB = 16 / sizeof(T);
    const unsigned g = grid_of(npix / (VECB / C), sm);
    if (C == 2) { if (un) alpha_mul_kernel<T, 2, true><<<g, 256, 0, st>>>(buf, npix); else alpha_mul_kernel<T, 2, false><<<g, 256, 0, st>>>(buf, npix); }
    else { if (un) alpha_mul_kernel<T, 4, true><<<g, 256, 0, st>>>(buf, npix); else alpha_mul_kernel<T, 4, false><<<g, 256, 0, st>>>(buf, npix); }
}

}  // namespace

// *d_flag (device int, zeroed here) becomes 1 when some alpha differs from the first pixel's
int alpha_scan_launch(const void *src, size_t npixels, int C, int dtype, int *d_flag, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    B200_CUDA(cudaMemsetAsync(d_flag, 0, sizeof(int), st));
    if ((C != 2 && C != 4) || npixels == 0) return B200IMG_OK;  // no alpha channel, or empty: constant
    switch (dtype) {
        case B200IMG_U8: scan_t((const uint8_t *)src, npixels, C, d_flag, ctx->sm_count, st); break;
        case B200IMG_U16: scan_t((const uint16_t *)src, npixels, C, d_flag, ctx->sm_count, st); break;
        default: scan_t((const float *)src, npixels, C, d_flag, ctx->sm_count, st); break;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

int alpha_mul_launch(void *buf, size_t npixels, int C, int dtype, bool un, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    if ((C != 2 && C != 4) || npixels == 0) return B200IMG_OK;  // resize.rs:105-129: Luma / Rgb variants are left alone
    switch (dtype) {
        case B200IMG_U8: mul_t((uint8_t *)buf, npixels, C, un, ctx->sm_count, st); break;
        case B200IMG_U16: mul_t((uint16_t *)buf, npixels, C, un, ctx->sm_count, st); break;
        default: mul_t((float *)buf, npixels, C, un, ctx->sm_count, st); break;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace b200
