// convert.cu -- pixel-format conversions and the BGR(A) swizzle.
//
// One persistent, HBM-bound streaming kernel: tiles of 1024 pixels are pulled
// into shared memory with TMA 1-D bulk copies (cp.async.bulk + mbarrier, two
// stages in flight), converted pixel-by-pixel out of shared memory, and pushed
// back with bulk stores.  The staging exists because 3-channel pixels are not
// 16-byte multiples: through shared memory every global access is a full-line
// bulk transfer whatever the pixel size.  Unaligned pointers and the ragged
// last tile take a cooperative element-wise path through the same buffers.
//
// Arithmetic follows (bit-exactly):
//   FromPrimitive                      src/color.rs:540-622
//   CicpRgb::cast_pixels (mode CAST)   src/metadata/cicp.rs:706-1033, fallback
//                                      782-887, expand/clamp 620-699, 1064-1108
//   FromColor (mode CONVERT)           src/color.rs:625-850
//   RgbToLuma                          src/primitive_sealed.rs:187-231
//   BgraSwizzle (mode SWIZZLE)         src/primitive_sealed.rs:34-67
#include "common.cuh"

namespace b200 {
namespace {

constexpr int TILE_PX = 1024;
constexpr int THREADS = 256;
enum { MODE_CAST = 0, MODE_CONVERT = 1, MODE_SWIZZLE = 2 };

// x / d for integer-valued x, d in {255, 65535}: q = x*r, then one FMA-exact
// residual correction.  Equal to the IEEE quotient for every u8 / u16 input
// (checked exhaustively by tests/test_convert.py through the kernel).
__device__ __forceinline__ float div_exact(float x, float d, float r) {
    float q = __fmul_rn(x, r);
    float e = __fmaf_rn(-q, d, x);
    return __fmaf_rn(e, r, q);
}

// ---- FromPrimitive, src/color.rs:540-622 ----
template <typename D, typename S> struct FromPrim;
template <typename T> struct FromPrim<T, T> { static __device__ __forceinline__ T f(T v) { return v; } };
template <> struct FromPrim<uint8_t, float> {
    static __device__ __forceinline__ uint8_t f(float v) {
        float x = fminf(v, 1.0f);  // NaN -> 1.0
        return (uint8_t)min(sat_u32_rz(__fadd_rn(__fmul_rn(x, 255.0f), 0.5f)), 255u);
    }
};
template <> struct FromPrim<uint16_t, float> {
    static __device__ __forceinline__ uint16_t f(float v) {
        float x = fminf(v, 1.0f);
        return (uint16_t)min(sat_u32_rz(__fadd_rn(__fmul_rn(x, 65535.0f), 0.5f)), 65535u);
    }
};
template <> struct FromPrim<uint8_t, uint16_t> {
    static __device__ __forceinline__ uint8_t f(uint16_t c) { return (uint8_t)(((uint32_t)c + 128u) / 257u); }
};
template <> struct FromPrim<float, uint16_t> {
    static __device__ __forceinline__ float f(uint16_t v) { return div_exact((float)v, 65535.0f, 1.0f / 65535.0f); }
};
template <> struct FromPrim<float, uint8_t> {
    static __device__ __forceinline__ float f(uint8_t v) { return div_exact((float)v, 255.0f, 1.0f / 255.0f); }
};
template <> struct FromPrim<uint16_t, uint8_t> {
    static __device__ __forceinline__ uint16_t f(uint8_t c) { return (uint16_t)(((uint32_t)c << 8) | c); }
};

template <typename T> struct MaxV;
template <> struct MaxV<uint8_t> { static __device__ __forceinline__ uint8_t v() { return 255; } };
template <> struct MaxV<uint16_t> { static __device__ __forceinline__ uint16_t v() { return 65535; } };
template <> struct MaxV<float> { static __device__ __forceinline__ float v() { return 1.0f; } };

// ---- ColorComponentForCicp, src/metadata/cicp.rs:1064-1108 ----
template <typename T> struct Cicp;
template <> struct Cicp<uint8_t> {
    static __device__ __forceinline__ float expand(uint8_t v) { return __fmul_rn((float)v, 1.0f / 255.0f); }
    static __device__ __forceinline__ uint8_t clamp(float v) { return (uint8_t)min(sat_u32_rz(__fadd_rn(__fmul_rn(v, 255.0f), 0.5f)), 255u); }
};
template <> struct Cicp<uint16_t> {
    static __device__ __forceinline__ float expand(uint16_t v) { return __fmul_rn((float)v, 1.0f / 65535.0f); }
    static __device__ __forceinline__ uint16_t clamp(float v) { return (uint16_t)min(sat_u32_rz(__fadd_rn(__fmul_rn(v, 65535.0f), 0.5f)), 65535u); }
};
template <> struct Cicp<float> {
    static __device__ __forceinline__ float expand(float v) { return v; }
    static __device__ __forceinline__ float clamp(float v) { return v; }
};

// ---- RgbToLuma, src/primitive_sealed.rs:187-231 ----
__device__ __forceinline__ uint8_t rgb_to_luma(uint8_t r, uint8_t g, uint8_t b) {
    // ceil(2^24 * {2126, 7152, 722} / 10000)
    return (uint8_t)(((uint32_t)r * 3566837u + (uint32_t)g * 11999065u + (uint32_t)b * 1211315u + 0x800000u) >> 24);
}
__device__ __forceinline__ uint16_t rgb_to_luma(uint16_t r, uint16_t g, uint16_t b) {
    return (uint16_t)((2126u * (uint32_t)r + 7152u * (uint32_t)g + 722u * (uint32_t)b) / 10000u);
}
__device__ __forceinline__ float rgb_to_luma(float r, float g, float b) {
    const float SR = 2126.0f / 10000.0f, SG = 7152.0f / 10000.0f, SB = 722.0f / 10000.0f;
    return __fadd_rn(__fadd_rn(__fmul_rn(SR, r), __fmul_rn(SG, g)), __fmul_rn(SB, b));
}

struct Coefs { float r, g, b; int fused_mac; };

template <int C> struct Layout {
    static constexpr bool luma = C <= 2;
    static constexpr bool alpha = (C == 2 || C == 4);
    static constexpr int colour = C <= 2 ? 1 : 3;
};

template <typename S, typename D, int CIN, int COUT, int MODE>
__device__ __forceinline__ void convert_px(const S (&s)[CIN], D (&d)[COUT], const Coefs &k) {
    using LI = Layout<CIN>;
    using LO = Layout<COUT>;
    if constexpr (MODE == MODE_SWIZZLE) {
#pragma unroll
        for (int c = 0; c < COUT; ++c) d[c] = (D)s[c == 0 ? 2 : (c == 2 ? 0 : c)];
    } else if constexpr (LI::luma == LO::luma) {
        // same family: per-channel FromPrimitive, alpha mapped or DEFAULT_MAX_VALUE
#pragma unroll
        for (int c = 0; c < LO::colour; ++c) d[c] = FromPrim<D, S>::f(s[c]);
        if constexpr (LO::alpha) {
            if constexpr (LI::alpha) d[LO::colour] = FromPrim<D, S>::f(s[LI::colour]);
            else d[LO::colour] = MaxV<D>::v();
        }
    } else if constexpr (MODE == MODE_CONVERT) {
        if constexpr (LI::luma) {  // Luma(A) -> Rgb(a): broadcast
            D g = FromPrim<D, S>::f(s[0]);
            d[0] = d[1] = d[2] = g;
            if constexpr (LO::alpha) {
                if constexpr (LI::alpha) d[3] = FromPrim<D, S>::f(s[1]);
                else d[3] = MaxV<D>::v();
            }
        } else {  // Rgb(a) -> Luma(A): luma in the source type first
            d[0] = FromPrim<D, S>::f(rgb_to_luma(s[0], s[1], s[2]));
            if constexpr (LO::alpha) {
                if constexpr (LI::alpha) d[1] = FromPrim<D, S>::f(s[3]);
                else d[1] = MaxV<D>::v();
            }
        }
    } else {
        // cast_pixels_by_fallback: through f32 RGBA
        float rgba[4];
        if constexpr (LI::luma) {
            float l = Cicp<S>::expand(s[0]);
            rgba[0] = rgba[1] = rgba[2] = l;
            if constexpr (LI::alpha) rgba[3] = Cicp<S>::expand(s[1]);
            else rgba[3] = 1.0f;
        } else {
            rgba[0] = Cicp<S>::expand(s[0]);
            rgba[1] = Cicp<S>::expand(s[1]);
            rgba[2] = Cicp<S>::expand(s[2]);
            if constexpr (LI::alpha) rgba[3] = Cicp<S>::expand(s[3]);
            else rgba[3] = 1.0f;
        }
        if constexpr (LO::luma) {
            float luma;
            if (k.fused_mac) {
                luma = __fmaf_rn(rgba[0], k.r, 0.0f);
                luma = __fmaf_rn(rgba[1], k.g, luma);
                luma = __fmaf_rn(rgba[2], k.b, luma);
            } else {
                luma = __fadd_rn(0.0f, __fmul_rn(rgba[0], k.r));
                luma = __fadd_rn(luma, __fmul_rn(rgba[1], k.g));
                luma = __fadd_rn(luma, __fmul_rn(rgba[2], k.b));
            }
            d[0] = Cicp<D>::clamp(luma);
            if constexpr (LO::alpha) d[1] = Cicp<D>::clamp(rgba[3]);
        } else {
#pragma unroll
            for (int c = 0; c < COUT; ++c) d[c] = Cicp<D>::clamp(rgba[c]);
        }
    }
}

// pixel <-> shared memory with the widest aligned access the pixel size allows
template <typename T, int C>
__device__ __forceinline__ void lds_px(const T *p, T (&v)[C]) {
    constexpr int B = sizeof(T) * C;
    if (B == 16) { *reinterpret_cast<uint4 *>(v) = *reinterpret_cast<const uint4 *>(p); }
    else if (B == 8) { *reinterpret_cast<uint2 *>(v) = *reinterpret_cast<const uint2 *>(p); }
    else if (B == 4) { *reinterpret_cast<uint32_t *>(v) = *reinterpret_cast<const uint32_t *>(p); }
    else if (B == 2) { *reinterpret_cast<uint16_t *>(v) = *reinterpret_cast<const uint16_t *>(p); }
    else {
#pragma unroll
        for (int c = 0; c < C; ++c) v[c] = p[c];
    }
}
template <typename T, int C>
__device__ __forceinline__ void sts_px(T *p, const T (&v)[C]) {
    constexpr int B = sizeof(T) * C;
    if (B == 16) { *reinterpret_cast<uint4 *>(p) = *reinterpret_cast<const uint4 *>(v); }
    else if (B == 8) { *reinterpret_cast<uint2 *>(p) = *reinterpret_cast<const uint2 *>(v); }
    else if (B == 4) { *reinterpret_cast<uint32_t *>(p) = *reinterpret_cast<const uint32_t *>(v); }
    else if (B == 2) { *reinterpret_cast<uint16_t *>(p) = *reinterpret_cast<const uint16_t *>(v); }
    else {
#pragma unroll
        for (int c = 0; c < C; ++c) p[c] = v[c];
    }
}

template <typename S, typename D, int CIN, int COUT, int MODE>
__global__ void __launch_bounds__(THREADS)
convert_kernel(const S *src, D *dst, size_t npixels, Coefs k, int bulk_ok) {
    constexpr uint32_t IN_BYTES = TILE_PX * CIN * sizeof(S);
    constexpr uint32_t OUT_BYTES = TILE_PX * COUT * sizeof(D);
    extern __shared__ __align__(128) unsigned char smem[];
    S *s_in[2] = {reinterpret_cast<S *>(smem), reinterpret_cast<S *>(smem + IN_BYTES)};
    D *s_out[2] = {reinterpret_cast<D *>(smem + 2 * IN_BYTES), reinterpret_cast<D *>(smem + 2 * IN_BYTES + OUT_BYTES)};
    __shared__ uint64_t full[2];

    const size_t n_tiles = (npixels + TILE_PX - 1) / TILE_PX;
    const size_t n_full = npixels / TILE_PX;  // tiles eligible for bulk copies
    const int tid = threadIdx.x;

    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        fence_mbar_init();
    }
    __syncthreads();

    // bulk prefetch of tile t into stage st (thread 0 only; full tiles only)
    auto prefetch = [&](size_t t, int st) {
        if (bulk_ok && t < n_full) {
            mbar_expect_tx(&full[st], IN_BYTES);
            bulk_g2s(s_in[st], src + t * (size_t)TILE_PX * CIN, IN_BYTES, &full[st]);
        }
    };

    size_t tile = blockIdx.x;
    if (tid == 0 && tile < n_tiles) prefetch(tile, 0);
    uint32_t phase[2] = {0, 0};
    int it = 0;
    for (; tile < n_tiles; tile += gridDim.x, ++it) {
        const int st = it & 1;
        const size_t next = tile + gridDim.x;
        if (tid == 0 && next < n_tiles) prefetch(next, st ^ 1);  // stage st^1 was released by the barrier ending the previous iteration

        const bool bulk = bulk_ok && tile < n_full;
        const int px_here = (int)min((size_t)TILE_PX, npixels - tile * TILE_PX);
        if (bulk) {
            mbar_wait(&full[st], phase[st]);
            phase[st] ^= 1;
        } else {
            const S *g = src + tile * (size_t)TILE_PX * CIN;
            for (int i = tid; i < px_here * CIN; i += THREADS) s_in[st][i] = g[i];
        }
        // the bulk store that last used s_out[st] (two tiles ago) must have drained
        if (tid == 0) bulk_wait_read<1>();
        __syncthreads();

#pragma unroll
        for (int j = 0; j < TILE_PX / THREADS; ++j) {
            const int p = tid + j * THREADS;
            if (p < px_here) {
                S sv[CIN];
                D dv[COUT];
                lds_px<S, CIN>(s_in[st] + p * CIN, sv);
                convert_px<S, D, CIN, COUT, MODE>(sv, dv, k);
                sts_px<D, COUT>(s_out[st] + p * COUT, dv);
            }
        }
        if (bulk) fence_async_smem();
        __syncthreads();

        D *g = dst + tile * (size_t)TILE_PX * COUT;
        if (bulk) {
            if (tid == 0) {
                bulk_s2g(g, s_out[st], OUT_BYTES);
                bulk_commit();
            }
        } else {
            for (int i = tid; i < px_here * COUT; i += THREADS) g[i] = s_out[st][i];
        }
    }
    if (tid == 0) bulk_wait<0>();
}

// The same conversion straight from / to global memory: a warp takes 128 consecutive pixels per trip, lane L the pixels
// L, L + 32, L + 64, L + 96 (so every access of the warp is one contiguous run), all loads of a trip first.  Pixels of 2, 4, 8
// or 16 bytes move as one access; 3-, 6- and 12-byte pixels element by element (the three accesses of a warp cover the same
// lines, L1 and L2 merge them).
template <typename S, typename D, int CIN, int COUT, int MODE>
__global__ void __launch_bounds__(256)
convert_direct_kernel(const S *__restrict__ src, D *__restrict__ dst, size_t npixels, Coefs k) {
    const int lane = threadIdx.x & 31;
    const size_t warps = (size_t)gridDim.x * 8;
    for (size_t base = ((size_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 128; base < npixels; base += warps * 128) {
        S sv[4][CIN];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const size_t p = base + lane + 32 * j;
            if (p < npixels) lds_px<S, CIN>(src + p * CIN, sv[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const size_t p = base + lane + 32 * j;
            if (p < npixels) {
                D dv[COUT];
                convert_px<S, D, CIN, COUT, MODE>(sv[j], dv, k);
                sts_px<D, COUT>(dst + p * COUT, dv);
            }
        }
    }
}

// Rgb8 sources with narrow destination pixels (to_luma8, to_luma_alpha8, to_rgba8, Luma16 / La8 ...): byte loads of 3-byte
// pixels are what holds the two kernels above back on these, so a thread takes 48 bytes = 16 pixels as three 128-bit loads,
// unpacks them in registers and writes its 16 destination pixels (1, 2 or 4 bytes each) as one, two or four 128-bit stores.
template <typename D, int COUT> __device__ __forceinline__ uint32_t px_bits(const D (&d)[COUT]) {
    uint32_t v = 0;
    if constexpr (sizeof(D) == 4) {
        v = __float_as_uint((float)d[0]);
    } else {
#pragma unroll
        for (int c = 0; c < COUT; ++c) v |= (uint32_t)d[c] << (8 * sizeof(D) * c);
    }
    return v;
}

template <typename D, int COUT, int MODE>
__global__ void __launch_bounds__(256)
convert_rgb8_chunks_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, size_t nchunks, Coefs k) {
    constexpr int BO = sizeof(D) * COUT;   // 1, 2 or 4
    const size_t stride = (size_t)gridDim.x * 256;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nchunks; i += stride) {
        const uint4 a = src[3 * i], b = src[3 * i + 1], c = src[3 * i + 2];
        const uint32_t w[12] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w, c.x, c.y, c.z, c.w};
        uint32_t o[16];
#pragma unroll
        for (int p = 0; p < 16; ++p) {
            uint8_t sv[3];
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const int byte = 3 * p + ch;
                sv[ch] = (uint8_t)(w[byte >> 2] >> (8 * (byte & 3)));
            }
            D dv[COUT];
            convert_px<uint8_t, D, 3, COUT, MODE>(sv, dv, k);
            o[p] = px_bits<D, COUT>(dv);
        }
        uint4 *out = dst + (size_t)BO * i;
        if constexpr (BO == 4) {
#pragma unroll
            for (int j = 0; j < 4; ++j) out[j] = make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        } else if constexpr (BO == 2) {
#pragma unroll
            for (int j = 0; j < 2; ++j)
                out[j] = make_uint4(o[8 * j] | o[8 * j + 1] << 16, o[8 * j + 2] | o[8 * j + 3] << 16, o[8 * j + 4] | o[8 * j + 5] << 16,
                                    o[8 * j + 6] | o[8 * j + 7] << 16);
        } else {
            out[0] = make_uint4(o[0] | o[1] << 8 | o[2] << 16 | o[3] << 24, o[4] | o[5] << 8 | o[6] << 16 | o[7] << 24,
                                o[8] | o[9] << 8 | o[10] << 16 | o[11] << 24, o[12] | o[13] << 8 | o[14] << 16 | o[15] << 24);
        }
    }
}

// Sub-word source pixels (3 or 6 bytes) with wide destination pixels (Rgb8 -> Rgba32F, Rgb16 -> Rgba16 ...): the source bytes of
// a warp's 128 pixels go through a warp-private, double-buffered slice of shared memory (coalesced 4-byte cp.async, no CTA
// barrier anywhere), every lane picks its pixels L, L + 32, L + 64, L + 96 out of it, converts, and stores them straight from
// registers: contiguous runs per warp, no staging of the (much larger) destination.
template <typename S, typename D, int CIN, int COUT, int MODE>
__global__ void __launch_bounds__(256)
convert_staged_in_kernel(const S *__restrict__ src, D *__restrict__ dst, size_t ngroups, Coefs k) {   // a group = 128 pixels
    constexpr int BI = sizeof(S) * CIN, GROUP_BYTES = 128 * BI, GROUP_WORDS = GROUP_BYTES / 4;   // 96 or 192 words
    __shared__ __align__(16) uint32_t s_buf[8][2][GROUP_WORDS];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const size_t warps = (size_t)gridDim.x * 8;
    size_t g = (size_t)blockIdx.x * 8 + wib;
    auto fetch = [&](size_t grp, int st) {
        const uint32_t *gw = reinterpret_cast<const uint32_t *>(reinterpret_cast<const unsigned char *>(src) + grp * GROUP_BYTES);
        const uint32_t d = (uint32_t)__cvta_generic_to_shared(&s_buf[wib][st][0]);
#pragma unroll
        for (int j = 0; j < GROUP_WORDS / 32; ++j)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d + 4u * (lane + 32 * j)), "l"(gw + lane + 32 * j) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (g < ngroups) fetch(g, 0);
    for (int it = 0; g < ngroups; g += warps, ++it) {
        const int st = it & 1;
        if (g + warps < ngroups) {
            fetch(g + warps, st ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncwarp();
        const S *sp = reinterpret_cast<const S *>(&s_buf[wib][st][0]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int q = lane + 32 * j;
            S sv[CIN];
#pragma unroll
            for (int c = 0; c < CIN; ++c) sv[c] = sp[q * CIN + c];
            D dv[COUT];
            convert_px<S, D, CIN, COUT, MODE>(sv, dv, k);
            sts_px<D, COUT>(dst + (g * 128 + q) * COUT, dv);
        }
        __syncwarp();   // the slice is free for the fetch after next
    }
}

template <typename S, typename D, int CIN, int COUT, int MODE>
int launch_direct(const void *src, void *dst, size_t npixels, const Coefs &k, cudaStream_t stream) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    const size_t want = (npixels + 1023) / 1024;
    const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)ctx->sm_count * 8);
    convert_direct_kernel<S, D, CIN, COUT, MODE><<<grid, 256, 0, stream>>>((const S *)src, (D *)dst, npixels, k);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename S, typename D, int CIN, int COUT, int MODE>
int launch(const void *src, void *dst, size_t npixels, const Coefs &k, cudaStream_t stream) {
    {
        // Which kernel: the direct one wherever the source pixel is one access (1 / 2 / 4 / 8 / 16 bytes, or f32 elements) or the
        // destination pixel is narrow; sub-word source pixels (Rgb8, Rgb16, La16 ...) feeding wide destination pixels stay on the
        // tile kernel, whose bulk copies read them better than byte loads do (8192^2, share of the HBM rate, tile / direct:
        // Rgba16 -> Rgb8 0.80 / 0.98, Rgba8 -> Rgb8 0.60 / 0.88, Rgba32F -> Rgba8 0.91 / 0.94, Luma8 -> Rgba8 0.44 / 0.66,
        // Rgb8 -> Rgba32F 0.82 / 0.81, Rgb8 -> Rgba8 0.62 / 0.57, Rgb16 -> Rgba16 0.80 / 0.78; tools/time_convert.py).
        // B200IMG_CONVERT_KERNEL = tile | direct forces one (tests run both); the direct kernel's vector accesses need
        // pixel-aligned buffers.
        constexpr size_t BI = sizeof(S) * CIN, BO = sizeof(D) * COUT;
        constexpr bool VI = BI == 16 || BI == 8 || BI == 4 || BI == 2 || BI == 1, VO = BO == 16 || BO == 8 || BO == 4 || BO == 2 || BO == 1;
        constexpr size_t AI = VI ? BI : sizeof(S), AO = VO ? BO : sizeof(D);
        const char *force = getenv("B200IMG_CONVERT_KERNEL");
        if constexpr (sizeof(S) == 1 && CIN == 3 && MODE != MODE_SWIZZLE && (BO == 1 || BO == 2 || BO == 4)) {
            // Rgb8 -> 1 / 2 / 4-byte pixels: whole 16-pixel chunks by the chunk kernel, the last pixels by the kernels below
            if (!force && npixels >= 4096 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0) {
                DeviceCtx *ctx = current_ctx();
                if (!ctx) return B200IMG_ERR_CUDA;
                const size_t chunks = npixels / 16;
                const unsigned grid = (unsigned)std::min<size_t>((chunks + 255) / 256, (size_t)ctx->sm_count * 16);
                convert_rgb8_chunks_kernel<D, COUT, MODE><<<grid, 256, 0, stream>>>((const uint4 *)src, (uint4 *)dst, chunks, k);
                B200_CUDA(cudaGetLastError());
                if (chunks * 16 == npixels) return B200IMG_OK;
                src = reinterpret_cast<const unsigned char *>(src) + chunks * 48;
                dst = reinterpret_cast<unsigned char *>(dst) + chunks * 16 * BO;
                npixels -= chunks * 16;
            }
        }
        if constexpr ((BI == 3 || BI == 6) && BO >= 4 && MODE != MODE_SWIZZLE) {
            constexpr size_t AOs = VO ? BO : sizeof(D);
            const bool want = force ? force[0] == 's' : true;
            if (want && npixels >= 4096 && (reinterpret_cast<uintptr_t>(src) & 3) == 0 && reinterpret_cast<uintptr_t>(dst) % AOs == 0) {
                DeviceCtx *ctx = current_ctx();
                if (!ctx) return B200IMG_ERR_CUDA;
                const size_t groups = npixels / 128;
                const unsigned grid = (unsigned)std::min<size_t>((groups + 7) / 8, (size_t)ctx->sm_count * 8);
                convert_staged_in_kernel<S, D, CIN, COUT, MODE><<<grid, 256, 0, stream>>>((const S *)src, (D *)dst, groups, k);
                B200_CUDA(cudaGetLastError());
                if (groups * 128 == npixels) return B200IMG_OK;
                src = reinterpret_cast<const unsigned char *>(src) + groups * 128 * BI;
                dst = reinterpret_cast<unsigned char *>(dst) + groups * 128 * BO;
                npixels -= groups * 128;
            }
        }
        bool direct = VI || sizeof(S) == 4 || BO < 4;
        if (force && force[0] == 't') direct = false;
        if (force && force[0] == 'd') direct = true;
        if (direct && MODE != MODE_SWIZZLE && reinterpret_cast<uintptr_t>(src) % AI == 0 && reinterpret_cast<uintptr_t>(dst) % AO == 0)
            return launch_direct<S, D, CIN, COUT, MODE>(src, dst, npixels, k, stream);
    }
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    auto kern = convert_kernel<S, D, CIN, COUT, MODE>;
    constexpr size_t smem = 2 * (size_t)TILE_PX * (CIN * sizeof(S) + COUT * sizeof(D));
    static thread_local int blocks_per_sm[64] = {0};
    int &bps = blocks_per_sm[ctx->device & 63];
    if (bps == 0) {
        B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, kern, THREADS, smem));
        if (bps < 1) bps = 1;
    }
    const size_t n_tiles = (npixels + TILE_PX - 1) / TILE_PX;
    const size_t grid = n_tiles < (size_t)ctx->sm_count * bps ? n_tiles : (size_t)ctx->sm_count * bps;
    const int bulk_ok = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
    kern<<<(unsigned)grid, THREADS, smem, stream>>>((const S *)src, (D *)dst, npixels, k, bulk_ok);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename S, typename D, int CIN, int COUT>
int launch_mode(const void *src, void *dst, size_t n, int mode, const Coefs &k, cudaStream_t st) {
    if (Layout<CIN>::luma == Layout<COUT>::luma || mode == B200IMG_CAST)
        return launch<S, D, CIN, COUT, MODE_CAST>(src, dst, n, k, st);
    return launch<S, D, CIN, COUT, MODE_CONVERT>(src, dst, n, k, st);
}

template <typename S, typename D, int CIN>
int launch_cout(const void *src, void *dst, int cout, size_t n, int mode, const Coefs &k, cudaStream_t st) {
    switch (cout) {
        case 1: return launch_mode<S, D, CIN, 1>(src, dst, n, mode, k, st);
        case 2: return launch_mode<S, D, CIN, 2>(src, dst, n, mode, k, st);
        case 3: return launch_mode<S, D, CIN, 3>(src, dst, n, mode, k, st);
        case 4: return launch_mode<S, D, CIN, 4>(src, dst, n, mode, k, st);
    }
    return fail(B200IMG_ERR_BAD_ARG, "convert: dst_channels must be 1..4");
}

template <typename S, typename D>
int launch_cin(const void *src, int cin, void *dst, int cout, size_t n, int mode, const Coefs &k, cudaStream_t st) {
    switch (cin) {
        case 1: return launch_cout<S, D, 1>(src, dst, cout, n, mode, k, st);
        case 2: return launch_cout<S, D, 2>(src, dst, cout, n, mode, k, st);
        case 3: return launch_cout<S, D, 3>(src, dst, cout, n, mode, k, st);
        case 4: return launch_cout<S, D, 4>(src, dst, cout, n, mode, k, st);
    }
    return fail(B200IMG_ERR_BAD_ARG, "convert: src_channels must be 1..4");
}

template <typename S>
int launch_dst(const void *src, int cin, void *dst, int dd, int cout, size_t n, int mode, const Coefs &k, cudaStream_t st) {
    switch (dd) {
        case B200IMG_U8: return launch_cin<S, uint8_t>(src, cin, dst, cout, n, mode, k, st);
        case B200IMG_U16: return launch_cin<S, uint16_t>(src, cin, dst, cout, n, mode, k, st);
        case B200IMG_F32: return launch_cin<S, float>(src, cin, dst, cout, n, mode, k, st);
    }
    return fail(B200IMG_ERR_BAD_ARG, "convert: bad dst_dtype");
}

}  // namespace

int convert_dev(const void *src, int sd, int cin, void *dst, int dd, int cout, size_t n, int mode,
                const float coefs[3], int fused_mac, cudaStream_t st) {
    if (mode != B200IMG_CAST && mode != B200IMG_CONVERT) return fail(B200IMG_ERR_BAD_ARG, "convert: bad mode %d", mode);
    if (n == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "convert: null buffer");
    Coefs k = {0.2126f, 0.7152f, 0.0722f, fused_mac};  // src/images/buffer.rs:1743
    if (coefs) { k.r = coefs[0]; k.g = coefs[1]; k.b = coefs[2]; }
    switch (sd) {
        case B200IMG_U8: return launch_dst<uint8_t>(src, cin, dst, dd, cout, n, mode, k, st);
        case B200IMG_U16: return launch_dst<uint16_t>(src, cin, dst, dd, cout, n, mode, k, st);
        case B200IMG_F32: return launch_dst<float>(src, cin, dst, dd, cout, n, mode, k, st);
    }
    return fail(B200IMG_ERR_BAD_ARG, "convert: bad src_dtype");
}

// Four-channel swizzle in place (bgra_to_rgba, src/primitive_sealed.rs:34-67): a pixel never straddles a 16-byte vector, so
// nothing needs staging: 128-bit loads (four per thread in flight), a byte / lane permute in registers, 128-bit stores to the
// same addresses.  (The tile kernel above moves the same bytes through shared memory with one tile of prefetch: 71 % of the
// HBM rate on 8192^2 BGRA8 against this kernel's figure in DESIGN.md.)
template <int ESZ> __device__ __forceinline__ uint4 swizzle4_vec(uint4 v) {
    if constexpr (ESZ == 1) {          // four pixels: bytes (b, g, r, a) -> (r, g, b, a) in every word
        v.x = __byte_perm(v.x, 0, 0x3012); v.y = __byte_perm(v.y, 0, 0x3012);
        v.z = __byte_perm(v.z, 0, 0x3012); v.w = __byte_perm(v.w, 0, 0x3012);
        return v;
    } else if constexpr (ESZ == 2) {   // two pixels: halves (c0 c1 | c2 c3) -> (c2 c1 | c0 c3)
        return make_uint4(__byte_perm(v.x, v.y, 0x3254), __byte_perm(v.x, v.y, 0x7610), __byte_perm(v.z, v.w, 0x3254), __byte_perm(v.z, v.w, 0x7610));
    } else {                           // one pixel
        return make_uint4(v.z, v.y, v.x, v.w);
    }
}

template <int ESZ>
__global__ void __launch_bounds__(256)
swizzle4_kernel(uint4 *buf, size_t nvec) {
    const size_t stride = (size_t)gridDim.x * 256;
    size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        const uint4 a = buf[i], b = buf[i + stride], c = buf[i + 2 * stride], d = buf[i + 3 * stride];
        buf[i] = swizzle4_vec<ESZ>(a);
        buf[i + stride] = swizzle4_vec<ESZ>(b);
        buf[i + 2 * stride] = swizzle4_vec<ESZ>(c);
        buf[i + 3 * stride] = swizzle4_vec<ESZ>(d);
    }
    for (; i < nvec; i += stride) buf[i] = swizzle4_vec<ESZ>(buf[i]);
}

// Three channels: 48 bytes (three vectors) hold a whole number of pixels for every element size (16 / 8 / 4 pixels), so a
// thread takes three consecutive vectors and permutes inside them; the byte patterns repeat every three words.
template <int ESZ> __device__ __forceinline__ void swizzle3_words(uint32_t &w0, uint32_t &w1, uint32_t &w2) {
    if constexpr (ESZ == 1) {          // bytes b g r b | g r b g | r b g r  ->  r g b r | g b r g | b r g b
        const uint32_t n0 = __byte_perm(w0, w1, 0x5012);
        const uint32_t t = __byte_perm(w0, w1, 0x7034);
        const uint32_t n1 = __byte_perm(t, w2, 0x3410);
        const uint32_t n2 = __byte_perm(w1, w2, 0x5672);
        w0 = n0; w1 = n1; w2 = n2;
    } else if constexpr (ESZ == 2) {   // halves c0 c1 | c2 c0' | c1' c2'  ->  c2 c1 | c0 c2' | c1' c0'
        const uint32_t n0 = __byte_perm(w0, w1, 0x3254), n1 = __byte_perm(w0, w2, 0x7610), n2 = __byte_perm(w2, w1, 0x7610);
        w0 = n0; w1 = n1; w2 = n2;
    } else {
        const uint32_t t = w0; w0 = w2; w2 = t;
    }
}

template <int ESZ>
__global__ void __launch_bounds__(256)
swizzle3_kernel(uint4 *buf, size_t nchunks) {   // a chunk = three vectors
    const size_t stride = (size_t)gridDim.x * 256;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < nchunks; i += stride) {
        uint4 a = buf[3 * i], b = buf[3 * i + 1], c = buf[3 * i + 2];
        swizzle3_words<ESZ>(a.x, a.y, a.z);
        swizzle3_words<ESZ>(a.w, b.x, b.y);
        swizzle3_words<ESZ>(b.z, b.w, c.x);
        swizzle3_words<ESZ>(c.y, c.z, c.w);
        buf[3 * i] = a; buf[3 * i + 1] = b; buf[3 * i + 2] = c;
    }
}

template <int ESZ>
int swizzle3_launch(void *buf, size_t nchunks, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    if (nchunks == 0) return B200IMG_OK;
    const unsigned grid = (unsigned)std::min<size_t>((nchunks + 255) / 256, (size_t)ctx->sm_count * 16);
    swizzle3_kernel<ESZ><<<grid, 256, 0, st>>>(reinterpret_cast<uint4 *>(buf), nchunks);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <int ESZ>
int swizzle4_launch(void *buf, size_t npixels, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    constexpr size_t PX_PER_VEC = 16 / (4 * ESZ);
    const size_t nvec = npixels / PX_PER_VEC;     // (npixels % PX_PER_VEC pixels are left to the caller)
    if (nvec == 0) return B200IMG_OK;
    const size_t want = (nvec + 4 * 256 - 1) / (4 * 256);
    const unsigned grid = (unsigned)std::min<size_t>(want, (size_t)ctx->sm_count * 8);
    swizzle4_kernel<ESZ><<<grid, 256, 0, st>>>(reinterpret_cast<uint4 *>(buf), nvec);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

int swizzle_dev(void *buf, int dtype, int channels, size_t nsub, cudaStream_t st) {
    if (channels != 3 && channels != 4) return fail(B200IMG_ERR_BAD_ARG, "swizzle: channels must be 3 or 4");
    size_t n = nsub / (size_t)channels;  // full chunks only (as_chunks_mut)
    if (n == 0) return B200IMG_OK;
    if (!buf) return fail(B200IMG_ERR_BAD_ARG, "swizzle: null buffer");
    Coefs k = {0, 0, 0, 0};
    if (channels == 4 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0 && n >= 4096) {
        // whole 16-byte vectors by the direct kernel, the last one to three pixels (u8 / u16 only) by the tile kernel
        const size_t esz = dtype_size(dtype), per_vec = 16 / (4 * esz), body = n / per_vec * per_vec;
        int rc = dtype == B200IMG_U8 ? swizzle4_launch<1>(buf, body, st) : dtype == B200IMG_U16 ? swizzle4_launch<2>(buf, body, st)
                                                                                                  : swizzle4_launch<4>(buf, body, st);
        if (rc || body == n) return rc;
        buf = reinterpret_cast<unsigned char *>(buf) + body * 4 * esz;
        n -= body;
    }
    if (channels == 3 && (reinterpret_cast<uintptr_t>(buf) & 15) == 0 && n >= 4096) {
        const size_t esz = dtype_size(dtype), per_chunk = 48 / (3 * esz), chunks = n / per_chunk, body = chunks * per_chunk;
        int rc = dtype == B200IMG_U8 ? swizzle3_launch<1>(buf, chunks, st) : dtype == B200IMG_U16 ? swizzle3_launch<2>(buf, chunks, st)
                                                                                                  : swizzle3_launch<4>(buf, chunks, st);
        if (rc || body == n) return rc;
        buf = reinterpret_cast<unsigned char *>(buf) + body * 3 * esz;
        n -= body;
    }
#define SWZ(T, C) launch<T, T, C, C, MODE_SWIZZLE>(buf, buf, n, k, st)
    switch (dtype) {
        case B200IMG_U8: return channels == 3 ? SWZ(uint8_t, 3) : SWZ(uint8_t, 4);
        case B200IMG_U16: return channels == 3 ? SWZ(uint16_t, 3) : SWZ(uint16_t, 4);
        case B200IMG_F32: return channels == 3 ? SWZ(float, 3) : SWZ(float, 4);
    }
#undef SWZ
    return fail(B200IMG_ERR_BAD_ARG, "swizzle: bad dtype");
}

}  // namespace b200
