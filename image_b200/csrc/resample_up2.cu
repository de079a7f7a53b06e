// resample_up2.cu -- the resampler for exact 2x upscales of four-channel images under B200IMG_FP_FAST:
// BASELINE config 4, the 32768^2 Rgba32F Lanczos3 upscale and its row bands.  Written out below for f32
// pixels and the support-3 filters (Lanczos3, Gaussian: seven taps); CatmullRom (five) and Triangle (three)
// are the same kernel with a shorter window, Rgba8 / Rgba16 the same kernel with an integer load in front
// (2^23 mantissa conversion) and the reference's rounding store behind it.
//
// At ratio 1/2 destination samples 2m and 2m+1 both tap the SAME seven source samples m-3 .. m+3
// (vertical_sample / horizontal_sample, src/imageops/sample.rs:494-562, 217-340: centre m + 0.25 resp.
// m + 0.75, support 3, so left = m - 3 and right = m + 4 for both), each with its own weight vector
// (one of whose ends is a zero of the filter -- kept, as the reference keeps it).  So:
//   vertical:   a thread owns one source pixel column (a float4) and streams down the source rows with a
//               window of seven rows in registers.  The row loop is unrolled over 14 rows, so the window is a
//               circular register file indexed at compile time (no moves); each step is one 128-bit load
//               issued seven rows ahead, 28 FFMA2 with weights in uniform registers, and two 128-bit stores
//               of the destination-row pair into a ring in shared memory.
//   horizontal: every seven steps (14 ring rows) the warps share out (pair of ring rows, 32 source positions)
//               items: lane = source position, fourteen 128-bit loads of neighbouring ring pixels, 56 FFMA2 in
//               eight independent chains for the destination pixels 2x and 2x+1 of both rows, stored as 32
//               contiguous bytes per lane and row (1 KB per warp and row).
// The ring is double buffered, so a phase costs one CTA barrier.  The job is bound by its stores (64 bytes
// written per 16 read); the arithmetic is ~160 thread instructions per source pixel.
// Borders: samples outside the image count as zero and the sum is multiplied by 1 / (weights inside), the
// reference's renormalisation of the clipped window up to f32 rounding (FAST contract: 1e-5 relative).  The
// accumulator's start value acc_init (sample.rs:547) is added once after the vertical pass.
#include "common.cuh"
#include "host_tables.h"
#include "packed.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200 {
namespace {

constexpr int U2_MAXT = 7;       // taps, at most: 2 * support + 1 (the kernel is a template on the tap count: 3, 5 or 7)
constexpr int U2_THREADS = 128;  // source pixel columns of a strip, halo included

struct Up2Args {
    const void *src;
    void *dst;
    size_t src_img_stride, dst_img_stride;  // bytes
    uint32_t w, h, nw, nh;
    uint32_t src_row0, dst_row0, dst_rows;
    int TWS;            // source pixels per strip (<= U2_THREADS - 6)
    int BANDS;          // source rows (= destination row pairs) per band, a multiple of 7
    int pitch;          // ring pixels per row (odd)
    float acc_init;
    f32x2 wv[2][U2_MAXT], wh[2][U2_MAXT];  // (w, w) pairs; [phase][tap]
};

__device__ __forceinline__ float up2_lo(f32x2 v) { float a, b; unpack2(v, a, b); return a; }

// 1 / (sum of the taps of the window [first, first + T) that fall inside [0, len)); only for clipped windows
__device__ __noinline__ float up2_border_scale(const f32x2 *w, int T, int first, int len) {
    const int k0 = max(0, -first), k1 = min(T, len - first);
    float s = 0.0f;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) s += up2_lo(w[k]);
    return 1.0f / s;
}

// One source pixel (four channels of E) as two packed pairs of floats (r, g) (b, a); integers by the 2^23 mantissa trick.
template <typename E> __device__ __forceinline__ ulonglong2 up2_load_px(const unsigned char *p) {
    if constexpr (sizeof(E) == 4) {
        return __ldg(reinterpret_cast<const ulonglong2 *>(p));
    } else if constexpr (sizeof(E) == 1) {
        float f[4];
        Word<uint8_t>::biased(__ldg(reinterpret_cast<const uint32_t *>(p)), f);
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        return make_ulonglong2(add2(pack2(f[0], f[1]), bias2), add2(pack2(f[2], f[3]), bias2));
    } else {
        const uint2 w = __ldg(reinterpret_cast<const uint2 *>(p));
        float f[2], g[2];
        Word<uint16_t>::biased(w.x, f);
        Word<uint16_t>::biased(w.y, g);
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        return make_ulonglong2(add2(pack2(f[0], f[1]), bias2), add2(pack2(g[0], g[1]), bias2));
    }
}

// Two neighbouring destination pixels (phase 0 and 1 of one source position) of E from their sums: clamp_nearest_from
// (src/primitive_sealed.rs:107-147) and one store of 2 x 4 x sizeof(E) bytes (f32: two 128-bit stores).
template <typename E> __device__ __forceinline__ void up2_store_pair(unsigned char *p, const f32x2 (&a)[2][2]) {
    float v[2][4];
#pragma unroll
    for (int ph = 0; ph < 2; ++ph) { unpack2(a[ph][0], v[ph][0], v[ph][1]); unpack2(a[ph][1], v[ph][2], v[ph][3]); }
    if constexpr (sizeof(E) == 4) {
#pragma unroll
        for (int ph = 0; ph < 2; ++ph)
            reinterpret_cast<float4 *>(p)[ph] = make_float4(Px<float>::clamp_nearest(v[ph][0]), Px<float>::clamp_nearest(v[ph][1]),
                                                            Px<float>::clamp_nearest(v[ph][2]), Px<float>::clamp_nearest(v[ph][3]));
    } else if constexpr (sizeof(E) == 1) {
        *reinterpret_cast<uint2 *>(p) = make_uint2(pack_u8x4(v[0][0], v[0][1], v[0][2], v[0][3]), pack_u8x4(v[1][0], v[1][1], v[1][2], v[1][3]));
    } else {
        *reinterpret_cast<uint4 *>(p) = make_uint4(pack_u16x2(v[0][0], v[0][1]), pack_u16x2(v[0][2], v[0][3]), pack_u16x2(v[1][0], v[1][1]),
                                                   pack_u16x2(v[1][2], v[1][3]));
    }
}

template <int T, typename E>
__global__ void __launch_bounds__(U2_THREADS, 3)
resize_up2_kernel(const __grid_constant__ Up2Args a) {
    constexpr int S = (T - 1) / 2;   // support
    constexpr int HR = 2 * T;        // ring rows per buffer = destination rows per phase (T source steps)
    extern __shared__ __align__(16) unsigned char smem[];
    ulonglong2 *s_ring = reinterpret_cast<ulonglong2 *>(smem);            // [2][HR][pitch] float4 pixels
    ulonglong2 *s_dump = s_ring + (size_t)2 * HR * a.pitch;               // one pixel
    float *s_vs = reinterpret_cast<float *>(s_dump + 1);                  // [2 BANDS + 2 HR]: border scale of local destination row n at [n + HR]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int nwarps = U2_THREADS / 32;
    const unsigned img = blockIdx.z;
    const int x0 = blockIdx.x * a.TWS;                 // first source pixel of the strip
    const int tws = min(a.TWS, (int)a.w - x0);
    const int m_lo = (int)a.dst_row0 >> 1;             // source row of the first destination pair this call touches
    const int mA = m_lo + blockIdx.y * a.BANDS;
    const int m_end = (int)((a.dst_row0 + a.dst_rows + 1) >> 1);
    const int nm = min(a.BANDS, m_end - mA);           // destination pairs of this band
    if (nm <= 0) return;

    constexpr size_t PXB = 4 * sizeof(E);                // bytes per pixel
    const unsigned char *src = reinterpret_cast<const unsigned char *>(a.src) + (size_t)img * a.src_img_stride;
    unsigned char *dst = reinterpret_cast<unsigned char *>(a.dst) + (size_t)img * a.dst_img_stride;
    const int p0 = x0 - S;                             // source pixel of ring column 0
    const int px = p0 + tid;                           // this thread's source pixel
    const bool active = tid < tws + 2 * S && px >= 0 && px < (int)a.w;
    const bool edge = p0 < 0 || x0 + tws + S > (int)a.w;

    // ---- vertical stream: step u consumes source row j0 + u and completes the pair m = j0 + u - 3; the first seven steps
    // only fill the window (their pairs lie in front of the band and are never read) ----
    const int j0 = mA + S - T;
    const int steps = T * (1 + (nm + T - 1) / T);
    const size_t gp = (size_t)a.w * PXB;               // bytes per source row
    const unsigned char *gptr = src + PXB * (size_t)(active ? px : max(min(px, (int)a.w - 1), 0)) + ((ptrdiff_t)j0 - (ptrdiff_t)a.src_row0) * (ptrdiff_t)gp;
    int jnext = j0;
    auto load_next = [&]() -> ulonglong2 {
        ulonglong2 v = make_ulonglong2(0ull, 0ull);
        // (only rows that a pair of this band taps: a band's buffer holds nothing else, and the stream starts one row earlier)
        if (jnext >= max(mA - S, 0) && jnext < min((int)a.h, mA + nm + S)) v = up2_load_px<E>(gptr);
        gptr += gp;
        ++jnext;
        return v;
    };
    ulonglong2 R[2 * T];   // circular window: row j0 + u lives in R[u % (2 T)]
#pragma unroll
    for (int k = 0; k < 2 * T; ++k) R[k] = make_ulonglong2(0ull, 0ull);
#pragma unroll
    for (int k = 0; k < T; ++k) R[k] = load_next();

    for (int i = tid; i < (2 * HR * a.pitch + 1) * 4; i += U2_THREADS) reinterpret_cast<float *>(s_ring)[i] = 0.0f;
    for (int i = tid; i < 2 * a.BANDS + 2 * HR; i += U2_THREADS) {
        const int n = i - HR, o = 2 * mA + n;          // local / global destination row
        const int first = (o >> 1) - S;                // its window (o >= 0 wherever it matters)
        float vs = 1.0f;
        if (n >= 0 && n < 2 * nm && (first < 0 || first + T > (int)a.h)) vs = up2_border_scale(a.wv[o & 1], T, first, (int)a.h);
        s_vs[i] = vs;
    }
    __syncthreads();

    const f32x2 init2 = pack2(a.acc_init, a.acc_init);
    const int pitch_eff = active ? a.pitch : 0;
    ulonglong2 *const tb0 = active ? s_ring + tid : s_dump;
    ulonglong2 *const tb1 = active ? tb0 + (size_t)HR * a.pitch : s_dump;
    ulonglong2 *tflush = tb0;
    int phase = 0;     // horizontal phases done
    
    auto horizontal = [&](int n0) {
        // local destination rows n0 .. n0 + 14 from the buffer just filled.  A warp item = one ring row x 32 source positions:
        // lane = source position, seven 128-bit loads (neighbouring lanes, no bank conflicts), both phases' pixels from them,
        // stored as 32 contiguous bytes per lane: a warp writes 1 KB of one destination row.
        __syncthreads();
        const int segs = (tws + 31) >> 5;
        const ulonglong2 *rbase = s_ring + (size_t)(phase & 1) * HR * a.pitch;
        const int row_lo = (int)a.dst_row0 - (2 * mA + n0), row_hi = (int)(a.dst_row0 + a.dst_rows) - (2 * mA + n0);  // ring rows to store
        int rp = 0, seg = warp;              // item = (pair of ring rows, segment of 32 source positions)
        while (seg >= segs) { seg -= segs; ++rp; }
        while (rp < T) {
            const int x = 32 * seg + lane;   // strip-local source position; ring column of x - 3
            const ulonglong2 *t = rbase + (size_t)(2 * rp) * a.pitch + min(x, tws - 1);
            f32x2 acc[2][2][2];              // [ring row of the pair][phase][half]
#pragma unroll
            for (int rr = 0; rr < 2; ++rr)
#pragma unroll
                for (int ph = 0; ph < 2; ++ph) { acc[rr][ph][0] = 0ull; acc[rr][ph][1] = 0ull; }
#pragma unroll
            for (int k = 0; k < T; ++k) {
#pragma unroll
                for (int rr = 0; rr < 2; ++rr) {
                    const ulonglong2 px4 = t[rr * a.pitch + k];
#pragma unroll
                    for (int ph = 0; ph < 2; ++ph) {
                        acc[rr][ph][0] = fma2(px4.x, a.wh[ph][k], acc[rr][ph][0]);
                        acc[rr][ph][1] = fma2(px4.y, a.wh[ph][k], acc[rr][ph][1]);
                    }
                }
            }
            if (edge) {   // strips on the image's left / right edge: some windows were clipped
                const int first = x0 + x - S;
                if (first < 0 || first + T > (int)a.w) {
#pragma unroll
                    for (int ph = 0; ph < 2; ++ph) {
                        const float hs = up2_border_scale(a.wh[ph], T, first, (int)a.w);
                        const f32x2 hs2 = pack2(hs, hs);
#pragma unroll
                        for (int rr = 0; rr < 2; ++rr) {
                            acc[rr][ph][0] = fma2(acc[rr][ph][0], hs2, 0ull);
                            acc[rr][ph][1] = fma2(acc[rr][ph][1], hs2, 0ull);
                        }
                    }
                }
            }
            if (x < tws) {
                unsigned char *gout = dst + ((size_t)(2 * rp - row_lo) * a.nw + 2 * (size_t)(x0 + x)) * PXB;
#pragma unroll
                for (int rr = 0; rr < 2; ++rr)
                    if (2 * rp + rr >= row_lo && 2 * rp + rr < row_hi) up2_store_pair<E>(gout + (size_t)rr * a.nw * PXB, acc[rr]);
            }
            seg += nwarps;
            while (seg >= segs) { seg -= segs; ++rp; }
        }
        ++phase;
        tflush = (phase & 1) ? tb1 : tb0;
    };

    for (int u0 = 0; u0 < steps; u0 += 2 * T) {
        const float *vsp = s_vs + 2 * u0;   // pair of step u: local rows 2 (u - 7), + 1 -> s_vs[2 u], [2 u + 1]
#pragma unroll
        for (int s = 0; s < 2 * T; ++s) {   // s is a compile-time constant after unrolling
            // the row loaded seven steps ago is in R[s]; fetch the one seven steps ahead into the slot that has just left the window
            R[(s + T) % (2 * T)] = load_next();
            f32x2 o[2][2];
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) { o[ph][0] = 0ull; o[ph][1] = 0ull; }
#pragma unroll
            for (int k = 0; k < T; ++k) {
                const ulonglong2 v = R[(s - (T - 1) + k + 2 * T) % (2 * T)];
#pragma unroll
                for (int ph = 0; ph < 2; ++ph) {
                    o[ph][0] = fma2(v.x, a.wv[ph][k], o[ph][0]);
                    o[ph][1] = fma2(v.y, a.wv[ph][k], o[ph][1]);
                }
            }
#pragma unroll
            for (int ph = 0; ph < 2; ++ph) {
                const float vs = vsp[2 * s + ph];
                const f32x2 vs2 = pack2(vs, vs);
                *tflush = make_ulonglong2(fma2(o[ph][0], vs2, init2), fma2(o[ph][1], vs2, init2));
                tflush += pitch_eff;
            }
            if (s == T - 1 || s == 2 * T - 1) {   // a phase of seven steps is complete
                const int u = u0 + s;
                if (u < T) {
                    tflush = tb0;                 // the window-filling phase
                } else if (u < steps) {
                    horizontal(2 * (u - (T - 1) - T));
                }
            }
        }
    }
}

int up2_env_int(const char *name, int dflt) {
    const char *e = std::getenv(name);
    return e && *e ? std::atoi(e) : dflt;
}

}  // namespace

// Returns 1 if this kernel took the job (and launched it), 0 if it does not apply, < 0 on error.
int resize_up2_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                   size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh, uint32_t n_images,
                   const AxisTable &vt, const AxisTable &ht, float acc_init, const DeviceCtx *ctx, cudaStream_t st) {
    if (C != 4) return 0;
    if (nw != 2 * w || nh != 2 * h || w < 16 || h < 16) return 0;
    const int T = vt.max_count;
    if (T != ht.max_count || (T != 3 && T != 5 && T != 7)) return 0;
    const int S = (T - 1) / 2;
    if ((reinterpret_cast<uintptr_t>(src) & 15) != 0 || (reinterpret_cast<uintptr_t>(dst) & 15) != 0 || src_img_stride % 16 != 0 ||
        dst_img_stride % 16 != 0)
        return 0;
    if (w > (1u << 28) || h > (1u << 28)) return 0;  // int arithmetic in the kernel
    // both axes follow the 2x rule (checked when the tables were built); whole windows in the middle give the weight vectors
    if (!vt.up2 || !ht.up2) return 0;
    const uint32_t vm = (nh / 2) & ~1u, hm = (nw / 2) & ~1u;
    const float *wv0 = &vt.w[(size_t)vm * vt.stride], *wv1 = &vt.w[(size_t)(vm + 1) * vt.stride];
    const float *wh0 = &ht.w[(size_t)hm * ht.stride], *wh1 = &ht.w[(size_t)(hm + 1) * ht.stride];

    Up2Args a;
    a.src = src; a.dst = dst; a.src_img_stride = src_img_stride; a.dst_img_stride = dst_img_stride;
    a.w = w; a.h = h; a.nw = nw; a.nh = nh; a.src_row0 = src_row0; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.acc_init = acc_init;
    // strip width: the widest that the block covers with its halo, narrowed so that the strips are of equal width (multiples of 4)
    int TWS = U2_THREADS - 2 * S;
    const int strips0 = (int)((w + TWS - 1) / TWS);
    TWS = std::min(TWS, (int)(((w + strips0 - 1) / strips0 + 3) & ~3u));
    TWS = std::min(TWS, up2_env_int("B200IMG_U2_TWS", TWS)) & ~3;
    if (TWS < 4) return 0;
    a.TWS = TWS;
    a.pitch = (4 * ((TWS + 3) / 4) + 2 * S) | 1;
    const int strips = (int)((w + TWS - 1) / TWS);
    const int pairs = (int)(((dst_row0 + dst_rows + 1) >> 1) - (dst_row0 >> 1));
    // rows per band: the job is bound by its stores and a CTA is a chain of short phases, so what matters is that the grid is
    // many times the resident CTAs (B200: 4096^2 -> 8192^2 0.34 ms with 21-row bands, 0.41 with 84; 8192^2 best at 63)
    int BANDS = 126;
    BANDS = BANDS / T * T;
    while (BANDS > 3 * T && (uint64_t)strips * ((pairs + BANDS - 1) / BANDS) * n_images < (uint64_t)ctx->sm_count * 48) BANDS -= T;
    {
        const int forced = up2_env_int("B200IMG_U2_BANDS", 0);
        if (forced > 0) BANDS = std::min((forced + T - 1) / T * T, 252 / T * T);
    }
    a.BANDS = BANDS;
    const int bands = (pairs + BANDS - 1) / BANDS;
    if (bands > 65535 || n_images > 65535) return 0;
    // small jobs: a CTA here is a chain of short phases with a window to fill and a ring to clear first, and the tile kernels
    // spread a small image better (1080p -> 2160p 65 us here, 56 us there; 4096 x 2048 and up are faster here)
    if ((uint64_t)w * pairs * n_images < (6u << 20) && !up2_env_int("B200IMG_U2_FORCE", 0)) return 0;
    for (int ph = 0; ph < 2; ++ph)
        for (int k = 0; k < U2_MAXT; ++k) {
            const float v = k < T ? (ph ? wv1 : wv0)[k] : 0.0f, hh = k < T ? (ph ? wh1 : wh0)[k] : 0.0f;
            uint32_t vb, hb;
            std::memcpy(&vb, &v, 4);
            std::memcpy(&hb, &hh, 4);
            a.wv[ph][k] = ((uint64_t)vb << 32) | vb;
            a.wh[ph][k] = ((uint64_t)hb << 32) | hb;
        }
    const int HR = 2 * T;
    const size_t smem = ((size_t)2 * HR * a.pitch + 1) * 16 + (size_t)(2 * BANDS + 2 * HR) * 4;
    if (smem > ctx->smem_optin) return 0;
    const dim3 grid((unsigned)strips, (unsigned)bands, n_images);
#define U2_GO(TT, EE)                                                                                                   \
    {                                                                                                                   \
        B200_CUDA(cudaFuncSetAttribute(resize_up2_kernel<TT, EE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        resize_up2_kernel<TT, EE><<<grid, U2_THREADS, smem, st>>>(a);                                                   \
    }
#define U2_GO_T(EE) { if (T == 7) U2_GO(7, EE) else if (T == 5) U2_GO(5, EE) else U2_GO(3, EE) }
    if (dtype == B200IMG_F32) U2_GO_T(float) else if (dtype == B200IMG_U8) U2_GO_T(uint8_t) else U2_GO_T(uint16_t)
#undef U2_GO_T
#undef U2_GO
    B200_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace b200
