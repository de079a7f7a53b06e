// resample_int.cu -- the resampler for downscales by an INTEGER ratio (2x, 4x; four-channel images also 3x) of u8
// images (and of four-channel 16-bit / f32 images) under B200IMG_FP_FAST: thumbnails and mip levels, BASELINE configs 0 (1080p -> 540p, Lanczos3) and 3
// (1024^2 -> 256^2, CatmullRom).
//
// For such a ratio R every output sample away from the borders has the SAME window shape: its window
// starts at source sample R * o + L0 and has NT = 2 * support * R taps whose weights depend on the tap
// index only (vertical_sample / horizontal_sample, src/imageops/sample.rs:494-562, 217-340: the centre
// (o + 0.5) * R lies on the same sub-sample phase for every o).  So both passes are strided
// convolutions with ONE weight vector, and everything the general streaming kernel
// (resample_stream.cu) spends on per-row weight tables in shared memory, accumulator-slot bookkeeping
// and register shifting goes away:
//   vertical:   a thread owns a 32-bit word column of the source strip (4 subpixels) and streams down
//               the rows once.  A = NT / R output rows are open at any time; with the row loop
//               unrolled over one period of A * R rows, which accumulator a row feeds and with which
//               tap is known at compile time: weights come from the kernel-parameter bank as
//               uniform operands of FFMA2, closed rows leave from a fixed accumulator.
//   horizontal: every HR finished rows, lane = row, a warp item = two 8-pixel blocks of one channel:
//               16 outputs per thread from ONE walk over the R * 7 + NT samples their windows span
//               (each sample is loaded once and feeds up to A outputs), weights again uniform.
// Borders: the reference clips a window at the image edge and renormalises the remaining weights.
// Here samples outside the image contribute zero and the result is multiplied by
// 1 / (sum of the weights that stayed inside) -- the same numbers up to f32 rounding, which the FAST
// contract (+-1 LSB) covers; STRICT runs never come here.  The accumulator's start value acc_init
// (sample.rs:547) is added once after the vertical pass (the horizontal weights sum to one).
#include "common.cuh"
#include "host_tables.h"
#include "packed.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>

namespace b200 {
namespace {

constexpr int RI_THREADS = 320;  // word columns per strip, at most (the block is as wide as the strip needs)
constexpr int RI_THREADS_PX = 224;  // ... of four-channel images (three CTAs per SM: 96 registers)
constexpr int RI_MAX_TAPS = 24;
constexpr int RI_MAX_BAND = 256;    // output rows per band, at most

struct IntArgs {
    const void *src;
    void *dst;
    size_t src_img_stride, dst_img_stride;  // bytes
    uint32_t w, h, nw, nh;
    uint32_t src_row0, dst_row0, dst_rows;  // the source buffer's first row; destination rows to produce
    int L0v, L0h;       // the window of output o starts at source sample R * o + L0 (before clipping)
    int TW, BAND;       // output pixels per strip (multiple of 16); output rows per band (multiple of HR)
    int tmp_pitch;      // floats per ring row (1 mod 32: lane = row reads hit 32 banks)
    int out_pitch;      // bytes per staged output row (4 * odd)
    float acc_init;
    int esize;          // four-channel kernel: bytes per channel (1, 2, 4)
    int vec_out;        // four-channel kernel: destination rows are 16-byte aligned (128-bit stores)
    f32x2 wv[RI_MAX_TAPS], wh[RI_MAX_TAPS];  // (w, w) pairs
};

__device__ __forceinline__ float lo_of(f32x2 v) { float a, b; unpack2(v, a, b); return a; }

// The weight vectors are symmetric (the host checked): tap k and tap NT - 1 - k share one entry, so that the NT / 2 distinct
// (w, w) pairs of an axis stay resident in uniform registers.
#define RI_SYM(w, k) ((w)[(k) < NT / 2 ? (k) : NT - 1 - (k)])

// 1 / (sum of the taps of a window starting at `first` that fall inside [0, len)).  Only windows that the image's edge clips
// come here (a few rows and columns per image): kept out of line so that the unrolled passes stay small.
template <int NT>
__device__ __noinline__ float border_scale(const f32x2 *w, int first, int len) {
    const int k0 = max(0, -first), k1 = min(NT, len - first);
    float s = 0.0f;
#pragma unroll 1
    for (int k = k0; k < k1; ++k) s += lo_of(RI_SYM(w, k));
    return 1.0f / s;
}

template <int C, int R, int A>
__global__ void __launch_bounds__(RI_THREADS)
resize_int_kernel(const __grid_constant__ IntArgs a) {
    constexpr int NT = A * R;            // taps
    constexpr int P = A * R;             // rows of one period of the vertical pass
    static_assert(C < 4, "four-channel images have their own kernel below");
    constexpr int HR = A == 6 ? 30 : 32; // ring rows = output rows per horizontal phase (a multiple of A); lane = row
    constexpr int PH = HR / A;           // periods per horizontal phase
    constexpr int D = P <= 16 ? P : P / 2;  // source rows in flight per thread (divides P: the row buffer is indexed statically)
    constexpr int SPAN = R * 7 + NT;     // samples the windows of 8 consecutive outputs span
    constexpr int PAD = 3;               // ring columns in front of the strip's first sample (the first word may start before it)
    static_assert(P % D == 0 && HR % A == 0, "period and phase geometry");
    extern __shared__ __align__(16) unsigned char smem[];
    float *s_tmp = reinterpret_cast<float *>(smem);                                    // [HR][tmp_pitch]
    uint8_t *s_out = reinterpret_cast<uint8_t *>(s_tmp + (size_t)HR * a.tmp_pitch);     // [HR][out_pitch]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x, nwarps = nthreads >> 5;
    const unsigned img = blockIdx.z;
    const int ox0 = blockIdx.x * a.TW;
    const int tw = min(a.TW, (int)a.nw - ox0);
    const int oyA = (int)a.dst_row0 + blockIdx.y * a.BAND;
    const int nout = min(a.BAND, (int)(a.dst_row0 + a.dst_rows) - oyA);
    if (nout <= 0) return;

    const uint8_t *src = reinterpret_cast<const uint8_t *>(a.src) + (size_t)img * a.src_img_stride;
    uint8_t *dst = reinterpret_cast<uint8_t *>(a.dst) + (size_t)img * a.dst_img_stride;
    const size_t gp = (size_t)a.w * C;           // source row pitch in bytes (a multiple of 4: the host checked)
    const int p0 = R * ox0 + a.L0h;              // source pixel of ring column 0 (may be negative)
    const int p1 = R * (ox0 + tw - 1) + a.L0h + NT;
    const int b0 = (max(p0, 0) * C) & ~3;        // first source byte of the strip's words
    const int b1 = min((min(p1, (int)a.w) * C + 3) & ~3, (int)gp);
    const int words = (b1 - b0) >> 2;            // <= blockDim.x (the host's choice of TW)
    const bool active = tid < words;
    const int col0 = b0 + 4 * tid - p0 * C + PAD;  // ring column of this thread's first subpixel (>= 0: b0 was rounded down by < 4)
    const bool edge = p0 < 0 || p1 > (int)a.w;   // only strips that touch the image's left / right edge have clipped windows


    // ---- vertical stream: row u of the stream is source row jbase + u, one virtual output in front of the band (so that
    // period m completes exactly the outputs (m - 1) A .. m A - 1); rows that no output of this band taps, and rows outside
    // the image, count as zero.  Which rows those are is the same for the whole CTA; threads beyond the strip's last word
    // follow word 0 (their sums are never stored). ----
    const int jbase = R * oyA + a.L0v - R;
    const int ulo = max(R * oyA + a.L0v, 0) - jbase;
    const unsigned ulen = (unsigned)(min((int)a.h, R * (oyA + nout - 1) + a.L0v + NT) - jbase - ulo);
    const uint8_t *gptr = src + b0 + 4 * (active ? tid : 0) + ((ptrdiff_t)jbase - (ptrdiff_t)a.src_row0) * (ptrdiff_t)gp;
    int unext = 0;  // the stream row gptr points at
    auto load_next = [&]() -> uint32_t {
        uint32_t v = 0u;
        if ((unsigned)(unext - ulo) < ulen) v = __ldg(reinterpret_cast<const uint32_t *>(gptr));
        gptr += gp;
        ++unext;
        return v;
    };
    f32x2 lo[A], hi[A];
#pragma unroll
    for (int s = 0; s < A; ++s) { lo[s] = 0ull; hi[s] = 0ull; }
    const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
    const f32x2 init2 = pack2(a.acc_init, a.acc_init);
    // ring stores of the generic layout: lanes 8 apart start on different subpixels of their word, so a store hits 32 banks
    const int rot = lane >> 3;
    const int ro0 = rot & 3, ro1 = (rot + 1) & 3, ro2 = (rot + 2) & 3, ro3 = (rot + 3) & 3;
    float *const tflush0 = s_tmp + col0;
    float *tflush = tflush0;
    const int periods = (nout + A - 1) / A + 1;
    const int rl = min(lane, HR - 1);  // lanes beyond the ring's last row recompute it (and store nothing)

    auto horizontal = [&](int n0, int rows_g) {
        // destination rows oyA + n0 .. + rows_g; their intermediates are ring rows 0 .. rows_g
        __syncthreads();
        const int nb = a.TW >> 3;                 // 8-pixel blocks per strip (even)
        {
            // a warp item = two blocks of 8 pixels, one channel: the two blocks' samples as packed pairs
            const int pairs = (nb >> 1) * C;
            const float *trow = s_tmp + (size_t)rl * a.tmp_pitch + PAD;
            for (int item = warp; item < pairs; item += nwarps) {
                const int pb = item / C, c = item - pb * C;
                const int xa = pb * 8, xb = (pb + (nb >> 1)) * 8;   // first output pixel of block A / block B (strip-local)
                const float *ta = trow + (R * xa) * C + c, *tb = trow + (R * xb) * C + c;
                f32x2 acc[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = 0ull;
#pragma unroll
                for (int e = 0; e < SPAN; ++e) {
                    const f32x2 x = pack2(ta[e * C], tb[e * C]);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int k = e - R * i;
                        if (k >= 0 && k < NT) acc[i] = fma2(x, RI_SYM(a.wh, k), acc[i]);
                    }
                }
                uint8_t *oa = s_out + (size_t)rl * a.out_pitch + xa * C + c, *ob = s_out + (size_t)rl * a.out_pitch + xb * C + c;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float va, vb;
                    unpack2(acc[i], va, vb);
                    if (edge) {
                        const int fa = R * (ox0 + xa + i) + a.L0h, fb = R * (ox0 + xb + i) + a.L0h;
                        if (fa < 0 || fa + NT > (int)a.w) va *= border_scale<NT>(a.wh, fa, (int)a.w);
                        if (fb < 0 || fb + NT > (int)a.w) vb *= border_scale<NT>(a.wh, fb, (int)a.w);
                    }
                    if (lane < rows_g && xa + i < tw) oa[i * C] = Px<uint8_t>::clamp_nearest(va);
                    if (lane < rows_g && xb + i < tw) ob[i * C] = Px<uint8_t>::clamp_nearest(vb);
                }
            }
        }
        __syncthreads();
        const size_t d_pitch = (size_t)a.nw * C;
        const int sub = tw * C;
        for (int rr = warp; rr < rows_g; rr += nwarps) {
            uint8_t *g = dst + (size_t)(oyA + n0 + rr - (int)a.dst_row0) * d_pitch + (size_t)ox0 * C;
            const uint8_t *sr = s_out + (size_t)rr * a.out_pitch;
            if ((sub & 3) == 0 && (reinterpret_cast<uintptr_t>(g) & 3) == 0) {
                for (int v = lane; v < (sub >> 2); v += 32) reinterpret_cast<uint32_t *>(g)[v] = reinterpret_cast<const uint32_t *>(sr)[v];
            } else {
                for (int v = lane; v < sub; v += 32) g[v] = sr[v];
            }
        }
        tflush = tflush0;
    };

    uint32_t buf[D];
#pragma unroll
    for (int k = 0; k < D; ++k) buf[k] = load_next();
    // a ring of zeros (while the first rows are on their way): the columns of samples outside the image are never written
    // and must contribute nothing
    for (int i = tid; i < HR * a.tmp_pitch; i += nthreads) s_tmp[i] = 0.0f;
    __syncthreads();
    for (int m = 0; m < periods; ++m) {
#pragma unroll
        for (int q = 0; q < P; ++q) {   // q is a compile-time constant after unrolling
            float f[4];
            Word<uint8_t>::biased(buf[q % D], f);
            buf[q % D] = load_next();
            const f32x2 f01 = add2(pack2(f[0], f[1]), bias2), f23 = add2(pack2(f[2], f[3]), bias2);
            const int fq = q / R, qr = q % R;
#pragma unroll
            for (int aa = 0; aa < A; ++aa) {
                const int slot = ((fq - aa) % A + A) % A, tap = qr + R * aa;
                lo[slot] = fma2(f01, RI_SYM(a.wv, tap), lo[slot]);
                hi[slot] = fma2(f23, RI_SYM(a.wv, tap), hi[slot]);
            }
            if (qr == R - 1) {   // an output row is complete: n = (m - 1) A + fq, in accumulator (fq + 1) mod A
                const int fs = (fq + 1) % A;
                const int n = (m - 1) * A + fq;
                if (n >= 0 && n < nout) {
                    float vs = 1.0f;
                    const int first = R * (oyA + n) + a.L0v;
                    if (first < 0 || first + NT > (int)a.h) vs = border_scale<NT>(a.wv, first, (int)a.h);
                    const f32x2 vs2 = pack2(vs, vs);
                    const f32x2 x = fma2(lo[fs], vs2, init2), y = fma2(hi[fs], vs2, init2);
                    if (active) {
                        {
                            float v0, v1, v2, v3;
                            unpack2(x, v0, v1);
                            unpack2(y, v2, v3);
                            // rotate the four values by `rot` so that store jj writes subpixel (jj + rot) & 3
                            const float r0 = (rot & 1) ? v1 : v0, r1 = (rot & 1) ? v2 : v1, r2 = (rot & 1) ? v3 : v2, r3 = (rot & 1) ? v0 : v3;
                            const float t0 = (rot & 2) ? r2 : r0, t1 = (rot & 2) ? r3 : r1, t2 = (rot & 2) ? r0 : r2, t3 = (rot & 2) ? r1 : r3;
                            tflush[ro0] = t0;
                            tflush[ro1] = t1;
                            tflush[ro2] = t2;
                            tflush[ro3] = t3;
                        }
                    }
                    tflush += a.tmp_pitch;
                }
                lo[fs] = 0ull;
                hi[fs] = 0ull;
            }
        }
        // periods 1 .. m have completed outputs 0 .. m A - 1: a horizontal phase every PH periods and at the band's end
        const int done = min(m * A, nout);
        if (m >= 1 && ((m % PH) == 0 || done == nout)) {
            const int n0 = (done - 1) / HR * HR;
            horizontal(n0, done - n0);
            if (done == nout) break;
        }
    }
}

// ---- four-channel images: a source word is one pixel ----
// The same two passes with a float4 ring: the vertical pass stores a pixel with one 128-bit store, the horizontal pass
// loads one with one 128-bit load as two packed pairs (r, g) (b, a), and writes its 8 output pixels (32 contiguous bytes per
// row) straight to global memory.  The ring is double buffered: a warp goes from its share of the horizontal pass of one
// phase straight into the vertical pass of the next, which fills the other buffer, so a phase costs one CTA barrier.
// Nothing in the row loop branches: border scales come from a small table in shared memory, threads beyond the strip's
// last word store to a dump slot, rows that no output needs are stored and never read.
// One source pixel (four channels of E) as loaded, and as two packed pairs of floats (r, g) (b, a).
template <typename E> struct RawPx;
template <> struct RawPx<uint8_t> {
    typedef uint32_t type;
    static __device__ __forceinline__ type zero() { return 0u; }
    static __device__ __forceinline__ void pairs(type w, f32x2 &rg, f32x2 &ba) {
        float f[4];
        Word<uint8_t>::biased(w, f);
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        rg = add2(pack2(f[0], f[1]), bias2);
        ba = add2(pack2(f[2], f[3]), bias2);
    }
};
template <> struct RawPx<uint16_t> {
    typedef uint2 type;
    static __device__ __forceinline__ type zero() { return make_uint2(0u, 0u); }
    static __device__ __forceinline__ void pairs(type w, f32x2 &rg, f32x2 &ba) {
        float f[2], g[2];
        Word<uint16_t>::biased(w.x, f);
        Word<uint16_t>::biased(w.y, g);
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        rg = add2(pack2(f[0], f[1]), bias2);
        ba = add2(pack2(g[0], g[1]), bias2);
    }
};
template <> struct RawPx<float> {
    typedef ulonglong2 type;
    static __device__ __forceinline__ type zero() { return make_ulonglong2(0ull, 0ull); }
    static __device__ __forceinline__ void pairs(type w, f32x2 &rg, f32x2 &ba) { rg = w.x; ba = w.y; }
};

// A destination pixel of E from its four sums (clamp_nearest_from, src/primitive_sealed.rs:107-147), as 32-bit words
template <typename E> __device__ __forceinline__ void out_words(float r, float g, float b, float al, uint32_t (&o)[sizeof(E)]) {
    if constexpr (sizeof(E) == 1) {
        o[0] = pack_u8x4(r, g, b, al);
    } else if constexpr (sizeof(E) == 2) {
        o[0] = pack_u16x2(r, g);
        o[1] = pack_u16x2(b, al);
    } else {
        o[0] = __float_as_uint(Px<float>::clamp_nearest(r)); o[1] = __float_as_uint(Px<float>::clamp_nearest(g));
        o[2] = __float_as_uint(Px<float>::clamp_nearest(b)); o[3] = __float_as_uint(Px<float>::clamp_nearest(al));
    }
}

// Geometry of the four-channel kernel, shared with the host's planning
__host__ __device__ constexpr int ri_ring_rows(int S) { return S == 6 ? 12 : S == 7 ? 14 : S == 5 ? 15 : S == 3 ? 15 : 16; }
__host__ __device__ constexpr int ri_item_px(int R) { return R == 2 ? 8 : R == 3 ? 5 : 4; }
__host__ __device__ constexpr int ri_rows_in_flight(int P, int most) {
    int d = 1;
    for (int k = 1; k <= most; ++k)
        if (P % k == 0) d = k;
    return d;
}

template <int R, int A, typename E>
__global__ void __maxnreg__(sizeof(E) == 4 ? 128 : 96)  // u8 / u16: three CTAs of RI_THREADS_PX threads per SM
resize_int4_kernel(const __grid_constant__ IntArgs a) {
    // An odd ratio puts the window's centre on a sample: NT = A R + 1 taps, and the row that closes output n (its tap NT - 1)
    // is also the row that opens output n + A + 1 (tap 0): S = A + 1 outputs are open on those rows, A on the others.
    constexpr int ODD = R & 1, NT = A * R + ODD, S = A + ODD, P = S * R;
    constexpr int HR = ri_ring_rows(S);    // ring rows per buffer = output rows per horizontal phase (a multiple of S); lane & 15 = row
    constexpr int PH = HR / S;
    constexpr int D = ri_rows_in_flight(P, sizeof(E) == 4 ? 8 : 16);   // source rows in flight per thread (f32 rows are four registers each)
    constexpr int PXB = 4 * sizeof(E);      // bytes per pixel
    constexpr int WPP = sizeof(E);          // 32-bit words per destination pixel
    typedef typename RawPx<E>::type Raw;
    constexpr int BLK = ri_item_px(R);      // output pixels per horizontal item: about two items per warp of thread columns
    constexpr int SPAN = R * (BLK - 1) + NT;
    static_assert(P % D == 0 && HR % S == 0 && HR <= 16, "period and phase geometry");
    extern __shared__ __align__(16) unsigned char smem[];
    float *s_ring = reinterpret_cast<float *>(smem);                 // [2][HR][tmp_pitch]
    float *s_dump = s_ring + (size_t)2 * HR * a.tmp_pitch;           // 4 floats
    float *s_vs = s_dump + 4;                                        // [BAND + 2 S]: vertical border scale of output n at [n + S]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nthreads = blockDim.x, nwarps = nthreads >> 5;
    const unsigned img = blockIdx.z;
    const int ox0 = blockIdx.x * a.TW;
    const int tw = min(a.TW, (int)a.nw - ox0);
    const int oyA = (int)a.dst_row0 + blockIdx.y * a.BAND;
    const int nout = min(a.BAND, (int)(a.dst_row0 + a.dst_rows) - oyA);
    if (nout <= 0) return;

    const uint8_t *src = reinterpret_cast<const uint8_t *>(a.src) + (size_t)img * a.src_img_stride;
    uint8_t *dst = reinterpret_cast<uint8_t *>(a.dst) + (size_t)img * a.dst_img_stride;
    const size_t gp = (size_t)a.w * PXB;
    const int p0 = R * ox0 + a.L0h;              // source pixel of ring column 0 (may be negative)
    const int p1 = R * (ox0 + tw - 1) + a.L0h + NT;
    const int q0 = max(p0, 0), q1 = min(p1, (int)a.w);
    const bool active = tid < q1 - q0;           // <= blockDim.x pixels (the host's choice of TW)
    const bool edge = p0 < 0 || p1 > (int)a.w;

    const int jbase = R * oyA + a.L0v - R;
    const int ulo = max(R * oyA + a.L0v, 0) - jbase;
    const unsigned ulen = (unsigned)(min((int)a.h, R * (oyA + nout - 1) + a.L0v + NT) - jbase - ulo);
    const uint8_t *gptr = src + PXB * (size_t)(q0 + (active ? tid : 0)) + ((ptrdiff_t)jbase - (ptrdiff_t)a.src_row0) * (ptrdiff_t)gp;
    int unext = 0;
    auto load_next = [&]() -> Raw {
        Raw v = RawPx<E>::zero();
        if ((unsigned)(unext - ulo) < ulen) v = __ldg(reinterpret_cast<const Raw *>(gptr));
        gptr += gp;
        ++unext;
        return v;
    };
    Raw buf[D];
#pragma unroll
    for (int k = 0; k < D; ++k) buf[k] = load_next();

    // while the first rows are on their way: a ring of zeros (the columns of samples outside the image are never written and
    // must contribute nothing) and the border scales of this band's rows
    for (int i = tid; i < 2 * HR * a.tmp_pitch + 4; i += nthreads) s_ring[i] = 0.0f;
    for (int i = tid; i < a.BAND + 2 * S; i += nthreads) {
        const int n = i - S, first = R * (oyA + n) + a.L0v;
        float vs = 1.0f;
        if (n >= 0 && n < nout && (first < 0 || first + NT > (int)a.h)) vs = border_scale<NT>(a.wv, first, (int)a.h);
        s_vs[i] = vs;
    }
    __syncthreads();

    f32x2 lo[S], hi[S];
#pragma unroll
    for (int s = 0; s < S; ++s) { lo[s] = 0ull; hi[s] = 0ull; }
    const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
    const f32x2 init2 = pack2(a.acc_init, a.acc_init);
    const int pitch_eff = active ? a.tmp_pitch : 0;
    float *const tb0 = active ? s_ring + 4 * (q0 - p0 + tid) : s_dump;
    float *const tb1 = active ? tb0 + (size_t)HR * a.tmp_pitch : s_dump;
    float *tflush = tb0;
    int phase = 0;                       // horizontal phases done: the vertical pass fills buffer phase & 1
    const int rlane = lane & 15, rl = min(rlane, HR - 1);
    const bool vec_out = a.vec_out != 0;

    auto horizontal = [&](int n0, int rows_g) {
        // destination rows oyA + n0 .. + rows_g from the buffer just filled
        __syncthreads();
        const int nb = (a.TW + BLK - 1) / BLK;
        const ulonglong2 *trow = reinterpret_cast<const ulonglong2 *>(s_ring + ((size_t)(phase & 1) * HR + rl) * a.tmp_pitch);
        uint8_t *grow = dst + (size_t)(oyA + n0 + rl - (int)a.dst_row0) * ((size_t)a.nw * PXB) + (size_t)ox0 * PXB;
        for (int item = 2 * warp + (lane >> 4); item < nb; item += 2 * nwarps) {
            const int xa = item * BLK;
            const ulonglong2 *t = trow + R * xa;
            f32x2 acc[BLK][2];
#pragma unroll
            for (int i = 0; i < BLK; ++i) { acc[i][0] = 0ull; acc[i][1] = 0ull; }
#pragma unroll
            for (int e = 0; e < SPAN; ++e) {
                const ulonglong2 px = t[e];
#pragma unroll
                for (int i = 0; i < BLK; ++i) {
                    const int k = e - R * i;
                    if (k >= 0 && k < NT) {
                        acc[i][0] = fma2(px.x, RI_SYM(a.wh, k), acc[i][0]);
                        acc[i][1] = fma2(px.y, RI_SYM(a.wh, k), acc[i][1]);
                    }
                }
            }
            if (edge) {   // strips on the image's left / right edge: some windows were clipped
#pragma unroll
                for (int i = 0; i < BLK; ++i) {
                    const int first = R * (ox0 + xa + i) + a.L0h;
                    if (first < 0 || first + NT > (int)a.w) {
                        const float hs = border_scale<NT>(a.wh, first, (int)a.w);
                        const f32x2 hs2 = pack2(hs, hs);
                        acc[i][0] = fma2(acc[i][0], hs2, 0ull);
                        acc[i][1] = fma2(acc[i][1], hs2, 0ull);
                    }
                }
            }
            uint32_t o[BLK][WPP];
#pragma unroll
            for (int i = 0; i < BLK; ++i) {
                float r, g, b, al;
                unpack2(acc[i][0], r, g);
                unpack2(acc[i][1], b, al);
                out_words<E>(r, g, b, al, o[i]);
            }
            if (rlane < rows_g) {
                uint32_t *gout = reinterpret_cast<uint32_t *>(grow) + xa * WPP;
                if ((BLK * PXB) % 16 == 0 && vec_out && xa + BLK <= tw) {
                    const uint32_t *ow = &o[0][0];
#pragma unroll
                    for (int i = 0; i < BLK * WPP; i += 4) reinterpret_cast<uint4 *>(gout + i)[0] = make_uint4(ow[i], ow[i + 1], ow[i + 2], ow[i + 3]);
                } else {
#pragma unroll
                    for (int i = 0; i < BLK; ++i)
                        if (xa + i < tw) {
#pragma unroll
                            for (int k = 0; k < WPP; ++k) gout[i * WPP + k] = o[i][k];
                        }
                }
            }
        }
        ++phase;
        tflush = (phase & 1) ? tb1 : tb0;
    };

    const int periods = (nout + S - 1) / S + 1;
    for (int m = 0; m < periods; ++m) {
        const float *vsp = s_vs + m * S;
#pragma unroll
        for (int q = 0; q < P; ++q) {   // q is a compile-time constant after unrolling
            f32x2 f01, f23;
            RawPx<E>::pairs(buf[q % D], f01, f23);
            buf[q % D] = load_next();
            const int fq = q / R, qr = q % R;
#pragma unroll
            for (int aa = 0; aa < S; ++aa) {
                const int slot = ((fq - aa) % S + S) % S, tap = qr + R * aa;
                if (tap < NT) {
                    lo[slot] = fma2(f01, RI_SYM(a.wv, tap), lo[slot]);
                    hi[slot] = fma2(f23, RI_SYM(a.wv, tap), hi[slot]);
                }
            }
            if (qr == (NT - 1) % R) {   // output row n = (m - 1) S + fq has had its last tap (from row group fq - (NT - 1) / R)
                const int fs = ((fq - (NT - 1) / R) % S + S) % S;
                const float vs = vsp[fq];
                const f32x2 vs2 = pack2(vs, vs);
                *reinterpret_cast<ulonglong2 *>(tflush) = make_ulonglong2(fma2(lo[fs], vs2, init2), fma2(hi[fs], vs2, init2));
                tflush += pitch_eff;
                lo[fs] = 0ull;
                hi[fs] = 0ull;
            }
        }
        if (m == 0) {
            tflush = tb0;  // period 0 completed only the virtual outputs in front of the band
        } else {
            // periods 1 .. m have completed outputs 0 .. m S - 1: a horizontal phase every PH periods and at the band's end
            const int done = min(m * S, nout);
            if ((m % PH) == 0 || done == nout) {
                const int n0 = (done - 1) / HR * HR;
                horizontal(n0, done - n0);
                if (done == nout) break;
            }
        }
    }
}

#undef RI_SYM

int round_up_i(int v, int m) { return (v + m - 1) / m * m; }

struct IntGeo {
    int TW = 0, threads = 0, strips = 0, tmp_pitch = 0, out_pitch = 0;
    size_t smem = 0;
    double cost = 0.0;  // issue slots per output row of the image, up to a constant
};

IntGeo int_geo(int TW, uint32_t w, uint32_t nw, int C, int R, int NT, int L0h, int HR) {
    IntGeo g;
    g.TW = TW;
    g.strips = (int)((nw + TW - 1) / TW);
    int words = 0;
    for (int s = 0; s < g.strips; ++s) {
        const int ox0 = s * TW, tw = std::min(TW, (int)nw - ox0);
        const int p0 = R * ox0 + L0h, p1 = R * (ox0 + tw - 1) + L0h + NT;
        const int b0 = (std::max(p0, 0) * C) & ~3, b1 = std::min((std::min(p1, (int)w) * C + 3) & ~3, (int)w * C);
        words = std::max(words, (b1 - b0) / 4);
    }
    g.threads = round_up_i(words, 32);
    int tp;
    if (C == 4) {  // float4 pixels: an odd number of pixels per row, so that the 128-bit loads of 8 rows hit 8 bank groups
        const int blk = ri_item_px(R);
        const int px = R * (round_up_i(TW, blk) - 1) + NT;  // the last item reads its whole span
        tp = 4 * (px | 1);
    } else {
        tp = 3 + (R * (TW - 1) + NT) * C + 3;
        tp = round_up_i(tp - 1, 32) + 1;  // 1 mod 32
    }
    g.tmp_pitch = tp;
    int op = round_up_i(TW * C, 4);
    if ((op / 4) % 2 == 0) op += 4;
    g.out_pitch = op;
    if (C == 4) g.smem = (size_t)2 * HR * tp * 4 + 16 + (size_t)(RI_MAX_BAND + 2 * 7) * 4;  // two ring buffers, dump slot, border scales
    else g.smem = (size_t)HR * tp * 4 + (size_t)HR * op;
    // every warp of the CTA walks HR * R source rows per phase, then the CTA's horizontal items go round its warps
    const int nwarps = g.threads / 32;
    if (C == 4) {
        const int blk = ri_item_px(R), nb = (TW + blk - 1) / blk, rounds = (nb + 2 * nwarps - 1) / (2 * nwarps);
        const int A = NT / R, item = (R * (blk - 1) + NT) + 2 * blk * NT + 22 * blk + 40;
        g.cost = (double)g.strips * nwarps * (HR * R * (10 + 2 * A + 12.0 / R) + rounds * item) / HR;
    } else {
        g.cost = (double)g.threads * g.strips;
    }
    return g;
}

template <int C, int R, int A>
int launch_ra(const IntArgs &a, dim3 grid, int threads, size_t smem, cudaStream_t st) {
    if constexpr (C == 4) {
#define RI_GO4(EE)                                                                              \
    {                                                                                           \
        auto k4 = resize_int4_kernel<R, A, EE>;                                                 \
        B200_CUDA(cudaFuncSetAttribute(k4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
        k4<<<grid, threads, smem, st>>>(a);                                                     \
    }
        if (a.esize == 1) RI_GO4(uint8_t) else if (a.esize == 2) RI_GO4(uint16_t) else RI_GO4(float)
#undef RI_GO4
        B200_CUDA(cudaGetLastError());
        return B200IMG_OK;
    }
    auto k = resize_int_kernel<C == 4 ? 3 : C, R, A>;
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, threads, smem, st>>>(a);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <int C, int R>
int launch_r(const IntArgs &a, int A, dim3 grid, int threads, size_t smem, cudaStream_t st) {
    switch (A) {
        case 2: return launch_ra<C, R, 2>(a, grid, threads, smem, st);
        case 4: return launch_ra<C, R, 4>(a, grid, threads, smem, st);
        default: return launch_ra<C, R, 6>(a, grid, threads, smem, st);
    }
}

template <int C>
int launch_c(const IntArgs &a, int R, int A, dim3 grid, int threads, size_t smem, cudaStream_t st) {
    if constexpr (C == 4) {
        if (R == 3) return launch_r<4, 3>(a, A, grid, threads, smem, st);
    }
    return R == 2 ? launch_r<C, 2>(a, A, grid, threads, smem, st) : launch_r<C, 4>(a, A, grid, threads, smem, st);
}

int env_int(const char *name, int dflt) {
    const char *e = std::getenv(name);
    return e && *e ? std::atoi(e) : dflt;
}

}  // namespace

// Returns 1 if this kernel took the job (and launched it), 0 if it does not apply, < 0 on error.
int resize_int_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                   size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh, uint32_t n_images,
                   const AxisTable &vt, const AxisTable &ht, float acc_init, const DeviceCtx *ctx, cudaStream_t st) {
    if (dtype != B200IMG_U8 && C != 4) return 0;   // 16-bit and f32 images: the four-channel kernel only
    if (nw == 0 || nh == 0 || w % nw != 0 || h % nh != 0 || w / nw != h / nh) return 0;
    const int R = (int)(w / nw);
    if (R != 2 && R != 4 && !(R == 3 && C == 4)) return 0;   // (3x: the four-channel kernel only)
    const size_t esize = dtype_size(dtype), gp = (size_t)w * C * esize;
    if (gp % 4 != 0 || (reinterpret_cast<uintptr_t>(src) & 3) != 0 || src_img_stride % 4 != 0) return 0;
    if ((uint64_t)w * C > (1u << 28) || h > (1u << 29)) return 0;  // int arithmetic in the kernel
    if (C == 4) {   // the four-channel kernel loads and stores whole pixels
        const size_t pxb = 4 * esize;
        if (reinterpret_cast<uintptr_t>(src) % pxb != 0 || src_img_stride % pxb != 0 || reinterpret_cast<uintptr_t>(dst) % pxb != 0 ||
            dst_img_stride % pxb != 0)
            return 0;
    }
    // both axes follow the fixed-window rule with symmetric weights (checked when the tables were built)
    if (vt.fixed_R != R || ht.fixed_R != R) return 0;
    const int NT = vt.max_count;
    if (NT != ht.max_count || NT % R != (R & 1) || NT > RI_MAX_TAPS) return 0;   // an odd ratio centres the window on a sample: one more tap
    const int A = NT / R, S = A + (R & 1);   // open outputs per row (odd ratios: S on the rows that close one)
    if (A != 2 && A != 4 && A != 6) return 0;
    // The one- to three-channel kernel (a 30 / 32-row scalar ring per CTA) pays only where there is work to amortise it: not on
    // two-tap-per-output filters, narrow rows or short images (measured against the streaming kernel, tools/check_choice.py:
    // Triangle 2 x 4096^2 Rgb8 1.15x slower, 64 x 256^2 Luma8 1.2-1.7x, 8192 x 128 Rgb8 1.7x).
    if (C != 4 && (A == 2 || (size_t)w * C < 1024 || dst_rows < 96)) return 0;
    const int L0v = vt.fixed_L0, L0h = ht.fixed_L0;
    const float *wvm = &vt.w[(size_t)(nh / 2) * vt.stride], *whm = &ht.w[(size_t)(nw / 2) * ht.stride];

    const int HR = C == 4 ? ri_ring_rows(S) : (A == 6 ? 30 : 32);
    const int max_threads = C == 4 ? RI_THREADS_PX : RI_THREADS;
    // strip width: the cheapest (fewest thread columns over the image's width; for four channels also whole rounds of
    // horizontal items) among the widths that leave room for two (three) CTAs per SM
    IntGeo best;
    const int forced_tw = env_int("B200IMG_RI_TW", 0);
    for (int pass = 0; pass < 2 && best.TW == 0; ++pass) {
        const size_t cap = pass == 0 ? (C == 4 ? (size_t)(233472 - 3072) / 3 : (size_t)(233472 - 2048) / 2) : ctx->smem_optin;
        const int step = C == 4 ? ri_item_px(R) : 16;  // whole horizontal items (and, at even ratios, 16-byte aligned output blocks)
        for (int TW = step; TW <= 256; TW += step) {
            if (forced_tw && TW != forced_tw) continue;
            const IntGeo g = int_geo(TW, w, nw, C, R, NT, L0h, HR);
            if (g.threads > max_threads || g.smem > std::min(cap, ctx->smem_optin)) continue;
            if (best.TW == 0 || g.cost < best.cost || (g.cost == best.cost && TW > best.TW)) best = g;
        }
    }
    if (best.TW == 0) return 0;
    const int ctas_per_sm = std::max<int>(1, std::min<size_t>(2048 / best.threads, (233472 - 1024) / (best.smem + 1024)));
    // bands of about 128 output rows (the window overlap between bands costs (NT - R) source rows each), of equal height, and
    // shorter ones when that leaves SMs without work
    int BAND;
    const int forced_band = env_int("B200IMG_RI_BAND", 0);
    if (forced_band > 0) {
        BAND = std::min(round_up_i(forced_band, HR), RI_MAX_BAND / HR * HR);
    } else {
        int nbands = (int)((dst_rows + 127) / 128);
        BAND = round_up_i((int)((dst_rows + nbands - 1) / nbands), HR);
        while (BAND > HR && (uint64_t)best.strips * ((dst_rows + BAND - 1) / BAND) * n_images < (uint64_t)ctx->sm_count * ctas_per_sm * 3) BAND -= HR;
    }
    const int bands = (int)((dst_rows + BAND - 1) / BAND);
    if (bands > 65535 || n_images > 65535) return 0;
    if ((uint64_t)best.strips * bands * n_images < (uint64_t)ctx->sm_count) return 0;  // a small single image: the tile kernels spread it better

    IntArgs a;
    a.src = src; a.dst = dst; a.src_img_stride = src_img_stride; a.dst_img_stride = dst_img_stride;
    a.w = w; a.h = h; a.nw = nw; a.nh = nh; a.src_row0 = src_row0; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.L0v = L0v; a.L0h = L0h; a.TW = best.TW; a.BAND = BAND; a.tmp_pitch = best.tmp_pitch; a.out_pitch = best.out_pitch;
    a.acc_init = acc_init;
    a.esize = (int)esize;
    a.vec_out = ((size_t)nw * C * esize) % 16 == 0 && dst_img_stride % 16 == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    for (int k = 0; k < RI_MAX_TAPS; ++k) {
        const float v = k < NT ? wvm[k] : 0.0f, hh = k < NT ? whm[k] : 0.0f;
        uint32_t vb, hb;
        std::memcpy(&vb, &v, 4);
        std::memcpy(&hb, &hh, 4);
        a.wv[k] = ((uint64_t)vb << 32) | vb;
        a.wh[k] = ((uint64_t)hb << 32) | hb;
    }
    const dim3 grid((unsigned)best.strips, (unsigned)bands, n_images);
    int rc;
    switch (C) {
        case 1: rc = launch_c<1>(a, R, A, grid, best.threads, best.smem, st); break;
        case 2: rc = launch_c<2>(a, R, A, grid, best.threads, best.smem, st); break;
        case 3: rc = launch_c<3>(a, R, A, grid, best.threads, best.smem, st); break;
        default: rc = launch_c<4>(a, R, A, grid, best.threads, best.smem, st); break;
    }
    return rc ? rc : 1;
}

}  // namespace b200
