// gauss_tiled.cu -- the fast row / column passes of the separable convolution engine.
//
// Used when reordering the sum is allowed: the f32 engine under B200IMG_FP_FAST
// (+-1 LSB contract) and the Q0.15 integer engine always (integer sums are exact
// in any order, so it stays bit-identical to src/imageops/filter_1d.rs:171-390).
//
// Both passes are register-tiled 1-D convolutions: a thread produces R = 8
// consecutive outputs along the filter axis from a sliding window of 15 staged
// inputs, so each shared-memory load feeds 8 multiply-adds (LDS : FFMA = 1 : 8;
// a per-output loop would be 1 : 1 and bound by the shared-memory pipe at a
// quarter of the FP32 rate).  Weights sit in the kernel-parameter constant bank
// and reach the FFMA/IMAD as uniform operands, costing no load slots.
//   row pass:    lanes = 32 image rows (odd row pitch: conflict-free), warps =
//                (8-pixel block, channel) pairs; results go through a staging
//                buffer so global stores are row-contiguous.
//   column pass: lanes = 32 consecutive subpixels of a row (coalesced loads and
//                stores, conflict-free), thread = one column x 8 rows.
// Clamp-to-edge (make_arena_row / make_columns_arenas, filter_1d.rs:60-130) is
// applied while staging.
#include "common.cuh"

#include <algorithm>
#include <cstring>
#include <type_traits>

namespace b200 {

constexpr int GT_R = 8;           // outputs per thread along the filter axis
constexpr int GT_MAX_TAPS = 384;  // longer kernels (sigma > 57) take the direct kernels in gauss.cu
constexpr int GT_THREADS = 256;

struct TapWeights { uint32_t w[GT_MAX_TAPS]; };  // f32 bits or Q0.15 integers, zero padded

namespace {

template <typename Acc> struct Arith;
template <> struct Arith<float> {
    static __device__ __forceinline__ float w(uint32_t b) { return __uint_as_float(b); }
    static __device__ __forceinline__ float mac(float acc, float a, float b) { return __fmaf_rn(a, b, acc); }
    static __device__ __forceinline__ uint32_t bits(float v) { return __float_as_uint(v); }
    static __device__ __forceinline__ float from_bits(uint32_t b) { return __uint_as_float(b); }
};
template <> struct Arith<uint32_t> {
    static __device__ __forceinline__ uint32_t w(uint32_t b) { return b; }
    static __device__ __forceinline__ uint32_t mac(uint32_t acc, uint32_t a, uint32_t b) { return acc + a * b; }
    static __device__ __forceinline__ uint32_t bits(uint32_t v) { return v; }
    static __device__ __forceinline__ uint32_t from_bits(uint32_t b) { return b; }
};

// ToStorage (filter_1d.rs:132-164) / final nearest_from (sample.rs:1596-1601)
template <typename Out, typename Acc> struct St;
template <> struct St<uint8_t, uint32_t> { static __device__ __forceinline__ uint8_t f(uint32_t a) { return (uint8_t)min((a + 16384u) >> 15, 255u); } };
template <> struct St<uint16_t, uint32_t> { static __device__ __forceinline__ uint16_t f(uint32_t a) { return (uint16_t)min((a + 16384u) >> 15, 65535u); } };
template <> struct St<float, float> { static __device__ __forceinline__ float f(float a) { return a; } };
template <> struct St<uint8_t, float> { static __device__ __forceinline__ uint8_t f(float a) { return Px<uint8_t>::nearest(a); } };
template <> struct St<uint16_t, float> { static __device__ __forceinline__ uint16_t f(float a) { return Px<uint16_t>::nearest(a); } };

// acc[r] = sum_k in[(r + k) * STRIDE] * w[k], k < nch * 8.  `in` points at the first staged
// input of output 0.  Window registers slide by 8 per chunk.  STRIDE is a compile-time
// constant so every LDS carries an immediate offset (no address arithmetic in the loop).
template <typename Acc, int STRIDE>
__device__ __forceinline__ void conv8(const uint32_t *in, int nch, const TapWeights &tw, Acc (&acc)[GT_R]) {
    using A = Arith<Acc>;
    Acc win[2 * GT_R - 1];
#pragma unroll
    for (int i = 0; i < GT_R - 1; ++i) win[i] = A::from_bits(in[i * STRIDE]);
    in += (GT_R - 1) * STRIDE;
    const uint2 *w2 = reinterpret_cast<const uint2 *>(tw.w);
#pragma unroll 1
    for (int ch = 0; ch < nch; ++ch) {
#pragma unroll
        for (int i = 0; i < GT_R; ++i) win[GT_R - 1 + i] = A::from_bits(in[i * STRIDE]);
        in += GT_R * STRIDE;
#pragma unroll
        for (int kp = 0; kp < GT_R / 2; ++kp) {
            const uint2 wv = w2[ch * (GT_R / 2) + kp];
            const Acc wa = A::w(wv.x), wb = A::w(wv.y);
#pragma unroll
            for (int r = 0; r < GT_R; ++r) acc[r] = A::mac(acc[r], win[r + 2 * kp], wa);
#pragma unroll
            for (int r = 0; r < GT_R; ++r) acc[r] = A::mac(acc[r], win[r + 2 * kp + 1], wb);
        }
#pragma unroll
        for (int i = 0; i < GT_R - 1; ++i) win[i] = win[GT_R + i];
    }
}

// ---- packed f32x2 variant (FFMA2): two independent convolutions per thread.  FFMA2 has the
// FLOP rate of FFMA (measured 126 vs 123 lanes/clk/SM) at half the issue slots, so the loads,
// constant fetches and loop overhead of the loop body fit beside a saturated FMA pipe. ----
typedef unsigned long long f32x2;
struct TapWeights2 { f32x2 w[GT_MAX_TAPS]; };  // (w, w) pairs, zero padded
__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

// acc[r] = (sum_k A[(r+k)*STRIDE] w[k], sum_k B[(r+k)*STRIDE] w[k]), k < nch * TC.  PAIRED: B = A + 1
// and A is 8-byte aligned, so one LDS.64 fetches both.  TC (taps per unrolled chunk) is chosen per
// launch from {5,6,7,8} to minimise zero-padded taps (65 taps = 13 x 5, 21 = 3 x 7).
// The window lives in a register RING of RS = TC * U entries (the smallest multiple of TC that
// holds GT_R - 1 + TC of them); the chunk loop is unrolled U times, so ring positions are
// compile-time constants and sliding the window by TC costs no register moves (sliding a linear
// window cost 14 moves per 40 FFMA2 in loops that are issue-bound).
// RING = false keeps a linear window and moves it (fewer registers, less code): measured faster for
// short kernels (21 taps), the ring for long ones (65 taps); the launcher picks by chunk count.
template <int STRIDE, bool PAIRED, int TC, bool RING>
__device__ __forceinline__ void conv8x2(const float *inA, const float *inB, int nch, const TapWeights2 &tw, f32x2 (&acc)[GT_R]) {
    if constexpr (!RING) {
        f32x2 win[GT_R - 1 + TC];
        auto ld = [&](int i) -> f32x2 {
            if constexpr (PAIRED) return *reinterpret_cast<const f32x2 *>(inA + i * STRIDE);
            else return pack2(inA[i * STRIDE], inB[i * STRIDE]);
        };
#pragma unroll
        for (int i = 0; i < GT_R - 1; ++i) win[i] = ld(i);
        inA += (GT_R - 1) * STRIDE;
        inB += (GT_R - 1) * STRIDE;
#pragma unroll 1
        for (int ch = 0; ch < nch; ++ch) {
#pragma unroll
            for (int i = 0; i < TC; ++i) win[GT_R - 1 + i] = ld(i);
            inA += TC * STRIDE;
            inB += TC * STRIDE;
#pragma unroll
            for (int kk = 0; kk < TC; ++kk) {
                const f32x2 wk = tw.w[ch * TC + kk];
#pragma unroll
                for (int r = 0; r < GT_R; ++r) acc[r] = fma2(win[r + kk], wk, acc[r]);
            }
#pragma unroll
            for (int i = 0; i < GT_R - 1; ++i) win[i] = win[TC + i];
        }
        return;
    }
    constexpr int U = (GT_R - 1 + TC + TC - 1) / TC;
    constexpr int RS = TC * U;
    f32x2 win[RS];
    auto ld = [&](int i) -> f32x2 {
        if constexpr (PAIRED) return *reinterpret_cast<const f32x2 *>(inA + i * STRIDE);
        else return pack2(inA[i * STRIDE], inB[i * STRIDE]);
    };
#pragma unroll
    for (int i = 0; i < GT_R - 1; ++i) win[i] = ld(i);
    inA += (GT_R - 1) * STRIDE;
    inB += (GT_R - 1) * STRIDE;
#pragma unroll 1
    for (int ch = 0; ch < nch; ch += U) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (ch + u < nch) {  // chunk ch + u: logical window entry i sits at ring position (u * TC + i) mod RS
#pragma unroll
                for (int i = 0; i < TC; ++i) win[(u * TC + GT_R - 1 + i) % RS] = ld(i);
                inA += TC * STRIDE;
                inB += TC * STRIDE;
#pragma unroll
                for (int kk = 0; kk < TC; ++kk) {
                    const f32x2 wk = tw.w[(ch + u) * TC + kk];
#pragma unroll
                    for (int r = 0; r < GT_R; ++r) acc[r] = fma2(win[(u * TC + r + kk) % RS], wk, acc[r]);
                }
            }
        }
    }
}

// u8 / u16 / f32 source subpixels of a row -> f32 words in shared memory.  Integer types go a
// 32-bit word at a time through the 2^23 trick (I2F would run on the quarter-rate XU pipe).
// `s` receives elements [0, n) of `g`; needs g's row pitch and base 4-byte aligned (checked by the caller).
template <typename In>
__device__ __forceinline__ void stage_convert_row(const In *g, float *s, int n, int lane) {
    if constexpr (sizeof(In) == 4) {
        for (int e = lane; e < n; e += 32) s[e] = g[e];
    } else {
        constexpr int PER = 4 / sizeof(In);
        const int phase = (int)((reinterpret_cast<uintptr_t>(g) & 3) / sizeof(In));  // elements before g in its word
        const uint32_t *gw = reinterpret_cast<const uint32_t *>(g - phase);
        const int words = (phase + n + PER - 1) / PER;
        for (int wi = lane; wi < words; wi += 32) {
            const uint32_t w = __ldg(gw + wi);
            const int e0 = wi * PER - phase;
#pragma unroll
            for (int j = 0; j < PER; ++j) {
                const int e = e0 + j;
                uint32_t m;
                if constexpr (PER == 4) m = __byte_perm(w, 0x4B000000u, 0x7540 + j);
                else m = __byte_perm(w, 0x4B000000u, j == 0 ? 0x7510 : 0x7532);
                if (e >= 0 && e < n) s[e] = __uint_as_float(m) - 8388608.0f;
            }
        }
    }
}

// The same for the NR rows a warp stages, with every global load of the warp issued before the first
// conversion: one row at a time leaves each load's DRAM latency exposed (the row pass spent 45 % of its
// stall samples in that loop).  Rows are `g_pitch` elements apart in global memory, `s_pitch` floats apart
// in shared memory; all rows have the same byte phase (row pitch is a multiple of 4 bytes).  Returns false
// (nothing done) when a row has more than 32 * MAXIT words.
template <typename In, int NR, int MAXIT>
__device__ __forceinline__ bool stage_convert_rows(const In *const (&g)[NR], float *const (&s)[NR], int n, int lane) {
    static_assert(sizeof(In) < 4, "integer subpixels only");
    constexpr int PER = 4 / sizeof(In);
    const int phase = (int)((reinterpret_cast<uintptr_t>(g[0]) & 3) / sizeof(In));
    const int words = (phase + n + PER - 1) / PER;
    if (words > 32 * MAXIT) return false;
    uint32_t wv[NR][MAXIT];
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const uint32_t *gw = reinterpret_cast<const uint32_t *>(g[r] - phase);
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int wi = it * 32 + lane;
            wv[r][it] = wi < words ? __ldg(gw + wi) : 0u;
        }
    }
#pragma unroll
    for (int r = 0; r < NR; ++r)
#pragma unroll
        for (int it = 0; it < MAXIT; ++it) {
            const int wi = it * 32 + lane;
            if (wi < words) {
                const int e0 = wi * PER - phase;
#pragma unroll
                for (int j = 0; j < PER; ++j) {
                    const int e = e0 + j;
                    uint32_t m;
                    if constexpr (PER == 4) m = __byte_perm(wv[r][it], 0x4B000000u, 0x7540 + j);
                    else m = __byte_perm(wv[r][it], 0x4B000000u, j == 0 ? 0x7510 : 0x7532);
                    if (e >= 0 && e < n) s[r][e] = __uint_as_float(m) - 8388608.0f;
                }
            }
        }
    return true;
}

constexpr int ROW_XT = 64;    // output pixels per CTA tile (row pass)
constexpr int ROW_ROWS = 32;  // rows per CTA tile = lanes

// Row pass.  total_rows = rows of the whole batch (rows are independent).
template <typename In, typename Mid, typename Acc, int C>
__global__ void __launch_bounds__(GT_THREADS)
row_tiled_kernel(const In *__restrict__ src, Mid *__restrict__ mid, uint32_t W, size_t total_rows, int half,
                 int shift, int nch, int in_pitch, const __grid_constant__ TapWeights tw) {
    using A = Arith<Acc>;
    constexpr int OUT_PITCH = (ROW_XT * C) | 1;
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t *s_in = sm;                                 // [ROW_ROWS][in_pitch]
    uint32_t *s_out = sm + (size_t)ROW_ROWS * in_pitch;  // [ROW_ROWS][OUT_PITCH]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int x0 = blockIdx.x * ROW_XT;
    const size_t row0 = (size_t)blockIdx.y * ROW_ROWS;
    const int sp = ROW_XT + nch * GT_R - 1;  // staged pixels per row
    const int row_elems = sp * C;
    const size_t g_pitch = (size_t)W * C;
    const int xs = x0 - half - shift;  // source pixel of staged pixel 0
    const bool interior = xs >= 0 && xs + sp <= (int)W;

    for (int r = warp; r < ROW_ROWS; r += GT_THREADS / 32) {
        const In *g = src + min(row0 + r, total_rows - 1) * g_pitch;
        uint32_t *s = s_in + r * in_pitch;
        if (interior) {
            g += (size_t)xs * C;
            for (int e = lane; e < row_elems; e += 32) s[e] = A::bits((Acc)g[e]);
        } else {
            for (int e = lane; e < row_elems; e += 32) {
                const int p = e / C, c = e - p * C;
                const int px = min(max(xs + p, 0), (int)W - 1);
                s[e] = A::bits((Acc)g[(size_t)px * C + c]);
            }
        }
    }
    __syncthreads();

    constexpr int ITEMS = (ROW_XT / GT_R) * C;
    for (int item = warp; item < ITEMS; item += GT_THREADS / 32) {
        const int xb = item / C, c = item - xb * C;
        Acc acc[GT_R];
#pragma unroll
        for (int r = 0; r < GT_R; ++r) acc[r] = (Acc)0;
        conv8<Acc, C>(s_in + lane * in_pitch + (xb * GT_R) * C + c, nch, tw, acc);
        uint32_t *o = s_out + lane * OUT_PITCH + (xb * GT_R) * C + c;
#pragma unroll
        for (int r = 0; r < GT_R; ++r) {
            if constexpr (sizeof(Mid) == 4) o[r * C] = A::bits(acc[r]);
            else o[r * C] = (uint32_t)St<Mid, Acc>::f(acc[r]);
        }
    }
    __syncthreads();

    const int out_elems = min(ROW_XT, (int)W - x0) * C;
    for (int r = warp; r < ROW_ROWS; r += GT_THREADS / 32) {
        const size_t row = row0 + r;
        if (row >= total_rows) break;
        Mid *g = mid + row * g_pitch + (size_t)x0 * C;
        const uint32_t *s = s_out + r * OUT_PITCH;
        for (int e = lane; e < out_elems; e += 32) {
            if constexpr (sizeof(Mid) == 4) reinterpret_cast<uint32_t *>(g)[e] = s[e];
            else g[e] = (Mid)s[e];
        }
    }
}

constexpr int COL_TH = 128;  // output rows per CTA tile (column pass)
constexpr int COL_TW = 64;   // subpixel columns per CTA tile

template <typename Mid, typename Out, typename Acc>
__global__ void __launch_bounds__(GT_THREADS)
col_tiled_kernel(const Mid *__restrict__ mid, Out *__restrict__ dst, uint32_t row_sub, uint32_t H, int tiles_y, int half,
                 int nch, int zero_first_row, const __grid_constant__ TapWeights tw) {
    using A = Arith<Acc>;
    extern __shared__ __align__(16) uint32_t sm[];  // [COL_TH + nch*8 - 1][COL_TW]
    const int tid = threadIdx.x;
    const uint32_t s0 = blockIdx.x * COL_TW;
    const uint32_t img = blockIdx.y / tiles_y;
    const int y0 = (int)(blockIdx.y % tiles_y) * COL_TH;
    const int sr = COL_TH + nch * GT_R - 1;
    const Mid *m = mid + (size_t)img * H * row_sub;

    {
        static_assert(GT_THREADS % COL_TW == 0, "one staging thread per column");
        const int s = tid % COL_TW;
        const uint32_t col = min(s0 + s, row_sub - 1);
        for (int r = tid / COL_TW; r < sr; r += GT_THREADS / COL_TW) {
            const int y = min(max(y0 - half + r, 0), (int)H - 1);
            sm[r * COL_TW + s] = A::bits((Acc)m[(size_t)y * row_sub + col]);
        }
    }
    __syncthreads();

    Out *d = dst + (size_t)img * H * row_sub;
    for (int item = tid; item < COL_TW * (COL_TH / GT_R); item += GT_THREADS) {
        const int s = item % COL_TW, rb = item / COL_TW;
        Acc acc[GT_R];
#pragma unroll
        for (int r = 0; r < GT_R; ++r) acc[r] = (Acc)0;
        conv8<Acc, COL_TW>(sm + (rb * GT_R) * COL_TW + s, nch, tw, acc);
        if (s0 + s < row_sub) {
#pragma unroll
            for (int r = 0; r < GT_R; ++r) {
                const int y = y0 + rb * GT_R + r;
                if (y < (int)H) d[(size_t)y * row_sub + s0 + s] = (zero_first_row && y == 0) ? (Out)0 : St<Out, Acc>::f(acc[r]);
            }
        }
    }
}

// ---- f32 engine, FFMA2 kernels ------------------------------------------------------------------
// They also run the Q0.15 engine for u8 pixels: weights <= 32767 and pixels <= 255 are integers,
// every product is < 2^23 and every partial sum <= 255 * 32768 < 2^24, so f32 arithmetic on them
// is exact in any order and the kernel is bit-identical to the reference's u32 accumulator
// (src/imageops/filter_1d.rs:132-150, 396-408).  (Not so for u16: 65535 * 32767 needs 31 bits.)
template <typename Out, bool Q15> struct StF;
template <typename Out> struct StF<Out, false> { static __device__ __forceinline__ Out f(float a) { return St<Out, float>::f(a); } };
template <> struct StF<uint8_t, true> {
    static __device__ __forceinline__ uint8_t f(float a) { return (uint8_t)min((__float2uint_rz(a) + 16384u) >> 15, 255u); }
};

// Row pass: lane = row, warp item = (pixel blocks xb and xb + 4, channel): 16 outputs per thread.
template <typename In, typename Mid, int C, bool Q15, int TC, bool RING>
__global__ void __launch_bounds__(GT_THREADS)
row_f32x2_kernel(const In *__restrict__ src, Mid *__restrict__ mid, uint32_t W, size_t total_rows, int half,
                 int shift, int nch, int in_pitch, const __grid_constant__ TapWeights2 tw) {
    constexpr int OUT_PITCH = (ROW_XT * C) | 1;
    extern __shared__ __align__(16) uint32_t sm[];
    float *s_in = reinterpret_cast<float *>(sm);          // [ROW_ROWS][in_pitch]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int x0 = blockIdx.x * ROW_XT;
    const size_t row0 = (size_t)blockIdx.y * ROW_ROWS;
    const int sp = ROW_XT + nch * TC - 1;
    const int row_elems = sp * C;
    const size_t g_pitch = (size_t)W * C;
    const int xs = x0 - half - shift;
    const bool interior = xs >= 0 && xs + sp <= (int)W;
    const bool words_ok = (g_pitch * sizeof(In)) % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0;

    bool staged = false;
    if constexpr (sizeof(In) < 4) {
        if (interior && words_ok) {  // the warp's four rows at once, all loads in flight together
            constexpr int NR = ROW_ROWS / (GT_THREADS / 32);
            const In *gr[NR];
            float *sr[NR];
#pragma unroll
            for (int k = 0; k < NR; ++k) {
                const int r = warp + k * (GT_THREADS / 32);
                gr[k] = src + min(row0 + r, total_rows - 1) * g_pitch + (size_t)xs * C;
                sr[k] = s_in + r * in_pitch;
            }
            staged = stage_convert_rows<In, NR, 5>(gr, sr, row_elems, lane);
        }
    }
    for (int r = warp; r < ROW_ROWS && !staged; r += GT_THREADS / 32) {
        const In *g = src + min(row0 + r, total_rows - 1) * g_pitch;
        float *s = s_in + r * in_pitch;
        if (interior && words_ok) {
            stage_convert_row<In>(g + (size_t)xs * C, s, row_elems, lane);
        } else {
            for (int e = lane; e < row_elems; e += 32) {
                const int p = e / C, c = e - p * C;
                const int px = min(max(xs + p, 0), (int)W - 1);
                s[e] = (float)g[(size_t)px * C + c];
            }
        }
    }
    __syncthreads();

    // Results stay in registers until every warp is done with s_in, then go out through the same
    // shared memory (s_out aliases s_in): 52 KB per CTA instead of 77 KB, i.e. 4 CTAs per SM.
    constexpr int HALF_BLOCKS = ROW_XT / GT_R / 2;  // 4: item pairs block xb with block xb + 4
    constexpr int ITEMS = HALF_BLOCKS * C;
    constexpr int PER_WARP = (ITEMS + GT_THREADS / 32 - 1) / (GT_THREADS / 32);
    f32x2 acc[PER_WARP][GT_R];
#pragma unroll
    for (int it = 0; it < PER_WARP; ++it) {
        const int item = warp + it * (GT_THREADS / 32);
#pragma unroll
        for (int r = 0; r < GT_R; ++r) acc[it][r] = 0ull;
        if (item < ITEMS) {
            const int xb = item / C, c = item - xb * C;
            const float *a = s_in + lane * in_pitch + (xb * GT_R) * C + c;
            conv8x2<C, false, TC, RING>(a, a + HALF_BLOCKS * GT_R * C, nch, tw, acc[it]);
        }
    }
    __syncthreads();
    float *s_out = s_in;  // [ROW_ROWS][OUT_PITCH]
#pragma unroll
    for (int it = 0; it < PER_WARP; ++it) {
        const int item = warp + it * (GT_THREADS / 32);
        if (item < ITEMS) {
            const int xb = item / C, c = item - xb * C;
            float *o = s_out + lane * OUT_PITCH + (xb * GT_R) * C + c;
#pragma unroll
            for (int r = 0; r < GT_R; ++r) unpack2(acc[it][r], o[r * C], o[(HALF_BLOCKS * GT_R + r) * C]);
        }
    }
    __syncthreads();

    const int out_elems = min(ROW_XT, (int)W - x0) * C;
    for (int r = warp; r < ROW_ROWS; r += GT_THREADS / 32) {
        const size_t row = row0 + r;
        if (row >= total_rows) break;
        Mid *g = mid + row * g_pitch + (size_t)x0 * C;
        const float *s = s_out + r * OUT_PITCH;
#pragma unroll 4
        for (int e = lane; e < out_elems; e += 32) g[e] = StF<Mid, Q15>::f(s[e]);
    }
}

// Column pass: thread = two adjacent subpixel columns x 8 rows; one LDS.64 per staged row.
template <typename Mid, typename Out, bool Q15, int TC, bool RING>
__global__ void __launch_bounds__(GT_THREADS)
col_f32x2_kernel(const Mid *__restrict__ mid, Out *__restrict__ dst, uint32_t row_sub, uint32_t H, int tiles_y, int half,
                 int nch, int zero_first_row, const __grid_constant__ TapWeights2 tw) {
    extern __shared__ __align__(16) uint32_t sm[];
    float *s = reinterpret_cast<float *>(sm);  // [COL_TH + nch*8 - 1][COL_TW]
    const int tid = threadIdx.x;
    const uint32_t s0 = blockIdx.x * COL_TW;
    const uint32_t img = blockIdx.y / tiles_y;
    const int y0 = (int)(blockIdx.y % tiles_y) * COL_TH;
    const int sr = COL_TH + nch * TC - 1;
    const Mid *m = mid + (size_t)img * H * row_sub;

    // stage [sr][COL_TW] with clamped rows; 16-byte async copies when the rows allow it
    const bool vec = sizeof(Mid) == 4 && (row_sub % 4 == 0) && (s0 + COL_TW <= row_sub) && ((reinterpret_cast<uintptr_t>(m) & 15) == 0);
    if (vec) {
        constexpr int VPR = COL_TW / 4;  // 16-byte vectors per staged row
        for (int i = tid; i < sr * VPR; i += GT_THREADS) {
            const int r = i / VPR, v = i - r * VPR;
            const int y = min(max(y0 - half + r, 0), (int)H - 1);
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(s + r * COL_TW + v * 4);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(m + (size_t)y * row_sub + s0 + v * 4) : "memory");
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else if (sizeof(Mid) == 1 && (row_sub % 4 == 0) && (s0 + COL_TW <= row_sub) && ((reinterpret_cast<uintptr_t>(m) & 3) == 0)) {
        // u8 intermediate (Q0.15 engine): a 32-bit word = four subpixels -> four floats through the
        // 2^23 trick, stored as one 16-byte vector; four words in flight per thread
        constexpr int WPR = COL_TW / 4;  // words per staged row
        const int total = sr * WPR;
        for (int i0 = tid; i0 < total; i0 += 4 * GT_THREADS) {
            uint32_t wv[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = i0 + k * GT_THREADS;
                const int r = i / WPR, v = i - r * WPR;
                const int y = min(max(y0 - half + r, 0), (int)H - 1);
                wv[k] = i < total ? __ldg(reinterpret_cast<const uint32_t *>(m + (size_t)y * row_sub + s0) + v) : 0u;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int i = i0 + k * GT_THREADS;
                if (i < total) {
                    float4 f;
                    f.x = __uint_as_float(__byte_perm(wv[k], 0x4B000000u, 0x7540)) - 8388608.0f;
                    f.y = __uint_as_float(__byte_perm(wv[k], 0x4B000000u, 0x7541)) - 8388608.0f;
                    f.z = __uint_as_float(__byte_perm(wv[k], 0x4B000000u, 0x7542)) - 8388608.0f;
                    f.w = __uint_as_float(__byte_perm(wv[k], 0x4B000000u, 0x7543)) - 8388608.0f;
                    *reinterpret_cast<float4 *>(s + i * 4) = f;   // i * 4 = r * COL_TW + v * 4
                }
            }
        }
    } else {
        const int c = tid % COL_TW;
        const uint32_t col = min(s0 + c, row_sub - 1);
        for (int r = tid / COL_TW; r < sr; r += GT_THREADS / COL_TW) {
            const int y = min(max(y0 - half + r, 0), (int)H - 1);
            s[r * COL_TW + c] = (float)m[(size_t)y * row_sub + col];
        }
    }
    __syncthreads();

    Out *d = dst + (size_t)img * H * row_sub;
    for (int item = tid; item < (COL_TW / 2) * (COL_TH / GT_R); item += GT_THREADS) {
        const int cp = item % (COL_TW / 2), rb = item / (COL_TW / 2);
        f32x2 acc[GT_R];
#pragma unroll
        for (int r = 0; r < GT_R; ++r) acc[r] = 0ull;
        const float *a = s + (rb * GT_R) * COL_TW + 2 * cp;
        conv8x2<COL_TW, true, TC, RING>(a, a + 1, nch, tw, acc);
        const uint32_t col = s0 + 2 * cp;
#pragma unroll
        for (int r = 0; r < GT_R; ++r) {
            const int y = y0 + rb * GT_R + r;
            if (y >= (int)H) break;
            float v0, v1;
            unpack2(acc[r], v0, v1);
            Out *o = d + (size_t)y * row_sub + col;
            const bool z = zero_first_row && y == 0;
            if (col < row_sub) o[0] = z ? (Out)0 : StF<Out, Q15>::f(v0);
            if (col + 1 < row_sub) o[1] = z ? (Out)0 : StF<Out, Q15>::f(v1);
        }
    }
}

}  // namespace

// Launches both passes.  rk / ck are the scanned kernels (f32 bits or Q0.15), lengths odd <= GT_MAX_TAPS.
// Returns 1 (nothing launched) when a pass's tile does not fit shared memory: the caller falls back.
template <typename T, typename Mid, typename Acc>
int gauss_tiled_run(const T *src, T *dst, Mid *mid, uint32_t w, uint32_t h, int C, uint32_t n_images,
                    const uint32_t *rk, int rlen, const uint32_t *ck, int clen, int row_shift, int zero_first_row,
                    cudaStream_t st) {
    // Q0.15 on u8: exact in f32 (see StF), so it rides the FFMA2 kernels with integer-valued weights
    constexpr bool Q15F = std::is_same<Acc, uint32_t>::value && std::is_same<T, uint8_t>::value;
    constexpr bool F32X2 = std::is_same<Acc, float>::value || Q15F;
    // taps per unrolled chunk: the FFMA2 kernels pick the chunk size with the least zero padding
    auto best_tc = [](int len) {
        int best = GT_R, waste = (len + GT_R - 1) / GT_R * GT_R - len;
        for (int tc = GT_R - 1; tc >= 5; --tc) {
            const int w_ = (len + tc - 1) / tc * tc - len;
            if (w_ < waste) { waste = w_; best = tc; }
        }
        return best;
    };
    // Measured on B200 (65 and 21 taps): the row pass is fastest with the least padding (13 x 5, 3 x 7),
    // the column pass with 8-tap chunks even when that pads (fewer, longer unrolled bodies).
    const int rtc = F32X2 ? best_tc(rlen) : GT_R, ctc = GT_R;
    const int rnch = (rlen + rtc - 1) / rtc, cnch = (clen + ctc - 1) / ctc;
    const size_t total_rows = (size_t)h * n_images;
    const uint32_t row_sub = w * (uint32_t)C;
    const int tiles_y = (int)((h + COL_TH - 1) / COL_TH);
    const int sp = ROW_XT + rnch * rtc - 1;
    const int in_pitch = (sp * C + 4) | 1, out_pitch = (ROW_XT * C) | 1;
    const size_t row_smem_int = (size_t)ROW_ROWS * (in_pitch + out_pitch) * 4;
    const size_t row_smem = (size_t)ROW_ROWS * std::max(in_pitch, out_pitch) * 4;  // f32 kernels alias the two
    const size_t col_smem = (size_t)(COL_TH + cnch * ctc - 1) * COL_TW * 4;
    constexpr size_t SMEM_MAX = 227 * 1024;  // per-CTA opt-in limit on sm_100
    if ((F32X2 ? row_smem : row_smem_int) > SMEM_MAX || col_smem > SMEM_MAX) return 1;
    const dim3 row_grid((w + ROW_XT - 1) / ROW_XT, (unsigned)((total_rows + ROW_ROWS - 1) / ROW_ROWS));
    const dim3 col_grid((row_sub + COL_TW - 1) / COL_TW, (unsigned)(tiles_y * n_images));

    constexpr int RING_MIN_CHUNKS = 6;  // ring-window kernels from this many tap chunks on (see conv8x2)
    if constexpr (F32X2) {
        TapWeights2 rwl, cwl;
        for (int i = 0; i < GT_MAX_TAPS; ++i) {
            uint32_t a = i < rlen ? rk[i] : 0u, b = i < clen ? ck[i] : 0u;
            if (Q15F) {  // integer weights -> the same integers as floats
                const float fa = (float)a, fb = (float)b;
                memcpy(&a, &fa, 4);
                memcpy(&b, &fb, 4);
            }
            rwl.w[i] = ((f32x2)a << 32) | a;
            cwl.w[i] = ((f32x2)b << 32) | b;
        }
#define ROW_GO2(CC, TCC)                                                                                             \
    {                                                                                                                \
        auto k = rnch >= RING_MIN_CHUNKS ? row_f32x2_kernel<T, Mid, CC, Q15F, TCC, true> : row_f32x2_kernel<T, Mid, CC, Q15F, TCC, false>; \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row_smem));              \
        k<<<row_grid, GT_THREADS, row_smem, st>>>(src, mid, w, total_rows, rlen / 2, row_shift, rnch, in_pitch, rwl); \
    }
#define ROW_GO(CC)                                                                                                   \
    switch (rtc) {                                                                                                   \
        case 5: ROW_GO2(CC, 5) break;                                                                                \
        case 6: ROW_GO2(CC, 6) break;                                                                                \
        case 7: ROW_GO2(CC, 7) break;                                                                                \
        default: ROW_GO2(CC, 8) break;                                                                               \
    }
        switch (C) {
            case 1: ROW_GO(1) break;
            case 2: ROW_GO(2) break;
            case 3: ROW_GO(3) break;
            default: ROW_GO(4) break;
        }
#undef ROW_GO
#undef ROW_GO2
#define COL_GO(TCC)                                                                                                  \
    {                                                                                                                \
        auto k = cnch >= RING_MIN_CHUNKS ? col_f32x2_kernel<Mid, T, Q15F, TCC, true> : col_f32x2_kernel<Mid, T, Q15F, TCC, false>; \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)col_smem));              \
        k<<<col_grid, GT_THREADS, col_smem, st>>>((const Mid *)mid, dst, row_sub, h, tiles_y, clen / 2, cnch, zero_first_row, cwl); \
    }
        switch (ctc) {
            case 5: COL_GO(5) break;
            case 6: COL_GO(6) break;
            case 7: COL_GO(7) break;
            default: COL_GO(8) break;
        }
#undef COL_GO
    } else {
        TapWeights rw, cw;
        for (int i = 0; i < GT_MAX_TAPS; ++i) { rw.w[i] = i < rlen ? rk[i] : 0u; cw.w[i] = i < clen ? ck[i] : 0u; }
#define ROW_GO(CC)                                                                                                   \
    {                                                                                                                \
        auto k = row_tiled_kernel<T, Mid, Acc, CC>;                                                                  \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)row_smem_int));          \
        k<<<row_grid, GT_THREADS, row_smem_int, st>>>(src, mid, w, total_rows, rlen / 2, row_shift, rnch, in_pitch, rw); \
    }
        switch (C) {
            case 1: ROW_GO(1) break;
            case 2: ROW_GO(2) break;
            case 3: ROW_GO(3) break;
            default: ROW_GO(4) break;
        }
#undef ROW_GO
        auto k = col_tiled_kernel<Mid, T, Acc>;
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)col_smem));
        k<<<col_grid, GT_THREADS, col_smem, st>>>(mid, dst, row_sub, h, tiles_y, clen / 2, cnch, zero_first_row, cw);
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

#define INST(T, Mid, Acc)                                                                                              \
    template int gauss_tiled_run<T, Mid, Acc>(const T *, T *, Mid *, uint32_t, uint32_t, int, uint32_t, const uint32_t *, \
                                              int, const uint32_t *, int, int, int, cudaStream_t);
INST(uint8_t, float, float)
INST(uint16_t, float, float)
INST(float, float, float)
INST(uint8_t, uint8_t, uint32_t)
INST(uint16_t, uint16_t, uint32_t)
#undef INST

}  // namespace b200
