// boxblur_u8.cu -- the u8 box-blur passes of fast_blur, word-at-a-time.
//
// Same arithmetic as the reference's u8 accumulator (src/imageops/fast_blur.rs:36-71: u32 window
// sums, store = ((sum + ks/2) * ceil(2^32/ks)) >> 32; horizontal pass 303-413, vertical pass
// 416-473) -- window sums of integers, so bit-identical however they are accumulated -- but laid
// out for the machine: the generic kernels in boxblur.cu spend ~20 instructions per subpixel on
// byte-granular loads, stores and index arithmetic; here every shared/global access is a 32-bit
// word and the per-byte work is two PRMTs, one IADD3 and one IMAD.HI.
//
//   vertical:   thread = one word column (4 subpixels) x a band of rows.  The four running sums
//               live in two registers as 16-bit lanes (sum + ks/2 <= 255*255 + 127 < 2^16), so the
//               sliding update is 4 PRMT + 2 IADD3 per row.
//   horizontal: a CTA stages 32 rows x 256 pixels (+ halo) word-aligned in shared memory (funnel
//               shift from the rows' global alignment); lane = row; a thread slides all C channel
//               sums along a 64-pixel chunk in steps of 4 pixels = C whole words in, C whole words
//               out, fully unrolled.
// Taken when ks <= 255 and rows are 4-byte aligned; boxblur.cu's generic kernels cover the rest.
#include "common.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>

namespace b200 {
namespace {

constexpr int V_THREADS = 256;
constexpr int V_BAND = 64;       // rows per CTA band of the vertical pass (halved for small jobs, see box_pass_u8)
constexpr int H_THREADS = 128;   // 4 warps = the 4 chunks of a tile; small jobs run 16-pixel chunks on 16 warps
constexpr int H_ROWS = 32;
constexpr int H_XT = 256;
constexpr int H_XC = 64;

struct U8Geom {
    uint32_t w, h;
    uint32_t r, ks, recip, bias;
};

__device__ __forceinline__ uint32_t even_lanes(uint32_t w) { return __byte_perm(w, 0, 0x4240); }  // [b0, 0, b2, 0]
__device__ __forceinline__ uint32_t odd_lanes(uint32_t w) { return __byte_perm(w, 0, 0x4341); }   // [b1, 0, b3, 0]
__device__ __forceinline__ uint32_t byte_of(uint32_t w, int j) { return __byte_perm(w, 0, 0x4440 + j); }
// c + sum of (unsigned bytes of a) x (signed bytes of b): with b = 1 << 8k it adds byte k of a, with 0xFF << 8k it subtracts it
__device__ __forceinline__ uint32_t dp4a_us(uint32_t a, int b, uint32_t c) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"((int)c));
    return (uint32_t)d;
}

// ---- vertical ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(V_THREADS)
box_v_u8_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, U8Geom g, uint32_t words, int bands, int vband) {
    const uint32_t wd = blockIdx.x * V_THREADS + threadIdx.x;
    if (wd >= words) return;
    const uint32_t img = blockIdx.y / bands, band = blockIdx.y % bands;
    const int y0 = (int)band * vband, y1 = min(y0 + vband, (int)g.h);
    const int r = (int)g.r, hmax = (int)g.h - 1;
    const uint32_t *p = src + (size_t)img * g.h * words + wd;
    uint32_t *d = dst + (size_t)img * g.h * words + wd;

    // lanes carry sum + ks/2
    uint32_t se = g.bias | (g.bias << 16), so = se;
    {
        const int lo = y0 - r, hi = y0 + r;
        const int nlo = lo < 0 ? -lo : 0, nhi = hi > hmax ? hi - hmax : 0;
        int y = max(lo, 0);
        const int yend = min(hi, hmax);
        constexpr int PB = 8;  // loads batched: a load-use-per-iteration loop serialises on DRAM latency
        for (; y + PB - 1 <= yend; y += PB) {
            uint32_t v[PB];
#pragma unroll
            for (int k = 0; k < PB; ++k) v[k] = __ldg(p + (size_t)(y + k) * words);
#pragma unroll
            for (int k = 0; k < PB; ++k) { se += even_lanes(v[k]); so += odd_lanes(v[k]); }
        }
        for (; y <= yend; ++y) { const uint32_t v = __ldg(p + (size_t)y * words); se += even_lanes(v); so += odd_lanes(v); }
        if (nlo) { const uint32_t v = __ldg(p); se += even_lanes(v) * (uint32_t)nlo; so += odd_lanes(v) * (uint32_t)nlo; }
        if (nhi) { const uint32_t v = __ldg(p + (size_t)hmax * words); se += even_lanes(v) * (uint32_t)nhi; so += odd_lanes(v) * (uint32_t)nhi; }
    }
    auto emit = [&](int y) {
        const uint32_t b0 = __umulhi(se & 0xffffu, g.recip), b2 = __umulhi(se >> 16, g.recip);
        const uint32_t b1 = __umulhi(so & 0xffffu, g.recip), b3 = __umulhi(so >> 16, g.recip);
        d[(size_t)y * words] = __byte_perm(__byte_perm(b0, b1, 0x0040), __byte_perm(b2, b3, 0x0040), 0x5410);
    };
    constexpr int U = 4;
    int y = y0;
    for (; y + U <= y1; y += U) {
        uint32_t a[U], b[U];
#pragma unroll
        for (int k = 0; k < U; ++k) {
            a[k] = __ldg(p + (size_t)min(y + k + r + 1, hmax) * words);
            b[k] = __ldg(p + (size_t)max(y + k - r, 0) * words);
        }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            emit(y + k);
            se = se + even_lanes(a[k]) - even_lanes(b[k]);
            so = so + odd_lanes(a[k]) - odd_lanes(b[k]);
        }
    }
    for (; y < y1; ++y) {
        emit(y);
        const uint32_t a = __ldg(p + (size_t)min(y + r + 1, hmax) * words), b = __ldg(p + (size_t)max(y - r, 0) * words);
        se = se + even_lanes(a) - even_lanes(b);
        so = so + odd_lanes(a) - odd_lanes(b);
    }
}

// ---- horizontal ----------------------------------------------------------------------------------
template <int C>
__global__ void __launch_bounds__(512)
box_h_u8_kernel(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, U8Geom g, size_t total_rows, int in_pitch,
                int out_pitch, int xc) {
    const int nwarps = blockDim.x >> 5;   // = H_XT / xc
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t *s_in = sm;                                 // [H_ROWS][in_pitch] words; staged pixel 0 = source pixel x0 - r
    uint32_t *s_out = sm + (size_t)H_ROWS * in_pitch;    // [H_ROWS][out_pitch] words; pixel 0 = x0
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int x0 = blockIdx.x * H_XT;
    const size_t row0 = (size_t)blockIdx.y * H_ROWS;
    const int r = (int)g.r, W = (int)g.w;
    const int tw = min(H_XT, W - x0);
    const int sp = tw + 2 * r;                           // staged pixels (clamped at the image edges)
    const int sbytes = sp * C;
    const size_t g_pitch = (size_t)W * C;                // bytes; multiple of 4 (checked by the host)
    const bool interior = x0 - r >= 0 && x0 + tw + r <= W;

    // Interior tiles are staged with asynchronous 4-byte copies of the ALIGNED global words (every
    // copy of the tile in flight at once); the rows' byte phase -- the same for all rows, the pitch
    // being a multiple of 4 -- is taken out by funnel shifts when the words are used.
    int phase = 0;
    if (interior) {
        const uint8_t *gb0 = src + (size_t)(x0 - r) * C;
        phase = (int)(reinterpret_cast<uintptr_t>(gb0) & 3);
    }
    const int swords = (phase + sbytes + 3) / 4;
    for (int rr = warp; rr < H_ROWS; rr += nwarps) {
        const uint8_t *grow = src + min(row0 + rr, total_rows - 1) * g_pitch;
        uint32_t *s = s_in + (size_t)rr * in_pitch;
        if (interior) {
            const uint32_t *gw = reinterpret_cast<const uint32_t *>(grow + (size_t)(x0 - r) * C - phase);
            for (int wi = lane; wi < swords; wi += 32) {
                const uint32_t d = (uint32_t)__cvta_generic_to_shared(s + wi);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d), "l"(gw + wi) : "memory");
            }
        } else {
            uint8_t *sb = reinterpret_cast<uint8_t *>(s);
            for (int e = lane; e < sbytes; e += 32) {
                const int p = e / C, c = e - p * C;
                sb[e] = grow[(size_t)min(max(x0 - r + p, 0), W - 1) * C + c];
            }
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();

    // lane = row; item = 64-pixel chunk; steps of 4 pixels = C words
    const int chunks = (tw + xc - 1) / xc;
    const uint32_t *in_row = s_in + (size_t)lane * in_pitch;
    uint32_t *out_row = s_out + (size_t)lane * out_pitch;
    const int pshift = phase * 8;                                // outgoing pixels: byte phase of the staged row
    const int nbyte = ((2 * r + 1) * C + phase) & 3;             // incoming pixels: phase + window offset
    for (int ck = warp; ck < chunks; ck += nwarps) {
        const int xa = ck * xc, xb = min(xa + xc, tw);
        uint32_t sum[C];
#pragma unroll
        for (int c = 0; c < C; ++c) sum[c] = g.bias;
        {
            const uint8_t *b = reinterpret_cast<const uint8_t *>(in_row) + phase + xa * C;
            for (int j = 0; j <= 2 * r; ++j) {
#pragma unroll
                for (int c = 0; c < C; ++c) sum[c] += b[j * C + c];
            }
        }
        const uint32_t *pw = in_row + (xa * C + phase) / 4;                    // outgoing pixels, aligned down
        const uint32_t *nw = in_row + ((xa + 2 * r + 1) * C + phase) / 4;      // incoming pixels, aligned down
        for (int x = xa; x < xb; x += 4) {
            uint32_t prev[C], next[C], praw[C + 1], nraw[C + 1], outw[C];
#pragma unroll
            for (int k = 0; k <= C; ++k) { praw[k] = pw[k]; nraw[k] = nw[k]; }
#pragma unroll
            for (int k = 0; k < C; ++k) {
                prev[k] = __funnelshift_r(praw[k], praw[k + 1], pshift);
                next[k] = __funnelshift_r(nraw[k], nraw[k + 1], nbyte * 8);
            }
            uint32_t ob[4 * C];
#pragma unroll
            for (int b = 0; b < 4 * C; ++b) {
                const int c = b % C;
                ob[b] = __umulhi(sum[c], g.recip);
                sum[c] = dp4a_us(next[b / 4], 1 << (8 * (b & 3)), sum[c]);
                sum[c] = dp4a_us(prev[b / 4], (int)(0xFFu << (8 * (b & 3))), sum[c]);
            }
#pragma unroll
            for (int k = 0; k < C; ++k)
                outw[k] = __byte_perm(__byte_perm(ob[4 * k], ob[4 * k + 1], 0x0040), __byte_perm(ob[4 * k + 2], ob[4 * k + 3], 0x0040), 0x5410);
            uint32_t *ow = out_row + (x * C) / 4;
#pragma unroll
            for (int k = 0; k < C; ++k) ow[k] = outw[k];
            pw += C;
            nw += C;
        }
    }
    __syncthreads();

    const int obytes = tw * C;
    for (int rr = warp; rr < H_ROWS; rr += nwarps) {
        const size_t row = row0 + rr;
        if (row >= total_rows) break;
        uint8_t *gout = dst + row * g_pitch + (size_t)x0 * C;  // 4-byte aligned: g_pitch % 4 == 0, x0 % 4 == 0
        const uint32_t *s = s_out + (size_t)rr * out_pitch;
        const int full = obytes / 4;
        for (int wi = lane; wi < full; wi += 32) reinterpret_cast<uint32_t *>(gout)[wi] = s[wi];
        if (lane < obytes - 4 * full) gout[4 * full + lane] = reinterpret_cast<const uint8_t *>(s)[4 * full + lane];
    }
}

int odd_words(int bytes) { return (((bytes + 3) / 4) | 1); }

// ---- horizontal + vertical pass of one box in one kernel ---------------------------------------------
// The two-kernel form writes the horizontally blurred image to HBM and reads it back: per box 4 bytes of
// traffic per subpixel for a 1-byte-in / 1-byte-out operator.  Here a CTA owns a strip of F_XT pixels and
// streams down a band of rows in chunks of 32: (1) the chunk's source rows (+ r pixels either side) are
// staged by asynchronous word copies; (2) the HORIZONTAL pass (lane = row, warp = chunk of F_XC pixels,
// sliding sums as above) writes its quantised u8 rows into a ring of rows in shared memory; (3) the
// VERTICAL pass (thread = one word column, four running sums updated with dp4a) takes
// the chunk's rows from the ring one by one: emit a destination row, add the incoming row, subtract the
// row 2r + 1 back.  Running sums persist from chunk to chunk, so inside a band nothing is computed twice;
// a band primes its window with 2r + 1 rows.  The quantised store between the passes is the reference's
// (fast_blur.rs:36-71), so the result stays bit-identical.  Rows above / below the image are the first /
// last row (fast_blur.rs:416-473): ring row q holds the horizontal pass of source row clamp(q).
// strip width in pixels: 128 for three and four channels (small enough for four or five CTAs per SM, which overlap one another's
// staging waits); one and two channels take wider strips so that the vertical pass (a thread per 32-bit word column) has all
// 128 threads busy and every warp gets several horizontal items per chunk
__host__ __device__ constexpr int f_xt(int C) { return C == 1 ? 512 : C == 2 ? 256 : 128; }
constexpr int F_XC = 32;       // pixels per horizontal-pass warp item
constexpr int F_WARPS = 4;
constexpr int F_THREADS = F_WARPS * 32;
constexpr int F_CH = 32;       // rows per chunk = lanes of the horizontal pass

struct FusedBoxArgs {
    const uint8_t *src;
    uint8_t *dst;
    U8Geom g;
    int bands, band_rows;      // rows of a band (the last may be shorter)
    int in_pitch, ring_pitch;  // words per staged row / ring row (odd)
    int ring_rows;             // multiple of 32, >= 32 + 2r + 1
};

template <int C>
__global__ void __launch_bounds__(F_THREADS)
box_fused_u8_kernel(const FusedBoxArgs a) {
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t *s_in = sm;                                    // [32][in_pitch]
    uint32_t *ring = sm + (size_t)F_CH * a.in_pitch;        // [ring_rows][ring_pitch]
    const uint32_t s_in_u32 = (uint32_t)__cvta_generic_to_shared(s_in);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const U8Geom &g = a.g;
    const int r = (int)g.r, W = (int)g.w, H = (int)g.h, ks = (int)g.ks;
    constexpr int F_XT = f_xt(C);
    const int x0 = blockIdx.x * F_XT;
    const int band = blockIdx.y % a.bands;
    const size_t img = blockIdx.y / a.bands;
    const int b0 = band * a.band_rows, b1 = min(b0 + a.band_rows, H);
    if (b0 >= b1) return;
    const int tw = min(F_XT, W - x0);
    const int sp = tw + 2 * r;
    const int sbytes = sp * C;
    const size_t g_pitch = (size_t)W * C;                   // bytes; multiple of 4 (checked by the host)
    const uint8_t *src = a.src + img * (size_t)H * g_pitch;
    uint8_t *dst = a.dst + img * (size_t)H * g_pitch;
    // A staged row = sp pixels from source pixel x0 - r, clamped at the image's edges.  The pixels inside the image are
    // one contiguous span of the source row: it is copied with asynchronous copies of ALIGNED global words (every copy
    // of the chunk in flight at once); the staged row's byte phase is chosen so that the span lands on the same byte
    // phase in shared memory, and is taken out by funnel shifts when the words are used.  The clamped pixels (edge
    // strips only) are filled from the span's first / last pixel afterwards.
    const int lo_px = max(x0 - r, 0), hi_px = min(x0 + tw + r, W);
    const int e_off = (lo_px - (x0 - r)) * C, n_in = (hi_px - lo_px) * C;     // the span's place and length in the staged row (bytes)
    const int a_g = (int)(reinterpret_cast<uintptr_t>(src + (size_t)lo_px * C) & 3);  // the span's phase in global memory (same for every row)
    const int phase = (a_g - e_off) & 3;                                        // staged byte i lives at shared byte phase + i
    const int span_words = (a_g + n_in + 3) / 4;
    const int span_w0 = (phase + e_off - a_g) / 4;                              // first shared word the span's words go to
    const int owords = (tw * C + 3) / 4;                    // destination words of a strip row (tw * C is a multiple of 4 unless the row ends)
    const int q0 = b0 - r;                                  // source row (before clamping) of ring row 0
    const int nq = (b1 - b0) + 2 * r + 1;                   // ring rows the band consumes (the last one only triggers the last emit)
    const int nchunks = (nq + F_CH - 1) / F_CH;
    const int pshift = phase * 8, nbyte = ((2 * r + 1) * C + phase) & 3;
    const int hchunks = (tw + F_XC - 1) / F_XC;

    // vertical state of this thread's word column
    uint32_t sv[4] = {g.bias, g.bias, g.bias, g.bias};   // running sums of the word's four subpixels
    uint32_t *dcol = reinterpret_cast<uint32_t *>(dst + (size_t)b0 * g_pitch + (size_t)x0 * C) + tid;
    const size_t dwords = g_pitch / 4;
    int primed = 0;     // rows added so far (the window is full at ks)
    int out_left = b1 - b0;

    // staging of chunk ci: asynchronous, in flight while the vertical pass of the chunk before it runs
    auto stage = [&](int ci) {
    {
        // a warp stages rows warp, warp + 4, ...: the eight row pointers first, then one sweep over the span's words with
        // the eight copies of a word column back to back (per-row address arithmetic once per chunk, not once per copy)
        constexpr int RPW = F_CH / F_WARPS;
        const uint8_t *grow[RPW];
#pragma unroll
        for (int i = 0; i < RPW; ++i) {
            const int y = min(max(q0 + ci * F_CH + warp + F_WARPS * i, 0), H - 1);
            grow[i] = src + (size_t)y * g_pitch + (size_t)lo_px * C - a_g;
        }
        const uint32_t d0 = s_in_u32 + (uint32_t)((warp * a.in_pitch + span_w0) * 4);
        const uint32_t dstep = (uint32_t)(F_WARPS * a.in_pitch * 4);
#pragma unroll 1
        for (int wi = lane; wi < span_words; wi += 32) {
#pragma unroll
            for (int i = 0; i < RPW; ++i)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d0 + i * dstep + 4u * wi), "l"(grow[i] + 4 * (size_t)wi) : "memory");
        }
    }
    };
    stage(0);
    for (int ci = 0; ci < nchunks; ++ci) {
        // ---- (1) the chunk's source rows have been requested (before the loop / during the previous chunk's vertical pass) ----
        asm volatile("cp.async.wait_all;" ::: "memory");
        if (n_in < sbytes) {   // an edge strip: replicate the span's first / last pixel outwards (each warp on its own rows)
            __syncwarp();
            for (int rr = warp; rr < F_CH; rr += F_WARPS) {
                uint8_t *sb = reinterpret_cast<uint8_t *>(s_in + (size_t)rr * a.in_pitch) + phase;
                for (int e = lane; e < e_off; e += 32) sb[e] = sb[e_off + e % C];
                const int right0 = e_off + n_in;
                for (int e = right0 + lane; e < sbytes; e += 32) sb[e] = sb[right0 - C + (e - right0) % C];
            }
        }
        __syncthreads();

        // ---- (2) horizontal pass: lane = row of the chunk, item = F_XC-pixel piece; output into the ring ----
        {
            const uint32_t *in_row = s_in + (size_t)lane * a.in_pitch;
            uint32_t *out_row = ring + (size_t)((ci * F_CH) % a.ring_rows + lane) * a.ring_pitch;
            for (int ck = warp; ck < hchunks; ck += F_WARPS) {
                const int xa = ck * F_XC, xb = min(xa + F_XC, tw);
                uint32_t sum[C];
#pragma unroll
                for (int c = 0; c < C; ++c) sum[c] = g.bias;
                {
                    const uint8_t *b = reinterpret_cast<const uint8_t *>(in_row) + phase + xa * C;
                    for (int j = 0; j <= 2 * r; ++j) {
#pragma unroll
                        for (int c = 0; c < C; ++c) sum[c] += b[j * C + c];
                    }
                }
                const uint32_t *pw = in_row + (xa * C + phase) / 4;                    // outgoing pixels, aligned down
                const uint32_t *nw = in_row + ((xa + 2 * r + 1) * C + phase) / 4;      // incoming pixels, aligned down
                for (int x = xa; x < xb; x += 4) {
                    uint32_t prev[C], next[C], praw[C + 1], nraw[C + 1];
#pragma unroll
                    for (int k = 0; k <= C; ++k) { praw[k] = pw[k]; nraw[k] = nw[k]; }
#pragma unroll
                    for (int k = 0; k < C; ++k) {
                        prev[k] = __funnelshift_r(praw[k], praw[k + 1], pshift);
                        next[k] = __funnelshift_r(nraw[k], nraw[k + 1], nbyte * 8);
                    }
                    uint32_t ob[4 * C];
#pragma unroll
                    for (int b = 0; b < 4 * C; ++b) {
                        const int c = b % C;
                        ob[b] = __umulhi(sum[c], g.recip);
                        sum[c] = dp4a_us(next[b / 4], 1 << (8 * (b & 3)), sum[c]);             // + byte (b & 3) of the incoming word
                        sum[c] = dp4a_us(prev[b / 4], (int)(0xFFu << (8 * (b & 3))), sum[c]);  // - byte (b & 3) of the outgoing word
                    }
                    uint32_t *ow = out_row + (x * C) / 4;
#pragma unroll
                    for (int k = 0; k < C; ++k)
                        ow[k] = __byte_perm(__byte_perm(ob[4 * k], ob[4 * k + 1], 0x0040), __byte_perm(ob[4 * k + 2], ob[4 * k + 3], 0x0040), 0x5410);
                    pw += C;
                    nw += C;
                }
            }
        }
        __syncthreads();
        if (ci + 1 < nchunks) stage(ci + 1);   // the staging rows are free: the next chunk's copies fly during the vertical pass

        // ---- (3) vertical pass over the chunk's ring rows ----
        if (tid < owords) {
            const int rows_here = min(F_CH, nq - ci * F_CH);
            const int slot = (ci * F_CH) % a.ring_rows;                 // ring row of the chunk's first incoming row
            // the row leaving the window when ring row q comes in is q - ks; while the window fills, nothing leaves
            const int fill = min(max(ks - primed, 0), rows_here);        // rows of this chunk that only fill the window
            int back = slot + fill - ks;
            if (back < 0) back += a.ring_rows;
            const uint32_t *rin = ring + (size_t)slot * a.ring_pitch + tid;
            const uint32_t *rout = ring + (size_t)back * a.ring_pitch + tid;
            const uint32_t *rend = ring + (size_t)a.ring_rows * a.ring_pitch + tid;
            const size_t ring_span = (size_t)a.ring_rows * a.ring_pitch;
            const int pitch = a.ring_pitch;
            auto emit = [&]() {
                const uint32_t o0 = __umulhi(sv[0], g.recip), o1 = __umulhi(sv[1], g.recip);
                const uint32_t o2 = __umulhi(sv[2], g.recip), o3 = __umulhi(sv[3], g.recip);
                *dcol = __byte_perm(__byte_perm(o0, o1, 0x0040), __byte_perm(o2, o3, 0x0040), 0x5410);  // tw * C is a multiple of 4: whole words only
                dcol += dwords;
            };
            // (a chunk's incoming rows never wrap: chunks and the ring are multiples of 32 rows)
            for (int k = 0; k < fill; ++k, ++primed) {   // filling the window (first chunk(s) of a band only)
                const uint32_t vin = *rin;
#pragma unroll
                for (int k = 0; k < 4; ++k) sv[k] = dp4a_us(vin, 1 << (8 * k), sv[k]);
                rin += pitch;
            }
            // steady state: every row emits a destination row (while the band has any left) and slides the window.
            // The outgoing rows wrap at most once inside the chunk: two segments, no per-row wrap test; eight rows
            // at a time with their sixteen loads issued first (a load-use-per-row loop serialises on the shared-
            // memory latency, which the few warps of the vertical pass cannot hide).
            int m = rows_here - fill;
            while (m > 0) {
                const int seg = min(m, (int)((rend - rout) / pitch));
                const int e = min(seg, out_left);
                int i = 0;
                constexpr int VU = 8;
                for (; i + VU <= e; i += VU) {
                    uint32_t vin[VU], vout[VU];
#pragma unroll
                    for (int u = 0; u < VU; ++u) { vin[u] = rin[u * pitch]; vout[u] = rout[u * pitch]; }
#pragma unroll
                    for (int u = 0; u < VU; ++u) {
                        emit();
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            sv[k] = dp4a_us(vin[u], 1 << (8 * k), sv[k]);
                            sv[k] = dp4a_us(vout[u], (int)(0xFFu << (8 * k)), sv[k]);
                        }
                    }
                    rin += VU * pitch;
                    rout += VU * pitch;
                }
                for (; i < seg; ++i) {
                    const uint32_t vin = *rin, vout = *rout;
                    if (i < e) emit();
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        sv[k] = dp4a_us(vin, 1 << (8 * k), sv[k]);
                        sv[k] = dp4a_us(vout, (int)(0xFFu << (8 * k)), sv[k]);
                    }
                    rin += pitch;
                    rout += pitch;
                }
                out_left -= e;
                m -= seg;
                if (rout >= rend) rout -= ring_span;
            }
        }
        // (the barrier that follows the next chunk's staging wait also keeps its horizontal pass off the ring rows this
        // vertical pass is still reading)
    }
}



}  // namespace

// One box (both passes in one kernel) on u8 images, src -> dst.  Returns 1 if done, 0 if the fused kernel does not
// apply (the caller then runs box_pass_u8 or the generic kernels for ALL three boxes), < 0 on error.
bool box_fused_u8_applies(uint32_t w, int C, uint32_t radius, const void *a, const void *b, const void *c, const DeviceCtx *ctx) {
    const uint32_t ks = 2 * radius + 1;
    if (ks > 255 || ((size_t)w * C) % 4 != 0) return false;
    if (((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 3) != 0) return false;
    const int F_XT = f_xt(C);
    const int in_pitch = odd_words((F_XT + 2 * (int)radius + 1 + 8) * C + 8), ring_pitch = odd_words(F_XT * C + 4);
    const int ring_rows = (F_CH + (int)ks + F_CH - 1) / F_CH * F_CH;
    return ((size_t)F_CH * in_pitch + (size_t)ring_rows * ring_pitch) * 4 <= std::min<size_t>(ctx->smem_optin, 110 * 1024);  // at least two CTAs per SM
}

int box_fused_u8(const uint8_t *src, uint8_t *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t radius,
                 const DeviceCtx *ctx, cudaStream_t st) {
    FusedBoxArgs a;
    a.src = src; a.dst = dst;
    a.g.w = w; a.g.h = h; a.g.r = radius; a.g.ks = 2 * radius + 1;
    a.g.recip = UINT32_MAX / a.g.ks + 1u;
    a.g.bias = a.g.ks / 2;
    const int F_XT = f_xt(C);
    a.in_pitch = odd_words((F_XT + 2 * (int)radius + 1 + 8) * C + 8);  // slack: the last 4-pixel step reads up to C + 1 words past the last incoming pixel
    a.ring_pitch = odd_words(F_XT * C + 4);
    a.ring_rows = (F_CH + (int)a.g.ks + F_CH - 1) / F_CH * F_CH;
    const size_t smem = ((size_t)F_CH * a.in_pitch + (size_t)a.ring_rows * a.ring_pitch) * 4;
    const unsigned strips = (w + F_XT - 1) / F_XT;
    // bands: each primes its window with 2r + 1 extra rows.  A CTA runs its chunks as a chain of short phases, so the
    // GPU needs many of them in flight: about ten per SM (measured on B200: 8 x 4K Rgb8 0.96 ms with 2 bands, 0.73 ms
    // with 6), while a band stays at least four windows (and 64 rows) long so the priming stays a small share.
    // Candidates up to ~20 CTAs per SM; among them the split whose last wave of CTAs is fullest, weighted by the share of a
    // band's rows that are not priming (8 x 4K Rgb8: 6 bands = 1440 CTAs = 1.95 waves of 740: 0.67 ms; 7 bands = 2.27: 0.71).
    int bands = 1;
    {
        const int min_rows = std::max(64, 4 * (int)a.g.ks);
        const uint64_t slots = (uint64_t)ctx->sm_count * std::max<size_t>(1, std::min<size_t>(16, (233472 - 1024) / (smem + 1024)));
        double best = -1.0;
        for (int b = 1; b <= 64 && (b == 1 || (int)(h / b) >= min_rows); ++b) {
            const uint64_t total = (uint64_t)strips * n_images * b;
            if (b > 1 && total > 20ull * ctx->sm_count && total > 2 * slots) break;
            const double waves = (double)total / (double)slots, rows = (double)h / b;
            const double score = waves / std::ceil(waves) * rows / (rows + a.g.ks) * (total >= slots ? 1.0 : 0.0 + waves);
            if (score > best) { best = score; bands = b; }
        }
    }
    if (const char *e = getenv("B200IMG_BOX_BANDS")) { const int v = atoi(e); if (v >= 1) bands = std::min<int>(v, (int)h); }
    a.band_rows = (int)((h + bands - 1) / bands);
    a.bands = (int)((h + a.band_rows - 1) / a.band_rows);
    if ((uint64_t)a.bands * n_images > 65535) return 0;
    const dim3 grid(strips, (unsigned)(a.bands * n_images));
#define F_GO(CC)                                                                                        \
    {                                                                                                   \
        auto k = box_fused_u8_kernel<CC>;                                                               \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
        k<<<grid, F_THREADS, smem, st>>>(a);                                                            \
    }
    switch (C) {
        case 1: F_GO(1) break;
        case 2: F_GO(2) break;
        case 3: F_GO(3) break;
        default: F_GO(4) break;
    }
#undef F_GO
    B200_CUDA(cudaGetLastError());
    return 1;
}

// One box pass (horizontal into tmp, vertical into dst) on u8 images.  Returns 1 if done, 0 if the
// fast path does not apply (caller uses the generic kernels), < 0 on error.
int box_pass_u8(const uint8_t *src, uint8_t *tmp, uint8_t *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t radius,
                const DeviceCtx *ctx, cudaStream_t st) {
    const uint32_t ks = 2 * radius + 1;
    const size_t row_bytes = (size_t)w * C;
    if (ks > 255 || row_bytes % 4 != 0) return 0;
    if (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(tmp) | reinterpret_cast<uintptr_t>(dst)) & 3) != 0) return 0;
    U8Geom g;
    g.w = w; g.h = h; g.r = radius; g.ks = ks;
    g.recip = UINT32_MAX / ks + 1u;
    g.bias = ks / 2;
    const size_t total_rows = (size_t)h * n_images;
    {
        // slack: the last 4-pixel step reads up to C + 1 words past the last needed incoming pixel
        const int in_pitch = odd_words((H_XT + 2 * (int)radius + 1 + 8) * C + 8);
        const int out_pitch = odd_words((H_XT + 4) * C);
        const size_t smem = (size_t)H_ROWS * (in_pitch + out_pitch) * 4;
        if (smem > ctx->smem_optin) return 0;
        const dim3 grid((w + H_XT - 1) / H_XT, (unsigned)((total_rows + H_ROWS - 1) / H_ROWS));
        // a job too small to give every SM two 4-warp CTAs runs 16-pixel chunks on 16 warps per CTA: each
        // chunk re-primes its window (2r + 1 pixels), which is cheaper than idle SMs
        const bool small = (uint64_t)grid.x * grid.y < 2ull * ctx->sm_count;
        const int xc = small ? 16 : 32;   // 32-pixel chunks on 8 warps: measured best for large batches (64: 0.98 ms, 32: 0.88 ms, 16: 0.99 ms)
        const int hthreads = 32 * (H_XT / xc);
#define H_GO(CC)                                                                                        \
    {                                                                                                   \
        auto k = box_h_u8_kernel<CC>;                                                                   \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));     \
        k<<<grid, hthreads, smem, st>>>(src, tmp, g, total_rows, in_pitch, out_pitch, xc);              \
    }
        switch (C) {
            case 1: H_GO(1) break;
            case 2: H_GO(2) break;
            case 3: H_GO(3) break;
            default: H_GO(4) break;
        }
#undef H_GO
    }
    {
        const uint32_t words = (uint32_t)(row_bytes / 4);
        // shorter bands (each re-primes its 2r + 1 row window) until every SM has two CTAs
        int vband = V_BAND;
        const uint64_t per_band = (uint64_t)((words + V_THREADS - 1) / V_THREADS) * n_images;
        while (vband > 8 && per_band * ((h + vband - 1) / vband) < 2ull * ctx->sm_count) vband /= 2;
        const int bands = (int)((h + vband - 1) / vband);
        const dim3 grid((words + V_THREADS - 1) / V_THREADS, (unsigned)(bands * n_images));
        box_v_u8_kernel<<<grid, V_THREADS, 0, st>>>(reinterpret_cast<const uint32_t *>(tmp), reinterpret_cast<uint32_t *>(dst), g,
                                                     words, bands, vband);
    }
    B200_CUDA(cudaGetLastError());
    return 1;
}

}  // namespace b200
