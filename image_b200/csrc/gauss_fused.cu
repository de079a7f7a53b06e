// gauss_fused.cu -- the Gaussian blur's row AND column pass in one kernel.
//
// filter_2d_separable (src/imageops/filter_1d.rs:615-761) writes a whole intermediate image between
// its passes; on the GPU that image costs 8 bytes of HBM traffic per subpixel (f32 out, f32 back)
// around a 1-byte-in / 1-byte-out operator.  Here the intermediate never leaves the SM:
//
//   A CTA owns a strip of TW output pixels and streams down a band of rows in chunks of 32.  Per
//   chunk: (1) 32 source rows x (TW + taps) pixels are staged as f32 in shared memory; (2) the ROW
//   pass (lane = row, a warp = two 8-pixel blocks of one channel, 16 outputs per thread on FFMA2)
//   appends 32 intermediate rows to a RING of RR rows in shared memory; (3) the COLUMN pass (thread
//   = two adjacent subpixel columns x 8 rows, one LDS.64 per ring row) produces the 32 output rows
//   whose windows the ring now holds.  FMA work is exactly taps_x + taps_y per subpixel (no vertical
//   recomputation inside a band); HBM traffic is the source strip (+ halo columns) and the result.
//
// Arithmetic: the f32 engine with fused multiply-adds in window order (B200IMG_FP_FAST: +-1 LSB), or
// the Q0.15 engine for u8 (exact in f32: every partial sum < 2^24, so bit-identical to the
// reference's u32 accumulator, filter_1d.rs:132-150).  Clamp-to-edge as make_arena_row /
// column clamping do it (filter_1d.rs:60-130): source pixels clamped in x while staging, source ROWS
// clamped in y before the row pass (a clamped row's intermediate is that row's intermediate).
//
// Both convolutions run "8 taps per chunk on a 16-entry register window (two halves that swap
// roles, so sliding costs no moves), remaining < 8 taps straight from shared memory"; the column
// window starts on a multiple of 8 ring rows (the kernel is padded with leading zeros to get
// there), so ring wrap-around is one pointer fix-up per 8-row block.
#include "common.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

namespace b200 {

constexpr int GF_MAX_TAPS = 136;  // per axis, padding included; longer kernels take the two-kernel path
constexpr int GF_CH = 32;         // rows per chunk = lanes of the row pass

typedef unsigned long long f32x2;
struct FusedWeights { f32x2 r[GF_MAX_TAPS]; f32x2 c[GF_MAX_TAPS]; };  // (w, w) pairs, zero padded

struct FusedArgs {
    const void *src;   // buffer row 0 = image row src_row0 (whole images: 0)
    void *dst;         // buffer row 0 = image row dst_row0
    uint32_t W, H;     // image size (rows are clamped to [0, H))
    uint32_t src_row0, src_rows_buf;  // rows the source buffer holds
    uint32_t dst_row0, dst_rows;      // rows to produce
    size_t src_img_stride, dst_img_stride;  // elements between images of a batch
    int strips, chunks;               // strips per image; output chunks (of 32 rows) per strip
    long long units;                  // strips x images x chunks: the work the persistent CTAs share out
    uint32_t wait_hint_ns;            // suspend-time hint of the barrier waits
    int half_r, pairs_r, rem_r;       // row kernel: centre offset, 16-tap steps, remaining (< 16) taps
    int lead_c, half_c, pairs_c, rem_c;   // column kernel (its lead_c leading zeros included)
    int RR;            // ring rows (multiple of 32)
    int win_chunks;    // ring chunks one output chunk reads (its window rounded up to whole chunks)
    int dbg;           // timing experiments only (B200IMG_GF_DBG): 1 = no staging work, 2 = no output stores, 4 = no ring stores, 8 = no multiply-adds
    int in_pitch;      // floats per staged row (odd)
    int raw_pitch;     // u8 only: words per row of the raw-byte buffers in front of the f32 staging (0 = not used)
    int sp;            // staged pixels per row
};

namespace {

// One CTA per SM, three kinds of warps that never meet at a CTA-wide barrier (mbarriers carry the hand-offs):
//   6 row warps    (two row-pass items each per chunk)     staging buffer -> ring
//   6 column warps (two column-pass items each per chunk)  ring -> destination
//   4 staging warps                                        source -> staging buffer (f32)
// so that one group's loads, conversions and stores overlap the other groups' multiply-adds; each SM
// sub-partition holds three compute warps (row and column warps mixed) and one staging warp.
constexpr int GF_COLS = 192;               // subpixel columns per strip, whatever the channel count
constexpr int GF_RP = GF_COLS + 2;         // ring row pitch: even (LDS.64), 2 mod 4 (2-way, not 4-way, on the row pass's stores)
constexpr int GF_CWARPS = 12;              // = row-pass items per chunk = (TW / 16) * C
constexpr int GF_CTHREADS = GF_CWARPS * 32;  // = 2 * COLS: one column-pass item (column pair x 8 rows) per thread
constexpr int GF_PWARPS = 4;
constexpr int GF_PTHREADS = GF_PWARPS * 32;
constexpr int GF_THREADS = GF_CTHREADS + GF_PTHREADS;
static_assert(GF_RP % 4 == 2, "ring pitch must be 2 mod 4");
template <int C> struct Geo {
    static constexpr int TW = GF_COLS / C;  // output pixels per strip: 192 / 96 / 64 / 48
    static_assert((TW / 16) * C == GF_CWARPS && TW % 16 == 0, "one row-pass item per compute warp");
};

constexpr int GF_RWARPS = GF_CWARPS / 2;   // row warps: warps [0, 6)
constexpr int GF_MAX_SLOTS = 8;            // ring chunks (RR / 32) the barrier arrays provide for

// Wait for the phase with this parity.  try_wait suspends the warp in hardware for up to the hinted time, so
// a waiting warp does not spin through the issue port the working warps need.
__device__ __forceinline__ void mbar_wait_parked(uint64_t *bar, uint32_t parity, uint32_t hint_ns = 1000000u) {
    if (hint_ns == 0) { mbar_wait(bar, parity); return; }
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(parity), "r"(hint_ns) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

// acc[r] += sum_{kk < 8} window[r + kk] * w[kk], window = lo[0..7] ++ hi[0..7]
__device__ __forceinline__ void taps8(const f32x2 (&lo)[8], const f32x2 (&hi)[8], const f32x2 *w, f32x2 (&acc)[8]) {
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
        const f32x2 wk = w[kk];
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r] = fma2(r + kk < 8 ? lo[r + kk] : hi[r + kk - 8], wk, acc[r]);
    }
}
// the first `t` (< 16, uniform) taps on the window a ++ b ++ c
__device__ __forceinline__ void taps_some(const f32x2 (&a)[8], const f32x2 (&b)[8], const f32x2 (&c)[8], const f32x2 *w, int t,
                                          f32x2 (&acc)[8]) {
#pragma unroll
    for (int kk = 0; kk < 15; ++kk) {
        if (kk < t) {
            const f32x2 wk = w[kk];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int i = r + kk;
                acc[r] = fma2(i < 8 ? a[i] : i < 16 ? b[i - 8] : c[i - 16], wk, acc[r]);
            }
        }
    }
}

// One thread's convolution: 8 outputs of two independent signals (a pair per f32x2).  Taps go 16 at a time
// (two chunks of 8 on a 16-entry register window whose halves swap roles, so sliding costs no moves); the
// remaining `rem` (< 16) taps run on the window that is then in registers plus up to two more blocks.
// `ld` delivers the signal in blocks of 8 samples: load(X) reads the current block, next() steps on.
template <typename Loader>
__device__ __forceinline__ void conv_blocks(Loader &ld, int npairs, int rem, const f32x2 *w, f32x2 (&acc)[8]) {
    f32x2 A[8], B[8];
    ld.load(A);
    // (the weights are indexed from the array's base with the uniform loop counter, not through a moving
    // pointer: that keeps their loads on the uniform datapath -- LDCU feeding FFMA2's uniform operand)
#pragma unroll 1
    for (int pr = 0; pr < npairs; ++pr) {
        ld.next(); ld.load(B);
        taps8(A, B, w + pr * 16, acc);
        ld.next(); ld.load(A);
        taps8(B, A, w + pr * 16 + 8, acc);
    }
    if (rem) {
        f32x2 Cc[8];
        if (rem >= 2) { ld.next(); ld.load(B); }
        if (rem >= 10) { ld.next(); ld.load(Cc); }
        taps_some(A, B, Cc, w + npairs * 16, rem, acc);
    }
}

// Row pass: samples of one channel are C floats apart in the staged row; a / b = the two pixel blocks.
template <int C> struct RowLoader {
    const float *pa, *pb;
    __device__ __forceinline__ void load(f32x2 (&x)[8]) const {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = pack2(pa[i * C], pb[i * C]);
    }
    __device__ __forceinline__ void next() { pa += 8 * C; pb += 8 * C; }
};
// Column pass: two adjacent columns of the ring; the window starts on a multiple of 8 ring rows, so it
// can only wrap between blocks.
template <int RP> struct ColLoader {
    const float *p, *first, *end;  // current block, ring row 0, one past the last ring row (of this column pair)
    __device__ __forceinline__ void load(f32x2 (&x)[8]) const {
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = *reinterpret_cast<const f32x2 *>(p + i * RP);
    }
    __device__ __forceinline__ void next() { p += 8 * RP; if (p >= end) p = first; }
};

// Store rules.  Q15: ToStorage (filter_1d.rs:132-150) on an exact integer held in f32.
template <typename T, bool Q15> struct Store {
    static __device__ __forceinline__ T f(float a) { return Px<T>::nearest(a); }
};
template <> struct Store<uint8_t, true> {
    static __device__ __forceinline__ uint8_t f(float a) { return (uint8_t)min((__float2uint_rz(a) + 16384u) >> 15, 255u); }
};

// Integer subpixels -> float without the conversion pipe (I2F issues at a quarter rate): byte j of
// a word goes into the mantissa of 2^23 and the 2^23 is subtracted.
__device__ __forceinline__ float byte_to_f32(uint32_t w, int j) { return __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540 + j)) - 8388608.0f; }
__device__ __forceinline__ float half_to_f32(uint32_t w, int j) { return __uint_as_float(__byte_perm(w, 0x4B000000u, j ? 0x7532 : 0x7510)) - 8388608.0f; }

constexpr int PF_ROWS = GF_CH / GF_PWARPS;  // staged rows per staging warp

// Work is a stream of "units" (one output chunk of one strip of one image); a CTA owns a contiguous range of
// units and walks it as segments (consecutive output chunks of one strip).  All three roles walk the same
// segments with their own running counters, so the hand-off barriers simply continue across segment boundaries:
// the row warps start on the next strip while the column warps finish the previous one.
struct Segment {
    int x0;           // first output pixel of the strip
    uint32_t img;
    int b0, b1;       // output rows [b0, b1) (image coordinates)
    int nj, ni;       // output chunks, row chunks (nj + ahead)
    int q0;           // image row of the segment's intermediate row 0 (may be negative: clamped)
};

template <typename In, int C, bool Q15>
__global__ void __launch_bounds__(GF_THREADS, 1)
gauss_fused_kernel(const __grid_constant__ FusedArgs a, const __grid_constant__ FusedWeights tw) {
    constexpr int RP = GF_RP, TW = Geo<C>::TW, COLS = GF_COLS;
    extern __shared__ __align__(16) float smf[];
    float *ring = smf;                                    // [RR][RP]
    float *stage0 = smf + (size_t)a.RR * RP;              // 2 x [32][in_pitch]
    const size_t stage_floats = (size_t)GF_CH * a.in_pitch;
    // hand-offs: staging buffer b filled / drained; ring chunk s written / read for the last time
    __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2], bar_rows[GF_MAX_SLOTS], bar_cols[GF_MAX_SLOTS];
    const int tid = threadIdx.x, lane = tid & 31;
    // broadcast from lane 0: tells the compiler the warp index (and so every role branch below) is warp-uniform,
    // which keeps the convolutions' weight loads on the uniform datapath (LDCU -> FFMA2's uniform operand)
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const int K = a.RR / GF_CH;          // ring chunks
    const int ahead = a.win_chunks - 1;  // output chunk j reads row chunks j .. j + ahead
    const size_t g_pitch = (size_t)a.W * C;
    const uint32_t hint = a.wait_hint_ns;
    const long long u_begin = a.units * (long long)blockIdx.x / gridDim.x, u_end = a.units * (long long)(blockIdx.x + 1) / gridDim.x;
    if (u_begin >= u_end) return;
    auto segment_at = [&](long long u, Segment &sg) -> long long {  // the segment starting at unit u; returns its last unit + 1
        const long long column = u / a.chunks;
        const int j0 = (int)(u - column * a.chunks);
        sg.nj = (int)min((long long)(a.chunks - j0), u_end - u);
        sg.img = (uint32_t)(column / a.strips);
        sg.x0 = (int)(column - (long long)sg.img * a.strips) * TW;
        sg.b0 = (int)a.dst_row0 + j0 * GF_CH;
        sg.b1 = min(sg.b0 + sg.nj * GF_CH, (int)(a.dst_row0 + a.dst_rows));
        sg.ni = sg.nj + ahead;
        sg.q0 = sg.b0 - a.half_c - a.lead_c;
        return u + sg.nj;
    };

    if (tid == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(&bar_full[b], GF_PTHREADS); mbar_init(&bar_empty[b], GF_RWARPS * 32); }
        for (int s = 0; s < GF_MAX_SLOTS; ++s) { mbar_init(&bar_rows[s], GF_RWARPS * 32); mbar_init(&bar_cols[s], GF_RWARPS * 32); }
        fence_mbar_init();
    }
    // a ring full of finite values from the start (belt and braces: no tap reads a row before it is written)
    for (int i = tid; i < a.RR * RP; i += GF_THREADS) ring[i] = 0.0f;
    __syncthreads();

    // Warp roles by index.  The staging warps take the HIGHEST indices, which the scheduler favours: they are few,
    // latency-bound and everything else waits for them (measured: with the lowest indices the kernel is 4 % slower).
    // Every SM sub-partition (warp % 4) gets one staging warp and three compute warps, row and column warps mixed.
    if (warp >= GF_CWARPS) {
        // ================= staging warps: row chunk g into buffer g & 1 =================
        // A staged row = sp pixels starting at image pixel xs (clamp-to-edge, make_arena_row filter_1d.rs:60-87).  The
        // pixels inside the image are one contiguous span of the source row: it is copied / converted with whole
        // 32-bit words (u8: four subpixels a word, u16: two, through the 2^23 trick; f32: cp.async), all loads of a
        // warp's rows in flight before the first conversion.  The clamped pixels left / right of it (edge strips only)
        // are then filled from the span's first / last pixel, which the same warp has just stored.
        const int pw = warp - GF_CWARPS;
        constexpr int EPW = 4 / (int)sizeof(In);                            // subpixels per 32-bit word
        constexpr int PF_W = sizeof(In) == 1 ? 4 : 8;                       // words per lane per row held in registers
        const int row_elems = a.sp * C;
        if constexpr (sizeof(In) == 1) {
            if (a.raw_pitch) {
                // ---- u8, word-aligned rows (the host checked): raw bytes first, conversion second ----
                // The warp's 8 rows of chunk g + 1 are on their way into a raw-byte buffer (cp.async, no registers
                // held) while its rows of chunk g are converted from the other raw buffer into the f32 staging
                // buffer: a chunk's DRAM latency hides behind the previous chunk's conversion, and the whole path is
                // two small rolled loops (the register-prefetch path below is 13 KB of straight-line code that the
                // staging warp drags through the sub-partition's instruction cache every chunk).
                uint32_t *raw0 = reinterpret_cast<uint32_t *>(stage0 + 2 * stage_floats);   // 2 x [32][raw_pitch] words
                const uint32_t raw_words = (uint32_t)GF_CH * a.raw_pitch;
                struct Cursor {
                    long long u_next;   // first unit after this segment
                    Segment sg;
                    int i;              // row chunk within the segment
                    const uint8_t *src; // image base
                    int lo_px, e_off, n_in, phase, words;
                    bool valid;
                };
                auto open_segment = [&](long long u, Cursor &c) {
                    c.valid = u < u_end;
                    if (!c.valid) return;
                    c.u_next = segment_at(u, c.sg);
                    c.i = 0;
                    c.src = reinterpret_cast<const uint8_t *>(a.src) + (size_t)c.sg.img * a.src_img_stride;
                    const int xs = c.sg.x0 - a.half_r;
                    c.lo_px = max(xs, 0);
                    const int hi_px = min(xs + a.sp, (int)a.W);
                    c.e_off = (c.lo_px - xs) * C;
                    c.n_in = (hi_px - c.lo_px) * C;
                    c.phase = (int)(((size_t)c.lo_px * C) & 3);
                    c.words = (c.phase + c.n_in + 3) >> 2;
                };
                auto advance = [&](Cursor &c) {
                    if (++c.i >= c.sg.ni) open_segment(c.u_next, c);
                };
                auto row_ptr = [&](const Cursor &c, int r) -> const uint8_t * {
                    int y = min(max(c.sg.q0 + c.i * GF_CH + r, 0), (int)a.H - 1);
                    y = min(max(y, (int)a.src_row0), (int)(a.src_row0 + a.src_rows_buf) - 1);
                    return c.src + (size_t)(y - (int)a.src_row0) * g_pitch + ((size_t)c.lo_px * C - c.phase);
                };
                auto issue_copies = [&](const Cursor &c, uint32_t *raw) {
#pragma unroll 2
                    for (int k = 0; k < PF_ROWS; ++k) {
                        const int r = pw + k * GF_PWARPS;
                        const uint32_t *gw = reinterpret_cast<const uint32_t *>(row_ptr(c, r)) + lane;
                        uint32_t d = smem_u32(raw + r * a.raw_pitch + lane);
                        for (int wi = lane; wi < c.words; wi += 32, gw += 32, d += 128)
                            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d), "l"(gw) : "memory");
                    }
                };
                Cursor cur, nxt;
                open_segment(u_begin, cur);
                if (cur.valid) issue_copies(cur, raw0);
                asm volatile("cp.async.commit_group;" ::: "memory");
                const int rot = lane >> 3;  // lanes 8 apart start on different bytes, so the four stores of a word hit 32 different banks
                for (uint32_t g = 0; cur.valid; ++g) {
                    nxt = cur;
                    advance(nxt);
                    if (nxt.valid && !(a.dbg & 1)) issue_copies(nxt, raw0 + ((g + 1) & 1) * raw_words);
                    asm volatile("cp.async.commit_group;" ::: "memory");
                    asm volatile("cp.async.wait_group 1;" ::: "memory");   // this chunk's bytes have landed (each lane converts the words it copied)
                    mbar_wait_parked(&bar_empty[g & 1], ((g >> 1) & 1) ^ 1, hint);  // the row pass two chunks back has left the f32 buffer
                    float *s_in = stage0 + (size_t)(g & 1) * stage_floats;
                    if (!(a.dbg & 1)) {
                        const uint32_t *raw = raw0 + (g & 1) * raw_words;
                        const int first_full = cur.phase ? 1 : 0, last_full = (cur.phase + cur.n_in) >> 2;  // whole words: [first_full, last_full)
#pragma unroll 1
                        for (int k = 0; k < PF_ROWS; ++k) {
                            const int r = pw + k * GF_PWARPS;
                            const uint32_t *rw = raw + r * a.raw_pitch;
                            float *sd = s_in + r * a.in_pitch + cur.e_off - cur.phase;   // word wi starts at sd[wi * 4]
#pragma unroll 2
                            for (int wi = lane; wi < cur.words; wi += 32) {
                                const uint32_t w = rw[wi];
                                if (wi >= first_full && wi < last_full) {
#pragma unroll
                                    for (int jj = 0; jj < 4; ++jj) {
                                        const int j = (jj + rot) & 3;
                                        sd[wi * 4 + j] = byte_to_f32(w, j);
                                    }
                                } else {
#pragma unroll
                                    for (int j = 0; j < 4; ++j) {
                                        const int e = wi * 4 + j - cur.phase;
                                        if (e >= 0 && e < cur.n_in) sd[wi * 4 + j] = byte_to_f32(w, j);
                                    }
                                }
                            }
                        }
                        if (cur.n_in < row_elems) {  // an edge strip: replicate the span's first / last pixel outwards
                            __syncwarp();
                            for (int r = pw; r < GF_CH; r += GF_PWARPS) {
                                float *sr = s_in + r * a.in_pitch;
                                for (int e = lane; e < cur.e_off; e += 32) sr[e] = sr[cur.e_off + e % C];
                                const int right0 = cur.e_off + cur.n_in;
                                for (int e = right0 + lane; e < row_elems; e += 32) sr[e] = sr[right0 - C + (e - right0) % C];
                            }
                        }
                    }
                    mbar_arrive(&bar_full[g & 1]);
                    cur = nxt;
                }
                return;
            }
        }
        uint32_t g = 0;  // row chunks staged so far
        Segment sg;
        for (long long u = u_begin; u < u_end;) {
            u = segment_at(u, sg);
            const In *src = reinterpret_cast<const In *>(a.src) + (size_t)sg.img * a.src_img_stride;
            const int xs = sg.x0 - a.half_r;                           // image pixel of staged pixel 0
            const int lo_px = max(xs, 0), hi_px = min(xs + a.sp, (int)a.W);     // span [lo_px, hi_px) of the source row (hi_px > lo_px: x0 < W)
            const int e_off = (lo_px - xs) * C, n_in = (hi_px - lo_px) * C;     // its place and length in the staged row
            const bool aligned = (g_pitch * sizeof(In)) % 4 == 0 && ((reinterpret_cast<uintptr_t>(src) & 3) == 0);
            const int phase = (int)(((size_t)lo_px * C) % EPW);                 // subpixels before the span in its first word
            const int words = (phase + n_in + EPW - 1) / EPW;
            const bool fast = aligned && (sizeof(In) == 4 || words <= 32 * PF_W);
            auto src_row = [&](int i, int r) -> const In * {  // source row of staged row r of the segment's row chunk i
                int y = min(max(sg.q0 + i * GF_CH + r, 0), (int)a.H - 1);  // clamp-to-edge: a clamped row's intermediate is that row's
                // a band's buffer ends where its windows end; rows past that are staged only because chunks are 32 rows
                // (no tap with a non-zero weight reads them): any row of the buffer will do
                y = min(max(y, (int)a.src_row0), (int)(a.src_row0 + a.src_rows_buf) - 1);
                return src + (size_t)(y - (int)a.src_row0) * g_pitch;
            };
            for (int i = 0; i < sg.ni; ++i, ++g) {
                float *s_in = stage0 + (size_t)(g & 1) * stage_floats;
                uint64_t *empty = &bar_empty[g & 1];
                const uint32_t empty_parity = ((g >> 1) & 1) ^ 1;  // passes at once the first time a buffer is used
                if (a.dbg & 1) {
                    mbar_wait_parked(empty, empty_parity, hint);
                } else if (fast) {
                    if constexpr (sizeof(In) == 4) {
                        mbar_wait_parked(empty, empty_parity, hint);  // the row pass two chunks back has left the buffer
#pragma unroll 2
                        for (int k = 0; k < PF_ROWS; ++k) {
                            const int r = pw + k * GF_PWARPS;
                            const In *gp = src_row(i, r) + (size_t)lo_px * C;
                            float *sd = s_in + r * a.in_pitch + e_off;
                            for (int e = lane; e < n_in; e += 32)
                                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(smem_u32(sd + e)), "l"(gp + e) : "memory");
                        }
                        asm volatile("cp.async.wait_all;" ::: "memory");
                    } else {
                        uint32_t pf[PF_ROWS][PF_W];
#pragma unroll
                        for (int k = 0; k < PF_ROWS; ++k) {
                            const uint32_t *gw = reinterpret_cast<const uint32_t *>(src_row(i, pw + k * GF_PWARPS) + ((size_t)lo_px * C - phase));
#pragma unroll
                            for (int it = 0; it < PF_W; ++it) {
                                const int wi = it * 32 + lane;
                                pf[k][it] = wi < words ? __ldg(gw + wi) : 0u;
                            }
                        }
                        mbar_wait_parked(empty, empty_parity, hint);  // the row pass two chunks back has left the buffer
                        const int rot = lane >> 3;  // u8: lanes 8 apart start on different bytes, so the four stores of a word hit 32 different banks
                        const int first_full = phase ? 1 : 0, last_full = (phase + n_in) / EPW;  // whole words: [first_full, last_full)
#pragma unroll
                        for (int k = 0; k < PF_ROWS; ++k) {
                            float *sd = s_in + (pw + k * GF_PWARPS) * a.in_pitch + e_off - phase;  // word wi starts at sd[wi * EPW]
#pragma unroll
                            for (int it = 0; it < PF_W; ++it) {
                                const int wi = it * 32 + lane;
                                if (wi >= first_full && wi < last_full) {
                                    if constexpr (sizeof(In) == 1) {
#pragma unroll
                                        for (int jj = 0; jj < 4; ++jj) {
                                            const int j = (jj + rot) & 3;
                                            sd[wi * 4 + j] = byte_to_f32(pf[k][it], j);
                                        }
                                    } else {
                                        sd[wi * 2] = half_to_f32(pf[k][it], 0);
                                        sd[wi * 2 + 1] = half_to_f32(pf[k][it], 1);
                                    }
                                }
                            }
                        }
                        // the (at most two) partial words of a row: a lane per subpixel, straight from global memory
                        const int head = first_full ? EPW - phase : 0, tail0 = max(last_full * EPW - phase, head);  // span elements [0, head) and [tail0, n_in)
                        const int odd = head + (n_in - tail0);
                        if (odd > 0) {
                            for (int t = lane; t < PF_ROWS * 8; t += 32) {
                                const int k = t >> 3, q = t & 7;
                                if (q < odd) {
                                    const int e = q < head ? q : tail0 + (q - head);
                                    const int r = pw + k * GF_PWARPS;
                                    s_in[r * a.in_pitch + e_off + e] = (float)__ldg(src_row(i, r) + (size_t)lo_px * C + e);
                                }
                            }
                        }
                    }
                } else {  // rows that are not word-aligned, rows too long for the register file: element by element
                    mbar_wait_parked(empty, empty_parity, hint);
                    for (int r = pw; r < GF_CH; r += GF_PWARPS) {
                        const In *gp = src_row(i, r) + (size_t)lo_px * C;
                        float *sd = s_in + r * a.in_pitch + e_off;
#pragma unroll 4
                        for (int e = lane; e < n_in; e += 32) sd[e] = (float)gp[e];
                    }
                }
                if (!(a.dbg & 1) && n_in < row_elems) {  // an edge strip: replicate the span's first / last pixel outwards
                    __syncwarp();
                    for (int r = pw; r < GF_CH; r += GF_PWARPS) {
                        float *s = s_in + r * a.in_pitch;
                        for (int e = lane; e < e_off; e += 32) s[e] = s[e_off + e % C];
                        const int right0 = e_off + n_in;
                        for (int e = right0 + lane; e < row_elems; e += 32) s[e] = s[right0 - C + (e - right0) % C];
                    }
                }
                mbar_arrive(&bar_full[g & 1]);
            }
        }
        return;
    }

    if (warp < GF_RWARPS) {
        // ================= row warps: items `rw` and `rw + 6` of every row chunk =================
        const int rw = warp;
        uint32_t g = 0;         // row chunks done so far
        int slot = 0;           // ring chunk of row chunk g, and the parity of its use count
        uint32_t use_parity = 0;
        Segment sg;
        for (long long u = u_begin; u < u_end;) {
            u = segment_at(u, sg);
            for (int i = 0; i < sg.ni; ++i, ++g) {
                mbar_wait_parked(&bar_full[g & 1], (g >> 1) & 1, hint);
#pragma unroll 1   // rolled: one copy of the convolution per role in the instruction cache (unrolled: 1.22 ms, rolled: 1.19 ms)
                for (int it = 0; it < 2; ++it) {
                    const int item = rw + it * GF_RWARPS;
                    const int xb = item / C, c = item - xb * C;  // pixel blocks xb and xb + TW/16, channel c
                    f32x2 acc[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) acc[r] = 0ull;
                    const float *pa = stage0 + (size_t)(g & 1) * stage_floats + lane * a.in_pitch + (xb * 8) * C + c;
                    RowLoader<C> ld{pa, pa + (TW / 2) * C};
                    if (!(a.dbg & 8)) conv_blocks(ld, a.pairs_r, a.rem_r, tw.r, acc);
                    // the column pass that read this ring chunk last (K row chunks ago) is done with it
                    if (it == 0) mbar_wait_parked(&bar_cols[slot], use_parity ^ 1, hint);
                    if (a.dbg & 4) continue;
                    float *o = ring + (size_t)(slot * GF_CH + lane) * RP + (xb * 8) * C + c;
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        float va, vb;
                        unpack2(acc[r], va, vb);
                        if constexpr (Q15) { va = (float)Store<uint8_t, true>::f(va); vb = (float)Store<uint8_t, true>::f(vb); }
                        o[r * C] = va;
                        o[(TW / 2 + r) * C] = vb;
                    }
                }
                mbar_arrive(&bar_empty[g & 1]);  // the staging warps may refill this buffer (two chunks on)
                mbar_arrive(&bar_rows[slot]);
                if (++slot == K) { slot = 0; use_parity ^= 1; }
            }
        }
        return;
    }

    // ================= column warps: output chunk j of a segment needs its row chunks j .. j + ahead =================
    {
        const int ct = tid - GF_RWARPS * 32;                       // 0 .. 191
        const int cp = ct % (COLS / 2), rb0 = ct / (COLS / 2);     // column pair; row blocks rb0 and rb0 + 2
        const int ring_floats = a.RR * RP;
        const float *col_first = ring + 2 * cp, *col_end = col_first + ring_floats;
        uint32_t rel = 0;       // row chunks released so far = index of the row chunk the next output chunk starts at
        int first = 0;          // its ring chunk
        Segment sg;
        for (long long u = u_begin; u < u_end;) {
            u = segment_at(u, sg);
            In *dst = reinterpret_cast<In *>(a.dst) + (size_t)sg.img * a.dst_img_stride;
            const int col = sg.x0 * C + 2 * cp;
            const bool col_ok = col < (int)g_pitch, two = col + 1 < (int)g_pitch;
            const bool vec = two && g_pitch % 2 == 0 && ((reinterpret_cast<uintptr_t>(dst) & (2 * sizeof(In) - 1)) == 0);
            In *out = dst + (size_t)(sg.b0 + rb0 * 8 - (int)a.dst_row0) * g_pitch + col;  // first output of item 0 of the next chunk
            int rows_left = sg.b1 - sg.b0 - rb0 * 8;                                      // rows from `out` to the end of the segment
            int last = (int)((rel + ahead) % K);  // ring chunk of row chunk rel + ahead, and the parity of its use count
            uint32_t last_parity = ((rel + ahead) / K) & 1;
            for (int j = 0; j < sg.nj; ++j) {
                mbar_wait_parked(&bar_rows[last], last_parity, hint);  // the rows of chunk j + ahead are in the ring (the row warps go in order: so are all before)
                const float *col_win = col_first + (first * GF_CH + rb0 * 8) * RP;   // first window block of item 0
#pragma unroll 1   // rolled: one copy of the convolution per role in the instruction cache (unrolled: 1.22 ms, rolled: 1.19 ms)
                for (int it = 0; it < 2; ++it) {
                    f32x2 acc[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) acc[r] = 0ull;
                    const float *win = col_win + it * 16 * RP;
                    if (win >= col_end) win -= ring_floats;
                    ColLoader<RP> ld{win, col_first, col_end};
                    if (!(a.dbg & 8)) conv_blocks(ld, a.pairs_c, a.rem_c, tw.c, acc);
                    const int left = rows_left - it * 16;
                    if (col_ok && left > 0 && !(a.dbg & 2)) {
                        In *op = out + (size_t)it * 16 * g_pitch;
                        if (vec && left >= 8) {
#pragma unroll
                            for (int r = 0; r < 8; ++r) {
                                float v0, v1;
                                unpack2(acc[r], v0, v1);
                                const In o0 = Store<In, Q15>::f(v0), o1 = Store<In, Q15>::f(v1);
                                if constexpr (sizeof(In) == 1) *reinterpret_cast<uint16_t *>(op) = (uint16_t)((uint32_t)o0 | ((uint32_t)o1 << 8));
                                else if constexpr (sizeof(In) == 2) *reinterpret_cast<uint32_t *>(op) = (uint32_t)o0 | ((uint32_t)o1 << 16);
                                else *reinterpret_cast<float2 *>(op) = make_float2(o0, o1);
                                op += g_pitch;
                            }
                        } else {
#pragma unroll
                            for (int r = 0; r < 8; ++r) {
                                if (r < left) {
                                    float v0, v1;
                                    unpack2(acc[r], v0, v1);
                                    op[0] = Store<In, Q15>::f(v0);
                                    if (two) op[1] = Store<In, Q15>::f(v1);
                                    op += g_pitch;
                                }
                            }
                        }
                    }
                }
                mbar_arrive(&bar_cols[first]);  // that ring chunk is not read again
                ++rel;
                if (++first == K) first = 0;
                if (++last == K) { last = 0; last_parity ^= 1; }
                out += (size_t)GF_CH * g_pitch;
                rows_left -= GF_CH;
            }
            // the segment's last `ahead` row chunks were read by its last output chunks only: release them too
            for (int k = 0; k < ahead; ++k) {
                mbar_arrive(&bar_cols[first]);
                ++rel;
                if (++first == K) first = 0;
            }
        }
    }
}

}  // namespace

// Returns 1 when this kernel does not apply (the caller runs the two-kernel path), 0 after a launch, < 0 on error.
// rk / ck: scanned kernels (f32 bits, or Q0.15 integers with q15 = true), odd lengths.
template <typename T>
int gauss_fused_run(const T *src, T *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t src_row0,
                    uint32_t src_rows_buf, uint32_t dst_row0, uint32_t dst_rows, const uint32_t *rk, int rlen,
                    const uint32_t *ck, int clen, bool q15, bool force, const DeviceCtx *ctx, cudaStream_t st) {
    if (rlen < 3 || clen < 3) return 1;  // the one-tap quirks live in the other kernels
    if (q15 && !std::is_same<T, uint8_t>::value) return 1;
    const int TW = GF_COLS / C, RP = GF_RP;
    FusedArgs a;
    a.half_r = rlen / 2;
    a.pairs_r = rlen / 16;
    a.rem_r = rlen % 16;
    a.half_c = clen / 2;
    a.lead_c = (8 - a.half_c % 8) % 8;
    const int cpad = a.lead_c + clen;
    a.pairs_c = cpad / 16;
    a.rem_c = cpad % 16;
    if (rlen > GF_MAX_TAPS || cpad > GF_MAX_TAPS) return 1;
    // blocks of 8 samples a window of `taps` taps reads: 8 outputs + taps - 1 samples, whole blocks, and the 16-tap
    // steps always read one block ahead
    auto blocks = [](int pairs, int rem) { return 1 + 2 * pairs + (rem >= 10 ? 2 : rem >= 2 ? 1 : 0); };
    // ring rows: a 32-row output chunk reads intermediate rows [0, 24 + window) of its own numbering
    const int window = 8 * blocks(a.pairs_c, a.rem_c);
    a.win_chunks = (24 + window + GF_CH - 1) / GF_CH;
    a.sp = TW - 8 + 8 * blocks(a.pairs_r, a.rem_r);  // the last pixel block's window, read in whole blocks
    a.in_pitch = (a.sp * C + 1) | 1;
    // ring: the window plus as many spare chunks (up to two) as shared memory has room for -- the row warps may
    // run that far ahead of the column warps; at least one, or the two groups would alternate instead of overlap
    // (two against one: 1.190 against 1.194 ms on the bench shape).  Where the raw-byte staging buffers are used (see
    // the staging warps) they take the second spare chunk's room.
    const size_t stage_bytes = 2 * (size_t)GF_CH * a.in_pitch * sizeof(float), chunk_bytes = (size_t)GF_CH * RP * sizeof(float);
    const size_t room = ctx->smem_optin > 1024 ? ctx->smem_optin - 1024 : 0;  // static shared memory: the barriers
    a.raw_pitch = 0;
    size_t raw_bytes = 0;
    // Measured on B200 (8 x 4K, sigma 10): Rgba8 1.59 ms with the raw buffers against 1.72 without (113 words a row leave the
    // register path's fourth load slot half empty); Rgb8 1.26 against 1.18, Luma8 0.43 against 0.42 -- so four channels only.
    const char *raw_env = getenv("B200IMG_GF_RAW");  // tuning aid: 0 / 1 force it off / on
    const bool want_raw = raw_env ? atoi(raw_env) != 0 : C == 4;
    if (want_raw && std::is_same<T, uint8_t>::value && ((size_t)w * C) % 4 == 0 && (reinterpret_cast<uintptr_t>(src) & 3) == 0 &&
        ((size_t)w * src_rows_buf * C) % 4 == 0) {
        a.raw_pitch = (a.sp * C + 3 + 3) / 4;   // the widest span plus its byte phase, in words
        raw_bytes = 2 * (size_t)GF_CH * a.raw_pitch * 4;
        if (stage_bytes + raw_bytes + (size_t)(a.win_chunks + 1) * chunk_bytes > room) { a.raw_pitch = 0; raw_bytes = 0; }
    }
    int spare = 2;
    if (const char *e = getenv("B200IMG_GF_SPARE")) spare = std::max(1, std::min(2, atoi(e)));  // tuning aid
    while (spare > 0 && stage_bytes + raw_bytes + (size_t)(a.win_chunks + spare) * chunk_bytes > room) --spare;
    if (spare == 0 || a.win_chunks + spare > GF_MAX_SLOTS) return 1;
    a.RR = (a.win_chunks + spare) * GF_CH;
    const size_t smem = stage_bytes + raw_bytes + (size_t)a.RR * RP * sizeof(float);
    if (dst_rows == 0 || n_images == 0) return 0;

    a.src = src; a.dst = dst; a.W = w; a.H = h;
    a.src_row0 = src_row0; a.src_rows_buf = src_rows_buf; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.src_img_stride = (size_t)w * src_rows_buf * C;
    a.dst_img_stride = (size_t)w * dst_rows * C;
    a.strips = (int)((w + TW - 1) / TW);
    a.chunks = (int)((dst_rows + GF_CH - 1) / GF_CH);
    a.units = (long long)a.strips * n_images * a.chunks;
    // Persistent CTAs, one per SM, each a contiguous range of the unit stream (equal shares: no wave quantisation).
    // A range that starts inside a strip recomputes the (win_chunks - 1) row chunks before its first output chunk --
    // the only redundant work there is.  Small jobs: not more CTAs than there are units.
    long long ctas = std::min<long long>(ctx->sm_count, a.units);
    // Where it pays (measured on B200, tools/time_gauss.py): long kernels on jobs that give every SM a good run of
    // chunks.  Short kernels leave too little multiply-add work per chunk to cover the per-chunk hand-offs (21 + 21
    // taps on 8 x 4K: 0.99 ms against 0.69 ms for the kernel pair), small jobs are mostly pipeline fill.
    if (!force && (rlen + clen < 64 || a.units < 8ll * ctx->sm_count)) return 1;
    a.wait_hint_ns = 1000;
    a.dbg = 0;
    if (const char *e = getenv("B200IMG_GF_DBG")) { a.dbg = atoi(e) & 255; }
    if (const char *e = getenv("B200IMG_GF_CTAS")) {  // tuning aid (tools/): force the number of CTAs
        const long long want = atoll(e);
        if (want >= 1) ctas = std::min<long long>(want, a.units);
    }

    FusedWeights wts;
    memset(&wts, 0, sizeof wts);
    auto put = [&](f32x2 *out, int at, const uint32_t *k, int len) {
        for (int i = 0; i < len; ++i) {
            uint32_t b = k[i];
            if (q15) { const float f = (float)b; memcpy(&b, &f, 4); }  // the same integers as floats
            out[at + i] = ((f32x2)b << 32) | b;
        }
    };
    put(wts.r, 0, rk, rlen);
    put(wts.c, a.lead_c, ck, clen);

    const dim3 grid((unsigned)ctas);
#define GF_GO(CC)                                                                                                   \
    {                                                                                                               \
        auto k = q15 ? gauss_fused_kernel<T, CC, std::is_same<T, uint8_t>::value> : gauss_fused_kernel<T, CC, false>; \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                 \
        k<<<grid, GF_THREADS, smem, st>>>(a, wts);                                                            \
    }
    switch (C) {
        case 1: GF_GO(1) break;
        case 2: GF_GO(2) break;
        case 3: GF_GO(3) break;
        default: GF_GO(4) break;
    }
#undef GF_GO
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template int gauss_fused_run<uint8_t>(const uint8_t *, uint8_t *, uint32_t, uint32_t, int, uint32_t, uint32_t, uint32_t, uint32_t,
                                      uint32_t, const uint32_t *, int, const uint32_t *, int, bool, bool, const DeviceCtx *, cudaStream_t);
template int gauss_fused_run<uint16_t>(const uint16_t *, uint16_t *, uint32_t, uint32_t, int, uint32_t, uint32_t, uint32_t, uint32_t,
                                       uint32_t, const uint32_t *, int, const uint32_t *, int, bool, bool, const DeviceCtx *, cudaStream_t);
template int gauss_fused_run<float>(const float *, float *, uint32_t, uint32_t, int, uint32_t, uint32_t, uint32_t, uint32_t,
                                    uint32_t, const uint32_t *, int, const uint32_t *, int, bool, bool, const DeviceCtx *, cudaStream_t);

}  // namespace b200
