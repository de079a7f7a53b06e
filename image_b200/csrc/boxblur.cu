// boxblur.cu -- imageops::fast_blur: three box blurs (Kovesi), each a horizontal
// then a vertical sliding-sum pass that stores back to the pixel type.
//
// Reference: src/imageops/fast_blur.rs -- fast_blur 111-191, boxes_for_gauss
// 201-226, box_blur_horizontal_pass 303-413, box_blur_vertical_pass 416-473,
// accumulators 36-94 (u8: u32 sums and an exact multiply-shift division; other
// types: f32 sums times 1/ks, then clamp_nearest_from).
//
// The reference slides one running sum along each whole row / column.  A window
// sum of integers is the same number however it is accumulated, so u8 (u32
// sums) and u16 while ks * 65535 < 2^24 (the reference's f32 sums of integers
// below 2^24 are exact) come out bit-identical even though the sums here are
// restarted per tile.  For f32 pixels, and u16 with kernels beyond 257 taps, the
// reference's rounding history along the whole row is not reproducible by any
// parallel schedule; those agree to f32 round-off.
//
//   vertical pass:   thread = one 32-bit word of a row (4 u8 / 2 u16 / 1 f32
//                    subpixels) x a band of rows; the running sums live in
//                    registers, loads and stores are fully coalesced.
//   horizontal pass: a CTA stages 32 rows x a span of pixels in shared memory;
//                    lane = row (row pitch = odd number of words: conflict-free),
//                    each lane slides its sums along a 64-pixel chunk; results go
//                    back through shared memory so global stores are contiguous.
#include "common.cuh"

#include <cstdlib>
#include "host_tables.h"

#include <mutex>
#include <type_traits>

namespace b200 {

// boxblur_u8.cu: word-at-a-time passes for u8 images (the common case)
int box_pass_u8(const uint8_t *src, uint8_t *tmp, uint8_t *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t radius,
                const DeviceCtx *ctx, cudaStream_t st);
// boxblur_u8.cu: both passes of a box in one kernel (the horizontally blurred rows stay in shared memory)
bool box_fused_u8_applies(uint32_t w, int C, uint32_t radius, const void *a, const void *b, const void *c, const DeviceCtx *ctx);
int box_fused_u8(const uint8_t *src, uint8_t *dst, uint32_t w, uint32_t h, int C, uint32_t n_images, uint32_t radius,
                 const DeviceCtx *ctx, cudaStream_t st);

namespace {

constexpr int THREADS = 256;

struct BoxGeom {
    uint32_t w, h, C;
    uint32_t r, ks, recip, bias;  // radius, 2r+1, ceil(2^32 / ks), ks / 2
    float inv;                    // 1.0f / ks
};

// per-type sum / store rules
template <typename T> struct Box;
template <> struct Box<uint8_t> {
    typedef uint32_t sum_t;
    static __device__ __forceinline__ uint32_t lift(uint8_t v) { return v; }
    // to_store, fast_blur.rs:66-70: ((acc + ks/2) * ceil(2^32/ks)) >> 32
    static __device__ __forceinline__ uint8_t store(uint32_t s, const BoxGeom &g) { return (uint8_t)__umulhi(s + g.bias, g.recip); }
};
template <> struct Box<uint16_t> {
    typedef uint32_t sum_t;  // integer sums: equal to the reference's f32 sums while those are exact
    static __device__ __forceinline__ uint32_t lift(uint16_t v) { return v; }
    static __device__ __forceinline__ uint16_t store(uint32_t s, const BoxGeom &g) {
        return Px<uint16_t>::clamp_nearest(__fmul_rn((float)s, g.inv));
    }
};
template <> struct Box<float> {
    typedef float sum_t;
    static __device__ __forceinline__ float lift(float v) { return v; }
    static __device__ __forceinline__ float store(float s, const BoxGeom &g) { return Px<float>::clamp_nearest(__fmul_rn(s, g.inv)); }
};

// ---- vertical pass ---------------------------------------------------------------------
constexpr int V_BAND = 64;  // rows per thread

template <typename T>
__global__ void __launch_bounds__(THREADS, 5)
box_v_kernel(const T *__restrict__ src, T *__restrict__ dst, BoxGeom g, uint32_t n_images, int bands) {
    typedef typename Box<T>::sum_t S;
    constexpr int N = 4 / sizeof(T);  // subpixels per 32-bit word
    const size_t row_sub = (size_t)g.w * g.C;
    const uint32_t words = (uint32_t)((row_sub + N - 1) / N);
    const uint32_t wd = blockIdx.x * THREADS + threadIdx.x;
    if (wd >= words) return;
    const uint32_t img = blockIdx.y / bands, band = blockIdx.y % bands;
    const int y0 = (int)band * V_BAND, y1 = min(y0 + V_BAND, (int)g.h);
    const int r = (int)g.r, hmax = (int)g.h - 1;
    const T *p = src + (size_t)img * g.h * row_sub;
    T *d = dst + (size_t)img * g.h * row_sub;
    const size_t s0 = (size_t)wd * N;
    const int n = (int)min((size_t)N, row_sub - s0);  // ragged last word
    // whole words only when every row start is word aligned
    const bool vec = n == N && (row_sub * sizeof(T)) % 4 == 0 && ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(d)) & 3) == 0;

    auto load = [&](int y, S (&v)[N]) {
        const T *q = p + (size_t)y * row_sub + s0;
        if (vec) {
            const uint32_t wv = *reinterpret_cast<const uint32_t *>(q);
            if constexpr (N == 4) { v[0] = wv & 0xff; v[1] = (wv >> 8) & 0xff; v[2] = (wv >> 16) & 0xff; v[3] = wv >> 24; }
            else if constexpr (N == 2) { v[0] = wv & 0xffff; v[1] = wv >> 16; }
            else v[0] = __uint_as_float(wv);
        } else {
#pragma unroll
            for (int j = 0; j < N; ++j) v[j] = j < n ? Box<T>::lift(q[j]) : (S)0;
        }
    };

    // sums for row y0: sum_{d=-r..r} p[clamp(y0 + d)]
    S sum[N];
#pragma unroll
    for (int j = 0; j < N; ++j) sum[j] = (S)0;
    {
        const int lo = y0 - r, hi = y0 + r;
        const int nlo = lo < 0 ? -lo : 0, nhi = hi > hmax ? hi - hmax : 0;
        S v[N];
        // loads batched 8 deep: a load-use-per-iteration loop would serialise on DRAM latency
        constexpr int PB = 8;
        int y = max(lo, 0);
        const int yend = min(hi, hmax);
        for (; y + PB - 1 <= yend; y += PB) {
            S vb[PB][N];
#pragma unroll
            for (int k = 0; k < PB; ++k) load(y + k, vb[k]);
#pragma unroll
            for (int k = 0; k < PB; ++k) {
#pragma unroll
                for (int j = 0; j < N; ++j) sum[j] += vb[k][j];
            }
        }
        for (; y <= yend; ++y) {
            load(y, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j];
        }
        if (nlo) { load(0, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j] * (S)nlo; }
        if (nhi) { load(hmax, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j] * (S)nhi; }
    }
    auto emit = [&](int y) {
        T o[N];
#pragma unroll
        for (int j = 0; j < N; ++j) o[j] = Box<T>::store(sum[j], g);
        T *q = d + (size_t)y * row_sub + s0;
        if (vec) {
            uint32_t wv;
            if constexpr (N == 4) wv = (uint32_t)o[0] | ((uint32_t)o[1] << 8) | ((uint32_t)o[2] << 16) | ((uint32_t)o[3] << 24);
            else if constexpr (N == 2) wv = (uint32_t)o[0] | ((uint32_t)o[1] << 16);
            else wv = __float_as_uint(o[0]);
            *reinterpret_cast<uint32_t *>(q) = wv;
        } else {
#pragma unroll
            for (int j = 0; j < N; ++j) if (j < n) q[j] = o[j];
        }
    };
    constexpr int U = 4;  // rows per iteration: 2*U independent loads in flight per thread
    int y = y0;
    for (; y + U <= y1; y += U) {
        S a[U][N], b[U][N];
#pragma unroll
        for (int k = 0; k < U; ++k) { load(min(y + k + r + 1, hmax), a[k]); load(max(y + k - r, 0), b[k]); }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            emit(y + k);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] = sum[j] + a[k][j] - b[k][j];  // the reference's order: add, then subtract
        }
    }
    for (; y < y1; ++y) {
        emit(y);
        S a[N], b[N];
        load(min(y + r + 1, hmax), a);
        load(max(y - r, 0), b);
#pragma unroll
        for (int j = 0; j < N; ++j) sum[j] = sum[j] + a[j] - b[j];
    }
}

// The same pass with a 16-byte column per thread (8 u16 or 4 f32 subpixels) for rows that are whole vectors: the
// word-per-thread kernel above keeps 32 bytes in flight per thread and runs at a quarter of the HBM rate on 16-bit and f32
// images; this one keeps 128.  Same sums, same order of the add and the subtract.
template <typename T>
__global__ void __launch_bounds__(THREADS)
box_v_wide_kernel(const T *__restrict__ src, T *__restrict__ dst, BoxGeom g, uint32_t n_images, int bands) {
    typedef typename Box<T>::sum_t S;
    constexpr int N = 16 / sizeof(T);
    const size_t row_sub = (size_t)g.w * g.C;
    const uint32_t vecs = (uint32_t)(row_sub / N);
    const uint32_t vi = blockIdx.x * THREADS + threadIdx.x;
    if (vi >= vecs) return;
    const uint32_t img = blockIdx.y / bands, band = blockIdx.y % bands;
    const int y0 = (int)band * V_BAND, y1 = min(y0 + V_BAND, (int)g.h);
    const int r = (int)g.r, hmax = (int)g.h - 1;
    const T *p = src + (size_t)img * g.h * row_sub + (size_t)vi * N;
    T *d = dst + (size_t)img * g.h * row_sub + (size_t)vi * N;

    auto load = [&](int y, S (&v)[N]) {
        const uint4 q = *reinterpret_cast<const uint4 *>(p + (size_t)y * row_sub);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < N; ++j) {
            if constexpr (sizeof(T) == 2) v[j] = (w[j >> 1] >> (16 * (j & 1))) & 0xffffu;
            else v[j] = __uint_as_float(w[j]);
        }
    };
    auto emit = [&](int y, const S (&sum)[N]) {
        uint32_t w[4];
        if constexpr (sizeof(T) == 2) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                w[j] = (uint32_t)Box<T>::store(sum[2 * j], g) | ((uint32_t)Box<T>::store(sum[2 * j + 1], g) << 16);
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = __float_as_uint(Box<T>::store(sum[j], g));
        }
        *reinterpret_cast<uint4 *>(d + (size_t)y * row_sub) = make_uint4(w[0], w[1], w[2], w[3]);
    };

    S sum[N];
#pragma unroll
    for (int j = 0; j < N; ++j) sum[j] = (S)0;
    {   // sums for row y0: sum_{d=-r..r} p[clamp(y0 + d)], in ascending row order like the word kernel
        const int lo = y0 - r, hi = y0 + r;
        const int nlo = lo < 0 ? -lo : 0, nhi = hi > hmax ? hi - hmax : 0;
        S v[N];
        constexpr int PB = 4;
        int y = max(lo, 0);
        const int yend = min(hi, hmax);
        for (; y + PB - 1 <= yend; y += PB) {
            S vb[PB][N];
#pragma unroll
            for (int k = 0; k < PB; ++k) load(y + k, vb[k]);
#pragma unroll
            for (int k = 0; k < PB; ++k)
#pragma unroll
                for (int j = 0; j < N; ++j) sum[j] += vb[k][j];
        }
        for (; y <= yend; ++y) {
            load(y, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j];
        }
        if (nlo) {
            load(0, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j] * (S)nlo;
        }
        if (nhi) {
            load(hmax, v);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] += v[j] * (S)nhi;
        }
    }
    constexpr int U = 4;
    int y = y0;
    for (; y + U <= y1; y += U) {
        S a[U][N], b[U][N];
#pragma unroll
        for (int k = 0; k < U; ++k) { load(min(y + k + r + 1, hmax), a[k]); load(max(y + k - r, 0), b[k]); }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            emit(y + k, sum);
#pragma unroll
            for (int j = 0; j < N; ++j) sum[j] = sum[j] + a[k][j] - b[k][j];  // the reference's order: add, then subtract
        }
    }
    for (; y < y1; ++y) {
        emit(y, sum);
        S a[N], b[N];
        load(min(y + r + 1, hmax), a);
        load(max(y - r, 0), b);
#pragma unroll
        for (int j = 0; j < N; ++j) sum[j] = sum[j] + a[j] - b[j];
    }
}

// ---- horizontal pass ---------------------------------------------------------------------
constexpr int H_ROWS = 32;  // rows per CTA tile = lanes
constexpr int H_XT = 256;   // output pixels per CTA tile
constexpr int H_XC = 64;    // pixels per thread chunk

// Copies `n` elements between a global row and a shared row that share the same 4-byte phase:
// scalar head up to the first word boundary, 32-bit body, scalar tail.  No over-read.  (Words,
// not uint4: shared rows must keep an odd word pitch for the lane-per-row passes, so they are
// only word aligned.)
template <typename T, bool TO_SHARED>
__device__ __forceinline__ void copy_row_phased(const T *from, T *to, int n, int lane) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(TO_SHARED ? (const void *)from : (const void *)to);  // the global side
    constexpr int PER = 4 / sizeof(T);
    int head = (int)(((4 - (a & 3)) & 3) / sizeof(T));
    if (head > n) head = n;
    const int body = (n - head) / PER;
    if (lane < head) to[lane] = from[lane];
    const uint32_t *vs = reinterpret_cast<const uint32_t *>(from + head);
    uint32_t *vd = reinterpret_cast<uint32_t *>(to + head);
    if constexpr (TO_SHARED) {
        // asynchronous 4-byte copies: every word of the row is in flight at once, no registers held
        for (int v = lane; v < body; v += 32) {
            const uint32_t d = (uint32_t)__cvta_generic_to_shared(vd + v);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d), "l"(vs + v) : "memory");
        }
    } else {
#pragma unroll 4
        for (int v = lane; v < body; v += 32) vd[v] = vs[v];
    }
    const int done = head + body * PER;
    if (lane < n - done) to[done + lane] = from[done + lane];
}

template <typename T, int C>
__global__ void __launch_bounds__(THREADS)
box_h_kernel(const T *__restrict__ src, T *__restrict__ dst, BoxGeom g, size_t total_rows, int in_pitch, int out_pitch, int xt,
             int xc) {
    typedef typename Box<T>::sum_t S;
    extern __shared__ __align__(16) unsigned char smem[];
    T *s_in = reinterpret_cast<T *>(smem);                              // [H_ROWS][in_pitch]   (pitch in elements; odd number of words)
    T *s_out = reinterpret_cast<T *>(smem) + (size_t)H_ROWS * in_pitch;  // [H_ROWS][out_pitch]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int x0 = blockIdx.x * xt;   // xt = output pixels per CTA tile, xc = pixels per warp chunk (host's choice)
    const size_t row0 = (size_t)blockIdx.y * H_ROWS;
    const int r = (int)g.r, W = (int)g.w;
    const int tw = min(xt, W - x0);
    const int sp = tw + 2 * r;  // staged pixels: x0 - r .. x0 + tw - 1 + r, clamped
    const size_t g_pitch = (size_t)W * C;
    const bool interior = x0 - r >= 0 && x0 + tw + r <= W;

    // a staged row starts at the same 4-byte phase as its global source, so the copy moves whole words
    auto in_phase = [&](size_t row) -> int {
        if (!interior) return 0;
        return (int)((reinterpret_cast<uintptr_t>(src + row * g_pitch + (size_t)(x0 - r) * C) & 3) / sizeof(T));
    };
    auto out_phase = [&](size_t row) -> int {
        return (int)((reinterpret_cast<uintptr_t>(dst + row * g_pitch + (size_t)x0 * C) & 3) / sizeof(T));
    };

    for (int rr = warp; rr < H_ROWS; rr += THREADS / 32) {
        const size_t row = min(row0 + rr, total_rows - 1);
        const T *grow = src + row * g_pitch;
        T *s = s_in + (size_t)rr * in_pitch + in_phase(row);
        if (interior) {
            copy_row_phased<T, true>(grow + (size_t)(x0 - r) * C, s, sp * C, lane);
        } else {
            for (int e = lane; e < sp * C; e += 32) {
                const int p = e / C, c = e - p * C;
                s[e] = grow[(size_t)min(max(x0 - r + p, 0), W - 1) * C + c];
            }
        }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();

    const int chunks = (tw + xc - 1) / xc;
    const size_t my_row = min(row0 + lane, total_rows - 1);
    const T *in_row = s_in + (size_t)lane * in_pitch + in_phase(my_row);
    T *out_row = s_out + (size_t)lane * out_pitch + out_phase(my_row);
    for (int item = warp; item < chunks * C; item += THREADS / 32) {
        const int ck = item / C, c = item - ck * C;
        const int xa = ck * xc, xb = min(xa + xc, tw);  // outputs [xa, xb) of this tile
        const T *in = in_row + c;                            // staged pixel j <-> source pixel x0 - r + j
        T *out = out_row + c;
        S sum = (S)0;
        for (int j = xa; j <= xa + 2 * r; ++j) sum += Box<T>::lift(in[j * C]);  // window of output xa
#pragma unroll 4
        for (int x = xa; x < xb; ++x) {
            out[x * C] = Box<T>::store(sum, g);
            sum = sum + Box<T>::lift(in[(x + 2 * r + 1) * C]) - Box<T>::lift(in[x * C]);
        }
    }
    __syncthreads();

    for (int rr = warp; rr < H_ROWS; rr += THREADS / 32) {
        const size_t row = row0 + rr;
        if (row >= total_rows) break;
        copy_row_phased<T, false>(s_out + (size_t)rr * out_pitch + out_phase(row), dst + row * g_pitch + (size_t)x0 * C, tw * C, lane);
    }
}

// Fallback for radii whose staged tile does not fit shared memory: direct window sum per output.
template <typename T>
__global__ void __launch_bounds__(THREADS)
box_h_direct_kernel(const T *__restrict__ src, T *__restrict__ dst, BoxGeom g, size_t total) {
    typedef typename Box<T>::sum_t S;
    const size_t row_sub = (size_t)g.w * g.C;
    const int wmax = (int)g.w - 1, r = (int)g.r;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const size_t row = idx / row_sub;
        const uint32_t s = (uint32_t)(idx - row * row_sub);
        const int x = (int)(s / g.C), c = (int)(s - (uint32_t)x * g.C);
        const T *p = src + row * row_sub + c;
        const int lo = x - r, hi = x + r;
        const int nlo = lo < 0 ? -lo : 0, nhi = hi > wmax ? hi - wmax : 0;
        S acc = (S)0;
        for (int xx = max(lo, 0); xx <= min(hi, wmax); ++xx) acc += Box<T>::lift(p[(size_t)xx * g.C]);
        if (nlo) acc += Box<T>::lift(p[0]) * (S)nlo;
        if (nhi) acc += Box<T>::lift(p[(size_t)wmax * g.C]) * (S)nhi;
        dst[idx] = Box<T>::store(acc, g);
    }
}

// Row pitch (in elements of T) for a staged row of n elements: room for a 4-byte phase shift,
// whole words, and an odd number of them so the 32 lanes (one row each) hit 32 different banks.
int staged_pitch(int n, int tsize) {
    int words = (n * tsize + 3 + 3) / 4;
    words |= 1;
    return words * 4 / tsize;
}

template <typename T>
int run(const void *src, void *dst, void *tmp, uint32_t w, uint32_t h, int C, uint32_t n_images, const uint32_t radii[3],
        const DeviceCtx *ctx, cudaStream_t st) {
    const size_t total_rows = (size_t)h * n_images;
    const T *cur = (const T *)src;
    if constexpr (std::is_same<T, uint8_t>::value) {
        // three fused kernels, src -> dst -> tmp -> dst: one read and one write of the image per box
        bool fused = getenv("B200IMG_BOX_UNFUSED") == nullptr && src != dst;  // (in place: the first box would read rows other CTAs overwrite)
        for (int pass = 0; pass < 3 && fused; ++pass) fused = box_fused_u8_applies(w, C, radii[pass], src, tmp, dst, ctx);
        if (fused) {
            const uint8_t *from[3] = {(const uint8_t *)src, (const uint8_t *)dst, (const uint8_t *)tmp};
            uint8_t *to[3] = {(uint8_t *)dst, (uint8_t *)tmp, (uint8_t *)dst};
            for (int pass = 0; pass < 3; ++pass) {
                const int rc = box_fused_u8(from[pass], to[pass], w, h, C, n_images, radii[pass], ctx, st);
                if (rc != 1) return rc < 0 ? rc : fail(B200IMG_ERR_CUDA, "fast_blur: fused box kernel declined after accepting");
            }
            return B200IMG_OK;
        }
    }
    for (int pass = 0; pass < 3; ++pass) {
        BoxGeom g;
        g.w = w; g.h = h; g.C = (uint32_t)C;
        g.r = radii[pass];
        g.ks = 2 * g.r + 1;
        g.recip = UINT32_MAX / g.ks + 1u;
        g.bias = g.ks / 2;
        g.inv = 1.0f / (float)g.ks;
        if constexpr (std::is_same<T, uint8_t>::value) {
            const int rc = box_pass_u8((const uint8_t *)cur, (uint8_t *)tmp, (uint8_t *)dst, w, h, C, n_images, g.r, ctx, st);
            if (rc < 0) return rc;
            if (rc == 1) { cur = (const T *)dst; continue; }
        }
        {
            // one staged element past the last window keeps the sliding update branch-free.  Tile width:
            // the widest of 256 / 128 / 64 / 32 pixels whose staged rows leave room for two or three CTAs per SM
            // (f32 x 4 channels at 256 pixels would not fit at all); 32-pixel chunks keep all 8 warps busy
            int xt = H_XT, in_pitch = 0, out_pitch = 0;
            size_t smem = 0;
            for (;; xt /= 2) {
                in_pitch = staged_pitch((xt + 2 * (int)g.r + 1) * C, sizeof(T));
                out_pitch = staged_pitch(xt * C, sizeof(T));
                smem = (size_t)H_ROWS * (in_pitch + out_pitch) * sizeof(T);
                if (smem <= 80 * 1024 || xt <= 32) break;   // (80 KB, not 100: 8 x 4K Rgb16 1.79 -> 1.63 ms with 128-pixel tiles)
            }
            const int xc = std::max(8, std::min(32, xt * C / 8));
            const dim3 grid((w + xt - 1) / xt, (unsigned)((total_rows + H_ROWS - 1) / H_ROWS));
            if (smem > ctx->smem_optin) {
                const size_t total = total_rows * w * C;
                const unsigned gd = (unsigned)std::min<size_t>((total + THREADS - 1) / THREADS, (size_t)ctx->sm_count * 32);
                box_h_direct_kernel<T><<<gd, THREADS, 0, st>>>(cur, (T *)tmp, g, total);
            } else
#define H_GO(CC)                                                                                            \
    {                                                                                                       \
        auto k = box_h_kernel<T, CC>;                                                                       \
        B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));         \
        k<<<grid, THREADS, smem, st>>>(cur, (T *)tmp, g, total_rows, in_pitch, out_pitch, xt, xc);          \
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
            constexpr int N = 4 / sizeof(T);
            const uint32_t words = (uint32_t)(((size_t)w * C + N - 1) / N);
            const int bands = (int)((h + V_BAND - 1) / V_BAND);
            const dim3 grid((words + THREADS - 1) / THREADS, (unsigned)(bands * n_images));
            bool wide = false;
            if constexpr (sizeof(T) == 2) {   // (f32 images measured slower this way: 2.53 against 2.36 ms for 8 x 4K Rgb32F)
                constexpr int NW = 16 / sizeof(T);
                wide = ((size_t)w * C) % NW == 0 && ((reinterpret_cast<uintptr_t>(tmp) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
                if (wide) {
                    const uint32_t vecs = (uint32_t)((size_t)w * C / NW);
                    const dim3 gridw((vecs + THREADS - 1) / THREADS, (unsigned)(bands * n_images));
                    box_v_wide_kernel<T><<<gridw, THREADS, 0, st>>>((const T *)tmp, (T *)dst, g, n_images, bands);
                }
            }
            if (!wide) box_v_kernel<T><<<grid, THREADS, 0, st>>>((const T *)tmp, (T *)dst, g, n_images, bands);
        }
        cur = (const T *)dst;
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace

int fast_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     float sigma, cudaStream_t st) {
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    uint32_t radii[3];
    boxes_for_gauss(sigma, radii);
    for (int i = 0; i < 3; ++i) {
        if (radii[i] >= (1u << 23))  // beyond here the reference's u8 path switches accumulators (fast_blur.rs:238-247)
            return fail(B200IMG_ERR_UNSUPPORTED, "fast_blur: box radius %u is beyond the supported range", radii[i]);
        // u16 running sums are u32 here (exact; the reference's are f32): they wrap once ks * 65535 >= 2^32
        if (dtype == B200IMG_U16 && 2ull * radii[i] + 1 > 65537ull)
            return fail(B200IMG_ERR_UNSUPPORTED, "fast_blur: box radius %u overflows the 32-bit sums of 16-bit images", radii[i]);
    }
    size_t bytes;
    if (!checked_mul((size_t)w * h, (size_t)C, dtype_size(dtype), n_images, &bytes))
        return fail(B200IMG_ERR_DIMENSION, "fast_blur: size overflows");
    void *tmp;
    int rc = scratch(SCRATCH_A, bytes, &tmp);
    if (rc) return rc;
    switch (dtype) {
        case B200IMG_U8: return run<uint8_t>(src, dst, tmp, w, h, C, n_images, radii, ctx, st);
        case B200IMG_U16: return run<uint16_t>(src, dst, tmp, w, h, C, n_images, radii, ctx, st);
        default: return run<float>(src, dst, tmp, w, h, C, n_images, radii, ctx, st);
    }
}


// ---- b200img_fast_blur_noalloc_dev: the box blur for callers that must not allocate ----------
//
// The ordinary entry points take the transient image of the blur (the horizontal pass writes
// it, the vertical pass reads it) from a cached cudaMalloc'd pool.  Some callers cannot let a
// library do that: a serving process that has handed the whole GPU to its own memory pool (a
// cudaMalloc from here would fail or eat into memory it has budgeted), or a latency-critical
// loop in which cudaMalloc's implicit device synchronisation and driver lock are not acceptable.
// For them the transient lives in this fixed array instead.  It is part of the library image:
// every GPU that loads libb200img holds these 32 MiB until the process exits.
namespace {

constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)32 << 20;  // fits 3840 x 2160 Rgba8 (31.6 MiB)
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address; the
// address is looked up once per device (with that device current) and remembered here.
constexpr int NOALLOC_MAX_DEVICES = 64;
void *g_noalloc_ptr[NOALLOC_MAX_DEVICES];
bool g_noalloc_ready[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mu;

int noalloc_scratch_ptr(int dev, void **out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ready[dev]) {
        void *p = nullptr;
        B200_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        g_noalloc_ptr[dev] = p;
        g_noalloc_ready[dev] = true;
    }
    *out = g_noalloc_ptr[dev];
    return B200IMG_OK;
}

}  // namespace
}  // namespace b200

extern "C" int b200img_fast_blur_noalloc_dev(const void *d_src, uint32_t w, uint32_t h, int channels, int dtype,
                                              void *d_dst, float sigma, void *stream) {
    using namespace b200;
    if (channels < 1 || channels > 4) return fail(B200IMG_ERR_BAD_ARG, "fast_blur_noalloc: channels must be 1..4 (got %d)", channels);
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "fast_blur_noalloc: bad dtype %d", dtype);
    size_t bytes;
    if (!checked_mul(w, h, (size_t)channels, dtype_size(dtype), &bytes)) return fail(B200IMG_ERR_DIMENSION, "fast_blur_noalloc: size overflows");
    if (bytes == 0) return B200IMG_OK;  // empty image: nothing to blur
    if (bytes > NOALLOC_SCRATCH_BYTES)
        return fail(B200IMG_ERR_UNSUPPORTED,
                    "fast_blur_noalloc: the image needs %zu bytes of transient storage but the built-in scratch holds %zu; "
                    "use b200img_fast_blur_dev (which allocates) for images this large", bytes, NOALLOC_SCRATCH_BYTES);
    if (!d_src || !d_dst) return fail(B200IMG_ERR_BAD_ARG, "fast_blur_noalloc: null buffer");
    uint32_t radii[3];
    boxes_for_gauss(sigma, radii);
    for (int i = 0; i < 3; ++i)
        if (radii[i] >= (1u << 23) || (dtype == B200IMG_U16 && 2ull * radii[i] + 1 > 65537ull))
            return fail(B200IMG_ERR_UNSUPPORTED, "fast_blur_noalloc: box radius %u is beyond the supported range", radii[i]);

    int dev;
    B200_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(B200IMG_ERR_UNSUPPORTED, "fast_blur_noalloc: device index %d not supported", dev);
    // device facts straight from the runtime: the shared per-device context lives on the heap
    DeviceCtx ctx;
    int sms = 0, smem = 0, major = 0;
    B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    B200_CUDA(cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
    B200_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
    if (major < 10) return fail(B200IMG_ERR_CUDA, "fast_blur_noalloc: device %d is not sm_100", dev);
    ctx.device = dev; ctx.sm_count = sms; ctx.smem_optin = (size_t)smem;

    void *tmp;
    int rc = noalloc_scratch_ptr(dev, &tmp);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    switch (dtype) {
        case B200IMG_U8: return run<uint8_t>(d_src, d_dst, tmp, w, h, channels, 1, radii, &ctx, st);
        case B200IMG_U16: return run<uint16_t>(d_src, d_dst, tmp, w, h, channels, 1, radii, &ctx, st);
        default: return run<float>(d_src, d_dst, tmp, w, h, channels, 1, radii, &ctx, st);
    }
}
