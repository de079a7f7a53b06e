// resample_group.cu -- the fused tile resampler with register reuse on both axes.
//
// Same structure as resize_kernel in resample.cu (one CTA per destination tile: stage the source
// region, vertical pass into an f32 shared-memory intermediate, horizontal pass out of it; the
// reference's order, vertical_sample then horizontal_sample, src/imageops/sample.rs:494-562 and
// 217-340), but every thread produces a GROUP of GR = 4 neighbouring destination samples along
// the filter axis and walks the UNION of their windows once:
//   * a staged sample is read from shared memory once per group instead of once per destination
//     sample -- the tile kernel is bound by the shared-memory pipe (8 bytes read per FFMA2 lane);
//     for a 2x Lanczos3 upscale the union is 9 taps against 4 x 7;
//   * the weights of the group are a dense [union][GR] table in shared memory, zero where a tap
//     lies outside a sample's own window, so the walk has no branches, and lanes are arranged so
//     that a warp reads ONE table row (broadcast): vertical pass lanes = neighbouring columns,
//     horizontal pass lanes = rows.
// A zero weight leaves the accumulator unchanged (acc + x*0 rounds to acc) for every finite x,
// so the sums are the reference's sums, tap for tap, also under B200IMG_FP_STRICT.  The one
// difference: a NON-FINITE f32 source sample also reaches the up to three neighbours that share
// its group (0 * inf = NaN); f32 images under FP_STRICT therefore stay on resize_kernel.
#include "common.cuh"
#include "host_tables.h"
#include "packed.cuh"

#include <list>
#include <memory>
#include <mutex>
#include <vector>

namespace b200 {

namespace {

constexpr int GK_THREADS = 256;
constexpr int GR = 4;  // destination samples per group, on either axis

struct GroupArgs {
    const void *src;
    void *dst;
    size_t src_img_stride, dst_img_stride;
    uint32_t w, nw;
    uint32_t src_row0, dst_row0, dst_rows;
    const int32_t *vleft, *vcount; const float *vw; int vstride;
    const int32_t *hleft, *hcount; const float *hw; int hstride;
    float acc_init;
    int TW, TH, log2TH;        // destination tile (powers of two, multiples of GR)
    int max_rows;              // staged source rows, at most
    int src_pitch;             // bytes per staged source row
    int src_bytes;             // bytes of the staged-source region (also holds the staged output tile)
    int tmp_pitch;             // floats per intermediate row
    int out_pitch;             // subpixels per staged output row
    int UV, UH;                // rows / pixels spanned by the windows of one group, at most
    // per-axis group tables, built once per (filter, lengths) on the host: weights [group][U][GR]
    // as (w, w) pairs, zero outside a sample's window; bu[group] = (first source index, span)
    const float2 *vwt, *hwt;
    const int2 *vbu, *hbu;
};

template <typename T, int C, bool STRICT>
__global__ void __launch_bounds__(GK_THREADS)
resize_group_kernel(const __grid_constant__ GroupArgs a) {
    constexpr int N = Word<T>::N;           // subpixels per 32-bit word
    constexpr bool PACKED = !STRICT;        // FFMA2 paths
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char *s_src = smem;
    float *s_tmp = reinterpret_cast<float *>(smem + a.src_bytes);                       // [TH][tmp_pitch]
    float2 *s_vwt = reinterpret_cast<float2 *>(s_tmp + a.TH * a.tmp_pitch);             // [TH/GR][UV][GR] (w, w)
    float2 *s_hwt = s_vwt + (a.TH / GR) * a.UV * GR;                                     // [TW/GR][UH][GR] (w, w)
    int2 *s_vbu = reinterpret_cast<int2 *>(s_hwt + (a.TW / GR) * a.UH * GR);            // [TH/GR] (first source row, rows in the span)
    int2 *s_hbu = s_vbu + a.TH / GR;                                                     // [TW/GR] (first source pixel, pixels in the span)
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const int ox0 = blockIdx.x * a.TW, oy0 = a.dst_row0 + blockIdx.y * a.TH;
    const int tw = min(a.TW, (int)a.nw - ox0);
    const int th = min(a.TH, (int)(a.dst_row0 + a.dst_rows) - oy0);
    const int col0 = a.hleft[ox0];
    const int col1 = a.hleft[ox0 + tw - 1] + a.hcount[ox0 + tw - 1];
    const int row0 = a.vleft[oy0];
    const int row1 = a.vleft[oy0 + th - 1] + a.vcount[oy0 + th - 1];
    const int rows = row1 - row0;
    const T *src = reinterpret_cast<const T *>(reinterpret_cast<const unsigned char *>(a.src) + (size_t)blockIdx.z * a.src_img_stride);
    T *dst = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(a.dst) + (size_t)blockIdx.z * a.dst_img_stride);

    // ---- stage rows [row0,row1) x bytes [b0,b1): one TMA bulk copy per row when rows are 16-byte aligned ----
    const size_t gp = sizeof(T) * (size_t)a.w * C;
    const unsigned char *gimg = reinterpret_cast<const unsigned char *>(src) + (size_t)(row0 - (int)a.src_row0) * gp;
    int b0 = col0 * C * (int)sizeof(T), b1 = col1 * C * (int)sizeof(T);
    const bool rows16 = (gp % 16 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    if (rows16) { b0 &= ~15; b1 = (b1 + 15) & ~15; }
    const int row_bytes = b1 - b0;
    const int sub0 = b0 / (int)sizeof(T);
    const int row_sub = row_bytes / (int)sizeof(T);
    if (rows16) {
        if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
        __syncthreads();
        if (tid == 0) mbar_expect_tx(&bar, (uint32_t)(rows * row_bytes));
        __syncthreads();
        for (int r = tid; r < rows; r += GK_THREADS)
            bulk_g2s(s_src + r * a.src_pitch, gimg + (size_t)r * gp + b0, (uint32_t)row_bytes, &bar);
    } else {
        const unsigned mis = (unsigned)(reinterpret_cast<uintptr_t>(gimg + b0) | gp | (unsigned)row_bytes | 16u);
        const int unit = (int)(mis & (~mis + 1u));  // widest power of two dividing row start, pitch and length
        const int lane = tid & 31;
        if (unit >= 4) {
            const int per_row = row_bytes / unit;
            for (int r = tid >> 5; r < rows; r += GK_THREADS / 32)
                for (int v = lane; v < per_row; v += 32)
                    cp_async(s_src + r * a.src_pitch + v * unit, gimg + (size_t)r * gp + b0 + v * unit, unit);
        } else {
            for (int r = tid >> 5; r < rows; r += GK_THREADS / 32) {
                const unsigned char *g = gimg + (size_t)r * gp + b0;
                unsigned char *sr = s_src + r * a.src_pitch;
#pragma unroll 4
                for (int v = lane; v < row_bytes; v += 32) sr[v] = __ldg(g + v);
            }
        }
    }

    // ---- this tile's slices of the group tables (16-byte async copies; groups are aligned to
    // multiples of GR from sample 0, and so are tiles) ----
    const int vgroups = (th + GR - 1) / GR, hgroups = (tw + GR - 1) / GR;
    {
        const uint4 *gv = reinterpret_cast<const uint4 *>(a.vwt + (size_t)(oy0 / GR) * a.UV * GR);
        const uint4 *gh = reinterpret_cast<const uint4 *>(a.hwt + (size_t)(ox0 / GR) * a.UH * GR);
        const int nv = vgroups * a.UV * (GR / 2), nh = hgroups * a.UH * (GR / 2);
        for (int i = tid; i < nv; i += GK_THREADS) cp_async(reinterpret_cast<uint4 *>(s_vwt) + i, gv + i, 16);
        for (int i = tid; i < nh; i += GK_THREADS) cp_async(reinterpret_cast<uint4 *>(s_hwt) + i, gh + i, 16);
        for (int g = tid; g < vgroups; g += GK_THREADS) s_vbu[g] = a.vbu[oy0 / GR + g];
        for (int g = tid; g < hgroups; g += GK_THREADS) s_hbu[g] = a.hbu[ox0 / GR + g];
    }
    cp_async_wait_all();
    if (rows16) mbar_wait(&bar, 0);
    __syncthreads();

    // ---- vertical pass: item = (row group g, unit x of a staged row); a warp's lanes are neighbouring
    // units of one group, so its weight reads are broadcasts.  Items are THREADS apart; the (g, x)
    // pair advances without a division. ----
    const f32x2 init2 = pack2(a.acc_init, a.acc_init);
    if (sizeof(T) == 4 && PACKED && row_bytes % 16 == 0) {
        const int units = row_bytes / 16;  // four f32 subpixels per unit
        const int dy = GK_THREADS / units, dx = GK_THREADS - dy * units;
        int g = tid / units, x = tid - g * units;
        while (g < vgroups) {
            const int2 bu = s_vbu[g];
            const int U = bu.y;
            const float2 *wt = s_vwt + g * a.UV * GR;
            const unsigned char *p = s_src + (bu.x - row0) * a.src_pitch + x * 16;
            f32x2 a01[GR], a23[GR];
#pragma unroll
            for (int k = 0; k < GR; ++k) { a01[k] = init2; a23[k] = init2; }
            for (int u = 0; u < U; ++u) {
                const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(p);
#pragma unroll
                for (int k = 0; k < GR; k += 2) {
                    const ulonglong2 ww = *reinterpret_cast<const ulonglong2 *>(wt + k);
                    a01[k] = fma2(v.x, ww.x, a01[k]);
                    a23[k] = fma2(v.y, ww.x, a23[k]);
                    a01[k + 1] = fma2(v.x, ww.y, a01[k + 1]);
                    a23[k + 1] = fma2(v.y, ww.y, a23[k + 1]);
                }
                p += a.src_pitch;
                wt += GR;
            }
#pragma unroll
            for (int k = 0; k < GR; ++k)
                if (g * GR + k < th)
                    *reinterpret_cast<ulonglong2 *>(s_tmp + (g * GR + k) * a.tmp_pitch + x * 4) = make_ulonglong2(a01[k], a23[k]);
            x += dx; g += dy;
            if (x >= units) { x -= units; ++g; }
        }
    } else {
        const int units = (row_sub + N - 1) / N;  // 32-bit words of a staged row
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        const int dy = GK_THREADS / units, dx = GK_THREADS - dy * units;
        int g = tid / units, x = tid - g * units;
        while (g < vgroups) {
            const int2 bu = s_vbu[g];
            const int U = bu.y;
            const float2 *wt = s_vwt + g * a.UV * GR;
            const unsigned char *p = s_src + (bu.x - row0) * a.src_pitch + x * 4;
            float *t = s_tmp + g * GR * a.tmp_pitch + x * N;
            if constexpr (PACKED && N == 4) {
                f32x2 a01[GR], a23[GR];
#pragma unroll
                for (int k = 0; k < GR; ++k) { a01[k] = init2; a23[k] = init2; }
                for (int u = 0; u < U; ++u) {
                    float f[4];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
                    const f32x2 f01 = add2(pack2(f[0], f[1]), bias2), f23 = add2(pack2(f[2], f[3]), bias2);
#pragma unroll
                    for (int k = 0; k < GR; k += 2) {
                        const ulonglong2 ww = *reinterpret_cast<const ulonglong2 *>(wt + k);
                        a01[k] = fma2(f01, ww.x, a01[k]);
                        a23[k] = fma2(f23, ww.x, a23[k]);
                        a01[k + 1] = fma2(f01, ww.y, a01[k + 1]);
                        a23[k + 1] = fma2(f23, ww.y, a23[k + 1]);
                    }
                    p += a.src_pitch;
                    wt += GR;
                }
#pragma unroll
                for (int k = 0; k < GR; ++k)
                    if (g * GR + k < th) *reinterpret_cast<ulonglong2 *>(t + k * a.tmp_pitch) = make_ulonglong2(a01[k], a23[k]);
            } else if constexpr (PACKED && N == 2) {
                f32x2 a01[GR];
#pragma unroll
                for (int k = 0; k < GR; ++k) a01[k] = init2;
                for (int u = 0; u < U; ++u) {
                    float f[2];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
                    const f32x2 f01 = add2(pack2(f[0], f[1]), bias2);
#pragma unroll
                    for (int k = 0; k < GR; k += 2) {
                        const ulonglong2 ww = *reinterpret_cast<const ulonglong2 *>(wt + k);
                        a01[k] = fma2(f01, ww.x, a01[k]);
                        a01[k + 1] = fma2(f01, ww.y, a01[k + 1]);
                    }
                    p += a.src_pitch;
                    wt += GR;
                }
#pragma unroll
                for (int k = 0; k < GR; ++k)
                    if (g * GR + k < th) *reinterpret_cast<f32x2 *>(t + k * a.tmp_pitch) = a01[k];
            } else {
                float acc[GR][N];
#pragma unroll
                for (int k = 0; k < GR; ++k)
#pragma unroll
                    for (int j = 0; j < N; ++j) acc[k][j] = a.acc_init;
                for (int u = 0; u < U; ++u) {
                    float f[N];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
#pragma unroll
                    for (int j = 0; j < N; ++j) f[j] = unbias<T>(f[j]);
#pragma unroll
                    for (int k = 0; k < GR; ++k) {
                        const float wgt = wt[k].x;
#pragma unroll
                        for (int j = 0; j < N; ++j) acc[k][j] = mac<STRICT>(acc[k][j], f[j], wgt);
                    }
                    p += a.src_pitch;
                    wt += GR;
                }
#pragma unroll
                for (int k = 0; k < GR; ++k)
                    if (g * GR + k < th) {
#pragma unroll
                        for (int j = 0; j < N; ++j) t[k * a.tmp_pitch + j] = acc[k][j];
                    }
            }
            x += dx; g += dy;
            if (x >= units) { x -= units; ++g; }
        }
    }
    __syncthreads();

    // ---- horizontal pass: item = (row, column group); a warp's lanes are rows of one group (weight
    // reads are broadcasts; the odd row pitches keep data reads and staged writes conflict-free).
    // The staged source region is free now and receives the tile. ----
    constexpr bool DIRECT = sizeof(T) * C == 16;
    T *s_out = reinterpret_cast<T *>(s_src);
    const int out_pitch = a.out_pitch;
    const size_t d_pitch = (size_t)a.nw * C;
    T *gdst = dst + (size_t)(oy0 - (int)a.dst_row0) * d_pitch + (size_t)ox0 * C;
    const bool vec_dst = (sizeof(T) * d_pitch) % 16 == 0 && (reinterpret_cast<uintptr_t>(gdst) & 15) == 0;
    for (int e0 = 0; e0 < (a.TW / GR) * a.TH; e0 += GK_THREADS) {   // same trip count for every thread (barriers inside)
        const int e = e0 + tid;
        const int row = e & (a.TH - 1), g = e >> a.log2TH;
        const bool valid = row < th && g < hgroups;
        if (!DIRECT && !valid) continue;
        const int2 bu = valid ? s_hbu[g] : make_int2(0, 0);
        const int U = bu.y;
        const float *t = s_tmp + row * a.tmp_pitch + (bu.x * C - sub0);
        const float2 *wt = s_hwt + g * a.UH * GR;
        float r[GR][C];
        if constexpr (PACKED && C == 4) {
            f32x2 a01[GR], a23[GR];
#pragma unroll
            for (int k = 0; k < GR; ++k) { a01[k] = 0ull; a23[k] = 0ull; }
            for (int u = 0; u < U; ++u) {
                const ulonglong2 q = *reinterpret_cast<const ulonglong2 *>(t);
#pragma unroll
                for (int k = 0; k < GR; k += 2) {
                    const ulonglong2 ww = *reinterpret_cast<const ulonglong2 *>(wt + k);
                    a01[k] = fma2(q.x, ww.x, a01[k]);
                    a23[k] = fma2(q.y, ww.x, a23[k]);
                    a01[k + 1] = fma2(q.x, ww.y, a01[k + 1]);
                    a23[k + 1] = fma2(q.y, ww.y, a23[k + 1]);
                }
                t += 4;
                wt += GR;
            }
#pragma unroll
            for (int k = 0; k < GR; ++k) {
                unpack2(a01[k], r[k][0], r[k][1 % C]);
                unpack2(a23[k], r[k][2 % C], r[k][3 % C]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < GR; ++k)
#pragma unroll
                for (int c = 0; c < C; ++c) r[k][c] = 0.0f;
            for (int u = 0; u < U; ++u) {
                float px[C];
#pragma unroll
                for (int c = 0; c < C; ++c) px[c] = t[c];
#pragma unroll
                for (int k = 0; k < GR; ++k) {
                    const float wgt = wt[k].x;
#pragma unroll
                    for (int c = 0; c < C; ++c) r[k][c] = mac<STRICT>(r[k][c], px[c], wgt);
                }
                t += C;
                wt += GR;
            }
        }
        constexpr int GBYTES = GR * C * (int)sizeof(T);
        if (DIRECT) {
            // f32 x 4 channels (16-byte pixels): the GK_THREADS items of this round are a block of
            // BG = GK_THREADS / TH neighbouring column groups x TH rows.  It is transposed through the
            // (now free) staged-source region so that a warp stores contiguous row segments: a
            // lane-per-row store would cost one memory transaction per pixel.
            const int BG = GK_THREADS >> a.log2TH;             // column groups in the block
            const int seg = BG * GR;                            // pixels per row of the block
            const int blk_pitch = seg * 16 + 16;                // bytes; odd number of 16-byte units
            const int g_first = e0 >> a.log2TH;
            if (valid) {
                unsigned char *o = s_src + row * blk_pitch + (g - g_first) * (GR * 16);
#pragma unroll
                for (int k = 0; k < GR; ++k) {
                    T px[C];
#pragma unroll
                    for (int c = 0; c < C; ++c) px[c] = Px<T>::clamp_nearest(r[k][c]);
                    *reinterpret_cast<uint4 *>(o + k * 16) = *reinterpret_cast<const uint4 *>(px);
                }
            }
            __syncthreads();
            // the block is TH x seg pixels = 1024 16-byte vectors: four per thread, same column, rows
            // GK_THREADS / seg apart
            const int lseg = 2 + 8 - a.log2TH;                  // log2(seg): GR = 4, GK_THREADS = 256
            const int cc = tid & (seg - 1), rstep = GK_THREADS >> lseg;
            const int px = g_first * GR + cc;                   // tile column of this thread's vectors
            if (px < tw) {
                int rr = tid >> lseg;
                const unsigned char *sp = s_src + rr * blk_pitch + cc * 16;
                T *o = gdst + (size_t)rr * d_pitch + (size_t)px * C;
                const size_t ostep = (size_t)rstep * d_pitch;
                if (vec_dst) {
#pragma unroll
                    for (int k = 0; k < 4; ++k, rr += rstep, sp += rstep * blk_pitch, o += ostep)
                        if (rr < th) stg_stream(reinterpret_cast<uint4 *>(o), *reinterpret_cast<const uint4 *>(sp));
                } else {
                    for (int k = 0; k < 4; ++k, rr += rstep, sp += rstep * blk_pitch, o += ostep)
                        if (rr < th) {
                            const uint4 q = *reinterpret_cast<const uint4 *>(sp);
                            o[0] = __uint_as_float(q.x); o[1 % C] = __uint_as_float(q.y); o[2 % C] = __uint_as_float(q.z); o[3 % C] = __uint_as_float(q.w);
                        }
                }
            }
            __syncthreads();
            continue;
        }
        T *o = s_out + row * out_pitch + g * GR * C;
        if (GBYTES % 16 == 0 && g * GR + GR <= tw) {
            union { T e[GR * C]; uint4 v[GBYTES % 16 == 0 ? GBYTES / 16 : 1]; } pk;
#pragma unroll
            for (int k = 0; k < GR; ++k)
#pragma unroll
                for (int c = 0; c < C; ++c) pk.e[k * C + c] = Px<T>::clamp_nearest(r[k][c]);
#pragma unroll
            for (int v = 0; v < GBYTES / 16; ++v) reinterpret_cast<uint4 *>(o)[v] = pk.v[v];
        } else {
#pragma unroll
            for (int k = 0; k < GR; ++k)
                if (g * GR + k < tw) {
#pragma unroll
                    for (int c = 0; c < C; ++c) o[k * C + c] = Px<T>::clamp_nearest(r[k][c]);
                }
        }
    }
    if (DIRECT) return;
    __syncthreads();

    // ---- store the tile, row-wise coalesced ----
    T *g0 = gdst;
    const int sub = tw * C;
    const bool vec_ok = (sizeof(T) * d_pitch) % 16 == 0 && (reinterpret_cast<uintptr_t>(g0) & 15) == 0 && (sizeof(T) * sub) % 16 == 0;
    const int lane = tid & 31;
    if (vec_ok) {
        const int v_per_row = (int)(sizeof(T) * sub / 16);
        for (int r = tid >> 5; r < th; r += GK_THREADS / 32)
            for (int v = lane; v < v_per_row; v += 32)
                stg_stream(reinterpret_cast<uint4 *>(g0 + (size_t)r * d_pitch) + v, reinterpret_cast<const uint4 *>(s_out + r * out_pitch)[v]);
    } else {
        for (int r = tid >> 5; r < th; r += GK_THREADS / 32)
            for (int v = lane; v < sub; v += 32) g0[(size_t)r * d_pitch + v] = s_out[r * out_pitch + v];
    }
}

int round_up_g(int v, int m) { return (v + m - 1) / m * m; }

// Largest staged extent over all tiles (tile = 1: over all groups) of `step` destination samples,
// where tiles start at multiples of `tile` from `first`.
int span_max(const AxisTable &t, uint32_t first, uint32_t count, int tile, int step) {
    int m = 1;
    for (uint32_t o0 = first; o0 < first + count; o0 += tile) {
        const uint32_t o1 = std::min<uint32_t>(o0 + tile, first + count);
        for (uint32_t g = o0; g < o1; g += step) {
            const uint32_t e = std::min<uint32_t>(g + step, o1) - 1;
            m = std::max(m, t.left[e] + t.count[e] - t.left[g]);
        }
    }
    return m;
}

struct GroupPlan { int TW, TH, max_rows, src_pitch, src_bytes, tmp_pitch, out_pitch, UV, UH; size_t smem; };

bool plan_group(const AxisTable &vt, const AxisTable &ht, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, int C, int tsize,
                int UV, int UH, size_t budget, GroupPlan *best) {
    static const int tws[] = {16, 32, 64, 128};
    static const int ths[] = {16, 32, 64};
    double best_cost = 1e300;
    bool found = false;
    for (int TW : tws)
        for (int TH : ths) {
            if ((TW / GR) * TH < GK_THREADS) continue;  // one horizontal item per thread at least
            GroupPlan p;
            p.TW = TW; p.TH = TH;
            const int max_cols = span_max(ht, 0, nw, TW, TW);
            p.max_rows = span_max(vt, dst_row0, dst_rows, TH, TH);
            p.UV = UV; p.UH = UH;
            int out_bytes_row = round_up_g(TW * C * tsize, 16);
            if ((out_bytes_row / 16) % 2 == 0) out_bytes_row += 16;  // odd number of 16-byte units: lanes = rows write conflict-free
            p.out_pitch = out_bytes_row / tsize;
            p.src_pitch = round_up_g(max_cols * C * tsize + 32, 16);  // +32: the span may be widened to 16-byte boundaries
            int tp = round_up_g(p.src_pitch / tsize, 4);
            if ((tp / 4) % 2 == 0) tp += 4;                            // same for the intermediate rows
            p.tmp_pitch = tp;
            const bool direct = C * tsize == 16;
            // the region also holds the staged output: the whole tile, or (16-byte pixels) one block of a round
            const int blk_bytes = TH * ((GK_THREADS / TH) * GR * 16 + 16);
            p.src_bytes = round_up_g(std::max(p.max_rows * p.src_pitch, direct ? blk_bytes : TH * out_bytes_row), 128);
            p.smem = (size_t)p.src_bytes + (size_t)TH * tp * 4 + (size_t)TH * p.UV * 8 + (size_t)TW * p.UH * 8 +
                     (size_t)(TH / GR + TW / GR) * 8 + 128;  // the last term: bu arrays
            if (p.smem > budget) continue;
            // cost per destination pixel: staged source bytes and vertical work (the intermediate is
            // max_cols wide), with a penalty when fewer than three CTAs fit an SM
            const double staged = (double)p.max_rows * max_cols / ((double)TW * TH);
            const double vwork = (double)max_cols / TW * p.UV / GR;
            const double hwork = (double)p.UH / GR;
            const double occ_pen = p.smem > 110 * 1024 ? 1.5 : (p.smem > 74 * 1024 ? 1.15 : 1.0);
            const double setup = 1200.0 / ((double)TW * TH);  // per-CTA table building and barriers, spread over the tile
            const double cost = (staged + vwork + hwork + setup) * occ_pen;
            if (cost < best_cost) { best_cost = cost; *best = p; found = true; }
        }
    return found;
}

// ---- group tables, cached per (device, axis table identity) -----------------------------------
// Built on the host from the axis table the caller already holds (same weights, regrouped):
// wt[g][u][k] = weight of source sample first(g) + u for destination sample g*GR + k, as a (w, w)
// pair; bu[g] = (first(g), span(g)).
struct GroupTable {
    int device = -1, filter = -1, axis = -1, U = 0;
    uint32_t src_len = 0, dst_len = 0;
    float2 *d_wt = nullptr;
    int2 *d_bu = nullptr;
    GroupTable() = default;
    GroupTable(const GroupTable &) = delete;
    GroupTable &operator=(const GroupTable &) = delete;
    ~GroupTable() {
        if (!d_wt && !d_bu) return;
        int cur = -1;
        cudaGetDevice(&cur);
        if (cur != device) cudaSetDevice(device);
        cudaFree(d_wt); cudaFree(d_bu);  // waits for queued kernels that still read them
        if (cur != device && cur >= 0) cudaSetDevice(cur);
    }
};
using GroupPtr = std::shared_ptr<const GroupTable>;  // a launch holds its tables until it is enqueued
std::mutex g_gt_mu;
std::list<std::shared_ptr<GroupTable>> g_gts;  // most recently used first; LRU per device
constexpr size_t MAX_GROUP_TABLES = 32;        // per device

int get_group_table(int device, int filter, int axis, uint32_t src_len, const AxisTable &t, uint32_t dst_len, int U,
                    GroupPtr *out) {
    {
        std::lock_guard<std::mutex> lk(g_gt_mu);
        for (auto it = g_gts.begin(); it != g_gts.end(); ++it) {
            const GroupTable &g = **it;
            if (g.device == device && g.filter == filter && g.axis == axis && g.src_len == src_len && g.dst_len == dst_len && g.U == U) {
                g_gts.splice(g_gts.begin(), g_gts, it);
                *out = g_gts.front();
                return B200IMG_OK;
            }
        }
    }
    const uint32_t groups = (dst_len + GR - 1) / GR;
    std::vector<float2> wt((size_t)groups * U * GR, make_float2(0.0f, 0.0f));
    std::vector<int2> bu(groups);
    for (uint32_t g = 0; g < groups; ++g) {
        const uint32_t first = g * GR, last = std::min<uint32_t>(first + GR, dst_len) - 1;
        bu[g] = make_int2(t.left[first], t.left[last] + t.count[last] - t.left[first]);
        for (uint32_t o = first; o <= last; ++o)
            for (int k = 0; k < t.count[o]; ++k) {
                const float w = t.w[(size_t)o * t.stride + k];
                wt[((size_t)g * U + (t.left[o] + k - t.left[first])) * GR + (o - first)] = make_float2(w, w);
            }
    }
    auto gt = std::make_shared<GroupTable>();
    gt->device = device; gt->filter = filter; gt->axis = axis; gt->src_len = src_len; gt->dst_len = dst_len; gt->U = U;
    if (cudaMalloc(&gt->d_wt, wt.size() * sizeof(float2)) != cudaSuccess || cudaMalloc(&gt->d_bu, bu.size() * sizeof(int2)) != cudaSuccess) {
        cudaGetLastError();
        return fail(B200IMG_ERR_OOM, "resize: cannot allocate group tables");
    }
    // synchronous copies: complete before any stream can use them
    B200_CUDA(cudaMemcpy(gt->d_wt, wt.data(), wt.size() * sizeof(float2), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(gt->d_bu, bu.data(), bu.size() * sizeof(int2), cudaMemcpyHostToDevice));
    std::shared_ptr<GroupTable> evicted;
    {
        std::lock_guard<std::mutex> lk(g_gt_mu);
        g_gts.push_front(gt);
        size_t mine = 0;
        for (auto it = g_gts.begin(); it != g_gts.end(); ++it)
            if ((*it)->device == device && ++mine > MAX_GROUP_TABLES) { evicted = *it; g_gts.erase(it); break; }
    }
    evicted.reset();
    *out = gt;
    return B200IMG_OK;
}

template <typename T, int C>
int launch_gc(const GroupArgs &a, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    auto k = strict ? resize_group_kernel<T, C, true> : resize_group_kernel<T, C, false>;
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, GK_THREADS, smem, st>>>(a);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename T>
int launch_gt(const GroupArgs &a, int C, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    switch (C) {
        case 1: return launch_gc<T, 1>(a, smem, grid, strict, st);
        case 2: return launch_gc<T, 2>(a, smem, grid, strict, st);
        case 3: return launch_gc<T, 3>(a, smem, grid, strict, st);
        default: return launch_gc<T, 4>(a, smem, grid, strict, st);
    }
}

}  // namespace

// Returns 1 if this kernel took the job (and launched it), 0 if it does not apply (the caller
// falls back to resize_kernel), < 0 on error.
void resize_group_release(int device) {
    std::list<std::shared_ptr<GroupTable>> dropped;
    {
        std::lock_guard<std::mutex> lk(g_gt_mu);
        for (auto it = g_gts.begin(); it != g_gts.end();) {
            if ((*it)->device == device) { dropped.push_back(*it); it = g_gts.erase(it); }
            else ++it;
        }
    }
}

int resize_group_try(int filter, uint32_t h, const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, int C, int dtype, void *dst,
                     size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t n_images,
                     const AxisTable &vt, const int32_t *d_vleft, const int32_t *d_vcount, const float *d_vw,
                     const AxisTable &ht, const int32_t *d_hleft, const int32_t *d_hcount, const float *d_hw,
                     float acc_init, bool strict, const DeviceCtx *ctx, cudaStream_t st) {
    if (dtype == B200IMG_F32 && strict) return 0;   // exact treatment of non-finite samples: resize_kernel
    if (vt.max_count < 2 && ht.max_count < 2) return 0;  // Nearest: one tap, nothing to share
    const int tsize = (int)dtype_size(dtype);
    // groups are aligned to multiples of GR from sample 0 (the tables are global): a band must start
    // on a group boundary and end on one or at the image's last row
    const uint32_t nh = (uint32_t)vt.left.size();
    if (dst_row0 % GR != 0 || ((dst_row0 + dst_rows) % GR != 0 && dst_row0 + dst_rows != nh)) return 0;
    const int UV = span_max(vt, 0, nh, nh, GR), UH = span_max(ht, 0, nw, nw, GR);
    // the union walk multiplies by zero outside a sample's own window: worth it only while the
    // unions stay close to the windows (upscales, moderate downscales of wide filters)
    if (UV * 10 > vt.max_count * 13 + 10 || UH * 10 > ht.max_count * 13 + 10) return 0;  // union <= 1.3 x window + 1
    GroupPlan p;
    const size_t budget = std::min<size_t>(ctx->smem_optin, 150 * 1024);
    if (!plan_group(vt, ht, dst_row0, dst_rows, nw, C, tsize, UV, UH, budget, &p)) return 0;
    GroupPtr gv, gh;  // held until the launch below has been enqueued
    int rc = get_group_table(ctx->device, filter, 0, h, vt, nh, UV, &gv);
    if (rc) return rc;
    const float2 *vwt = gv->d_wt; const int2 *vbu = gv->d_bu;
    if ((rc = get_group_table(ctx->device, filter, 1, w, ht, nw, UH, &gh))) return rc;
    const int tiles_x = (int)((nw + p.TW - 1) / p.TW), tiles_y = (int)((dst_rows + p.TH - 1) / p.TH);
    if (tiles_y > 65535 || n_images > 65535) return 0;

    GroupArgs a;
    a.src = src; a.dst = dst; a.src_img_stride = src_img_stride; a.dst_img_stride = dst_img_stride;
    a.w = w; a.nw = nw; a.src_row0 = src_row0; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.vleft = d_vleft; a.vcount = d_vcount; a.vw = d_vw; a.vstride = vt.stride;
    a.hleft = d_hleft; a.hcount = d_hcount; a.hw = d_hw; a.hstride = ht.stride;
    a.acc_init = acc_init;
    a.TW = p.TW; a.TH = p.TH;
    a.log2TH = 0;
    while ((1 << a.log2TH) < p.TH) ++a.log2TH;
    a.max_rows = p.max_rows; a.src_pitch = p.src_pitch; a.src_bytes = p.src_bytes; a.tmp_pitch = p.tmp_pitch;
    a.out_pitch = p.out_pitch; a.UV = p.UV; a.UH = p.UH;
    a.vwt = vwt; a.vbu = vbu; a.hwt = gh->d_wt; a.hbu = gh->d_bu;
    const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y, n_images);
    if (dtype == B200IMG_U8) rc = launch_gt<uint8_t>(a, C, p.smem, grid, strict, st);
    else if (dtype == B200IMG_U16) rc = launch_gt<uint16_t>(a, C, p.smem, grid, strict, st);
    else rc = launch_gt<float>(a, C, p.smem, grid, strict, st);
    return rc ? rc : 1;
}

}  // namespace b200
