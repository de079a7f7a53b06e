// resample_stream.cu -- the resampler for DOWNSCALES of integer-subpixel images (u8 / u16),
// where the tile kernel in resample.cu spends most of its issue slots re-converting the same
// source bytes for every output row whose window covers them.
//
// Here a thread OWNS one 32-bit word column of a source strip (4 u8 or 2 u16 subpixels) and
// streams down the source rows of a band exactly once: each word is loaded (coalesced, straight
// from global memory -- no staging), converted once, and multiplied into the accumulators of all
// output rows whose vertical window is open (vertical_sample, src/imageops/sample.rs:494-562:
// accumulator starts at acc_init, taps ascending -- the streaming order IS the reference order).
// Source rows are taken in chunks of 4.  Within a chunk the destination rows whose windows are
// still open occupy accumulator slots 0, 1, .. in order (slot s = row base(chunk) + s, base = rows
// already finished); the per-source-row weights of the A slots (zero where a slot has no open
// window) are laid out densely in shared memory so the inner loop has no branches.  After a chunk
// the rows whose windows closed are flushed from slot 0 and the accumulators shift down, so
// slot numbers never need run-time register indexing.  Flushed rows go to an 8-row ring of the
// f32 intermediate in shared memory; every 8 finished rows the CTA runs
// the horizontal pass (horizontal_sample, sample.rs:217-340) on them and stores 8 destination rows.
// The intermediate never leaves the SM; DRAM sees the source once and the destination once.
#include "common.cuh"
#include "host_tables.h"
#include "packed.cuh"

namespace b200 {
namespace {

constexpr int ST_THREADS = 256;  // = word columns per strip
constexpr int ST_TH = 8;         // destination rows per horizontal phase
constexpr int ST_PF = 4;         // source rows prefetched per thread

}  // namespace

struct StreamArgs {
    const void *src;
    void *dst;
    size_t src_img_stride, dst_img_stride;
    uint32_t w, nw;
    uint32_t src_row0, dst_row0, dst_rows;
    const int32_t *vleft, *vcount; const float *vw; int vstride;
    const int32_t *hleft, *hcount; const float *hw; int hstride;
    float acc_init;
    int TW, BAND;        // destination columns per strip, destination rows per band (multiple of 8)
    int max_band_rows;   // source rows a band taps, at most
    int tmp_pitch;       // floats per intermediate row
};

namespace {

template <typename T, int C, int A, bool STRICT>
__global__ void __launch_bounds__(ST_THREADS)
resize_stream_kernel(const StreamArgs a) {
    constexpr int N = Word<T>::N;
    constexpr bool PACKED = !STRICT;  // FFMA2 path
    extern __shared__ __align__(16) unsigned char smem[];
    float2 *s_ws = reinterpret_cast<float2 *>(smem);                           // [max_band_rows + ST_PF][A] (w, w)
    float *s_tmp = reinterpret_cast<float *>(s_ws + (size_t)(a.max_band_rows + ST_PF) * A);  // [ST_TH][tmp_pitch]
    float2 *s_hw = reinterpret_cast<float2 *>(s_tmp + (size_t)ST_TH * a.tmp_pitch);  // [hstride][TW] (w, w)
    int *s_hl = reinterpret_cast<int *>(s_hw + (size_t)a.TW * a.hstride);        // [TW]
    int *s_end = s_hl + a.TW;                                                    // [BAND] window end rows
    int *s_base = s_end + a.BAND;                                                // [chunks + 1] rows finished before a chunk
    T *s_out = reinterpret_cast<T *>(s_base + (a.max_band_rows / ST_PF + 3));    // [ST_TH][TW * C]

    const int tid = threadIdx.x;
    const unsigned img = blockIdx.z;
    const int ox0 = blockIdx.x * a.TW;
    const int tw = min(a.TW, (int)a.nw - ox0);
    const int oyA = a.dst_row0 + blockIdx.y * a.BAND;
    const int nout = min(a.BAND, (int)(a.dst_row0 + a.dst_rows) - oyA);
    const int rowA = a.vleft[oyA];
    const int rowB = a.vleft[oyA + nout - 1] + a.vcount[oyA + nout - 1];
    const int col0 = a.hleft[ox0];
    const int col1 = a.hleft[ox0 + tw - 1] + a.hcount[ox0 + tw - 1];

    const T *src = reinterpret_cast<const T *>(reinterpret_cast<const unsigned char *>(a.src) + (size_t)img * a.src_img_stride);
    T *dst = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(a.dst) + (size_t)img * a.dst_img_stride);
    const size_t gp = sizeof(T) * (size_t)a.w * C;  // row pitch in bytes; a multiple of 4 (checked by the host)
    const int b0 = (col0 * C * (int)sizeof(T)) & ~3;
    const int b1 = (col1 * C * (int)sizeof(T) + 3) & ~3;
    const int words = (b1 - b0) / 4;  // <= ST_THREADS (host's choice of TW)
    const int sub0 = b0 / (int)sizeof(T);

    // ---- set-up: horizontal weights of the strip, slot weights and window ends of the band ----
    for (int i = tid; i < tw * a.hstride; i += ST_THREADS) {
        const int oxl = i / a.hstride, k = i - oxl * a.hstride;
        const float w = a.hw[(size_t)(ox0 + oxl) * a.hstride + k];
        s_hw[k * a.TW + oxl] = make_float2(w, w);
    }
    for (int i = tid; i < tw; i += ST_THREADS) s_hl[i] = a.hleft[ox0 + i] * C - sub0;
    for (int i = tid; i < nout; i += ST_THREADS) s_end[i] = a.vleft[oyA + i] + a.vcount[oyA + i];
    for (int i = tid; i < (rowB - rowA + ST_PF) * A; i += ST_THREADS) s_ws[i] = make_float2(0.0f, 0.0f);
    const int chunks = (rowB - rowA + ST_PF - 1) / ST_PF;
    for (int c = tid; c <= chunks; c += ST_THREADS) {  // base(c) = rows of the band whose window ends at or before the chunk starts
        const int start = rowA + c * ST_PF;
        int lo_ = 0, hi_ = nout;  // window ends are non-decreasing: binary search
        while (lo_ < hi_) {
            const int mid = (lo_ + hi_) >> 1;
            if (a.vleft[oyA + mid] + a.vcount[oyA + mid] <= start) lo_ = mid + 1; else hi_ = mid;
        }
        s_base[c] = lo_;
    }
    __syncthreads();
    for (int i = tid; i < nout * a.vstride; i += ST_THREADS) {
        const int o = i / a.vstride, k = i - o * a.vstride;
        if (k < a.vcount[oyA + o]) {
            const float w = a.vw[(size_t)(oyA + o) * a.vstride + k];
            const int jr = a.vleft[oyA + o] + k - rowA;
            s_ws[(size_t)jr * A + (o - s_base[jr / ST_PF])] = make_float2(w, w);
        }
    }
    __syncthreads();

    // ---- accumulators: within a chunk, slot s holds destination row base(chunk) + s ----
    f32x2 lo[A], hi[A];   // PACKED: (sub 0,1) and (sub 2,3)
    float acc[A][N];      // scalar path
    const f32x2 init2 = pack2(a.acc_init, a.acc_init);
#pragma unroll
    for (int s = 0; s < A; ++s) {
        lo[s] = init2; hi[s] = init2;
#pragma unroll
        for (int j = 0; j < N; ++j) acc[s][j] = a.acc_init;
    }
    const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
    const bool active = tid < words;
    const unsigned char *gcol = reinterpret_cast<const unsigned char *>(src) + (size_t)(rowA - (int)a.src_row0) * gp + b0 + 4 * (active ? tid : 0);

    auto horizontal = [&](int g0, int rows_g) {
        // destination rows oyA + g0 .. + rows_g; their intermediates are ring rows 0 .. rows_g
        __syncthreads();
        for (int e = tid; e < a.TW * ST_TH; e += ST_THREADS) {
            const int oyl = e & 7, oxl = e >> 3;
            if (oyl >= rows_g || oxl >= tw) continue;
            const int n = a.hcount[ox0 + oxl];
            const float *t = s_tmp + (size_t)oyl * a.tmp_pitch + s_hl[oxl];
            const float2 *wh = s_hw + oxl;
            float r[C];
            if constexpr (PACKED && C == 4) {
                f32x2 a01 = 0ull, a23 = 0ull;
                for (int i = 0; i < n; ++i) {
                    const ulonglong2 q = *reinterpret_cast<const ulonglong2 *>(t);
                    const f32x2 ww = *reinterpret_cast<const f32x2 *>(wh + i * a.TW);
                    a01 = fma2(q.x, ww, a01);
                    a23 = fma2(q.y, ww, a23);
                    t += 4;
                }
                unpack2(a01, r[0], r[1 % C]);
                unpack2(a23, r[2 % C], r[3 % C]);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c) r[c] = 0.0f;
                for (int i = 0; i < n; ++i) {
                    const float wgt = wh[i * a.TW].x;
#pragma unroll
                    for (int c = 0; c < C; ++c) r[c] = mac<STRICT>(r[c], t[c], wgt);
                    t += C;
                }
            }
            T *o = s_out + (size_t)oyl * (a.TW * C) + oxl * C;
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = Px<T>::clamp_nearest(r[c]);
        }
        __syncthreads();
        const size_t d_pitch = (size_t)a.nw * C;
        const int sub = tw * C;
        const int lane = tid & 31;
        for (int rr = tid >> 5; rr < rows_g; rr += ST_THREADS / 32) {
            T *g = dst + (size_t)(oyA + g0 + rr - (int)a.dst_row0) * d_pitch + (size_t)ox0 * C;
            const T *sr = s_out + (size_t)rr * (a.TW * C);
            if ((sizeof(T) * sub) % 4 == 0 && ((reinterpret_cast<uintptr_t>(g) | (sizeof(T) * a.TW * C)) & 3) == 0) {
                for (int v = lane; v < (int)(sizeof(T) * sub / 4); v += 32)
                    reinterpret_cast<uint32_t *>(g)[v] = reinterpret_cast<const uint32_t *>(sr)[v];
            } else {
                for (int v = lane; v < sub; v += 32) g[v] = sr[v];
            }
        }
    };

    int next_out = 0;

    auto flush_one = [&]() {
        // the oldest open window (slot 0, destination row oyA + next_out) has closed
        float *t = s_tmp + (size_t)(next_out & (ST_TH - 1)) * a.tmp_pitch + tid * N;
        if (active) {
            if constexpr (PACKED && N == 4) *reinterpret_cast<ulonglong2 *>(t) = make_ulonglong2(lo[0], hi[0]);
            else if constexpr (PACKED && N == 2) *reinterpret_cast<f32x2 *>(t) = lo[0];
            else {
#pragma unroll
                for (int j = 0; j < N; ++j) t[j] = acc[0][j];
            }
        }
#pragma unroll
        for (int s = 0; s + 1 < A; ++s) {
            lo[s] = lo[s + 1]; hi[s] = hi[s + 1];
#pragma unroll
            for (int j = 0; j < N; ++j) acc[s][j] = acc[s + 1][j];
        }
        lo[A - 1] = init2; hi[A - 1] = init2;
#pragma unroll
        for (int j = 0; j < N; ++j) acc[A - 1][j] = a.acc_init;
        ++next_out;
        if ((next_out & (ST_TH - 1)) == 0 || next_out == nout) {
            const int rows_g = (next_out & (ST_TH - 1)) ? (next_out & (ST_TH - 1)) : ST_TH;
            horizontal(next_out - rows_g, rows_g);
        }
    };

    auto accumulate = [&](uint32_t wv, const float2 *ws) {
        if constexpr (PACKED && N == 4) {
            float f[4];
            Word<T>::biased(wv, f);
            const f32x2 f01 = add2(pack2(f[0], f[1]), bias2), f23 = add2(pack2(f[2], f[3]), bias2);
#pragma unroll
            for (int s = 0; s < A; ++s) {
                const f32x2 ww = *reinterpret_cast<const f32x2 *>(ws + s);
                lo[s] = fma2(f01, ww, lo[s]);
                hi[s] = fma2(f23, ww, hi[s]);
            }
        } else if constexpr (PACKED && N == 2) {
            float f[2];
            Word<T>::biased(wv, f);
            const f32x2 f01 = add2(pack2(f[0], f[1]), bias2);
#pragma unroll
            for (int s = 0; s < A; ++s) lo[s] = fma2(f01, *reinterpret_cast<const f32x2 *>(ws + s), lo[s]);
        } else {
            float f[N];
            Word<T>::biased(wv, f);
#pragma unroll
            for (int j2 = 0; j2 < N; ++j2) f[j2] = unbias<T>(f[j2]);
#pragma unroll
            for (int s = 0; s < A; ++s) {
                const float wgt = ws[s].x;
#pragma unroll
                for (int j2 = 0; j2 < N; ++j2) acc[s][j2] = mac<STRICT>(acc[s][j2], f[j2], wgt);
            }
        }
    };

    // ---- stream down the band's source rows in chunks of ST_PF; the next chunk's loads are in
    // flight while this one is accumulated.  Windows that closed inside a chunk are flushed after
    // it (the rows in between carry zero weights for their slots). ----
    const unsigned char *p = gcol;
    uint32_t cur[ST_PF], nxt[ST_PF];
#pragma unroll
    for (int k = 0; k < ST_PF; ++k) {
        cur[k] = 0;
        if (active && rowA + k < rowB) cur[k] = __ldg(reinterpret_cast<const uint32_t *>(p));
        p += gp;
    }
    for (int j = rowA; j < rowB; j += ST_PF) {
#pragma unroll
        for (int k = 0; k < ST_PF; ++k) {
            nxt[k] = 0;
            if (active && j + ST_PF + k < rowB) nxt[k] = __ldg(reinterpret_cast<const uint32_t *>(p));
            p += gp;
        }
        const float2 *ws = s_ws + (size_t)(j - rowA) * A;  // rows past rowB have zeroed weights (see set-up)
#pragma unroll
        for (int k = 0; k < ST_PF; ++k) accumulate(cur[k], ws + k * A);
        for (int f = s_base[(j - rowA) / ST_PF + 1] - next_out; f > 0; --f) flush_one();
#pragma unroll
        for (int k = 0; k < ST_PF; ++k) cur[k] = nxt[k];
    }
}

int round_up_i(int v, int m) { return (v + m - 1) / m * m; }

// Accumulator slots the kernel needs: the largest number of destination rows with a window
// overlapping one 4-row chunk, over all bands.  Rounded up to a compiled size; 0 if > 12 (upscales).
int slots_needed(const AxisTable &vt, uint32_t first, uint32_t count, int band) {
    int need = 1;
    for (uint32_t o0 = first; o0 < first + count; o0 += band) {
        const uint32_t o1 = std::min<uint32_t>(o0 + band, first + count);
        const int rowA = vt.left[o0], rowB = vt.left[o1 - 1] + vt.count[o1 - 1];
        uint32_t base = o0, open_end = o0;  // [base, open_end): rows whose window overlaps the chunk
        for (int start = rowA; start < rowB; start += ST_PF) {
            while (base < o1 && vt.left[base] + vt.count[base] <= start) ++base;
            while (open_end < o1 && vt.left[open_end] < start + ST_PF) ++open_end;
            need = std::max(need, (int)(open_end - base));
            if (need > 12) return 0;
        }
    }
    static const int cand[] = {4, 6, 8, 10, 12};
    for (int A : cand)
        if (need <= A) return A;
    return 0;
}

int strip_words(const AxisTable &ht, uint32_t nw, int TW, int C, int tsize) {
    int m = 0;
    for (uint32_t o = 0; o < nw; o += TW) {
        const uint32_t e = std::min<uint32_t>(o + TW, nw) - 1;
        const int b0 = (ht.left[o] * C * tsize) & ~3, b1 = ((ht.left[e] + ht.count[e]) * C * tsize + 3) & ~3;
        m = std::max(m, (b1 - b0) / 4);
    }
    return m;
}

template <typename T, int C, int A>
int launch_a(const StreamArgs &a, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    auto k = strict ? resize_stream_kernel<T, C, A, true> : resize_stream_kernel<T, C, A, false>;
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, ST_THREADS, smem, st>>>(a);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename T, int C>
int launch_c(const StreamArgs &a, int A, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    switch (A) {
        case 4: return launch_a<T, C, 4>(a, smem, grid, strict, st);
        case 6: return launch_a<T, C, 6>(a, smem, grid, strict, st);
        case 8: return launch_a<T, C, 8>(a, smem, grid, strict, st);
        case 10: return launch_a<T, C, 10>(a, smem, grid, strict, st);
        default: return launch_a<T, C, 12>(a, smem, grid, strict, st);
    }
}

template <typename T>
int launch_t(const StreamArgs &a, int C, int A, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    switch (C) {
        case 1: return launch_c<T, 1>(a, A, smem, grid, strict, st);
        case 2: return launch_c<T, 2>(a, A, smem, grid, strict, st);
        case 3: return launch_c<T, 3>(a, A, smem, grid, strict, st);
        default: return launch_c<T, 4>(a, A, smem, grid, strict, st);
    }
}

}  // namespace

// Returns 1 if the streaming kernel took the job (and launched it), 0 if it does not apply
// (caller falls back to the tile kernel), < 0 on error.
int resize_stream_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, int C, int dtype, void *dst,
                      size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t n_images,
                      const AxisTable &vt, const int32_t *d_vleft, const int32_t *d_vcount, const float *d_vw,
                      const AxisTable &ht, const int32_t *d_hleft, const int32_t *d_hcount, const float *d_hw,
                      float acc_init, bool strict, const DeviceCtx *ctx, cudaStream_t st) {
    if (dtype == B200IMG_F32) return 0;  // closed-window slots see 0 * x: not neutral for non-finite f32 pixels
    const int tsize = (int)dtype_size(dtype);
    const size_t gp = (size_t)w * C * tsize;
    if (gp % 4 != 0 || (reinterpret_cast<uintptr_t>(src) & 3) != 0 || src_img_stride % 4 != 0) return 0;
    if (vt.max_count < 2) return 0;  // Nearest: nothing to share between rows; the tile kernel is fine
    // vertical upscales belong to the tile kernels: many destination rows per source row would fill the
    // accumulator slots with windows that overlap almost entirely (measured: Triangle 1080 -> 2000 rows
    // took 83 us here against 49 us in the tile kernel)
    {
        const size_t nh_all = vt.left.size();
        if (nh_all && (size_t)(vt.left[nh_all - 1] + vt.count[nh_all - 1]) < nh_all) return 0;
    }
    // a quick reject for upscales before the exact count below
    if ((uint64_t)vt.left.size() > 0 && dst_rows > 12 && vt.left[dst_row0 + 12] - vt.left[dst_row0] < 1) return 0;

    int TW = 0;
    for (int cand = 128; cand >= 8; cand -= 8)
        if (strip_words(ht, nw, cand, C, tsize) <= ST_THREADS) { TW = cand; break; }
    if (TW == 0) return 0;
    const int strips = (int)((nw + TW - 1) / TW);

    // rows per band: enough CTAs to fill the GPU a few times over, bands not so short that the
    // window overlap between bands (re-read source rows) matters
    int BAND = 128;
    while (BAND > 16 && (uint64_t)strips * ((dst_rows + BAND - 1) / BAND) * n_images < (uint64_t)ctx->sm_count * 6) BAND /= 2;
    const int bands = (int)((dst_rows + BAND - 1) / BAND);
    if (bands > 65535 || n_images > 65535) return 0;
    // a single small image does not give this kernel one CTA per SM even with 16-row bands: the tile
    // kernels spread it better (1080p -> 400x300: 23-26 us here, 15-23 us there)
    if ((uint64_t)strips * bands * n_images < (uint64_t)ctx->sm_count) return 0;
    const int A = slots_needed(vt, dst_row0, dst_rows, BAND);
    if (A == 0) return 0;
    int max_band_rows = 1;
    for (uint32_t o = dst_row0; o < dst_row0 + dst_rows; o += BAND) {
        const uint32_t e = std::min<uint32_t>(o + BAND, dst_row0 + dst_rows) - 1;
        max_band_rows = std::max(max_band_rows, vt.left[e] + vt.count[e] - vt.left[o]);
    }

    StreamArgs a;
    a.src = src; a.dst = dst; a.src_img_stride = src_img_stride; a.dst_img_stride = dst_img_stride;
    a.w = w; a.nw = nw; a.src_row0 = src_row0; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.vleft = d_vleft; a.vcount = d_vcount; a.vw = d_vw; a.vstride = vt.stride;
    a.hleft = d_hleft; a.hcount = d_hcount; a.hw = d_hw; a.hstride = ht.stride;
    a.acc_init = acc_init; a.TW = TW; a.BAND = BAND; a.max_band_rows = max_band_rows;
    int tp = round_up_i(ST_THREADS * (4 / tsize), 4);
    if ((tp / 4) % 2 == 0) tp += 4;  // odd number of 16-byte units: the 8-row lane pattern of the horizontal pass
    a.tmp_pitch = tp;
    const size_t smem = (size_t)(max_band_rows + ST_PF) * A * 8 + (size_t)ST_TH * tp * 4 + (size_t)TW * ht.stride * 8 + (size_t)TW * 4 +
                        (size_t)BAND * 4 + (size_t)(max_band_rows / ST_PF + 3) * 4 + (size_t)ST_TH * TW * C * tsize + 64;
    if (smem > ctx->smem_optin) return 0;
    const dim3 grid((unsigned)strips, (unsigned)bands, n_images);
    int rc;
    if (dtype == B200IMG_U8) rc = launch_t<uint8_t>(a, C, A, smem, grid, strict, st);
    else rc = launch_t<uint16_t>(a, C, A, smem, grid, strict, st);
    return rc ? rc : 1;
}

}  // namespace b200
