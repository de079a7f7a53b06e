// resample.cu -- imageops::resize as ONE fused kernel.
//
// The reference runs vertical_sample (src/imageops/sample.rs:494-562) into a
// full Rgba32F intermediate (16 B per pixel in DRAM) and then horizontal_sample
// (217-340).  Here a CTA owns a TW x TH tile of the destination: it stages the
// source rows/columns that tile taps (with their halos) in shared memory, runs
// the vertical pass into an f32 shared-memory intermediate, runs the horizontal
// pass out of it, and stores the tile.  The intermediate never touches DRAM, so
// the kernel's DRAM traffic is the algorithmic minimum (source once through L2,
// destination once).  The arithmetic order is the reference's: vertical first
// (accumulator starts at acc_init, taps ascending), then horizontal (starts at
// 0, taps ascending), then the store rule of src/primitive_sealed.rs:107-147.
//
// All reductions run on the FMA pipes; the filter is a banded stencil, not a
// dense contraction, so tensor cores are not used.
#include "common.cuh"
#include "host_tables.h"
#include "packed.cuh"

#include <list>
#include <memory>
#include <mutex>
#include <tuple>

namespace b200 {

// resample_stream.cu: column-streaming kernel for integer-subpixel downscales
int resize_stream_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, int C, int dtype, void *dst,
                      size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t n_images,
                      const AxisTable &vt, const int32_t *d_vleft, const int32_t *d_vcount, const float *d_vw,
                      const AxisTable &ht, const int32_t *d_hleft, const int32_t *d_hcount, const float *d_hw,
                      float acc_init, bool strict, const DeviceCtx *ctx, cudaStream_t st);

// resample_int.cu: strided-convolution kernel for u8 downscales by an even integer ratio (FAST arithmetic only)
int resize_int_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                   size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh, uint32_t n_images,
                   const AxisTable &vt, const AxisTable &ht, float acc_init, const DeviceCtx *ctx, cudaStream_t st);

// resample_up2.cu: exact 2x upscales of four-channel f32 images with a support-3 filter (FAST arithmetic only)
int resize_up2_try(const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                   size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh, uint32_t n_images,
                   const AxisTable &vt, const AxisTable &ht, float acc_init, const DeviceCtx *ctx, cudaStream_t st);

// resample_group.cu: tile kernel with four-sample groups on both axes (upscales, f32 images)
void resize_group_release(int device);
int resize_group_try(int filter, uint32_t h, const void *src, size_t src_img_stride, uint32_t src_row0, uint32_t w, int C, int dtype, void *dst,
                     size_t dst_img_stride, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t n_images,
                     const AxisTable &vt, const int32_t *d_vleft, const int32_t *d_vcount, const float *d_vw,
                     const AxisTable &ht, const int32_t *d_hleft, const int32_t *d_hcount, const float *d_hw,
                     float acc_init, bool strict, const DeviceCtx *ctx, cudaStream_t st);

namespace {

constexpr int THREADS = 256;

struct ResizeArgs {
    const void *src;
    void *dst;
    size_t src_img_stride, dst_img_stride;  // bytes between images of a batch
    uint32_t w, C;                          // source width (pixels), channels
    uint32_t nw;                            // destination width
    uint32_t src_row0;                      // first source row present in `src` (band mode)
    uint32_t dst_row0, dst_rows;            // destination rows computed by this launch
    const int32_t *vleft, *vcount; const float *vw; int vstride;
    const int32_t *hleft, *hcount; const float *hw; int hstride;
    float acc_init;
    int TW, TH;            // tile of the destination (powers of two)
    int log2TW;
    int tiles_x, tiles_y;  // per image
    int max_cols, max_rows;     // bound of the staged source region (pixels, rows)
    int src_pitch;              // bytes per staged source row
    int tmp_pitch;              // floats per intermediate row
    int out_pitch;              // subpixels per staged output row
};

// One CTA = one TW x TH tile of the destination of one image (grid: x = tile column, y = tile
// row, z = image).
template <typename T, int C, bool STRICT>
__global__ void __launch_bounds__(THREADS)
resize_kernel(const ResizeArgs a) {
    extern __shared__ __align__(128) unsigned char smem[];
    // layout: [src region | reused as output staging] [tmp f32] [vw pairs] [hw pairs (transposed)] [hleft] [mbarrier]
    const int src_bytes = a.max_rows * a.src_pitch;
    unsigned char *s_src = smem;
    float *s_tmp = reinterpret_cast<float *>(smem + src_bytes);
    float2 *s_vw = reinterpret_cast<float2 *>(s_tmp + (size_t)a.TH * a.tmp_pitch);  // [TH][vstride] (w, w)
    float2 *s_hw = s_vw + a.TH * a.vstride;                                          // [hstride][TW] (w, w)
    int *s_hl = reinterpret_cast<int *>(s_hw + a.TW * a.hstride);   // [TW] first float of a column's window in a tmp row
    int *s_hn = s_hl + a.TW;                                        // [TW] taps of a column
    int *s_vl = s_hn + a.TW;                                        // [TH] first staged row of a row's window
    int *s_vn = s_vl + a.TH;                                        // [TH] taps of a row
    __shared__ __align__(8) uint64_t bar;

    const int tid = threadIdx.x;
    const unsigned img = blockIdx.z;
    const int tx = blockIdx.x, ty = blockIdx.y;
    const int ox0 = tx * a.TW, oy0 = a.dst_row0 + ty * a.TH;
    const int tw = min(a.TW, (int)a.nw - ox0);
    const int th = min(a.TH, (int)(a.dst_row0 + a.dst_rows) - oy0);

    const int col0 = a.hleft[ox0];
    const int col1 = a.hleft[ox0 + tw - 1] + a.hcount[ox0 + tw - 1];
    const int row0 = a.vleft[oy0];
    const int row1 = a.vleft[oy0 + th - 1] + a.vcount[oy0 + th - 1];
    const int rows = row1 - row0;

    const T *src = reinterpret_cast<const T *>(reinterpret_cast<const unsigned char *>(a.src) + (size_t)img * a.src_img_stride);
    T *dst = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(a.dst) + (size_t)img * a.dst_img_stride);

    // ---- the source region in bytes: rows [row0,row1) x bytes [b0,b1) of each row.  When the
    // image rows are 16-byte aligned the span is widened to 16-byte boundaries so that every row
    // is one TMA bulk copy. ----
    const size_t gp = sizeof(T) * (size_t)a.w * C;  // global row pitch, bytes
    const unsigned char *gimg = reinterpret_cast<const unsigned char *>(src) + (size_t)(row0 - (int)a.src_row0) * gp;
    int b0 = col0 * C * (int)sizeof(T), b1 = col1 * C * (int)sizeof(T);
    const bool rows16 = (gp % 16 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    if (rows16) { b0 &= ~15; b1 = (b1 + 15) & ~15; }   // b1 <= gp because gp is a multiple of 16
    const int row_bytes = b1 - b0;
    const int sub0 = b0 / (int)sizeof(T);               // first staged subpixel of each row
    const int row_sub = row_bytes / (int)sizeof(T);      // staged subpixels per row

    if (rows16) {
        if (tid == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
        __syncthreads();
        if (tid == 0) mbar_expect_tx(&bar, (uint32_t)(rows * row_bytes));
        __syncthreads();
        for (int r = tid; r < rows; r += THREADS)
            bulk_g2s(s_src + (size_t)r * a.src_pitch, gimg + (size_t)r * gp + b0, (uint32_t)row_bytes, &bar);
    } else {
        // widest power-of-two unit that divides the row start, the row pitch and the row length
        const unsigned mis = (unsigned)(reinterpret_cast<uintptr_t>(gimg + b0) | gp | (unsigned)row_bytes | 16u);
        const int unit = (int)(mis & (~mis + 1u));
        const int lane = tid & 31;
        if (unit >= 4) {
            const int per_row = row_bytes / unit;
            for (int r = tid >> 5; r < rows; r += THREADS / 32)
                for (int v = lane; v < per_row; v += 32)
                    cp_async(s_src + (size_t)r * a.src_pitch + v * unit, gimg + (size_t)r * gp + b0 + v * unit, unit);
        } else if (unit == 2) {
            for (int r = tid >> 5; r < rows; r += THREADS / 32) {
                const uint16_t *g = reinterpret_cast<const uint16_t *>(gimg + (size_t)r * gp + b0);
                uint16_t *sr = reinterpret_cast<uint16_t *>(s_src + (size_t)r * a.src_pitch);
#pragma unroll 4
                for (int v = lane; v < row_bytes / 2; v += 32) sr[v] = __ldg(g + v);
            }
        } else {
            for (int r = tid >> 5; r < rows; r += THREADS / 32) {
                const unsigned char *g = gimg + (size_t)r * gp + b0;
                unsigned char *sr = s_src + (size_t)r * a.src_pitch;
#pragma unroll 4
                for (int v = lane; v < row_bytes; v += 32) sr[v] = __ldg(g + v);
            }
        }
    }

    // ---- stage weights while the copies fly ----
    for (int i = tid; i < th * a.vstride; i += THREADS) { const float w = a.vw[(size_t)oy0 * a.vstride + i]; s_vw[i] = make_float2(w, w); }
    for (int i = tid; i < tw * a.hstride; i += THREADS) {
        const int oxl = i / a.hstride, k = i - oxl * a.hstride;
        const float w = a.hw[(size_t)(ox0 + oxl) * a.hstride + k];
        s_hw[k * a.TW + oxl] = make_float2(w, w);
    }
    for (int i = tid; i < tw; i += THREADS) {
        s_hl[i] = a.hleft[ox0 + i] * C - sub0;  // subpixel offset into a tmp row
        s_hn[i] = a.hcount[ox0 + i];
    }
    for (int i = tid; i < th; i += THREADS) {
        s_vl[i] = a.vleft[oy0 + i] - row0;
        s_vn[i] = a.vcount[oy0 + i];
    }

    if (rows16) mbar_wait(&bar, 0);
    else cp_async_wait_all();
    __syncthreads();

    // ---- vertical pass: s_tmp[oyl][s] = acc_init + sum_i src[vleft+i][s] * vw[i] ----
    if (sizeof(T) == 4 && !STRICT && row_bytes % 16 == 0) {
        // f32 pixels: four subpixels per thread and tap (LDS.128 + 2 FFMA2)
        // item = (row oyl, quad q), items walked THREADS apart without a division per item
        const int quads = row_bytes / 16;
        const f32x2 init2 = pack2(a.acc_init, a.acc_init);
        const int dy = THREADS / quads, dq = THREADS - dy * quads;
        int oyl = tid / quads, q = tid - oyl * quads;
        while (oyl < th) {
            const int n = s_vn[oyl];
            const float2 *wv = s_vw + oyl * a.vstride;
            const unsigned char *p = s_src + s_vl[oyl] * a.src_pitch + q * 16;
            f32x2 a01 = init2, a23 = init2;
            for (int i = 0; i < n; ++i) {
                const ulonglong2 v = *reinterpret_cast<const ulonglong2 *>(p);
                const f32x2 ww = *reinterpret_cast<const f32x2 *>(wv + i);
                a01 = fma2(v.x, ww, a01);
                a23 = fma2(v.y, ww, a23);
                p += a.src_pitch;
            }
            *reinterpret_cast<ulonglong2 *>(s_tmp + oyl * a.tmp_pitch + q * 4) = make_ulonglong2(a01, a23);
            q += dq; oyl += dy;
            if (q >= quads) { q -= quads; ++oyl; }
        }
    } else {
        constexpr int N = Word<T>::N;
        const int words = (row_sub + N - 1) / N;
        const f32x2 bias2 = pack2(-8388608.0f, -8388608.0f);
        const int dy = THREADS / words, dw = THREADS - dy * words;
        int oyl = tid / words, wd = tid - oyl * words;
        while (oyl < th) {
            const int n = s_vn[oyl];
            const float2 *wv = s_vw + oyl * a.vstride;
            const unsigned char *p = s_src + s_vl[oyl] * a.src_pitch + wd * 4;
            float *t = s_tmp + oyl * a.tmp_pitch + wd * N;
            if constexpr (!STRICT && N == 4) {
                f32x2 a01 = pack2(a.acc_init, a.acc_init), a23 = a01;
                for (int i = 0; i < n; ++i) {
                    float f[4];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
                    const f32x2 ww = *reinterpret_cast<const f32x2 *>(wv + i);
                    a01 = fma2(add2(pack2(f[0], f[1]), bias2), ww, a01);
                    a23 = fma2(add2(pack2(f[2], f[3]), bias2), ww, a23);
                    p += a.src_pitch;
                }
                *reinterpret_cast<ulonglong2 *>(t) = make_ulonglong2(a01, a23);
            } else if constexpr (!STRICT && N == 2) {
                f32x2 a01 = pack2(a.acc_init, a.acc_init);
                for (int i = 0; i < n; ++i) {
                    float f[2];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
                    a01 = fma2(add2(pack2(f[0], f[1]), bias2), *reinterpret_cast<const f32x2 *>(wv + i), a01);
                    p += a.src_pitch;
                }
                *reinterpret_cast<f32x2 *>(t) = a01;
            } else {
                float acc[N];
#pragma unroll
                for (int j = 0; j < N; ++j) acc[j] = a.acc_init;
                for (int i = 0; i < n; ++i) {
                    float f[N];
                    Word<T>::biased(*reinterpret_cast<const uint32_t *>(p), f);
                    const float wgt = wv[i].x;
#pragma unroll
                    for (int j = 0; j < N; ++j) acc[j] = mac<STRICT>(acc[j], unbias<T>(f[j]), wgt);
                    p += a.src_pitch;
                }
                if (N == 4) *reinterpret_cast<float4 *>(t) = make_float4(acc[0], acc[1 % N], acc[2 % N], acc[3 % N]);
                else if (N == 2) *reinterpret_cast<float2 *>(t) = make_float2(acc[0], acc[1 % N]);
                else t[0] = acc[0];
            }
            wd += dw; oyl += dy;
            if (wd >= words) { wd -= words; ++oyl; }
        }
    }
    __syncthreads();

    // ---- horizontal pass out of s_tmp.  Results are staged in s_src (now free) and stored row-wise,
    // except 16-byte pixels (f32 x 4): a warp's 8-row x 4-column patch already writes full 64-byte
    // runs, so those go straight to global memory. ----
    constexpr bool DIRECT = sizeof(T) * C == 16;
    T *s_out = reinterpret_cast<T *>(s_src);
    const int out_pitch = a.out_pitch;  // subpixels per staged output row: 16-byte multiple, odd number of 16-byte units
    const size_t d_pitch = (size_t)a.nw * C;
    T *g0 = dst + (size_t)(oy0 - (int)a.dst_row0) * d_pitch + (size_t)ox0 * C;
    {
        const int items = a.TW * a.TH;  // lane = 8 rows x 4 columns: distinct banks for the float4 reads
        for (int e = tid; e < items; e += THREADS) {
            const int oyl = ((e >> (3 + a.log2TW)) << 3) | (e & 7);   // TW is a power of two
            const int oxl = (e >> 3) & (a.TW - 1);
            if (oyl >= th || oxl >= tw) continue;
            const int n = s_hn[oxl];
            const float *t = s_tmp + oyl * a.tmp_pitch + s_hl[oxl];
            const float2 *wh = s_hw + oxl;
            float acc[C];
            if constexpr (!STRICT && C == 4) {
                f32x2 a01 = 0ull, a23 = 0ull;  // (+0.0f, +0.0f)
                for (int i = 0; i < n; ++i) {
                    const ulonglong2 q = *reinterpret_cast<const ulonglong2 *>(t);
                    const f32x2 ww = *reinterpret_cast<const f32x2 *>(wh + i * a.TW);
                    a01 = fma2(q.x, ww, a01);
                    a23 = fma2(q.y, ww, a23);
                    t += 4;
                }
                unpack2(a01, acc[0], acc[1 % C]);
                unpack2(a23, acc[2 % C], acc[3 % C]);
            } else {
#pragma unroll
                for (int c = 0; c < C; ++c) acc[c] = 0.0f;
                for (int i = 0; i < n; ++i) {
                    const float wgt = wh[i * a.TW].x;
                    float v[C];
                    if (C == 4) { const float4 q = *reinterpret_cast<const float4 *>(t); v[0] = q.x; v[1 % C] = q.y; v[2 % C] = q.z; v[3 % C] = q.w; }
                    else if (C == 2) { const float2 q = *reinterpret_cast<const float2 *>(t); v[0] = q.x; v[1 % C] = q.y; }
                    else {
#pragma unroll
                        for (int c = 0; c < C; ++c) v[c] = t[c];
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) acc[c] = mac<STRICT>(acc[c], v[c], wgt);
                    t += C;
                }
            }
            if constexpr (DIRECT) {
                T *o = g0 + (size_t)oyl * d_pitch + oxl * C;
                T r[C];
#pragma unroll
                for (int c = 0; c < C; ++c) r[c] = Px<T>::clamp_nearest(acc[c]);
                if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) *reinterpret_cast<uint4 *>(o) = *reinterpret_cast<const uint4 *>(r);
                else {
#pragma unroll
                    for (int c = 0; c < C; ++c) o[c] = r[c];
                }
            } else {
                T *o = s_out + (size_t)oyl * out_pitch + oxl * C;
#pragma unroll
                for (int c = 0; c < C; ++c) o[c] = Px<T>::clamp_nearest(acc[c]);
            }
        }
    }
    if constexpr (DIRECT) return;
    __syncthreads();

    // ---- store the tile, row-wise coalesced ----
    {
        const int sub = tw * C;
        const bool vec_ok = (sizeof(T) * d_pitch) % 16 == 0 && (reinterpret_cast<uintptr_t>(g0) & 15) == 0 && (sizeof(T) * sub) % 16 == 0;
        const int lane = tid & 31;
        if (vec_ok) {
            const int v_per_row = (int)(sizeof(T) * sub / 16);
            for (int r = tid >> 5; r < th; r += THREADS / 32)
                for (int v = lane; v < v_per_row; v += 32)
                    reinterpret_cast<uint4 *>(g0 + (size_t)r * d_pitch)[v] = reinterpret_cast<const uint4 *>(s_out + (size_t)r * out_pitch)[v];
        } else {
            for (int r = tid >> 5; r < th; r += THREADS / 32)
                for (int v = lane; v < sub; v += 32) g0[(size_t)r * d_pitch + v] = s_out[(size_t)r * out_pitch + v];
        }
    }
}

// ---- device-side weight tables, cached per (device, filter, lengths, axis) --------
// Tables are handed out as shared_ptr: a launch keeps its tables alive until it has been enqueued,
// whatever other host threads do to the cache meanwhile; the device memory goes when the last
// holder lets go (cudaFree waits for queued kernels that still read it).
struct DevTable {
    int device = -1, filter = -1, axis = -1;
    uint32_t src_len = 0, dst_len = 0;
    AxisTable host;  // left/count kept on the host for tile planning
    int32_t *d_left = nullptr, *d_count = nullptr;
    float *d_w = nullptr;
    DevTable() = default;
    DevTable(const DevTable &) = delete;
    DevTable &operator=(const DevTable &) = delete;
    ~DevTable() {
        if (!d_left && !d_count && !d_w) return;
        int cur = -1;
        cudaGetDevice(&cur);
        if (cur != device) cudaSetDevice(device);
        cudaFree(d_left); cudaFree(d_count); cudaFree(d_w);
        if (cur != device && cur >= 0) cudaSetDevice(cur);
    }
};
using TablePtr = std::shared_ptr<const DevTable>;
std::mutex g_tab_mu;
std::list<std::shared_ptr<DevTable>> g_tabs;  // most recently used first; LRU per device
constexpr size_t MAX_TABLES = 32;             // per device

int get_table(int device, int filter, uint32_t src_len, uint32_t dst_len, int axis, TablePtr *out) {
    {
        std::lock_guard<std::mutex> lk(g_tab_mu);
        for (auto it = g_tabs.begin(); it != g_tabs.end(); ++it) {
            const DevTable &t = **it;
            if (t.device == device && t.filter == filter && t.axis == axis && t.src_len == src_len && t.dst_len == dst_len) {
                g_tabs.splice(g_tabs.begin(), g_tabs, it);
                *out = g_tabs.front();
                return B200IMG_OK;
            }
        }
    }
    // build outside the lock (host arithmetic + synchronous uploads); two threads racing on the same
    // key both build, the second insert simply shadows the first
    auto t = std::make_shared<DevTable>();
    t->device = device; t->filter = filter; t->axis = axis; t->src_len = src_len; t->dst_len = dst_len;
    build_axis_table(filter, src_len, dst_len, axis, t->host);
    const size_t n = dst_len;
    if (cudaMalloc(&t->d_left, n * 4) != cudaSuccess || cudaMalloc(&t->d_count, n * 4) != cudaSuccess ||
        cudaMalloc(&t->d_w, t->host.w.size() * 4) != cudaSuccess) {
        cudaGetLastError();
        return fail(B200IMG_ERR_OOM, "resize: cannot allocate weight tables");
    }
    // synchronous copies: the table is complete before any stream can use it
    B200_CUDA(cudaMemcpy(t->d_left, t->host.left.data(), n * 4, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(t->d_count, t->host.count.data(), n * 4, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(t->d_w, t->host.w.data(), t->host.w.size() * 4, cudaMemcpyHostToDevice));
    std::shared_ptr<DevTable> evicted;
    {
        std::lock_guard<std::mutex> lk(g_tab_mu);
        g_tabs.push_front(t);
        size_t mine = 0;
        for (auto it = g_tabs.begin(); it != g_tabs.end(); ++it)
            if ((*it)->device == device && ++mine > MAX_TABLES) { evicted = *it; g_tabs.erase(it); break; }
    }
    evicted.reset();  // freed here (outside the lock) unless a launch in flight on another thread still holds it
    *out = t;
    return B200IMG_OK;
}

// Largest staged extent over all tiles of size `tile` along one axis.
int max_extent(const AxisTable &t, uint32_t first, uint32_t count, int tile) {
    int m = 1;
    for (uint32_t o = first; o < first + count; o += tile) {
        const uint32_t e = std::min<uint32_t>(o + tile, first + count) - 1;
        m = std::max(m, t.left[e] + t.count[e] - t.left[o]);
    }
    return m;
}

struct Plan { int TW, TH, max_cols, max_rows, src_pitch, tmp_pitch, out_pitch; size_t smem; };

int round_up(int v, int m) { return (v + m - 1) / m * m; }

// Pick the destination tile: smallest staged-source footprint per output pixel
// that still fits the shared-memory budget and keeps >= 2 CTAs per SM.
bool plan_tiles(const AxisTable &vt, const AxisTable &ht, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw,
                int C, int tsize, size_t smem_budget, Plan *best) {
    static const int tws[] = {8, 16, 32, 64, 128};
    static const int ths[] = {8, 16, 32, 64};
    double best_cost = 1e300;
    bool found = false;
    for (int TW : tws) {
        for (int TH : ths) {
            if (TW * TH < 256 && !(TW == 8 && TH == 8)) continue;  // keep every thread busy in the horizontal pass
            Plan p;
            p.TW = TW; p.TH = TH;
            p.max_cols = max_extent(ht, 0, nw, TW);
            p.max_rows = max_extent(vt, dst_row0, dst_rows, TH);
            // staged output rows: a 16-byte multiple with an odd number of 16-byte units, so the 8 rows a
            // warp writes in the horizontal pass fall into different bank groups
            int out_bytes_row = round_up(TW * C * tsize, 16);
            if ((out_bytes_row / 16) % 2 == 0) out_bytes_row += 16;
            p.out_pitch = out_bytes_row / tsize;
            const bool direct = C * tsize == 16;  // f32 x 4: no output staging
            // +32: the span may be widened to 16-byte boundaries on both sides
            p.src_pitch = round_up(p.max_cols * C * tsize + 32, 16);
            int tp = round_up(p.src_pitch / tsize, 4);
            if ((tp / 4) % 2 == 0) tp += 4;  // odd number of 16-byte units: the 8-row lane pattern hits 8 bank groups
            p.tmp_pitch = tp;
            if (!direct) p.max_rows = std::max(p.max_rows, (TH * out_bytes_row + p.src_pitch - 1) / p.src_pitch);
            p.smem = (size_t)p.max_rows * p.src_pitch + (size_t)TH * p.tmp_pitch * 4 + (size_t)TH * vt.stride * 8 +
                     (size_t)TW * ht.stride * 8 + (size_t)(2 * TW + 2 * TH) * 4 + 128;
            if (p.smem > smem_budget) continue;
            // cost: staged bytes + intermediate work per destination pixel, with a mild
            // penalty for tiles so large that few CTAs fit per SM
            const double staged = (double)p.max_rows * p.max_cols / ((double)TW * TH);
            const double vwork = (double)p.max_cols / TW;
            const double occ_pen = p.smem > 100 * 1024 ? 1.3 : (p.smem > 72 * 1024 ? 1.1 : 1.0);
            const double cost = (staged + 2.0 * vwork) * occ_pen;
            if (cost < best_cost) { best_cost = cost; *best = p; found = true; }
        }
    }
    return found;
}

template <typename T, int C>
int launch_tc(const ResizeArgs &a, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    auto k = strict ? resize_kernel<T, C, true> : resize_kernel<T, C, false>;
    if (smem > 48 * 1024) B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, THREADS, smem, st>>>(a);
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

template <typename T>
int launch_t(const ResizeArgs &a, size_t smem, dim3 grid, bool strict, cudaStream_t st) {
    switch (a.C) {
        case 1: return launch_tc<T, 1>(a, smem, grid, strict, st);
        case 2: return launch_tc<T, 2>(a, smem, grid, strict, st);
        case 3: return launch_tc<T, 3>(a, smem, grid, strict, st);
        case 4: return launch_tc<T, 4>(a, smem, grid, strict, st);
    }
    return fail(B200IMG_ERR_BAD_ARG, "resize: channels must be 1..4");
}

// ---- unfused fallback for extreme downscales (the tile's source footprint does not fit
// shared memory): the reference's own structure, an f32 intermediate in global scratch.
template <typename T, bool STRICT>
__global__ void __launch_bounds__(THREADS)
vertical_kernel(const T *src, float *tmp, size_t row_sub, uint32_t src_row0, uint32_t dst_row0, uint32_t dst_rows,
                const int32_t *vleft, const int32_t *vcount, const float *vw, int vstride, float acc_init) {
    const size_t total = (size_t)dst_rows * row_sub;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const uint32_t oyl = (uint32_t)(idx / row_sub);
        const size_t s = idx - (size_t)oyl * row_sub;
        const uint32_t oy = dst_row0 + oyl;
        const int l = vleft[oy], n = vcount[oy];
        const float *w = vw + (size_t)oy * vstride;
        const T *p = src + (size_t)(l - (int)src_row0) * row_sub + s;
        float acc = acc_init;
        for (int i = 0; i < n; ++i) { acc = mac<STRICT>(acc, Px<T>::to_f32(__ldg(p)), w[i]); p += row_sub; }
        tmp[idx] = acc;
    }
}

template <typename T, int C, bool STRICT>
__global__ void __launch_bounds__(THREADS)
horizontal_kernel(const float *tmp, T *dst, uint32_t w, uint32_t nw, uint32_t dst_rows,
                  const int32_t *hleft, const int32_t *hcount, const float *hw, int hstride) {
    const size_t total = (size_t)dst_rows * nw;
    for (size_t idx = (size_t)blockIdx.x * THREADS + threadIdx.x; idx < total; idx += (size_t)gridDim.x * THREADS) {
        const uint32_t oyl = (uint32_t)(idx / nw);
        const uint32_t ox = (uint32_t)(idx - (size_t)oyl * nw);
        const int l = hleft[ox], n = hcount[ox];
        const float *wgt = hw + (size_t)ox * hstride;
        const float *t = tmp + ((size_t)oyl * w + l) * C;
        float acc[C];
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = 0.0f;
        for (int i = 0; i < n; ++i) {
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] = mac<STRICT>(acc[c], t[c], wgt[i]);
            t += C;
        }
        T *o = dst + idx * C;
#pragma unroll
        for (int c = 0; c < C; ++c) o[c] = Px<T>::clamp_nearest(acc[c]);
    }
}

template <typename T, int C>
int unfused_tc(const void *src, void *dst, float *tmp, const ResizeArgs &a, uint32_t n_images, int sm_count,
               bool strict, cudaStream_t st) {
    const size_t row_sub = (size_t)a.w * C;
    const unsigned grid = (unsigned)std::min<size_t>(((size_t)a.dst_rows * row_sub + THREADS - 1) / THREADS, (size_t)sm_count * 16);
    const unsigned grid_h = (unsigned)std::min<size_t>(((size_t)a.dst_rows * a.nw + THREADS - 1) / THREADS, (size_t)sm_count * 16);
    for (uint32_t img = 0; img < n_images; ++img) {
        const T *s = reinterpret_cast<const T *>(reinterpret_cast<const unsigned char *>(src) + (size_t)img * a.src_img_stride);
        T *d = reinterpret_cast<T *>(reinterpret_cast<unsigned char *>(dst) + (size_t)img * a.dst_img_stride);
        if (strict) {
            vertical_kernel<T, true><<<grid, THREADS, 0, st>>>(s, tmp, row_sub, a.src_row0, a.dst_row0, a.dst_rows, a.vleft, a.vcount, a.vw, a.vstride, a.acc_init);
            horizontal_kernel<T, C, true><<<grid_h, THREADS, 0, st>>>(tmp, d, a.w, a.nw, a.dst_rows, a.hleft, a.hcount, a.hw, a.hstride);
        } else {
            vertical_kernel<T, false><<<grid, THREADS, 0, st>>>(s, tmp, row_sub, a.src_row0, a.dst_row0, a.dst_rows, a.vleft, a.vcount, a.vw, a.vstride, a.acc_init);
            horizontal_kernel<T, C, false><<<grid_h, THREADS, 0, st>>>(tmp, d, a.w, a.nw, a.dst_rows, a.hleft, a.hcount, a.hw, a.hstride);
        }
        B200_CUDA(cudaGetLastError());
    }
    return B200IMG_OK;
}

template <typename T>
int unfused_t(const void *src, void *dst, float *tmp, const ResizeArgs &a, uint32_t n, int sms, bool strict, cudaStream_t st) {
    switch (a.C) {
        case 1: return unfused_tc<T, 1>(src, dst, tmp, a, n, sms, strict, st);
        case 2: return unfused_tc<T, 2>(src, dst, tmp, a, n, sms, strict, st);
        case 3: return unfused_tc<T, 3>(src, dst, tmp, a, n, sms, strict, st);
        default: return unfused_tc<T, 4>(src, dst, tmp, a, n, sms, strict, st);
    }
}

}  // namespace

int resize_band_src_rows(uint32_t h, uint32_t nh, int filter, uint32_t dst_row0, uint32_t dst_rows,
                         uint32_t *src_row0, uint32_t *src_rows) {
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize: unknown filter %d", filter);
    if (h == 0 || nh == 0 || dst_rows == 0 || dst_row0 + (uint64_t)dst_rows > nh)
        return fail(B200IMG_ERR_BAD_ARG, "resize band: bad row range");
    AxisTable t;
    build_axis_table(filter, h, nh, 0, t);
    const uint32_t e = dst_row0 + dst_rows - 1;
    *src_row0 = (uint32_t)t.left[dst_row0];
    *src_rows = (uint32_t)(t.left[e] + t.count[e] - t.left[dst_row0]);
    return B200IMG_OK;
}

// Core launcher.  src holds source rows [src_row0, src_row0 + src_rows) of a w x h image (per
// batch entry); dst receives destination rows [dst_row0, dst_row0 + dst_rows).
int resize_launch(const void *src, size_t src_stride, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h,
                  int C, int dtype, void *dst, size_t dst_stride, uint32_t dst_row0, uint32_t dst_rows,
                  uint32_t nw, uint32_t nh, uint32_t n_images, int filter, float acc_init, int flags,
                  cudaStream_t st) {
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize: unknown filter %d", filter);
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "resize: bad dtype %d", dtype);
    if (C < 1 || C > 4) return fail(B200IMG_ERR_BAD_ARG, "resize: channels must be 1..4 (the reference is unimplemented beyond 4)");
    if (n_images == 0 || dst_rows == 0 || nw == 0) return B200IMG_OK;
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    const int tsize = (int)dtype_size(dtype);

    TablePtr vt, ht;  // held until this launch has been enqueued
    int rc = get_table(ctx->device, filter, h, nh, 0, &vt);
    if (rc) return rc;
    rc = get_table(ctx->device, filter, w, nw, 1, &ht);
    if (rc) return rc;
    const uint32_t e = dst_row0 + dst_rows - 1;
    if ((uint32_t)vt->host.left[dst_row0] < src_row0 ||
        (uint32_t)(vt->host.left[e] + vt->host.count[e]) > src_row0 + src_rows)
        return fail(B200IMG_ERR_BAD_ARG, "resize band: source band does not cover the rows the destination band taps");

    ResizeArgs a;
    a.src = src; a.dst = dst;
    a.src_img_stride = src_stride ? src_stride : (size_t)w * src_rows * C * tsize;
    a.dst_img_stride = dst_stride ? dst_stride : (size_t)nw * dst_rows * C * tsize;
    a.w = w; a.C = (uint32_t)C; a.nw = nw;
    a.src_row0 = src_row0; a.dst_row0 = dst_row0; a.dst_rows = dst_rows;
    a.vleft = vt->d_left; a.vcount = vt->d_count; a.vw = vt->d_w; a.vstride = vt->host.stride;
    a.hleft = ht->d_left; a.hcount = ht->d_count; a.hw = ht->d_w; a.hstride = ht->host.stride;
    a.acc_init = acc_init;
    const bool strict = (flags & B200IMG_FP_STRICT) != 0;

    if (!(flags & (B200IMG_RESIZE_UNFUSED | B200IMG_RESIZE_TILED))) {
        if (!strict && !(flags & B200IMG_RESIZE_GENERAL)) {
            rc = resize_int_try(src, a.src_img_stride, src_row0, w, h, C, dtype, dst, a.dst_img_stride, dst_row0, dst_rows, nw, nh,
                                n_images, vt->host, ht->host, acc_init, ctx, st);
            if (rc != 0) return rc < 0 ? rc : B200IMG_OK;
            rc = resize_up2_try(src, a.src_img_stride, src_row0, w, h, C, dtype, dst, a.dst_img_stride, dst_row0, dst_rows, nw, nh,
                                n_images, vt->host, ht->host, acc_init, ctx, st);
            if (rc != 0) return rc < 0 ? rc : B200IMG_OK;
        }
        rc = resize_stream_try(src, a.src_img_stride, src_row0, w, C, dtype, dst, a.dst_img_stride, dst_row0, dst_rows, nw,
                               n_images, vt->host, vt->d_left, vt->d_count, vt->d_w, ht->host, ht->d_left, ht->d_count,
                               ht->d_w, acc_init, strict, ctx, st);
        if (rc != 0) return rc < 0 ? rc : B200IMG_OK;
        rc = resize_group_try(filter, h, src, a.src_img_stride, src_row0, w, C, dtype, dst, a.dst_img_stride, dst_row0, dst_rows, nw,
                              n_images, vt->host, vt->d_left, vt->d_count, vt->d_w, ht->host, ht->d_left, ht->d_count,
                              ht->d_w, acc_init, strict, ctx, st);
        if (rc != 0) return rc < 0 ? rc : B200IMG_OK;
    }

    Plan p;
    const size_t budget = std::min<size_t>(ctx->smem_optin, 160 * 1024);
    const bool fused = !(flags & B200IMG_RESIZE_UNFUSED) &&
                       plan_tiles(vt->host, ht->host, dst_row0, dst_rows, nw, C, tsize, budget, &p);
    if (!fused) {
        size_t tmp_bytes;
        if (!checked_mul(dst_rows, w, (size_t)C, sizeof(float), &tmp_bytes))
            return fail(B200IMG_ERR_DIMENSION, "resize: intermediate size overflows");
        void *tmp;
        rc = scratch(SCRATCH_A, tmp_bytes, &tmp);
        if (rc) return rc;
        switch (dtype) {
            case B200IMG_U8: return unfused_t<uint8_t>(src, dst, (float *)tmp, a, n_images, ctx->sm_count, strict, st);
            case B200IMG_U16: return unfused_t<uint16_t>(src, dst, (float *)tmp, a, n_images, ctx->sm_count, strict, st);
            default: return unfused_t<float>(src, dst, (float *)tmp, a, n_images, ctx->sm_count, strict, st);
        }
    }

    a.TW = p.TW; a.TH = p.TH;
    a.log2TW = 0;
    while ((1 << a.log2TW) < p.TW) ++a.log2TW;
    a.tiles_x = (int)((nw + p.TW - 1) / p.TW);
    a.tiles_y = (int)((dst_rows + p.TH - 1) / p.TH);
    a.max_cols = p.max_cols; a.max_rows = p.max_rows; a.src_pitch = p.src_pitch; a.tmp_pitch = p.tmp_pitch; a.out_pitch = p.out_pitch;
    if (a.tiles_y > 65535 || n_images > 65535) return fail(B200IMG_ERR_DIMENSION, "resize: too many tiles / images for one launch");
    const dim3 grid((unsigned)a.tiles_x, (unsigned)a.tiles_y, n_images);
    switch (dtype) {
        case B200IMG_U8: return launch_t<uint8_t>(a, p.smem, grid, strict, st);
        case B200IMG_U16: return launch_t<uint16_t>(a, p.smem, grid, strict, st);
        default: return launch_t<float>(a, p.smem, grid, strict, st);
    }
}

void resize_release_tables(int device) {
    resize_group_release(device);
    std::list<std::shared_ptr<DevTable>> dropped;
    {
        std::lock_guard<std::mutex> lk(g_tab_mu);
        for (auto it = g_tabs.begin(); it != g_tabs.end();) {
            if ((*it)->device == device) { dropped.push_back(*it); it = g_tabs.erase(it); }
            else ++it;
        }
    }
}

}  // namespace b200
