// multi.cu -- one image split into row bands over several GPUs, driven from ONE host thread.
//
// The path has no reduction across shards, so there is no collective and no NCCL here:
// a band needs a few rows of its neighbours' bands (the rows its vertical taps reach, or the blur
// radius), fetched once before the kernels with cudaMemcpyPeerAsync over NVLink ("ghost rows"),
// after which every device runs the ordinary single-device kernels on its band.  Bands are never
// split in x: the horizontal passes need whole rows (src/imageops/sample.rs:217-340).
//
// What a band must reproduce (reference): the vertical taps of src/imageops/sample.rs:494-562 decide
// the resampler's ghost rows; the Gaussian column pass clamps to the IMAGE's first / last row
// (src/imageops/filter_1d.rs:102-130), never to a band's; the box blur's three passes compose, so a
// band needs r1 + r2 + r3 rows beyond its own (src/imageops/fast_blur.rs:303-473).
//
// The blurs run the whole-image kernels on [band - halo, band + halo) clipped to the image and keep
// the band's own rows: a row that is `halo` or more rows away from an artificial edge never sees
// that edge's clamping, and an edge that is the image's edge clamps exactly as the reference does.
#include "common.cuh"
#include "host_tables.h"

#include <algorithm>
#include <map>
#include <mutex>
#include <vector>

namespace b200 {

int resize_launch(const void *src, size_t src_stride, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h,
                  int C, int dtype, void *dst, size_t dst_stride, uint32_t dst_row0, uint32_t dst_rows,
                  uint32_t nw, uint32_t nh, uint32_t n_images, int filter, float acc_init, int flags,
                  cudaStream_t st);
int resize_band_src_rows(uint32_t h, uint32_t nh, int filter, uint32_t dst_row0, uint32_t dst_rows,
                         uint32_t *src_row0, uint32_t *src_rows);
int gaussian_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                         const b200img_gauss_params *p, int path, int flags, cudaStream_t st, bool image_top = true);
int fast_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     float sigma, cudaStream_t st);
int gaussian_blur_rows_launch(const void *src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h, int C, int dtype,
                              void *dst_band, uint32_t dst_row0, uint32_t dst_rows, const b200img_gauss_params *p, int path, int flags,
                              cudaStream_t st);

namespace {

// rows beyond a band that a blur of it reads
uint32_t gauss_halo(const b200img_gauss_params *p) { return p->y_kernel_size / 2; }
uint32_t fast_halo(float sigma) {
    uint32_t r[3];
    boxes_for_gauss(sigma, r);
    return r[0] + r[1] + r[2];
}

void blur_need(uint32_t h, uint32_t halo, uint32_t row0, uint32_t rows, uint32_t *s0, uint32_t *sn) {
    const uint32_t lo = row0 > halo ? row0 - halo : 0;
    const uint64_t hi = std::min<uint64_t>((uint64_t)row0 + rows + halo, h);
    *s0 = lo;
    *sn = (uint32_t)(hi - lo);
}

int check_geom(const char *op, int C, int dtype, uint32_t w, uint32_t h, size_t *row_bytes) {
    if (C < 1 || C > 4) return fail(B200IMG_ERR_BAD_ARG, "%s: channels must be 1..4 (got %d)", op, C);
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "%s: bad dtype %d", op, dtype);
    size_t total;
    if (!checked_mul(w, h, (size_t)C, dtype_size(dtype), &total)) return fail(B200IMG_ERR_DIMENSION, "%s: image size overflows", op);
    *row_bytes = (size_t)w * C * dtype_size(dtype);
    return B200IMG_OK;
}

// One band of a blur on the current device.  `blur(src, rows, dst, image_top)` runs the whole-image
// kernels on `rows` rows.
template <typename Blur>
int blur_band(const char *op, const void *src_band, uint32_t src_row0, uint32_t src_rows, uint32_t h, size_t row_bytes,
              void *dst_band, uint32_t dst_row0, uint32_t dst_rows, uint32_t halo, cudaStream_t st, Blur blur) {
    if (dst_rows == 0 || row_bytes == 0) return B200IMG_OK;
    if (dst_row0 + (uint64_t)dst_rows > h || src_row0 + (uint64_t)src_rows > h)
        return fail(B200IMG_ERR_BAD_ARG, "%s: band outside the image", op);
    if (!src_band || !dst_band) return fail(B200IMG_ERR_BAD_ARG, "%s: null buffer", op);
    uint32_t lo, n;
    blur_need(h, halo, dst_row0, dst_rows, &lo, &n);
    if (lo < src_row0 || lo + (uint64_t)n > src_row0 + (uint64_t)src_rows)
        return fail(B200IMG_ERR_BAD_ARG, "%s: the source band [%u, %u) does not cover rows [%u, %u) that destination rows [%u, %u) read",
                    op, src_row0, src_row0 + src_rows, lo, lo + n, dst_row0, dst_row0 + dst_rows);
    const char *sub = (const char *)src_band + (size_t)(lo - src_row0) * row_bytes;
    if (lo == dst_row0 && n == dst_rows) return blur(sub, n, dst_band, lo == 0);  // the band reads nothing beyond itself
    void *tmp;
    int rc = scratch(SCRATCH_B, (size_t)n * row_bytes, &tmp);
    if (rc) return rc;
    if ((rc = blur(sub, n, tmp, lo == 0))) return rc;
    B200_CUDA(cudaMemcpyAsync(dst_band, (const char *)tmp + (size_t)(dst_row0 - lo) * row_bytes, (size_t)dst_rows * row_bytes,
                              cudaMemcpyDeviceToDevice, st));
    return B200IMG_OK;
}

// ---- the multi-device driver ---------------------------------------------------------------------
struct DevStreams {
    std::mutex mu;
    std::map<int, cudaStream_t> st;
    std::map<std::pair<int, int>, bool> peer;  // (device, peer) -> access enabled or known impossible
} g_multi;

int internal_stream(int dev, cudaStream_t *out) {  // device `dev` is current
    std::lock_guard<std::mutex> lk(g_multi.mu);
    auto it = g_multi.st.find(dev);
    if (it == g_multi.st.end()) {
        cudaStream_t s;
        B200_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        it = g_multi.st.emplace(dev, s).first;
    }
    *out = it->second;
    return B200IMG_OK;
}

// Direct NVLink / PCIe peer copies need peer access from the destination's context; without it
// cudaMemcpyPeerAsync still works (staged through the host), so failure to enable is not an error.
void enable_peer(int dev, int peer) {  // device `dev` is current
    if (dev == peer) return;
    std::lock_guard<std::mutex> lk(g_multi.mu);
    auto key = std::make_pair(dev, peer);
    if (g_multi.peer.count(key)) return;
    int can = 0;
    if (cudaDeviceCanAccessPeer(&can, dev, peer) == cudaSuccess && can) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
        if (e != cudaSuccess) cudaGetLastError();  // already enabled by the caller (or refused): both fine
    } else {
        cudaGetLastError();
    }
    g_multi.peer[key] = true;
}

struct DeviceGuard {
    int saved = -1;
    DeviceGuard() { cudaGetDevice(&saved); }
    ~DeviceGuard() { if (saved >= 0) cudaSetDevice(saved); }
};

int check_bands(const char *op, const b200img_band *bands, int n, uint32_t h, uint32_t dst_h) {
    if (n < 1 || !bands) return fail(B200IMG_ERR_BAD_ARG, "%s: no bands", op);
    int count = 0;
    B200_CUDA(cudaGetDeviceCount(&count));
    for (int k = 0; k < n; ++k) {
        const b200img_band &b = bands[k];
        if (b.device < 0 || b.device >= count) return fail(B200IMG_ERR_BAD_ARG, "%s: band %d names device %d (the process sees %d)", op, k, b.device, count);
        if (b.src_row0 + (uint64_t)b.src_rows > h || b.dst_row0 + (uint64_t)b.dst_rows > dst_h)
            return fail(B200IMG_ERR_BAD_ARG, "%s: band %d lies outside the image", op, k);
        if (b.own_rows && (b.own_row0 < b.src_row0 || b.own_row0 + (uint64_t)b.own_rows > b.src_row0 + (uint64_t)b.src_rows))
            return fail(B200IMG_ERR_BAD_ARG, "%s: band %d owns rows outside its source buffer", op, k);
        if ((b.src_rows && !b.d_src) || (b.dst_rows && !b.d_dst)) return fail(B200IMG_ERR_BAD_ARG, "%s: band %d has a null buffer", op, k);
    }
    return B200IMG_OK;
}

// Fill every band's ghost rows (the rows of its source buffer it does not own) from the bands that
// own them; enqueued on the destination band's stream after the owner's stream has reached `ready`.
int exchange(const char *op, const b200img_band *bands, int n, size_t row_bytes, const std::vector<cudaStream_t> &st,
             const std::vector<cudaEvent_t> &ready) {
    for (int k = 0; k < n; ++k) {
        const b200img_band &b = bands[k];
        if (b.src_rows == 0) continue;
        const uint32_t own0 = b.own_rows ? b.own_row0 : b.src_row0, own1 = own0 + b.own_rows;
        const uint32_t ghost[2][2] = {{b.src_row0, own0}, {own1, b.src_row0 + b.src_rows}};
        B200_CUDA(cudaSetDevice(b.device));
        for (const auto &g : ghost) {
            uint32_t covered = g[0];
            while (covered < g[1]) {  // walk the range in owner order
                int owner = -1;
                uint32_t hi = 0;
                for (int j = 0; j < n; ++j) {
                    const b200img_band &o = bands[j];
                    if (j == k || !o.own_rows) continue;
                    if (o.own_row0 <= covered && covered < o.own_row0 + o.own_rows) { owner = j; hi = std::min(g[1], o.own_row0 + o.own_rows); break; }
                }
                if (owner < 0) return fail(B200IMG_ERR_BAD_ARG, "%s: no band owns source row %u, which band %d reads", op, covered, k);
                const b200img_band &o = bands[owner];
                enable_peer(b.device, o.device);
                B200_CUDA(cudaStreamWaitEvent(st[k], ready[owner], 0));
                B200_CUDA(cudaMemcpyPeerAsync((char *)b.d_src + (size_t)(covered - b.src_row0) * row_bytes, b.device,
                                              (const char *)o.d_src + (size_t)(covered - o.src_row0) * row_bytes, o.device,
                                              (size_t)(hi - covered) * row_bytes, st[k]));
                covered = hi;
            }
        }
    }
    return B200IMG_OK;
}

// Common driver: streams (the caller's or internal ones), ready events, ghost rows, per-band work.
// `work(band, stream)` runs with the band's device current.
template <typename Work>
int run_bands(const char *op, const b200img_band *bands, int n, size_t row_bytes, void *const *streams, Work work) {
    DeviceGuard guard;
    std::vector<cudaStream_t> st(n);
    std::vector<cudaEvent_t> ready(n, nullptr);
    int rc = B200IMG_OK;
    auto cleanup = [&]() {
        for (int k = 0; k < n; ++k)
            if (ready[k]) { cudaSetDevice(bands[k].device); cudaEventDestroy(ready[k]); }
    };
    for (int k = 0; k < n && !rc; ++k) {
        if (cudaSetDevice(bands[k].device) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "cudaSetDevice"); break; }
        if (streams) st[k] = (cudaStream_t)streams[k];
        else rc = internal_stream(bands[k].device, &st[k]);
        if (!rc && cudaEventCreateWithFlags(&ready[k], cudaEventDisableTiming) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventCreate");
        // the band's own rows are whatever its stream has produced up to here
        if (!rc && cudaEventRecord(ready[k], st[k]) != cudaSuccess) rc = cuda_fail(cudaGetLastError(), "cudaEventRecord");
    }
    if (!rc) rc = exchange(op, bands, n, row_bytes, st, ready);
    for (int k = 0; k < n && !rc; ++k) {
        if (cudaSetDevice(bands[k].device) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "cudaSetDevice"); break; }
        ScratchScope scope(st[k]);
        rc = work(bands[k], st[k]);
    }
    if (!streams || rc) {  // internal streams: the call is synchronous; on failure nothing stays in flight either way
        for (int k = 0; k < n; ++k) {
            if (cudaSetDevice(bands[k].device) != cudaSuccess) continue;
            const cudaError_t e = cudaStreamSynchronize(st[k]);
            if (e != cudaSuccess && !rc) rc = cuda_fail(e, "band stream");
        }
    }
    cleanup();
    return rc;
}

// own ranges from need ranges: a partition of the rows that at least one band reads, band k's share
// lying inside what it reads itself
void assign_owners(uint32_t h, int n, b200img_band *bands) {
    std::vector<uint32_t> cut(n + 1, 0);
    cut[n] = h;
    for (int k = 1; k < n; ++k) {
        const uint32_t prev_end = bands[k - 1].src_row0 + bands[k - 1].src_rows, next_begin = bands[k].src_row0;
        const uint32_t even = (uint32_t)((uint64_t)h * k / n);
        if (next_begin <= prev_end) cut[k] = std::min(std::max(even, next_begin), prev_end);
        else cut[k] = next_begin;  // nobody reads the rows in between
        cut[k] = std::max(cut[k], cut[k - 1]);
    }
    for (int k = 0; k < n; ++k) {
        const uint32_t lo = std::max(cut[k], bands[k].src_row0), hi = std::min(cut[k + 1], bands[k].src_row0 + bands[k].src_rows);
        bands[k].own_row0 = lo;
        bands[k].own_rows = hi > lo ? hi - lo : 0;
    }
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

// ---- single-device band entry points of the blurs ------------------------------------------------
int b200img_gaussian_blur_band_src_rows(uint32_t h, const b200img_gauss_params *p, uint32_t dst_row0, uint32_t dst_rows,
                                        uint32_t *src_row0, uint32_t *src_rows) {
    if (!p || !src_row0 || !src_rows) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur_band_src_rows: null pointer");
    if (dst_rows == 0 || dst_row0 + (uint64_t)dst_rows > h) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur band: bad row range");
    blur_need(h, gauss_halo(p), dst_row0, dst_rows, src_row0, src_rows);
    return B200IMG_OK;
}

int b200img_fast_blur_band_src_rows(uint32_t h, float sigma, uint32_t dst_row0, uint32_t dst_rows, uint32_t *src_row0,
                                    uint32_t *src_rows) {
    if (!src_row0 || !src_rows) return fail(B200IMG_ERR_BAD_ARG, "fast_blur_band_src_rows: null pointer");
    if (dst_rows == 0 || dst_row0 + (uint64_t)dst_rows > h) return fail(B200IMG_ERR_BAD_ARG, "fast_blur band: bad row range");
    blur_need(h, fast_halo(sigma), dst_row0, dst_rows, src_row0, src_rows);
    return B200IMG_OK;
}

int b200img_gaussian_blur_band_dev(const void *d_src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h,
                                   int C, int dtype, void *d_dst_band, uint32_t dst_row0, uint32_t dst_rows,
                                   const b200img_gauss_params *p, int path, int flags, void *stream) {
    size_t row_bytes;
    int rc = check_geom("gaussian_blur_band", C, dtype, w, h, &row_bytes);
    if (rc) return rc;
    if (!p) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur_band: null params");
    if (path != B200IMG_BLUR_GENERIC && path != B200IMG_BLUR_DYN) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur_band: bad path %d", path);
    if ((p->x_kernel_size & 1u) == 0 || (p->y_kernel_size & 1u) == 0)
        return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur_band: kernel sizes must be odd (%u, %u)", p->x_kernel_size, p->y_kernel_size);
    const cudaStream_t st = (cudaStream_t)stream;
    ScratchScope scope(st);
    // large bands of long kernels: the fused kernel computes exactly the band's rows from the band buffer (it clamps
    // at the IMAGE's edges itself); everything else blurs the haloed band as an image and keeps the band's rows
    if (dst_rows && row_bytes && d_src_band && d_dst_band && dst_row0 + (uint64_t)dst_rows <= h && src_row0 + (uint64_t)src_rows <= h &&
        (uint64_t)dst_rows * w >= (1u << 20) && p->x_kernel_size + p->y_kernel_size >= 64) {
        uint32_t lo, n;
        blur_need(h, gauss_halo(p), dst_row0, dst_rows, &lo, &n);
        if (lo >= src_row0 && lo + (uint64_t)n <= src_row0 + (uint64_t)src_rows) {
            rc = gaussian_blur_rows_launch(d_src_band, src_row0, src_rows, w, h, C, dtype, d_dst_band, dst_row0, dst_rows, p, path, flags, st);
            if (rc != 1) return rc;
        }
    }
    return blur_band("gaussian_blur_band", d_src_band, src_row0, src_rows, h, row_bytes, d_dst_band, dst_row0, dst_rows, gauss_halo(p), st,
                     [&](const void *s, uint32_t rows, void *d, bool top) {
                         return gaussian_blur_launch(s, w, rows, C, dtype, d, 1, p, path, flags, st, top);
                     });
}

int b200img_fast_blur_band_dev(const void *d_src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h, int C,
                               int dtype, void *d_dst_band, uint32_t dst_row0, uint32_t dst_rows, float sigma, void *stream) {
    size_t row_bytes;
    int rc = check_geom("fast_blur_band", C, dtype, w, h, &row_bytes);
    if (rc) return rc;
    const cudaStream_t st = (cudaStream_t)stream;
    ScratchScope scope(st);
    return blur_band("fast_blur_band", d_src_band, src_row0, src_rows, h, row_bytes, d_dst_band, dst_row0, dst_rows, fast_halo(sigma), st,
                     [&](const void *s, uint32_t rows, void *d, bool) { return fast_blur_launch(s, w, rows, C, dtype, d, 1, sigma, st); });
}

// ---- band plans --------------------------------------------------------------------------------
int b200img_plan_resize_bands(uint32_t h, uint32_t nh, int filter, int n_bands, b200img_band *bands) {
    if (n_bands < 1 || !bands) return fail(B200IMG_ERR_BAD_ARG, "plan_resize_bands: no bands");
    if (h == 0 || nh == 0) return fail(B200IMG_ERR_BAD_ARG, "plan_resize_bands: empty image");
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "plan_resize_bands: unknown filter %d", filter);
    AxisTable t;
    build_axis_table(filter, h, nh, 0, t);
    for (int k = 0; k < n_bands; ++k) {
        // destination bands start on multiples of four rows: the grouped resampler's unit
        auto cut = [&](int i) { return i >= n_bands ? nh : (uint32_t)((uint64_t)nh * i / n_bands) & ~3u; };
        const uint32_t d0 = cut(k), d1 = cut(k + 1);
        bands[k].dst_row0 = d0;
        bands[k].dst_rows = d1 - d0;
        if (d1 > d0) {
            bands[k].src_row0 = (uint32_t)t.left[d0];
            bands[k].src_rows = (uint32_t)(t.left[d1 - 1] + t.count[d1 - 1] - t.left[d0]);
        } else {  // more bands than rows: an empty band, placed where it keeps the ranges monotone
            bands[k].src_row0 = k ? bands[k - 1].src_row0 + bands[k - 1].src_rows : 0;
            bands[k].src_rows = 0;
        }
    }
    assign_owners(h, n_bands, bands);
    return B200IMG_OK;
}

static int plan_blur(const char *op, uint32_t h, uint32_t halo, int n_bands, b200img_band *bands) {
    if (n_bands < 1 || !bands) return fail(B200IMG_ERR_BAD_ARG, "%s: no bands", op);
    if (h == 0) return fail(B200IMG_ERR_BAD_ARG, "%s: empty image", op);
    for (int k = 0; k < n_bands; ++k) {
        const uint32_t d0 = (uint32_t)((uint64_t)h * k / n_bands), d1 = (uint32_t)((uint64_t)h * (k + 1) / n_bands);
        bands[k].dst_row0 = d0;
        bands[k].dst_rows = d1 - d0;
        bands[k].own_row0 = d0;  // a blur's band owns exactly the rows it produces
        bands[k].own_rows = d1 - d0;
        if (d1 > d0) blur_need(h, halo, d0, d1 - d0, &bands[k].src_row0, &bands[k].src_rows);
        else { bands[k].src_row0 = d0; bands[k].src_rows = 0; }
    }
    return B200IMG_OK;
}

int b200img_plan_gaussian_blur_bands(uint32_t h, const b200img_gauss_params *p, int n_bands, b200img_band *bands) {
    if (!p) return fail(B200IMG_ERR_BAD_ARG, "plan_gaussian_blur_bands: null params");
    return plan_blur("plan_gaussian_blur_bands", h, gauss_halo(p), n_bands, bands);
}

int b200img_plan_fast_blur_bands(uint32_t h, float sigma, int n_bands, b200img_band *bands) {
    return plan_blur("plan_fast_blur_bands", h, fast_halo(sigma), n_bands, bands);
}

// ---- multi-device calls ----------------------------------------------------------------------------
int b200img_exchange_ghost_rows(const b200img_band *bands, int n_bands, uint32_t h, size_t row_bytes, void *const *streams) {
    int rc = check_bands("exchange_ghost_rows", bands, n_bands, h, UINT32_MAX);
    if (rc) return rc;
    return run_bands("exchange_ghost_rows", bands, n_bands, row_bytes, streams,
                     [](const b200img_band &, cudaStream_t) { return (int)B200IMG_OK; });
}

int b200img_resize_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int C, int dtype, uint32_t nw,
                         uint32_t nh, int filter, float acc_init, int flags, void *const *streams) {
    size_t row_bytes, dst_row_bytes;
    int rc = check_geom("resize_bands", C, dtype, w, h, &row_bytes);
    if (rc) return rc;
    if ((rc = check_geom("resize_bands", C, dtype, nw, nh, &dst_row_bytes))) return rc;
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize_bands: unknown filter %d", filter);
    if (w == 0 || h == 0 || nw == 0 || nh == 0) return fail(B200IMG_ERR_BAD_ARG, "resize_bands: empty image");
    if ((rc = check_bands("resize_bands", bands, n_bands, h, nh))) return rc;
    return run_bands("resize_bands", bands, n_bands, row_bytes, streams, [&](const b200img_band &b, cudaStream_t st) {
        if (b.dst_rows == 0) return (int)B200IMG_OK;
        return b200img_resize_band_dev(b.d_src, b.src_row0, b.src_rows, w, h, C, dtype, b.d_dst, b.dst_row0, b.dst_rows, nw, nh,
                                       filter, acc_init, flags, st);
    });
}

int b200img_gaussian_blur_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int C, int dtype,
                                const b200img_gauss_params *p, int path, int flags, void *const *streams) {
    size_t row_bytes;
    int rc = check_geom("gaussian_blur_bands", C, dtype, w, h, &row_bytes);
    if (rc) return rc;
    if (!p) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur_bands: null params");
    if ((rc = check_bands("gaussian_blur_bands", bands, n_bands, h, h))) return rc;
    return run_bands("gaussian_blur_bands", bands, n_bands, row_bytes, streams, [&](const b200img_band &b, cudaStream_t st) {
        if (b.dst_rows == 0) return (int)B200IMG_OK;
        return b200img_gaussian_blur_band_dev(b.d_src, b.src_row0, b.src_rows, w, h, C, dtype, b.d_dst, b.dst_row0, b.dst_rows, p,
                                              path, flags, st);
    });
}

int b200img_fast_blur_bands(const b200img_band *bands, int n_bands, uint32_t w, uint32_t h, int C, int dtype, float sigma,
                            void *const *streams) {
    size_t row_bytes;
    int rc = check_geom("fast_blur_bands", C, dtype, w, h, &row_bytes);
    if (rc) return rc;
    if ((rc = check_bands("fast_blur_bands", bands, n_bands, h, h))) return rc;
    return run_bands("fast_blur_bands", bands, n_bands, row_bytes, streams, [&](const b200img_band &b, cudaStream_t st) {
        if (b.dst_rows == 0) return (int)B200IMG_OK;
        return b200img_fast_blur_band_dev(b.d_src, b.src_row0, b.src_rows, w, h, C, dtype, b.d_dst, b.dst_row0, b.dst_rows, sigma, st);
    });
}

}  // extern "C"
