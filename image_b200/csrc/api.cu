// api.cu -- the extern "C" surface of libb200img (include/b200img.h): argument
// checking in the reference's terms, per-device context and scratch, and the
// host-pointer wrappers (H2D -> kernels -> D2H) around the device entry points.
#include "common.cuh"
#include "host_tables.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <tuple>
#include <vector>

namespace b200 {

// implemented in the kernel files
int convert_dev(const void *src, int sd, int cin, void *dst, int dd, int cout, size_t n, int mode,
                const float coefs[3], int fused_mac, cudaStream_t st);
int swizzle_dev(void *buf, int dtype, int channels, size_t nsub, cudaStream_t st);
int resize_launch(const void *src, size_t src_stride, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h,
                  int C, int dtype, void *dst, size_t dst_stride, uint32_t dst_row0, uint32_t dst_rows,
                  uint32_t nw, uint32_t nh, uint32_t n_images, int filter, float acc_init, int flags,
                  cudaStream_t st);
int resize_band_src_rows(uint32_t h, uint32_t nh, int filter, uint32_t dst_row0, uint32_t dst_rows,
                         uint32_t *src_row0, uint32_t *src_rows);
void resize_release_tables(int device);
int gaussian_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                         const b200img_gauss_params *p, int path, int flags, cudaStream_t st, bool image_top = true);
int fast_blur_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     float sigma, cudaStream_t st);
int thumbnail_check(uint32_t w, uint32_t h, int dtype, uint32_t nw, uint32_t nh);
int thumbnail_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                     uint32_t n_images, int path, cudaStream_t st);
int colorop_launch(int op, const void *src, void *dst, size_t npixels, int C, int dtype, int32_t ivalue, float fvalue,
                   cudaStream_t st);
int alpha_scan_launch(const void *src, size_t npixels, int C, int dtype, int *d_flag, cudaStream_t st);
int alpha_mul_launch(void *buf, size_t npixels, int C, int dtype, bool un, cudaStream_t st);
int filter3x3_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images,
                     const float kernel[9], cudaStream_t st);
int unsharpen_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n_images, float sigma,
                     int threshold, int flags, cudaStream_t st);

// ---- errors -------------------------------------------------------------------
static thread_local char g_err[512] = "";

int fail(int status, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return status;
}

int cuda_fail(cudaError_t e, const char *what) {
    snprintf(g_err, sizeof g_err, "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    cudaGetLastError();  // clear the sticky-less error state
    return e == cudaErrorMemoryAllocation ? B200IMG_ERR_OOM : B200IMG_ERR_CUDA;
}

bool checked_mul(size_t a, size_t b, size_t c, size_t d, size_t *out) {
    size_t r;
    if (__builtin_mul_overflow(a, b, &r) || __builtin_mul_overflow(r, c, &r) || __builtin_mul_overflow(r, d, &r)) return false;
    *out = r;
    return true;
}

// ---- per-device context -------------------------------------------------------
static std::mutex g_ctx_mu;
static std::map<int, DeviceCtx> g_ctx;

DeviceCtx *current_ctx() {
    int dev;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaGetDevice (is there a GPU? libb200img has no CPU fallback)");
        return nullptr;
    }
    std::lock_guard<std::mutex> lk(g_ctx_mu);
    auto it = g_ctx.find(dev);
    if (it != g_ctx.end()) return &it->second;
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, dev);
    if (e != cudaSuccess) {
        cuda_fail(e, "cudaGetDeviceProperties");
        return nullptr;
    }
    if (prop.major < 10) {
        fail(B200IMG_ERR_CUDA, "device %d is sm_%d%d; libb200img is built for sm_100a only", dev, prop.major, prop.minor);
        return nullptr;
    }
    DeviceCtx c;
    c.device = dev;
    c.sm_count = prop.multiProcessorCount;
    c.smem_optin = prop.sharedMemPerBlockOptin;
    return &(g_ctx[dev] = c);
}

// ---- scratch pool: buffers per (device, stream, slot), grown on demand -------------------
// A buffer is LEASED to the host thread that asked for it until that thread's entry point returns
// (ScratchScope), so another host thread working on the same (device, stream, slot) can neither free
// nor share it between the first thread's scratch() and its launches: it gets a buffer of its own.
// Between leases, re-use is ordered by the stream the key names.  Growth is stream-ordered too
// (cudaFreeAsync / cudaMallocAsync on that stream): a `_dev` call never synchronises the device.
struct ScratchBuf { void *p = nullptr; size_t bytes = 0; bool leased = false; };
using ScratchKey = std::tuple<int, cudaStream_t, int>;
static std::mutex g_scr_mu;
static std::map<ScratchKey, std::vector<ScratchBuf>> g_scr;
static thread_local cudaStream_t g_scratch_stream = nullptr;
static thread_local std::vector<std::pair<ScratchKey, void *>> g_leases;

void set_scratch_stream(cudaStream_t st) { g_scratch_stream = st; }

void release_scratch_leases() {
    if (g_leases.empty()) return;
    std::lock_guard<std::mutex> lk(g_scr_mu);
    for (auto &l : g_leases) {
        auto it = g_scr.find(l.first);
        if (it == g_scr.end()) continue;
        for (ScratchBuf &b : it->second)
            if (b.p == l.second) b.leased = false;
    }
    g_leases.clear();
}

int scratch(ScratchSlot slot, size_t bytes, void **out) {
    int dev;
    B200_CUDA(cudaGetDevice(&dev));
    const cudaStream_t st = g_scratch_stream;
    const ScratchKey key = std::make_tuple(dev, st, (int)slot);
    std::lock_guard<std::mutex> lk(g_scr_mu);
    for (auto &l : g_leases)  // asked twice within one call: the same buffer if it is large enough
        if (l.first == key)
            for (ScratchBuf &b : g_scr[key])
                if (b.p == l.second && b.bytes >= bytes) { *out = b.p; return B200IMG_OK; }
    std::vector<ScratchBuf> &bufs = g_scr[key];
    ScratchBuf *pick = nullptr;
    for (ScratchBuf &b : bufs)
        if (!b.leased && b.bytes >= bytes) { pick = &b; break; }
    if (!pick) {
        for (ScratchBuf &b : bufs)
            if (!b.leased) { pick = &b; break; }  // too small: replaced below
        if (!pick) { bufs.emplace_back(); pick = &bufs.back(); }
        if (pick->p) cudaFreeAsync(pick->p, st);  // after the work already queued on `st` that uses it
        pick->p = nullptr; pick->bytes = 0;
        const size_t want = (bytes + ((size_t)1 << 20) - 1) & ~(((size_t)1 << 20) - 1);
        cudaError_t e = cudaMallocAsync(&pick->p, want, st);
        if (e != cudaSuccess) {
            pick->p = nullptr;
            cudaGetLastError();
            return fail(B200IMG_ERR_OOM, "cannot allocate %zu bytes of device scratch", want);
        }
        pick->bytes = want;
    }
    pick->leased = true;
    g_leases.emplace_back(key, pick->p);
    *out = pick->p;
    return B200IMG_OK;
}

static void release_scratch(int dev) {  // caller has synchronised the device
    std::lock_guard<std::mutex> lk(g_scr_mu);
    for (auto it = g_scr.begin(); it != g_scr.end();) {
        if (std::get<0>(it->first) == dev) {
            for (ScratchBuf &b : it->second) cudaFree(b.p);
            it = g_scr.erase(it);
        } else ++it;
    }
}

// Host-pointer entry points share the stream-0 scratch and the staging pipe of their device: one
// call at a time per DEVICE (not per process -- one process can feed every GPU of the box at once).
static std::mutex g_host_mu_map_mu;
static std::map<int, std::unique_ptr<std::mutex>> g_host_mu;
static std::mutex &host_mutex_of_current_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(g_host_mu_map_mu);
    auto &m = g_host_mu[dev];
    if (!m) m.reset(new std::mutex);
    return *m;
}

// RAII: host-pointer call = lock, device buffers, copy in ... copy out, sync
struct HostCall {
    std::lock_guard<std::mutex> lk;
    ScratchScope scope;
    HostCall() : lk(host_mutex_of_current_device()), scope(nullptr) {}
    int in(const void *h, size_t bytes, void **d) {
        int rc = scratch(SCRATCH_IO_SRC, bytes ? bytes : 1, d);
        if (rc) return rc;
        if (bytes) B200_CUDA(cudaMemcpyAsync(*d, h, bytes, cudaMemcpyHostToDevice, 0));
        return B200IMG_OK;
    }
    int out_buf(size_t bytes, void **d) { return scratch(SCRATCH_IO_DST, bytes ? bytes : 1, d); }
    int out(void *h, const void *d, size_t bytes) {
        if (bytes) B200_CUDA(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, 0));
        B200_CUDA(cudaStreamSynchronize(0));
        return B200IMG_OK;
    }
    // The caller's buffers must be quiescent when an entry point returns, also when it fails midway.
    ~HostCall() { cudaStreamSynchronize(0); }
};

// ---- host-resident batches: H2D, kernels and D2H of consecutive images overlap on three streams ----
struct HostPipe {
    static constexpr int SLOTS = 3;
    cudaStream_t h2d = nullptr, exec = nullptr, d2h = nullptr;
    cudaEvent_t e_in[SLOTS] = {}, e_run[SLOTS] = {}, e_out[SLOTS] = {};
    void *src[SLOTS] = {}, *dst[SLOTS] = {};
    size_t src_bytes = 0, dst_bytes = 0;
    bool ready = false;
};
static std::mutex g_pipes_mu;            // guards the map; a pipe itself is used under its device's host mutex
static std::map<int, HostPipe> g_pipes;  // std::map: references stay valid across inserts

static void pipe_free_buffers(HostPipe &p) {
    for (int i = 0; i < HostPipe::SLOTS; ++i) {
        cudaFree(p.src[i]); cudaFree(p.dst[i]);
        p.src[i] = p.dst[i] = nullptr;
    }
    p.src_bytes = p.dst_bytes = 0;
}

static void release_pipe(int dev) {  // caller holds the device's host mutex and has synchronised the device
    std::lock_guard<std::mutex> lk(g_pipes_mu);
    auto it = g_pipes.find(dev);
    if (it == g_pipes.end()) return;
    HostPipe &p = it->second;
    pipe_free_buffers(p);
    if (p.ready) {
        for (int i = 0; i < HostPipe::SLOTS; ++i) { cudaEventDestroy(p.e_in[i]); cudaEventDestroy(p.e_run[i]); cudaEventDestroy(p.e_out[i]); }
        cudaStreamDestroy(p.h2d); cudaStreamDestroy(p.exec); cudaStreamDestroy(p.d2h);
    }
    g_pipes.erase(it);
}

static int pipe_get(size_t sb, size_t db, HostPipe **out) {
    int dev;
    B200_CUDA(cudaGetDevice(&dev));
    HostPipe *pp;
    {
        std::lock_guard<std::mutex> lk(g_pipes_mu);
        pp = &g_pipes[dev];
    }
    HostPipe &p = *pp;
    if (!p.ready) {
        B200_CUDA(cudaStreamCreateWithFlags(&p.h2d, cudaStreamNonBlocking));
        B200_CUDA(cudaStreamCreateWithFlags(&p.exec, cudaStreamNonBlocking));
        B200_CUDA(cudaStreamCreateWithFlags(&p.d2h, cudaStreamNonBlocking));
        for (int i = 0; i < HostPipe::SLOTS; ++i) {
            B200_CUDA(cudaEventCreateWithFlags(&p.e_in[i], cudaEventDisableTiming));
            B200_CUDA(cudaEventCreateWithFlags(&p.e_run[i], cudaEventDisableTiming));
            B200_CUDA(cudaEventCreateWithFlags(&p.e_out[i], cudaEventDisableTiming));
        }
        p.ready = true;
    }
    if (p.src_bytes < sb || p.dst_bytes < db) {
        // the pipe is idle here: every pipe_run drains its streams before it returns
        pipe_free_buffers(p);
        for (int i = 0; i < HostPipe::SLOTS; ++i) {
            if (cudaMalloc(&p.src[i], sb) != cudaSuccess || cudaMalloc(&p.dst[i], db) != cudaSuccess) {
                cudaGetLastError();
                pipe_free_buffers(p);
                return fail(B200IMG_ERR_OOM, "cannot allocate the host-batch staging buffers (%zu + %zu bytes x %d)", sb, db, HostPipe::SLOTS);
            }
        }
        p.src_bytes = sb; p.dst_bytes = db;
    }
    *out = &p;
    return B200IMG_OK;
}

// op(d_src, d_dst, stream) enqueues the device work for one image.
template <typename Op>
static int pipe_enqueue(HostPipe *p, const void *hsrc, size_t sb, void *hdst, size_t db, uint32_t n, Op op) {
    int rc;
    for (uint32_t i = 0; i < n; ++i) {
        const int s = (int)(i % HostPipe::SLOTS);
        if (i >= (uint32_t)HostPipe::SLOTS) {
            B200_CUDA(cudaStreamWaitEvent(p->h2d, p->e_run[s], 0));   // the slot's previous kernels have consumed src[s]
            B200_CUDA(cudaStreamWaitEvent(p->exec, p->e_out[s], 0));  // and its previous result has left dst[s]
        }
        B200_CUDA(cudaMemcpyAsync(p->src[s], (const char *)hsrc + (size_t)i * sb, sb, cudaMemcpyHostToDevice, p->h2d));
        B200_CUDA(cudaEventRecord(p->e_in[s], p->h2d));
        B200_CUDA(cudaStreamWaitEvent(p->exec, p->e_in[s], 0));
        if ((rc = op(p->src[s], p->dst[s], p->exec))) return rc;
        B200_CUDA(cudaEventRecord(p->e_run[s], p->exec));
        B200_CUDA(cudaStreamWaitEvent(p->d2h, p->e_run[s], 0));
        B200_CUDA(cudaMemcpyAsync((char *)hdst + (size_t)i * db, p->dst[s], db, cudaMemcpyDeviceToHost, p->d2h));
        B200_CUDA(cudaEventRecord(p->e_out[s], p->d2h));
    }
    return B200IMG_OK;
}

template <typename Op>
static int pipe_run(const void *hsrc, size_t sb, void *hdst, size_t db, uint32_t n, Op op) {
    std::lock_guard<std::mutex> lk(host_mutex_of_current_device());
    HostPipe *p;
    int rc = pipe_get(sb, db, &p);
    if (rc) return rc;
    ScratchScope scope(p->exec);
    rc = pipe_enqueue(p, hsrc, sb, hdst, db, n, op);
    // Drain all three streams whatever happened: copies queued for earlier images still read `hsrc` and
    // write `hdst`, and the caller may free or reuse both as soon as this call has returned.
    const cudaError_t e1 = cudaStreamSynchronize(p->h2d), e2 = cudaStreamSynchronize(p->exec), e3 = cudaStreamSynchronize(p->d2h);
    if (rc) return rc;
    if (e1 != cudaSuccess) return cuda_fail(e1, "host batch: upload stream");
    if (e2 != cudaSuccess) return cuda_fail(e2, "host batch: kernel stream");
    if (e3 != cudaSuccess) return cuda_fail(e3, "host batch: download stream");
    return B200IMG_OK;
}

static int check_image(const char *op, int channels, int dtype, uint32_t w, uint32_t h, size_t *bytes) {
    if (channels < 1 || channels > 4) return fail(B200IMG_ERR_BAD_ARG, "%s: channels must be 1..4 (got %d)", op, channels);
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "%s: bad dtype %d", op, dtype);
    if (!checked_mul(w, h, (size_t)channels, dtype_size(dtype), bytes))
        return fail(B200IMG_ERR_DIMENSION, "%s: image size overflows", op);
    return B200IMG_OK;
}

// imageops::resize front end (src/imageops/sample.rs:970-991): the two early-outs
static int resize_front(const void *src, size_t src_stride, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                        size_t dst_stride, uint32_t nw, uint32_t nh, uint32_t n, int filter, float acc_init,
                        int flags, cudaStream_t st) {
    size_t sb, db;
    int rc = check_image("resize", C, dtype, w, h, &sb);
    if (rc) return rc;
    rc = check_image("resize", C, dtype, nw, nh, &db);
    if (rc) return rc;
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize: unknown filter %d", filter);
    if (n == 0 || db == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "resize: null buffer");
    if (!src_stride) src_stride = sb;
    if (!dst_stride) dst_stride = db;
    if (w == 0 || h == 0) {  // nothing to sample from: zero-filled buffer
        B200_CUDA(cudaMemset2DAsync(dst, dst_stride, 0, db, n, st));
        return B200IMG_OK;
    }
    if (nw == w && nh == h) {  // same dimensions: a copy
        B200_CUDA(cudaMemcpy2DAsync(dst, dst_stride, src, src_stride, sb, n, cudaMemcpyDeviceToDevice, st));
        return B200IMG_OK;
    }
    return resize_launch(src, src_stride, 0, h, w, h, C, dtype, dst, dst_stride, 0, nh, nw, nh, n, filter, acc_init, flags, st);
}

}  // namespace b200

using namespace b200;

extern "C" {

const char *b200img_last_error(void) { return g_err; }
int b200img_version(void) { return 100; }

int b200img_device_count(int *count) {
    if (!count) return fail(B200IMG_ERR_BAD_ARG, "device_count: null pointer");
    B200_CUDA(cudaGetDeviceCount(count));
    return B200IMG_OK;
}

int b200img_pinned_alloc(size_t bytes, void **out) {
    if (!out) return fail(B200IMG_ERR_BAD_ARG, "pinned_alloc: null pointer");
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(B200IMG_ERR_OOM, "pinned_alloc: %s", cudaGetErrorString(e)); }
    return B200IMG_OK;
}
int b200img_pinned_free(void *p) {
    if (p) B200_CUDA(cudaFreeHost(p));
    return B200IMG_OK;
}

int b200img_release_caches(void) {
    int dev;
    B200_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(host_mutex_of_current_device());  // no host-pointer call is using the pipe
    release_scratch_leases();
    B200_CUDA(cudaDeviceSynchronize());
    release_scratch(dev);
    release_pipe(dev);
    resize_release_tables(dev);
    return B200IMG_OK;
}

// ---- host tables --------------------------------------------------------------
float b200img_filter_support(int filter) { return filter_support(filter); }
float b200img_filter_kernel(int filter, float x) { return filter_kernel(filter, x); }

int b200img_resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len) {
    if (filter_support(filter) < 0 || dst_len == 0) return fail(B200IMG_ERR_BAD_ARG, "resize_max_taps: bad argument");
    return resize_max_taps(filter, src_len, dst_len);
}

int b200img_build_weights(int filter, uint32_t src_len, uint32_t dst_len, int axis, int32_t *left, int32_t *count,
                          float *weights, int stride) {
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "build_weights: unknown filter %d", filter);
    if (src_len == 0) return fail(B200IMG_ERR_BAD_ARG, "build_weights: empty source axis");
    if (axis != 0 && axis != 1) return fail(B200IMG_ERR_BAD_ARG, "build_weights: axis must be 0 (vertical) or 1 (horizontal)");
    if (dst_len == 0) return B200IMG_OK;
    if (!left || !count || !weights) return fail(B200IMG_ERR_BAD_ARG, "build_weights: null pointer");
    AxisTable t;
    build_axis_table(filter, src_len, dst_len, axis, t);
    if (t.max_count > stride) return fail(B200IMG_ERR_BAD_ARG, "build_weights: stride %d < %d taps", stride, t.max_count);
    for (uint32_t o = 0; o < dst_len; ++o) {
        left[o] = t.left[o];
        count[o] = t.count[o];
        for (int i = 0; i < stride; ++i) weights[(size_t)o * stride + i] = i < t.count[o] ? t.w[(size_t)o * t.stride + i] : 0.0f;
    }
    return B200IMG_OK;
}

int b200img_resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, int fill, uint32_t *ow, uint32_t *oh) {
    if (!ow || !oh) return fail(B200IMG_ERR_BAD_ARG, "resize_dimensions: null pointer");
    resize_dimensions(w, h, nw, nh, fill != 0, ow, oh);
    return B200IMG_OK;
}

#define PARAMS_FN(name, call)                                                                          \
    {                                                                                                  \
        if (!out) return fail(B200IMG_ERR_BAD_ARG, name ": null pointer");                             \
        int rc = call;                                                                                 \
        if (rc) return fail(rc, name ": invalid argument (the reference asserts here)");               \
        return B200IMG_OK;                                                                             \
    }
int b200img_gauss_params_from_radius(float r, b200img_gauss_params *out) PARAMS_FN("gauss_params_from_radius", gauss_params_from_radius(r, out))
int b200img_gauss_params_from_kernel_size(float k, b200img_gauss_params *out) PARAMS_FN("gauss_params_from_kernel_size", gauss_params_from_kernel_size(k, out))
int b200img_gauss_params_from_anisotropic(float x, float y, b200img_gauss_params *out) PARAMS_FN("gauss_params_from_anisotropic", gauss_params_from_anisotropic(x, y, out))
int b200img_gauss_params_from_sigma(float s, b200img_gauss_params *out) PARAMS_FN("gauss_params_from_sigma", gauss_params_from_sigma(s, out))
#undef PARAMS_FN

int b200img_gaussian_kernel_1d(uint32_t width, float sigma, float *out) {
    if (!out && width) return fail(B200IMG_ERR_BAD_ARG, "gaussian_kernel_1d: null pointer");
    std::vector<float> k = gaussian_kernel_1d(width, sigma);
    memcpy(out, k.data(), k.size() * sizeof(float));
    return B200IMG_OK;
}

int b200img_boxes_for_gauss(float sigma, uint32_t out[3]) {
    if (!out) return fail(B200IMG_ERR_BAD_ARG, "boxes_for_gauss: null pointer");
    boxes_for_gauss(sigma, out);
    return B200IMG_OK;
}

// ---- resampler ----------------------------------------------------------------
int b200img_resize_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                       int filter, float acc_init, int flags, void *stream) {
    ScratchScope scope((cudaStream_t)stream);
    return resize_front(src, 0, w, h, C, dtype, dst, 0, nw, nh, 1, filter, acc_init, flags, (cudaStream_t)stream);
}

int b200img_resize_batch_dev(const void *src, size_t src_stride, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                             size_t dst_stride, uint32_t nw, uint32_t nh, uint32_t n, int filter, float acc_init,
                             int flags, void *stream) {
    ScratchScope scope((cudaStream_t)stream);
    return resize_front(src, src_stride, w, h, C, dtype, dst, dst_stride, nw, nh, n, filter, acc_init, flags, (cudaStream_t)stream);
}

int b200img_resize(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                   int filter, float acc_init, int flags) {
    size_t sb, db;
    int rc = check_image("resize", C, dtype, w, h, &sb);
    if (rc) return rc;
    rc = check_image("resize", C, dtype, nw, nh, &db);
    if (rc) return rc;
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize: unknown filter %d", filter);
    if (db == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "resize: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, sb, &ds))) return rc;
    if ((rc = hc.out_buf(db, &dd))) return rc;
    if ((rc = resize_front(ds, 0, w, h, C, dtype, dd, 0, nw, nh, 1, filter, acc_init, flags, 0))) return rc;
    return hc.out(dst, dd, db);
}

int b200img_resize_band_src_rows(uint32_t h, uint32_t nh, int filter, uint32_t dst_row0, uint32_t dst_rows,
                                 uint32_t *src_row0, uint32_t *src_rows) {
    if (!src_row0 || !src_rows) return fail(B200IMG_ERR_BAD_ARG, "resize_band_src_rows: null pointer");
    return resize_band_src_rows(h, nh, filter, dst_row0, dst_rows, src_row0, src_rows);
}

int b200img_resize_band_dev(const void *src_band, uint32_t src_row0, uint32_t src_rows, uint32_t w, uint32_t h, int C,
                            int dtype, void *dst_band, uint32_t dst_row0, uint32_t dst_rows, uint32_t nw, uint32_t nh,
                            int filter, float acc_init, int flags, void *stream) {
    size_t sb, db;
    int rc = check_image("resize_band", C, dtype, w, h, &sb);
    if (rc) return rc;
    rc = check_image("resize_band", C, dtype, nw, nh, &db);
    if (rc) return rc;
    if (w == 0 || h == 0 || nw == 0 || nh == 0) return fail(B200IMG_ERR_BAD_ARG, "resize_band: empty image");
    if (dst_rows == 0) return B200IMG_OK;
    if (dst_row0 + (uint64_t)dst_rows > nh || src_row0 + (uint64_t)src_rows > h)
        return fail(B200IMG_ERR_BAD_ARG, "resize_band: band outside the image");
    if (!src_band || !dst_band) return fail(B200IMG_ERR_BAD_ARG, "resize_band: null buffer");
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize_band: unknown filter %d", filter);
    if (nw == w && nh == h) {  // same dimensions: imageops::resize copies (sample.rs:987-991), so does every band of it
        if (dst_row0 < src_row0 || dst_row0 + (uint64_t)dst_rows > src_row0 + (uint64_t)src_rows)
            return fail(B200IMG_ERR_BAD_ARG, "resize band: source band does not cover the rows the destination band copies");
        const size_t row_bytes = sb / h;
        B200_CUDA(cudaMemcpyAsync(dst_band, (const char *)src_band + (size_t)(dst_row0 - src_row0) * row_bytes,
                                  (size_t)dst_rows * row_bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        return B200IMG_OK;
    }
    ScratchScope scope((cudaStream_t)stream);
    return resize_launch(src_band, 0, src_row0, src_rows, w, h, C, dtype, dst_band, 0, dst_row0, dst_rows, nw, nh, 1,
                         filter, acc_init, flags, (cudaStream_t)stream);
}

// ---- Gaussian blur ------------------------------------------------------------
static int gauss_check(const b200img_gauss_params *p, int path) {
    if (!p) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: null params");
    if (path != B200IMG_BLUR_GENERIC && path != B200IMG_BLUR_DYN) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: bad path %d", path);
    // filter_2d_separable asserts odd lengths (src/imageops/filter_1d.rs:649-654)
    if ((p->x_kernel_size & 1u) == 0 || (p->y_kernel_size & 1u) == 0)
        return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: kernel sizes must be odd (%u, %u)", p->x_kernel_size, p->y_kernel_size);
    return B200IMG_OK;
}

int b200img_gaussian_blur_batch_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n,
                                    const b200img_gauss_params *p, int path, int flags, void *stream) {
    size_t bytes;
    int rc = check_image("gaussian_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if ((rc = gauss_check(p, path))) return rc;
    if (bytes == 0 || n == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: null buffer");
    ScratchScope scope((cudaStream_t)stream);
    return gaussian_blur_launch(src, w, h, C, dtype, dst, n, p, path, flags, (cudaStream_t)stream);
}

int b200img_gaussian_blur_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                              const b200img_gauss_params *p, int path, int flags, void *stream) {
    return b200img_gaussian_blur_batch_dev(src, w, h, C, dtype, dst, 1, p, path, flags, stream);
}

int b200img_gaussian_blur(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                          const b200img_gauss_params *p, int path, int flags) {
    size_t bytes;
    int rc = check_image("gaussian_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if ((rc = gauss_check(p, path))) return rc;
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    if ((rc = hc.out_buf(bytes, &dd))) return rc;
    if ((rc = gaussian_blur_launch(ds, w, h, C, dtype, dd, 1, p, path, flags, 0))) return rc;
    return hc.out(dst, dd, bytes);
}

int b200img_gaussian_blur_batch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n,
                                const b200img_gauss_params *p, int path, int flags) {
    size_t bytes;
    int rc = check_image("gaussian_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if ((rc = gauss_check(p, path))) return rc;
    if (bytes == 0 || n == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "gaussian_blur: null buffer");
    return pipe_run(src, bytes, dst, bytes, n, [&](void *ds, void *dd, cudaStream_t st) {
        return gaussian_blur_launch(ds, w, h, C, dtype, dd, 1, p, path, flags, st);
    });
}

int b200img_fast_blur_batch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n, float sigma) {
    size_t bytes;
    int rc = check_image("fast_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0 || n == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "fast_blur: null buffer");
    return pipe_run(src, bytes, dst, bytes, n, [&](void *ds, void *dd, cudaStream_t st) {
        return fast_blur_launch(ds, w, h, C, dtype, dd, 1, sigma, st);
    });
}

int b200img_resize_batch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                         uint32_t n, int filter, float acc_init, int flags) {
    size_t sb, db;
    int rc = check_image("resize", C, dtype, w, h, &sb);
    if (rc) return rc;
    if ((rc = check_image("resize", C, dtype, nw, nh, &db))) return rc;
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "resize: unknown filter %d", filter);
    if (db == 0 || n == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "resize: null buffer");
    return pipe_run(src, sb ? sb : 1, dst, db, n, [&](void *ds, void *dd, cudaStream_t st) {
        return resize_front(ds, 0, w, h, C, dtype, dd, 0, nw, nh, 1, filter, acc_init, flags, st);
    });
}

// ---- fast_blur ----------------------------------------------------------------
int b200img_fast_blur_batch_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t n,
                                float sigma, void *stream) {
    size_t bytes;
    int rc = check_image("fast_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0 || n == 0) return B200IMG_OK;  // empty image: clone (fast_blur.rs:117-119)
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "fast_blur: null buffer");
    ScratchScope scope((cudaStream_t)stream);
    return fast_blur_launch(src, w, h, C, dtype, dst, n, sigma, (cudaStream_t)stream);
}

int b200img_fast_blur_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float sigma, void *stream) {
    return b200img_fast_blur_batch_dev(src, w, h, C, dtype, dst, 1, sigma, stream);
}

int b200img_fast_blur(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float sigma) {
    size_t bytes;
    int rc = check_image("fast_blur", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "fast_blur: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    if ((rc = hc.out_buf(bytes, &dd))) return rc;
    if ((rc = fast_blur_launch(ds, w, h, C, dtype, dd, 1, sigma, 0))) return rc;
    return hc.out(dst, dd, bytes);
}

// ---- thumbnail ------------------------------------------------------------
static int thumbnail_front(uint32_t w, uint32_t h, int C, int dtype, uint32_t nw, uint32_t nh, size_t *sb, size_t *db) {
    int rc = check_image("thumbnail", C, dtype, w, h, sb);
    if (rc) return rc;
    if ((rc = check_image("thumbnail", C, dtype, nw, nh, db))) return rc;
    return thumbnail_check(w, h, dtype, nw, nh);
}

int b200img_thumbnail_batch_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                                uint32_t n_images, int flags, void *stream) {
    size_t sb, db;
    int rc = thumbnail_front(w, h, C, dtype, nw, nh, &sb, &db);
    if (rc) return rc;
    if (db == 0 || n_images == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "thumbnail: null buffer");
    return thumbnail_launch(src, w, h, C, dtype, dst, nw, nh, n_images, (flags & B200IMG_THUMBNAIL_PER_PIXEL) ? 1 : 0,
                            (cudaStream_t)stream);
}

int b200img_thumbnail_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                          int flags, void *stream) {
    return b200img_thumbnail_batch_dev(src, w, h, C, dtype, dst, nw, nh, 1, flags, stream);
}

int b200img_thumbnail(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh, int flags) {
    size_t sb, db;
    int rc = thumbnail_front(w, h, C, dtype, nw, nh, &sb, &db);
    if (rc) return rc;
    if (db == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "thumbnail: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, sb, &ds))) return rc;
    if ((rc = hc.out_buf(db, &dd))) return rc;
    if ((rc = thumbnail_launch(ds, w, h, C, dtype, dd, nw, nh, 1, (flags & B200IMG_THUMBNAIL_PER_PIXEL) ? 1 : 0, 0))) return rc;
    return hc.out(dst, dd, db);
}

// ---- filter3x3 / unsharpen ---------------------------------------------
int b200img_filter3x3_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, const float kernel[9], void *stream) {
    size_t bytes;
    int rc = check_image("filter3x3", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (!kernel) return fail(B200IMG_ERR_BAD_ARG, "filter3x3: null kernel");
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "filter3x3: null buffer");
    return filter3x3_launch(src, w, h, C, dtype, dst, 1, kernel, (cudaStream_t)stream);
}

int b200img_filter3x3(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, const float kernel[9]) {
    size_t bytes;
    int rc = check_image("filter3x3", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (!kernel) return fail(B200IMG_ERR_BAD_ARG, "filter3x3: null kernel");
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "filter3x3: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    if ((rc = hc.out_buf(bytes, &dd))) return rc;
    if ((rc = filter3x3_launch(ds, w, h, C, dtype, dd, 1, kernel, 0))) return rc;
    return hc.out(dst, dd, bytes);
}

int b200img_unsharpen_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float sigma, int threshold,
                          int flags, void *stream) {
    size_t bytes;
    int rc = check_image("unsharpen", C, dtype, w, h, &bytes);
    if (rc) return rc;
    b200img_gauss_params p;
    if (gauss_params_from_sigma(sigma, &p)) return fail(B200IMG_ERR_BAD_ARG, "unsharpen: sigma must be non-negative and finite");
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "unsharpen: null buffer");
    ScratchScope scope((cudaStream_t)stream);
    return unsharpen_launch(src, w, h, C, dtype, dst, 1, sigma, threshold, flags, (cudaStream_t)stream);
}

int b200img_unsharpen(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float sigma, int threshold, int flags) {
    size_t bytes;
    int rc = check_image("unsharpen", C, dtype, w, h, &bytes);
    if (rc) return rc;
    b200img_gauss_params p;
    if (gauss_params_from_sigma(sigma, &p)) return fail(B200IMG_ERR_BAD_ARG, "unsharpen: sigma must be non-negative and finite");
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "unsharpen: null buffer");
    HostCall hc;
    void *ds, *dd;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    if ((rc = hc.out_buf(bytes, &dd))) return rc;
    if ((rc = unsharpen_launch(ds, w, h, C, dtype, dd, 1, sigma, threshold, flags, 0))) return rc;
    return hc.out(dst, dd, bytes);
}

// ---- colorops --------------------------------------------------------
static int colorop_dev(const char *name, int op, const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                       int32_t ivalue, float fvalue, void *stream) {
    size_t bytes;
    int rc = check_image(name, C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "%s: null buffer", name);
    return colorop_launch(op, src, dst, (size_t)w * h, C, dtype, ivalue, fvalue, (cudaStream_t)stream);
}

static int colorop_host(const char *name, int op, const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst,
                        int32_t ivalue, float fvalue) {
    size_t bytes;
    int rc = check_image(name, C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "%s: null buffer", name);
    HostCall hc;
    void *ds;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    if ((rc = colorop_launch(op, ds, ds, (size_t)w * h, C, dtype, ivalue, fvalue, 0))) return rc;  // in place on the device copy
    return hc.out(dst, ds, bytes);
}

int b200img_invert(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst) {
    return colorop_host("invert", 0, src, w, h, C, dtype, dst, 0, 0.0f);
}
int b200img_invert_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, void *stream) {
    return colorop_dev("invert", 0, src, w, h, C, dtype, dst, 0, 0.0f, stream);
}
int b200img_brighten(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int32_t value) {
    return colorop_host("brighten", 1, src, w, h, C, dtype, dst, value, 0.0f);
}
int b200img_brighten_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int32_t value, void *stream) {
    return colorop_dev("brighten", 1, src, w, h, C, dtype, dst, value, 0.0f, stream);
}
int b200img_contrast(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float contrast) {
    return colorop_host("contrast", 2, src, w, h, C, dtype, dst, 0, contrast);
}
int b200img_contrast_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, float contrast, void *stream) {
    return colorop_dev("contrast", 2, src, w, h, C, dtype, dst, 0, contrast, stream);
}
int b200img_huerotate(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int32_t degrees) {
    return colorop_host("huerotate", 3, src, w, h, C, dtype, dst, degrees, 0.0f);
}
int b200img_huerotate_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int32_t degrees, void *stream) {
    return colorop_dev("huerotate", 3, src, w, h, C, dtype, dst, degrees, 0.0f, stream);
}
// grayscale / grayscale_alpha: pixel.to_luma() / to_luma_alpha() = FromColor in the source subpixel type
int b200img_grayscale(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int keep_alpha) {
    size_t bytes;
    int rc = check_image("grayscale", C, dtype, w, h, &bytes);
    if (rc) return rc;
    return b200img_convert(src, dtype, C, dst, dtype, keep_alpha ? 2 : 1, (size_t)w * h, B200IMG_CONVERT, nullptr);
}
int b200img_grayscale_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, int keep_alpha, void *stream) {
    size_t bytes;
    int rc = check_image("grayscale", C, dtype, w, h, &bytes);
    if (rc) return rc;
    return b200img_convert_dev(src, dtype, C, dst, dtype, keep_alpha ? 2 : 1, (size_t)w * h, B200IMG_CONVERT, nullptr, stream);
}

// ---- either side of DynamicImage::resize (src/imageops/resize.rs) ------------
static int constant_alpha_on(const void *d_src, size_t npix, int C, int dtype, int *result, cudaStream_t st) {
    void *flag;
    int rc = scratch(SCRATCH_FLAG, sizeof(int), &flag);
    if (rc) return rc;
    if ((rc = alpha_scan_launch(d_src, npix, C, dtype, (int *)flag, st))) return rc;
    int differs = 0;
    B200_CUDA(cudaMemcpyAsync(&differs, flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    *result = differs ? 0 : 1;
    return B200IMG_OK;
}

int b200img_has_constant_alpha_dev(const void *src, uint32_t w, uint32_t h, int C, int dtype, int *result, void *stream) {
    size_t bytes;
    int rc = check_image("has_constant_alpha", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (!result) return fail(B200IMG_ERR_BAD_ARG, "has_constant_alpha: null result");
    *result = 1;
    if (bytes == 0 || (C != 2 && C != 4)) return B200IMG_OK;  // resize.rs:155-171: no alpha channel / empty image
    if (!src) return fail(B200IMG_ERR_BAD_ARG, "has_constant_alpha: null buffer");
    ScratchScope scope((cudaStream_t)stream);
    return constant_alpha_on(src, (size_t)w * h, C, dtype, result, (cudaStream_t)stream);
}

int b200img_has_constant_alpha(const void *src, uint32_t w, uint32_t h, int C, int dtype, int *result) {
    size_t bytes;
    int rc = check_image("has_constant_alpha", C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (!result) return fail(B200IMG_ERR_BAD_ARG, "has_constant_alpha: null result");
    *result = 1;
    if (bytes == 0 || (C != 2 && C != 4)) return B200IMG_OK;
    if (!src) return fail(B200IMG_ERR_BAD_ARG, "has_constant_alpha: null buffer");
    HostCall hc;
    void *ds;
    if ((rc = hc.in(src, bytes, &ds))) return rc;
    return constant_alpha_on(ds, (size_t)w * h, C, dtype, result, 0);
}

static int alpha_mul_dev(const char *name, void *buf, uint32_t w, uint32_t h, int C, int dtype, bool un, void *stream) {
    size_t bytes;
    int rc = check_image(name, C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0 || (C != 2 && C != 4)) return B200IMG_OK;
    if (!buf) return fail(B200IMG_ERR_BAD_ARG, "%s: null buffer", name);
    return alpha_mul_launch(buf, (size_t)w * h, C, dtype, un, (cudaStream_t)stream);
}

static int alpha_mul_host(const char *name, void *buf, uint32_t w, uint32_t h, int C, int dtype, bool un) {
    size_t bytes;
    int rc = check_image(name, C, dtype, w, h, &bytes);
    if (rc) return rc;
    if (bytes == 0 || (C != 2 && C != 4)) return B200IMG_OK;
    if (!buf) return fail(B200IMG_ERR_BAD_ARG, "%s: null buffer", name);
    HostCall hc;
    void *d;
    if ((rc = hc.in(buf, bytes, &d))) return rc;
    if ((rc = alpha_mul_launch(d, (size_t)w * h, C, dtype, un, 0))) return rc;
    return hc.out(buf, d, bytes);
}

int b200img_premultiply_alpha(void *buf, uint32_t w, uint32_t h, int C, int dtype) {
    return alpha_mul_host("premultiply_alpha", buf, w, h, C, dtype, false);
}
int b200img_premultiply_alpha_dev(void *buf, uint32_t w, uint32_t h, int C, int dtype, void *stream) {
    return alpha_mul_dev("premultiply_alpha", buf, w, h, C, dtype, false, stream);
}
int b200img_unpremultiply_alpha(void *buf, uint32_t w, uint32_t h, int C, int dtype) {
    return alpha_mul_host("unpremultiply_alpha", buf, w, h, C, dtype, true);
}
int b200img_unpremultiply_alpha_dev(void *buf, uint32_t w, uint32_t h, int C, int dtype, void *stream) {
    return alpha_mul_dev("unpremultiply_alpha", buf, w, h, C, dtype, true, stream);
}

// decode -> device hand-off: one upload, DynamicImage::to_* on the device, imageops::resize, one download
int b200img_convert_resize(const void *src, int src_dtype, int src_channels, uint32_t w, uint32_t h, void *dst, int dst_dtype,
                           int dst_channels, uint32_t nw, uint32_t nh, int filter, float acc_init, int flags) {
    size_t sb, mb, db;
    int rc = check_image("convert_resize", src_channels, src_dtype, w, h, &sb);
    if (rc) return rc;
    if ((rc = check_image("convert_resize", dst_channels, dst_dtype, w, h, &mb))) return rc;
    if ((rc = check_image("convert_resize", dst_channels, dst_dtype, nw, nh, &db))) return rc;
    if (filter_support(filter) < 0) return fail(B200IMG_ERR_BAD_ARG, "convert_resize: unknown filter %d", filter);
    if (db == 0) return B200IMG_OK;
    if (!dst || (sb && !src)) return fail(B200IMG_ERR_BAD_ARG, "convert_resize: null buffer");
    HostCall hc;
    void *ds, *dm, *dd;
    if ((rc = hc.in(src, sb, &ds))) return rc;
    if ((rc = scratch(SCRATCH_MID, mb ? mb : 1, &dm))) return rc;
    if ((rc = hc.out_buf(db, &dd))) return rc;
    if (sb && (rc = convert_dev(ds, src_dtype, src_channels, dm, dst_dtype, dst_channels, (size_t)w * h, 0, nullptr, 0, 0))) return rc;
    if ((rc = resize_front(dm, 0, w, h, dst_channels, dst_dtype, dd, 0, nw, nh, 1, filter, acc_init, flags, 0))) return rc;
    return hc.out(dst, dd, db);
}

// ---- conversions and swizzle ----------------------------------------------------
static int convert_check(int sd, int sc, int dd, int dc, size_t n, size_t *sb, size_t *db) {
    if (!valid_dtype(sd) || !valid_dtype(dd)) return fail(B200IMG_ERR_BAD_ARG, "convert: bad dtype");
    if (sc < 1 || sc > 4 || dc < 1 || dc > 4) return fail(B200IMG_ERR_BAD_ARG, "convert: channels must be 1..4");
    if (!checked_mul(n, (size_t)sc, dtype_size(sd), 1, sb) || !checked_mul(n, (size_t)dc, dtype_size(dd), 1, db))
        return fail(B200IMG_ERR_DIMENSION, "convert: buffer size overflows");
    return B200IMG_OK;
}

int b200img_convert_dev(const void *src, int sd, int sc, void *dst, int dd, int dc, size_t n, int mode,
                        const float coefs[3], void *stream) {
    size_t sb, db;
    int rc = convert_check(sd, sc, dd, dc, n, &sb, &db);
    if (rc) return rc;
    return convert_dev(src, sd, sc, dst, dd, dc, n, mode & 1, coefs, (mode & 0x100) ? 1 : 0, (cudaStream_t)stream);
}

int b200img_convert(const void *src, int sd, int sc, void *dst, int dd, int dc, size_t n, int mode, const float coefs[3]) {
    size_t sb, db;
    int rc = convert_check(sd, sc, dd, dc, n, &sb, &db);
    if (rc) return rc;
    if (n == 0) return B200IMG_OK;
    if (!src || !dst) return fail(B200IMG_ERR_BAD_ARG, "convert: null buffer");
    HostCall hc;
    void *ds, *ddv;
    if ((rc = hc.in(src, sb, &ds))) return rc;
    if ((rc = hc.out_buf(db, &ddv))) return rc;
    if ((rc = convert_dev(ds, sd, sc, ddv, dd, dc, n, mode & 1, coefs, (mode & 0x100) ? 1 : 0, 0))) return rc;
    return hc.out(dst, ddv, db);
}

int b200img_swizzle_02_dev(void *buf, int dtype, int channels, size_t nsub, void *stream) {
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "swizzle: bad dtype %d", dtype);
    return swizzle_dev(buf, dtype, channels, nsub, (cudaStream_t)stream);
}

int b200img_swizzle_02(void *buf, int dtype, int channels, size_t nsub) {
    if (!valid_dtype(dtype)) return fail(B200IMG_ERR_BAD_ARG, "swizzle: bad dtype %d", dtype);
    if (channels != 3 && channels != 4) return fail(B200IMG_ERR_BAD_ARG, "swizzle: channels must be 3 or 4");
    size_t bytes;
    if (!checked_mul(nsub, dtype_size(dtype), 1, 1, &bytes)) return fail(B200IMG_ERR_DIMENSION, "swizzle: buffer size overflows");
    if (nsub < (size_t)channels) return B200IMG_OK;
    if (!buf) return fail(B200IMG_ERR_BAD_ARG, "swizzle: null buffer");
    HostCall hc;
    void *d;
    int rc;
    if ((rc = hc.in(buf, bytes, &d))) return rc;
    if ((rc = swizzle_dev(d, dtype, channels, nsub, 0))) return rc;
    return hc.out(buf, d, bytes);
}

}  // extern "C"
