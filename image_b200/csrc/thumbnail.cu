// thumbnail.cu -- imageops::thumbnail (src/imageops/sample.rs:564-886), SURVEY.md section 8f rank 1.
//
// Every output pixel owns the source window [left, right) x [bottom, top) that the reference derives
// in f32 (sample.rs:611-626) and takes one of four forms:
//   block      (both extents >= 1)  integer average with round-half-up, sums in S::Larger
//                                   (u32 for u8/u16, f64 for f32); windows of more than 2^16 pixels
//                                   are averaged per 256x256 sub-block first and then divided by the
//                                   reference's literal count (w/256)*(h/256)            :680-742
//   fraction_horizontal / _vertical / _both   2-column, 2-row or 2x2 interpolation in f32   :745-886
// The span arithmetic is the same IEEE f32 sequence as the reference (mul, add, ceil, saturating
// cast; the library is built without FP contraction), so windows agree exactly.
//
// Kernels:
//   thumb_cols_kernel   pure downscale, word-aligned rows: one CTA per (x segment, output row);
//                       threads stream the source rows as 32-bit words, keep column sums in
//                       registers, park them in shared memory, then one thread per output
//                       subpixel adds its columns.  The source is read exactly once, coalesced:
//                       HBM-bound at (src + dst) bytes.
//   thumb_pixel_kernel  everything else, literal per-pixel evaluation: one thread per output pixel
//                       when windows are small (upscales), one warp per output pixel when large.
#include "common.cuh"

#include <set>

namespace b200 {
namespace {

struct Span { uint32_t lo, hi; float lof, hif; };

// sample.rs:615-626 for one axis (o = output index, len = source length >= 1)
__host__ __device__ __forceinline__ Span thumb_span(uint32_t o, float ratio, uint32_t len) {
    Span s;
#ifdef __CUDA_ARCH__
    s.lof = __fmul_rn((float)o, ratio);
    s.hif = __fadd_rn(s.lof, ratio);
    uint32_t a = __float2uint_rz(ceilf(s.lof)), b = __float2uint_rz(ceilf(s.hif));  // saturating, NaN -> 0
#else
    volatile float lof = (float)o * ratio;  // volatile: no contraction whatever the host flags are
    volatile float hif = lof + ratio;
    s.lof = lof; s.hif = hif;
    auto sat = [](float v) -> uint32_t { return !(v > 0.0f) ? 0u : (v >= 4294967296.0f ? 0xFFFFFFFFu : (uint32_t)v); };
    uint32_t a = sat(ceilf(s.lof)), b = sat(ceilf(s.hif));
#endif
    s.lo = a > len - 1 ? len - 1 : a;
    s.hi = b < s.lo ? s.lo : (b > len ? len : b);
    return s;
}

template <typename T> struct Lg { typedef uint32_t type; };
template <> struct Lg<float> { typedef double type; };

// Enlargeable::clamp_from (src/traits.rs:79-87)
__device__ __forceinline__ uint8_t clamp_from(uint32_t n, uint8_t) { return (uint8_t)min(n, 255u); }
__device__ __forceinline__ uint16_t clamp_from(uint32_t n, uint16_t) { return (uint16_t)min(n, 65535u); }
__device__ __forceinline__ float clamp_from(double n, float) {
    const double mx = 3.4028234663852886e38;
    if (n > mx) return 3.4028234663852886e38f;
    if (n < -mx) return -3.4028234663852886e38f;
    return (float)n;  // cvt.rn.f32.f64; NaN stays NaN
}

// (sum + n/2) / n for the integer types, sum / n for f32 (sample.rs:729-739)
__device__ __forceinline__ uint32_t avg_of(uint32_t sum, uint64_t count) {
    const uint32_t n = (uint32_t)count;
    return n ? (sum + n / 2) / n : 0u;  // n == 0 is refused on the host (the reference panics)
}
__device__ __forceinline__ double avg_of(double sum, uint64_t count) { return sum / (double)count; }

// The same averages for many sums over one count (<= 2^16, sums + count/2 < 2^32): the quotient
// estimate from one f32 reciprocal is within 1 of the truth (relative error < 2^-22 on a
// quotient < 2^16), and one remainder check on each side makes it exact.
template <typename L> struct Rounder;
template <> struct Rounder<uint32_t> {
    uint32_t n, half;
    float rcp;
    __device__ __forceinline__ explicit Rounder(uint32_t count) : n(count), half(count / 2), rcp(1.0f / (float)count) {}
    __device__ __forceinline__ uint32_t avg(uint32_t sum) const {
        const uint32_t v = sum + half;
        uint32_t q = __float2uint_rz((float)v * rcp);
        int32_t r = (int32_t)(v - q * n);  // |true remainder error| < 2n, fits
        if (r < 0) { --q; r += (int32_t)n; }
        if (r >= (int32_t)n) ++q;
        return q;
    }
};
template <> struct Rounder<double> {
    double n;
    __device__ __forceinline__ explicit Rounder(uint32_t count) : n((double)count) {}
    __device__ __forceinline__ double avg(double sum) const { return sum / n; }
};

__device__ __forceinline__ float fract(float v) { return __fsub_rn(v, truncf(v)); }

// ---- literal per-pixel evaluation ------------------------------------------------------
// Sum of the window [x0,x1) x [y0,y1) per channel.  WARP = the 32 lanes share the window (lanes
// tile it gx wide, 32/gx high) and every lane ends with the full sum; otherwise one thread walks it
// in the reference's order (rows outer, columns inner), which keeps the f64 sums of f32 images
// bit-identical.
template <typename T, bool WARP>
__device__ __forceinline__ void window_sum(const T *__restrict__ img, uint32_t w, int C, uint32_t x0, uint32_t x1, uint32_t y0,
                                           uint32_t y1, unsigned lane, typename Lg<T>::type s[4]) {
    typedef typename Lg<T>::type L;
    uint32_t gx = 1, gy = 1, lx = 0, ly = 0;
    if (WARP) {
        const uint32_t bw = x1 - x0;
        gx = bw >= 32 ? 32u : (1u << (31 - __clz((int)max(bw, 1u))));
        gy = 32 / gx;
        lx = lane % gx;
        ly = lane / gx;
    }
    for (uint32_t y = y0 + ly; y < y1; y += gy) {
        const T *row = img + (size_t)y * w * C;
        for (uint32_t x = x0 + lx; x < x1; x += gx) {
            const T *p = row + (size_t)x * C;
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (c < C) s[c] += (L)p[c];
        }
        if (y + gy < y) break;  // y1 close to 2^32
    }
    if (WARP) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (c < C)
#pragma unroll
                for (int m = 16; m >= 1; m >>= 1) s[c] += __shfl_xor_sync(0xFFFFFFFFu, s[c], m);
    }
}

template <typename T, bool WARP>
__device__ __forceinline__ void thumb_pixel(const T *__restrict__ img, uint32_t w, int C, const Span xs, const Span ys,
                                            unsigned lane, T res[4]) {
    typedef typename Lg<T>::type L;
    L s[4] = {0, 0, 0, 0};
    if (ys.lo != ys.hi && xs.lo != xs.hi) {
        // thumbnail_sample_block, sample.rs:680-742
        uint64_t count = (uint64_t)(xs.hi - xs.lo) * (uint64_t)(ys.hi - ys.lo);
        if (count > 256 * 256) {
            for (uint64_t y = ys.lo; y < ys.hi; y += 256)
                for (uint64_t x = xs.lo; x < xs.hi; x += 256) {
                    const uint32_t xe = (uint32_t)min(x + 256, (uint64_t)xs.hi), ye = (uint32_t)min(y + 256, (uint64_t)ys.hi);
                    L t[4] = {0, 0, 0, 0};
                    window_sum<T, WARP>(img, w, C, (uint32_t)x, xe, (uint32_t)y, ye, lane, t);
                    const uint64_t cnt = (uint64_t)(xe - (uint32_t)x) * (uint64_t)(ye - (uint32_t)y);
#pragma unroll
                    for (int c = 0; c < 4; ++c) s[c] += (L)clamp_from(avg_of(t[c], cnt), T());
                }
            count = ((uint64_t)(xs.hi - xs.lo) / 256) * ((uint64_t)(ys.hi - ys.lo) / 256);  // the reference's literal count
        } else {
            window_sum<T, WARP>(img, w, C, xs.lo, xs.hi, ys.lo, ys.hi, lane, s);
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) res[c] = clamp_from(avg_of(s[c], count), T());
    } else if (ys.lo != ys.hi) {
        // thumbnail_sample_fraction_horizontal, sample.rs:745-789: columns right-1 and right
        const uint32_t left = xs.hi - 1;
        const float fr = (fract(xs.lof) + fract(xs.hif)) / 2.0f;
        L r[4] = {0, 0, 0, 0};
        window_sum<T, WARP>(img, w, C, left, left + 1, ys.lo, ys.hi, lane, s);
        window_sum<T, WARP>(img, w, C, left + 1, left + 2, ys.lo, ys.hi, lane, r);
        const float n = (float)(ys.hi - ys.lo);
        const float fact_right = fr / n, fact_left = (1.0f - fr) / n;
#pragma unroll
        for (int c = 0; c < 4; ++c) res[c] = Px<T>::nearest(fact_left * (float)s[c] + fact_right * (float)r[c]);
    } else if (xs.lo != xs.hi) {
        // thumbnail_sample_fraction_vertical, sample.rs:792-836: rows top-1 and top
        const uint32_t bottom = ys.hi - 1;
        const float fr = (fract(ys.hif) + fract(ys.lof)) / 2.0f;
        L t[4] = {0, 0, 0, 0};
        window_sum<T, WARP>(img, w, C, xs.lo, xs.hi, bottom, bottom + 1, lane, s);
        window_sum<T, WARP>(img, w, C, xs.lo, xs.hi, bottom + 1, bottom + 2, lane, t);
        const float n = (float)(xs.hi - xs.lo);
        const float fact_top = fr / n, fact_bot = (1.0f - fr) / n;
#pragma unroll
        for (int c = 0; c < 4; ++c) res[c] = Px<T>::nearest(fact_bot * (float)s[c] + fact_top * (float)t[c]);
    } else {
        // thumbnail_sample_fraction_both, sample.rs:839-886 (its two fraction names are crossed at
        // the call site; fv comes from the rows, fh from the columns)
        const uint32_t left = xs.hi - 1, bottom = ys.hi - 1;
        const float fv = (fract(ys.hif) + fract(ys.lof)) / 2.0f;
        const float fh = (fract(xs.lof) + fract(xs.hif)) / 2.0f;
        const float fact_tr = fv * fh, fact_tl = fv * (1.0f - fh), fact_br = (1.0f - fv) * fh, fact_bl = (1.0f - fv) * (1.0f - fh);
        const T *bl = img + ((size_t)bottom * w + left) * C, *tl = bl + (size_t)w * C, *br = bl + C, *tr = tl + C;
#pragma unroll
        for (int c = 0; c < 4; ++c)
            if (c < C)
                res[c] = Px<T>::nearest(fact_br * Px<T>::to_f32(br[c]) + fact_tr * Px<T>::to_f32(tr[c]) +
                                        fact_bl * Px<T>::to_f32(bl[c]) + fact_tl * Px<T>::to_f32(tl[c]));
    }
}

// grid.x = output rows of all images; threads (or warps) stride along the row
template <typename T, bool WARP>
__global__ void __launch_bounds__(256)
thumb_pixel_kernel(const T *__restrict__ src, T *__restrict__ dst, uint32_t w, uint32_t h, int C, uint32_t nw, uint32_t nh,
                   float x_ratio, float y_ratio) {
    const unsigned lane = threadIdx.x & 31;
    const uint32_t oy = blockIdx.x % nh, z = blockIdx.x / nh;
    const T *img = src + (size_t)z * w * h * C;
    T *drow = dst + (size_t)blockIdx.x * nw * C;
    const Span ys = thumb_span(oy, y_ratio, h);
    const uint32_t first = WARP ? threadIdx.x / 32 : threadIdx.x, step = WARP ? 8u : 256u;
    for (uint32_t ox = first; ox < nw; ox += step) {
        T res[4];
        thumb_pixel<T, WARP>(img, w, C, thumb_span(ox, x_ratio, w), ys, WARP ? lane : 0u, res);
        if (!WARP || lane == 0) {
            T *d = drow + (size_t)ox * C;
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (c < C) d[c] = res[c];
        }
    }
}

// ---- pure downscale: column sums --------------------------------------------------------
constexpr int TC_THREADS = 256;
template <typename T> struct ColCfg;
template <> struct ColCfg<uint8_t> { static constexpr int K = 4, EPW = 4; };   // words per thread, elements per word
template <> struct ColCfg<uint16_t> { static constexpr int K = 8, EPW = 2; };
template <> struct ColCfg<float> { static constexpr int K = 8, EPW = 1; };

__device__ __forceinline__ uint32_t ldg_word(const uint32_t *p) {
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}

template <typename T> __device__ __forceinline__ void add_word(typename Lg<T>::type *a, uint32_t v);
template <> __device__ __forceinline__ void add_word<uint8_t>(uint32_t *a, uint32_t v) {
    a[0] += v & 0xFFu; a[1] += (v >> 8) & 0xFFu; a[2] += (v >> 16) & 0xFFu; a[3] += v >> 24;
}
template <> __device__ __forceinline__ void add_word<uint16_t>(uint32_t *a, uint32_t v) { a[0] += v & 0xFFFFu; a[1] += v >> 16; }
template <> __device__ __forceinline__ void add_word<float>(double *a, uint32_t v) { a[0] += (double)__uint_as_float(v); }

// grid (x segments, ceil(nh / rows_per_cta), images).  Needs: both ratios give only block windows
// of <= 2^16 pixels, row pitch and base multiples of 4 bytes, segment source span <= K*256 words
// (host checks all).  The x windows of the segment are computed once per CTA and reused for its
// rows_per_cta output rows.
template <typename T>
__global__ void __launch_bounds__(TC_THREADS)
thumb_cols_kernel(const T *__restrict__ src, T *__restrict__ dst, uint32_t w, uint32_t h, int C, uint32_t nw, uint32_t nh,
                  float x_ratio, float y_ratio, uint32_t seg_n, uint32_t rows_per_cta) {
    typedef typename Lg<T>::type L;
    constexpr int K = ColCfg<T>::K, EPW = ColCfg<T>::EPW;
    __shared__ L colsum[K * TC_THREADS * EPW];
    __shared__ uint32_t xwin[K * TC_THREADS * EPW];  // per output pixel: first column (relative) | extent << 16

    const uint32_t ox0 = blockIdx.x * seg_n, ox1 = min(nw, ox0 + seg_n);
    const uint32_t x0 = thumb_span(ox0, x_ratio, w).lo, x1 = thumb_span(ox1 - 1, x_ratio, w).hi;
    const uint32_t ea = (x0 * (uint32_t)C) & ~(uint32_t)(EPW - 1);        // first element of the first word
    const uint32_t nwords = (x1 * (uint32_t)C - ea + EPW - 1) / EPW;
    const size_t pitch_w = (size_t)w * C / EPW;                          // row pitch in words
    const uint32_t *img = reinterpret_cast<const uint32_t *>(src + (size_t)blockIdx.z * w * h * C + ea);

    for (uint32_t j = threadIdx.x; j < ox1 - ox0; j += TC_THREADS) {
        const Span xs = thumb_span(ox0 + j, x_ratio, w);
        xwin[j] = (xs.lo * (uint32_t)C - ea) | ((xs.hi - xs.lo) << 16);
    }

    const uint32_t oy0 = blockIdx.y * rows_per_cta, oy1 = min(nh, oy0 + rows_per_cta);
    for (uint32_t oy = oy0; oy < oy1; ++oy) {
        const Span ys = thumb_span(oy, y_ratio, h);
        const uint32_t *base = img + (size_t)ys.lo * pitch_w;
        L acc[K][EPW];
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int e = 0; e < EPW; ++e) acc[k][e] = 0;

        constexpr int RU = 4;  // rows in flight per thread
        uint32_t y = ys.lo;
        for (; y + RU <= ys.hi; y += RU, base += RU * pitch_w) {
            uint32_t v[RU][K];
#pragma unroll
            for (int r = 0; r < RU; ++r)
#pragma unroll
                for (int k = 0; k < K; ++k) {
                    const uint32_t i = k * TC_THREADS + threadIdx.x;
                    v[r][k] = i < nwords ? ldg_word(base + r * pitch_w + i) : 0u;
                }
#pragma unroll
            for (int r = 0; r < RU; ++r)
#pragma unroll
                for (int k = 0; k < K; ++k) add_word<T>(acc[k], v[r][k]);
        }
        for (; y < ys.hi; ++y, base += pitch_w) {
            uint32_t v[K];
#pragma unroll
            for (int k = 0; k < K; ++k) {
                const uint32_t i = k * TC_THREADS + threadIdx.x;
                v[k] = i < nwords ? ldg_word(base + i) : 0u;
            }
#pragma unroll
            for (int k = 0; k < K; ++k) add_word<T>(acc[k], v[k]);
        }
#pragma unroll
        for (int k = 0; k < K; ++k)
#pragma unroll
            for (int e = 0; e < EPW; ++e) colsum[(k * TC_THREADS + threadIdx.x) * EPW + e] = acc[k][e];
        __syncthreads();  // also orders xwin on the first row

        T *drow = dst + (((size_t)blockIdx.z * nh + oy) * nw + ox0) * C;
        const uint32_t yext = ys.hi - ys.lo;
        for (uint32_t j = threadIdx.x; j < ox1 - ox0; j += TC_THREADS) {
            const uint32_t t = xwin[j], off = t & 0xFFFFu, n = t >> 16;
            const uint32_t cnt = n * yext;  // <= 2^16 on this path
            L sum[4] = {0, 0, 0, 0};
            for (uint32_t x = 0; x < n; ++x)
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    if (c < C) sum[c] += colsum[off + x * (uint32_t)C + c];
            const Rounder<L> rd(cnt);
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (c < C) drow[j * (uint32_t)C + c] = clamp_from(rd.avg(sum[c]), T());
        }
        __syncthreads();
    }
}

template <typename T>
int launch_t(const T *src, uint32_t w, uint32_t h, int C, T *dst, uint32_t nw, uint32_t nh, uint32_t n_images, bool cols_ok,
             uint32_t seg_n, bool big_windows, float xr, float yr, int sm_count, cudaStream_t st) {
    if (cols_ok) {
        // enough output rows per CTA to stream about 16 source rows
        const uint32_t rows = yr >= 16.0f ? 1u : std::max(1u, std::min(8u, (uint32_t)(16.0f / yr)));
        dim3 grid((nw + seg_n - 1) / seg_n, (nh + rows - 1) / rows, n_images);
        thumb_cols_kernel<T><<<grid, TC_THREADS, 0, st>>>(src, dst, w, h, C, nw, nh, xr, yr, seg_n, rows);
    } else {
        (void)sm_count;
        if ((size_t)nh * n_images > 0x7FFFFFFFull) return fail(B200IMG_ERR_DIMENSION, "thumbnail: too many output rows in one call");
        const unsigned g = (unsigned)((size_t)nh * n_images);
        if (big_windows) thumb_pixel_kernel<T, true><<<g, 256, 0, st>>>(src, dst, w, h, C, nw, nh, xr, yr);
        else thumb_pixel_kernel<T, false><<<g, 256, 0, st>>>(src, dst, w, h, C, nw, nh, xr, yr);
    }
    B200_CUDA(cudaGetLastError());
    return B200IMG_OK;
}

}  // namespace

namespace {
struct Extents { uint32_t min_x, max_x, min_y, max_y; bool int_div_by_zero; };

// window extents per axis, by the same arithmetic the kernels use (w, h, nw, nh all >= 1)
Extents extents_of(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh) {
    const float xr = (float)w / (float)nw, yr = (float)h / (float)nh;
    std::set<uint32_t> ex, ey;
    for (uint32_t o = 0; o < nw; ++o) { const Span s = thumb_span(o, xr, w); ex.insert(s.hi - s.lo); }
    for (uint32_t o = 0; o < nh; ++o) { const Span s = thumb_span(o, yr, h); ey.insert(s.hi - s.lo); }
    Extents e = {*ex.begin(), *ex.rbegin(), *ey.begin(), *ey.rbegin(), false};
    if ((uint64_t)e.max_x * e.max_y > 65536)
        for (uint32_t a : ex)
            for (uint32_t b : ey)
                if ((uint64_t)a * b > 65536 && (a / 256) * (uint64_t)(b / 256) == 0) e.int_div_by_zero = true;
    return e;
}
}  // namespace

// The one input the reference panics on (no GPU needed): an integer image with a window of more
// than 2^16 pixels that is under 256 wide or high makes sample.rs:716 count zero samples and
// sample.rs:739 divide by it.
int thumbnail_check(uint32_t w, uint32_t h, int dtype, uint32_t nw, uint32_t nh) {
    if (w == 0 || h == 0 || nw == 0 || nh == 0 || dtype == B200IMG_F32) return B200IMG_OK;
    if (extents_of(w, h, nw, nh).int_div_by_zero)
        return fail(B200IMG_ERR_BAD_ARG,
                    "thumbnail: a window of more than 65536 pixels is under 256 wide or high; "
                    "the reference divides by zero here (sample.rs:716,739)");
    return B200IMG_OK;
}

// path: 0 = choose, 1 = force the per-pixel kernels (tests)
int thumbnail_launch(const void *src, uint32_t w, uint32_t h, int C, int dtype, void *dst, uint32_t nw, uint32_t nh,
                     uint32_t n_images, int path, cudaStream_t st) {
    if (nw == 0 || nh == 0 || n_images == 0) return B200IMG_OK;
    int rc = thumbnail_check(w, h, dtype, nw, nh);
    if (rc) return rc;
    DeviceCtx *ctx = current_ctx();
    if (!ctx) return B200IMG_ERR_CUDA;
    const size_t esz = dtype == B200IMG_U8 ? 1 : (dtype == B200IMG_U16 ? 2 : 4);
    if (w == 0 || h == 0) {  // sample.rs:607: the zero-filled buffer comes back
        B200_CUDA(cudaMemsetAsync(dst, 0, (size_t)nw * nh * C * esz * n_images, st));
        return B200IMG_OK;
    }
    const float xr = (float)w / (float)nw, yr = (float)h / (float)nh;
    const Extents e = extents_of(w, h, nw, nh);
    const uint32_t min_x = e.min_x, max_x = e.max_x, min_y = e.min_y, max_y = e.max_y;
    const uint64_t max_px = (uint64_t)max_x * max_y;
    const int K = dtype == B200IMG_U8 ? 4 : 8, EPW = (int)(4 / esz);
    const size_t row_bytes = (size_t)w * C * esz;
    // outputs per segment so that the source span (plus one word of alignment slack) fits K*256 words
    const double cap_px = ((double)K * TC_THREADS * EPW - EPW) / C - 2.0;
    const uint32_t seg_n = cap_px / (double)xr >= 1.0 ? (uint32_t)std::min<double>(cap_px / (double)xr, (double)nw) : 0u;
    bool cols_ok = path == 0 && min_x >= 1 && min_y >= 1 && max_px <= 65536 && row_bytes % 4 == 0 &&
                   ((uintptr_t)src % 4) == 0 && seg_n >= 1 && (uint64_t)w * C < (1ull << 31) && nh <= 65535 && n_images <= 65535;
    if (cols_ok) {
        // the span bound above is a real-number estimate; confirm it with the kernel's arithmetic
        for (uint32_t o = 0; o < nw && cols_ok; o += seg_n) {
            const uint32_t last = std::min(nw, o + seg_n) - 1;
            const uint32_t x0 = thumb_span(o, xr, w).lo, x1 = thumb_span(last, xr, w).hi;
            const uint32_t ea = (x0 * (uint32_t)C) & ~(uint32_t)(EPW - 1);
            if ((x1 * (uint32_t)C - ea + EPW - 1) / EPW > (uint32_t)(K * TC_THREADS)) cols_ok = false;
        }
    }
    const bool big = max_px >= 256;
    switch (dtype) {
        case B200IMG_U8: return launch_t((const uint8_t *)src, w, h, C, (uint8_t *)dst, nw, nh, n_images, cols_ok, seg_n, big, xr, yr, ctx->sm_count, st);
        case B200IMG_U16: return launch_t((const uint16_t *)src, w, h, C, (uint16_t *)dst, nw, nh, n_images, cols_ok, seg_n, big, xr, yr, ctx->sm_count, st);
        default: return launch_t((const float *)src, w, h, C, (float *)dst, nw, nh, n_images, cols_ok, seg_n, big, xr, yr, ctx->sm_count, st);
    }
}

}  // namespace b200
