/*
 * gauss.cpp -- oracle restatement of the Gaussian blur
 * (TEST INFRASTRUCTURE ONLY; see oracle.h).
 *
 * Follows src/imageops/sample.rs: get_gaussian_kernel_1d 1057-1077,
 * GaussianBlurParameters 1081-1295, gaussian_blur_dyn_image 1297-1490,
 * gaussian_blur_indirect(_impl) 1492-1604; and src/imageops/filter_1d.rs:
 * make_arena_row 60-87, make_columns_arenas 102-130, ToStorage 132-164,
 * filter_symmetric_column 171-316, filter_symmetric_row 323-390,
 * KernelTransformer 392-420, prepare_symmetric_kernel 423-437, ring queue
 * 490-600, filter_2d_separable 615-761.  Both the ring-queue (< 55 column
 * taps) and the transient-image variants are restated literally, including
 * the 1-tap quirks (row shift; row 0 unwritten in the ring variant).
 */
#include "oracle.h"
#include "rust_sem.h"

#include <cstring>
#include <vector>

namespace {

const float PI_F32 = 3.14159265358979323846264338327950288f;

// sample.rs:1252-1269
uint32_t round_to_nearest_odd(float x) {
    uint32_t n = rs::sat_cast<uint32_t>(roundf(x));
    if (n % 2 != 0) return n;
    uint32_t lower = n - 1, upper = n + 1;
    float dist_lower = fabsf(x - (float)lower);
    float dist_upper = fabsf(x - (float)upper);
    return dist_lower <= dist_upper ? lower : upper;
}

// sample.rs:1271-1277
float sigma_from_kernel_size(float kernel_size) {
    float safe = kernel_size <= 1.0f ? 0.8f : kernel_size;
    return 0.3f * ((safe - 1.0f) * 0.5f - 1.0f) + 0.8f;
}

// sample.rs:1279-1294
uint32_t kernel_size_from_sigma(float sigma) {
    return rs::sat_cast<uint32_t>(fmaxf(6.66666f * sigma - 2.3333333f, 3.0f)) | 1u;
}

bool is_normal(float v) { return std::fpclassify(v) == FP_NORMAL; }

// ---- ToStorage, filter_1d.rs:132-164 ----
template <typename T, typename I> struct Store;
template <> struct Store<uint8_t, uint32_t> {
    static uint8_t to(uint32_t v) { uint32_t r = (v + (1u << 14)) >> 15; return (uint8_t)(r < 255u ? r : 255u); }
};
template <> struct Store<uint16_t, uint32_t> {
    static uint16_t to(uint32_t v) { uint32_t r = (v + (1u << 14)) >> 15; return (uint16_t)(r < 65535u ? r : 65535u); }
};
template <> struct Store<float, float> {
    static float to(float v) { return v; }
};

// make_arena_row, filter_1d.rs:60-87
template <typename T>
void make_arena_row(const T *image, std::vector<T> &row_buffer, size_t source_y, size_t width,
                    int N, size_t kernel_w) {
    size_t pad_w = kernel_w / 2 > 1 ? kernel_w / 2 : 1;
    const T *source_row = image + source_y * width * N;
    memcpy(&row_buffer[pad_w * N], source_row, width * N * sizeof(T));
    for (size_t p = 0; p < pad_w; ++p)
        for (int c = 0; c < N; ++c) {
            row_buffer[p * N + c] = source_row[c];
            row_buffer[(width + pad_w + p) * N + c] = source_row[(width - 1) * N + c];
        }
}

// filter_symmetric_row, filter_1d.rs:323-390.  (The 4-wide and scalar loops do
// the same per-subpixel arithmetic, so one loop restates both.)
template <typename T, typename I>
void filter_symmetric_row(const T *arena, T *dst_row, size_t dst_len, const std::vector<I> &k, int N) {
    const size_t length = k.size(), half_len = length / 2;
    const I hc = k[half_len];
    for (size_t x = 0; x < dst_len; ++x) {
        const T *src = arena + x;
        I k0 = (I)src[half_len * N] * hc;
        for (size_t i = 0; i < half_len; ++i) {
            size_t other = length - i - 1;
            k0 = orc_mac(k0, (I)src[i * N] + (I)src[other * N], k[i]);
        }
        dst_row[x] = Store<T, I>::to(k0);
    }
}

// filter_symmetric_column, filter_1d.rs:171-316 (32/16/4/1-wide chunks share
// the per-subpixel arithmetic).
template <typename T, typename I>
void filter_symmetric_column(const std::vector<const T *> &rows, T *dst_row, size_t dst_stride,
                             const std::vector<I> &k) {
    const size_t length = k.size(), half_len = length / 2;
    const I coeff = k[half_len];
    for (size_t x = 0; x < dst_stride; ++x) {
        I k0 = (I)rows[half_len][x] * coeff;
        for (size_t i = 0; i < half_len; ++i) {
            size_t other = length - i - 1;
            k0 = orc_mac(k0, (I)rows[i][x] + (I)rows[other][x], k[i]);
        }
        dst_row[x] = Store<T, I>::to(k0);
    }
}

// ---- the same two filters with the reference's own chunk structure, for the TIMED baseline -----------
// The reference writes its loops over fixed-size chunks so that rustc vectorises them: rows in chunks
// of 4 consecutive subpixels (filter_1d.rs:340-370), columns in chunks of 32 (one-byte subpixels only),
// 16, 4 and 1 (171-316).  Restated chunk for chunk, with the multiply_accumulate flavour a template
// parameter (src/math/utils.rs:7-47 picks it at compile time too), g++ -O3 emits packed AVX2 for them;
// per lane the arithmetic and its order are those of the scalar checker above, so the results are
// bit-identical (tests/test_oracle_kat.py asserts it).  The scalar functions stay the checker; these are
// what bench.py times, so that the CPU baseline is not slower than the code it stands in for.
template <bool FUSED> static inline float macc(float acc, float a, float b) { return FUSED ? fmaf(a, b, acc) : acc + a * b; }
template <bool FUSED> static inline uint32_t macc(uint32_t acc, uint32_t a, uint32_t b) { return acc + a * b; }

template <typename T, typename I, bool FUSED>
void filter_symmetric_row_chunked(const T *arena, T *dst_row, size_t dst_len, const std::vector<I> &k, int N) {
    const size_t length = k.size(), half_len = length / 2;
    const I hc = k[half_len];
    const I *kk = k.data();
    size_t x = 0;
    for (; x + 4 <= dst_len; x += 4) {  // filter_1d.rs:340-370
        const T *src = arena + x;
        I acc[4];
        for (int l = 0; l < 4; ++l) acc[l] = (I)src[half_len * N + l] * hc;
        for (size_t i = 0; i < half_len; ++i) {
            const T *fw = src + i * N, *bw = src + (length - i - 1) * N;
            const I c = kk[i];
            for (int l = 0; l < 4; ++l) acc[l] = macc<FUSED>(acc[l], (I)fw[l] + (I)bw[l], c);
        }
        for (int l = 0; l < 4; ++l) dst_row[x + l] = Store<T, I>::to(acc[l]);
    }
    for (; x < dst_len; ++x) {  // 374-389
        const T *src = arena + x;
        I k0 = (I)src[half_len * N] * hc;
        for (size_t i = 0; i < half_len; ++i) k0 = macc<FUSED>(k0, (I)src[i * N] + (I)src[(length - i - 1) * N], kk[i]);
        dst_row[x] = Store<T, I>::to(k0);
    }
}

template <typename T, typename I, bool FUSED, int W>
static inline void column_chunk(const T *const *rows, T *dst, size_t cx, const I *kk, size_t length, size_t half_len) {
    I acc[W];
    const T *mid = rows[half_len] + cx;
    const I coeff = kk[half_len];
    for (int l = 0; l < W; ++l) acc[l] = (I)mid[l] * coeff;
    for (size_t i = 0; i < half_len; ++i) {
        const T *fw = rows[i] + cx, *bw = rows[length - i - 1] + cx;
        const I c = kk[i];
        for (int l = 0; l < W; ++l) acc[l] = macc<FUSED>(acc[l], (I)fw[l] + (I)bw[l], c);
    }
    for (int l = 0; l < W; ++l) dst[cx + l] = Store<T, I>::to(acc[l]);
}

template <typename T, typename I, bool FUSED>
void filter_symmetric_column_chunked(const std::vector<const T *> &rows, T *dst_row, size_t dst_stride, const std::vector<I> &k) {
    const size_t length = k.size(), half_len = length / 2;
    const T *const *r = rows.data();
    const I *kk = k.data();
    size_t cx = 0;
    if (sizeof(T) == 1)  // filter_1d.rs:197-243 (two stores of 16)
        for (; cx + 32 <= dst_stride; cx += 32) {
            column_chunk<T, I, FUSED, 16>(r, dst_row, cx, kk, length, half_len);
            column_chunk<T, I, FUSED, 16>(r, dst_row, cx + 16, kk, length, half_len);
        }
    for (; cx + 16 <= dst_stride; cx += 16) column_chunk<T, I, FUSED, 16>(r, dst_row, cx, kk, length, half_len);  // 245-271
    for (; cx + 4 <= dst_stride; cx += 4) column_chunk<T, I, FUSED, 4>(r, dst_row, cx, kk, length, half_len);     // 273-297
    for (; cx < dst_stride; ++cx) column_chunk<T, I, FUSED, 1>(r, dst_row, cx, kk, length, half_len);             // 299-315
}

// which pair of filters a pass of filter_2d_separable runs: 0 = scalar checker, 1 = chunked, 2 = chunked + fused mac
template <typename T, typename I>
static inline void run_row(int mode, const T *arena, T *dst_row, size_t dst_len, const std::vector<I> &k, int N) {
    if (mode == 0) filter_symmetric_row<T, I>(arena, dst_row, dst_len, k, N);
    else if (mode == 1) filter_symmetric_row_chunked<T, I, false>(arena, dst_row, dst_len, k, N);
    else filter_symmetric_row_chunked<T, I, true>(arena, dst_row, dst_len, k, N);
}
template <typename T, typename I>
static inline void run_col(int mode, const std::vector<const T *> &rows, T *dst_row, size_t dst_stride, const std::vector<I> &k) {
    if (mode == 0) filter_symmetric_column<T, I>(rows, dst_row, dst_stride, k);
    else if (mode == 1) filter_symmetric_column_chunked<T, I, false>(rows, dst_row, dst_stride, k);
    else filter_symmetric_column_chunked<T, I, true>(rows, dst_row, dst_stride, k);
}
thread_local int t_filter_mode = 0;  // set by the *_chunked entry points around their calls

// KernelTransformer, filter_1d.rs:392-420
template <typename I> I transform(float w);
template <> uint32_t transform<uint32_t>(float w) {
    return rs::sat_cast<uint32_t>(fmaxf(fminf(w * 32768.0f, 32767.0f), 0.0f));
}
template <> float transform<float>(float w) { return w; }

// prepare_symmetric_kernel, filter_1d.rs:423-437
template <typename I>
std::vector<I> prepare_symmetric_kernel(std::vector<I> k) {
    while (k.size() > 2 && k.back() == (I)0 && k.front() == (I)0) {
        k.erase(k.begin());
        k.pop_back();
    }
    return k;
}

// filter_2d_separable_ring_queue, filter_1d.rs:490-600
template <typename T, typename I>
void ring_queue(const T *image, T *destination, size_t width, size_t height, int N,
                const std::vector<I> &row_kernel, const std::vector<I> &column_kernel) {
    const size_t pad_w = row_kernel.size() / 2 > 1 ? row_kernel.size() / 2 : 1;
    std::vector<T> row_buffer(width * N + pad_w * 2 * N);
    const size_t full_width = width * N;
    std::vector<T> buffer(full_width * column_kernel.size());
    const size_t column_kernel_len = column_kernel.size();
    const size_t half_kernel = column_kernel_len / 2;

    make_arena_row(image, row_buffer, 0, width, N, row_kernel.size());
    run_row<T, I>(t_filter_mode, row_buffer.data(), buffer.data(), full_width, row_kernel, N);
    for (size_t r = 0; r < half_kernel && (r + 1) * full_width < buffer.size(); ++r)
        memcpy(&buffer[(r + 1) * full_width], &buffer[0], full_width * sizeof(T));

    size_t start_ky = column_kernel_len / 2 + 1;
    start_ky %= column_kernel_len;

    for (size_t y = 1; y < height + half_kernel; ++y) {
        size_t new_y = y < height ? y : (y < height - 1 ? y : height - 1);
        make_arena_row(image, row_buffer, new_y, width, N, row_kernel.size());
        run_row<T, I>(t_filter_mode, row_buffer.data(), &buffer[start_ky * full_width], full_width, row_kernel, N);
        if (y >= half_kernel) {
            std::vector<const T *> rows(column_kernel_len);
            for (size_t i = 0; i < column_kernel_len; ++i) {
                size_t ky = (i + start_ky + 1) % column_kernel_len;
                rows[i] = &buffer[ky * full_width];
            }
            size_t dy = y - half_kernel;
            run_col<T, I>(t_filter_mode, rows, destination + dy * full_width, full_width, column_kernel);
        }
        start_ky += 1;
        start_ky %= column_kernel_len;
    }
}

// filter_2d_separable, filter_1d.rs:615-761.  `destination` must arrive zeroed
// (callers allocate vec![0; ..], sample.rs:1318.., 1536).
template <typename T, typename I>
int filter_2d_separable(const T *image, T *destination, size_t width, size_t height, int N,
                        const float *row_kernel_f, size_t nrow, const float *col_kernel_f,
                        size_t ncol) {
    if ((nrow & 1) == 0 || (ncol & 1) == 0) return -1;  // asserts 649-654
    std::vector<I> rk(nrow), ck(ncol);
    for (size_t i = 0; i < nrow; ++i) rk[i] = transform<I>(row_kernel_f[i]);
    for (size_t i = 0; i < ncol; ++i) ck[i] = transform<I>(col_kernel_f[i]);
    rk = prepare_symmetric_kernel(rk);
    ck = prepare_symmetric_kernel(ck);
    // (the is_empty() copy branch, 667-670, is unreachable: lengths stay >= 1)

    if (ncol < 55) {  // RING_QUEUE_CIRCULAR_CUTOFF on the *unscanned* length, 673
        ring_queue<T, I>(image, destination, width, height, N, rk, ck);
        return 0;
    }

    const size_t pad_w = rk.size() / 2 > 1 ? rk.size() / 2 : 1;
    std::vector<T> row_buffer(width * N + pad_w * 2 * N);
    std::vector<T> transient(width * height * N);
    const size_t src_stride = width * N;
    for (size_t y = 0; y < height; ++y) {
        make_arena_row(image, row_buffer, y, width, N, rk.size());
        run_row<T, I>(t_filter_mode, row_buffer.data(), &transient[y * src_stride], src_stride, rk, N);
    }
    // virtual column padding, 102-130 and 743-755: rows above/below are row 0 / row H-1
    const size_t pad_h = ck.size() / 2;
    std::vector<const T *> rows(ck.size());
    for (size_t y = 0; y < height; ++y) {
        for (size_t k = 0; k < ck.size(); ++k) {
            int64_t fy = (int64_t)y - (int64_t)pad_h + (int64_t)k;
            if (fy < 0) fy = 0;
            else if ((size_t)fy >= height) fy = (int64_t)height - 1;
            rows[k] = &transient[(size_t)fy * src_stride];
        }
        run_col<T, I>(t_filter_mode, rows, destination + y * src_stride, src_stride, ck);
    }
    return 0;
}

// gaussian_blur_indirect_impl, sample.rs:1505-1604
template <typename T>
int blur_generic(const T *src, uint32_t w, uint32_t h, int C, T *dst, const std::vector<float> &kx,
                 const std::vector<float> &ky) {
    const size_t n = (size_t)w * h * C;
    std::vector<float> transient(n), transient_dst(n, 0.0f);
    for (size_t i = 0; i < n; ++i) transient[i] = (float)src[i];
    int rc = filter_2d_separable<float, float>(transient.data(), transient_dst.data(), w, h, C,
                                               kx.data(), kx.size(), ky.data(), ky.size());
    if (rc) return rc;
    for (size_t i = 0; i < n; ++i) dst[i] = rs::nearest_from<T>(transient_dst[i]);
    return 0;
}

}  // namespace

extern "C" void orc_gaussian_kernel_1d(uint32_t width, float sigma, float *out) {
    // sample.rs:1057-1077.  powf(d, 2.0) lowers to d*d (LLVM SimplifyLibCalls).
    float sum_norm = 0.0f;
    float scale = 1.0f / (sqrtf(2.0f * PI_F32) * sigma);
    float mean = (float)(width / 2);
    for (uint32_t x = 0; x < width; ++x) {
        float d = ((float)x - mean) / sigma;
        float nw = expf(-0.5f * (d * d)) * scale;
        out[x] = nw;
        sum_norm += nw;
    }
    if (sum_norm != 0.0f) {
        float sum_scale = 1.0f / sum_norm;
        for (uint32_t x = 0; x < width; ++x) out[x] = out[x] * sum_scale;
    }
}

extern "C" int orc_gauss_params_from_kernel_size(float ks, orc_gauss_params *out) {
    // sample.rs:1148-1166
    if (!(ks > 0.0f) || !is_normal(ks)) return -1;
    uint32_t i = round_to_nearest_odd(ks);
    if (i % 2 == 0) return -1;
    float s = sigma_from_kernel_size(ks);
    *out = {i, s, i, s};
    return 0;
}

extern "C" int orc_gauss_params_from_radius(float radius, orc_gauss_params *out) {
    // sample.rs:1132-1142
    if (!(radius >= 0.0f)) return -1;
    if (radius != 0.0f && !is_normal(radius)) return -1;
    return orc_gauss_params_from_kernel_size(radius * 2.0f + 1.0f, out);
}

extern "C" int orc_gauss_params_from_anisotropic(float xks, float yks, orc_gauss_params *out) {
    // sample.rs:1172-1204 (the x size is only checked for > 0; y for normal, sic)
    if (!(xks > 0.0f) || !(yks > 0.0f) || !is_normal(yks)) return -1;
    uint32_t xi = round_to_nearest_odd(xks), yi = round_to_nearest_odd(yks);
    if (xi % 2 == 0 || yi % 2 == 0) return -1;
    *out = {xi, sigma_from_kernel_size(xks), yi, sigma_from_kernel_size(yks)};
    return 0;
}

extern "C" int orc_gauss_params_from_sigma(float sigma, orc_gauss_params *out) {
    // sample.rs:1210-1249
    if (!(sigma >= 0.0f) || !std::isfinite(sigma)) return -1;
    if (sigma < 0.16986436f) {
        *out = {1, 1.0f, 1, 1.0f};
        return 0;
    }
    uint32_t k = kernel_size_from_sigma(sigma);
    *out = {k, sigma, k, sigma};
    return 0;
}

extern "C" int orc_q15_kernel(const float *k, int n, uint32_t *out) {
    std::vector<uint32_t> q(n);
    for (int i = 0; i < n; ++i) q[i] = transform<uint32_t>(k[i]);
    q = prepare_symmetric_kernel(q);
    for (size_t i = 0; i < q.size(); ++i) out[i] = q[i];
    return (int)q.size();
}

extern "C" int orc_gaussian_blur_generic(const void *src, uint32_t w, uint32_t h, int channels,
                                         int dtype, void *dst, const orc_gauss_params *p) {
    if (channels < 1 || channels > 4) return -1;
    if (w == 0 || h == 0) return 0;
    std::vector<float> kx(p->x_kernel_size), ky(p->y_kernel_size);
    orc_gaussian_kernel_1d(p->x_kernel_size, p->x_sigma, kx.data());
    orc_gaussian_kernel_1d(p->y_kernel_size, p->y_sigma, ky.data());
    switch (dtype) {
        case ORC_U8: return blur_generic((const uint8_t *)src, w, h, channels, (uint8_t *)dst, kx, ky);
        case ORC_U16: return blur_generic((const uint16_t *)src, w, h, channels, (uint16_t *)dst, kx, ky);
        case ORC_F32: return blur_generic((const float *)src, w, h, channels, (float *)dst, kx, ky);
    }
    return -1;
}

extern "C" int orc_gaussian_blur_dyn(const void *src, uint32_t w, uint32_t h, int channels,
                                     int dtype, void *dst, const orc_gauss_params *p) {
    // gaussian_blur_dyn_image, sample.rs:1297-1490
    if (channels < 1 || channels > 4) return -1;
    if (w == 0 || h == 0) return 0;
    std::vector<float> kx(p->x_kernel_size), ky(p->y_kernel_size);
    orc_gaussian_kernel_1d(p->x_kernel_size, p->x_sigma, kx.data());
    orc_gaussian_kernel_1d(p->y_kernel_size, p->y_sigma, ky.data());
    const size_t n = (size_t)w * h * channels;
    switch (dtype) {
        case ORC_U8:
            memset(dst, 0, n);
            return filter_2d_separable<uint8_t, uint32_t>((const uint8_t *)src, (uint8_t *)dst, w, h,
                                                          channels, kx.data(), kx.size(), ky.data(), ky.size());
        case ORC_U16:
            memset(dst, 0, n * 2);
            return filter_2d_separable<uint16_t, uint32_t>((const uint16_t *)src, (uint16_t *)dst, w, h,
                                                           channels, kx.data(), kx.size(), ky.data(), ky.size());
        case ORC_F32:
            memset(dst, 0, n * 4);
            return filter_2d_separable<float, float>((const float *)src, (float *)dst, w, h, channels,
                                                     kx.data(), kx.size(), ky.data(), ky.size());
    }
    return -1;
}

// The timed baseline's entry points: the same calls with the chunked (vectorisable) filters.
extern "C" int orc_gaussian_blur_generic_chunked(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                                                 const orc_gauss_params *p) {
    t_filter_mode = g_orc_fused_mac ? 2 : 1;
    const int rc = orc_gaussian_blur_generic(src, w, h, channels, dtype, dst, p);
    t_filter_mode = 0;
    return rc;
}

extern "C" int orc_gaussian_blur_dyn_chunked(const void *src, uint32_t w, uint32_t h, int channels, int dtype, void *dst,
                                             const orc_gauss_params *p) {
    t_filter_mode = g_orc_fused_mac ? 2 : 1;
    const int rc = orc_gaussian_blur_dyn(src, w, h, channels, dtype, dst, p);
    t_filter_mode = 0;
    return rc;
}
