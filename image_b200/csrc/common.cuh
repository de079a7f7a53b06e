// common.cuh -- shared pieces of libb200img: error plumbing, per-device context
// (SM count, scratch pool), and the device-side restatement of the reference's
// numeric store rules.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/b200img.h"

namespace b200 {

// ---- errors -----------------------------------------------------------------
int fail(int status, const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);
#define B200_CUDA(expr)                                             \
    do {                                                            \
        cudaError_t _e = (expr);                                    \
        if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #expr); \
    } while (0)

// ---- per-device context -------------------------------------------------------
struct DeviceCtx {
    int device = -1;
    int sm_count = 0;
    size_t smem_optin = 0;  // max dynamic shared memory per block
};
// Context of the current device (created on first use).  nullptr + error set on failure.
DeviceCtx *current_ctx();

// Cached scratch allocation on the current device, at least `bytes` large.
// Slots are independent buffers; contents are undefined on return.
enum ScratchSlot { SCRATCH_TABLES = 0, SCRATCH_A = 1, SCRATCH_B = 2, SCRATCH_IO_SRC = 3, SCRATCH_IO_DST = 4, SCRATCH_MID = 5, SCRATCH_FLAG = 6, SCRATCH_SLOTS = 7 };
int scratch(ScratchSlot slot, size_t bytes, void **out);
// Scratch is keyed by (device, stream) and leased to the calling host thread until its entry point
// returns: every entry point that may take scratch opens a ScratchScope first.
void set_scratch_stream(cudaStream_t st);
void release_scratch_leases();
struct ScratchScope {
    explicit ScratchScope(cudaStream_t st) { release_scratch_leases(); set_scratch_stream(st); }
    ~ScratchScope() { release_scratch_leases(); }
    ScratchScope(const ScratchScope &) = delete;
    ScratchScope &operator=(const ScratchScope &) = delete;
};

inline size_t dtype_size(int dtype) { return dtype == B200IMG_U8 ? 1 : dtype == B200IMG_U16 ? 2 : 4; }
inline bool valid_dtype(int d) { return d == B200IMG_U8 || d == B200IMG_U16 || d == B200IMG_F32; }
// checked a*b*c*d into size_t (ImageError::Limits(DimensionError) when it overflows)
bool checked_mul(size_t a, size_t b, size_t c, size_t d, size_t *out);

// ---- device-side numeric rules ------------------------------------------------
#ifdef __CUDACC__

// Rust `v as u8 / u16` from f32: truncate toward zero, saturate, NaN -> 0.
__device__ __forceinline__ uint32_t sat_u32_rz(float v) { return __float2uint_rz(v); }  // cvt.rzi.u32.f32 saturates, NaN -> 0

template <typename T> struct Px;
template <> struct Px<uint8_t> {
    static constexpr uint32_t MAXI = 255u;
    // NearestFrom<f32>::(clamp_)nearest_from, src/primitive_sealed.rs:110-120: (v + 0.5) as u8
    static __device__ __forceinline__ uint8_t nearest(float v) { return (uint8_t)min(sat_u32_rz(__fadd_rn(v, 0.5f)), 255u); }
    static __device__ __forceinline__ uint8_t clamp_nearest(float v) { return nearest(v); }
    static __device__ __forceinline__ float to_f32(uint8_t v) { return (float)v; }
};
template <> struct Px<uint16_t> {
    static constexpr uint32_t MAXI = 65535u;
    static __device__ __forceinline__ uint16_t nearest(float v) { return (uint16_t)min(sat_u32_rz(__fadd_rn(v, 0.5f)), 65535u); }
    static __device__ __forceinline__ uint16_t clamp_nearest(float v) { return nearest(v); }
    static __device__ __forceinline__ float to_f32(uint16_t v) { return (float)v; }
};
template <> struct Px<float> {
    // src/primitive_sealed.rs:127-137: nearest_from = identity; clamp_nearest_from = max(0).min(1), NaN -> 0
    static __device__ __forceinline__ float nearest(float v) { return v; }
    static __device__ __forceinline__ float clamp_nearest(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }
    static __device__ __forceinline__ float to_f32(float v) { return v; }
};

// acc + a*b under the two arithmetic flags
template <bool STRICT>
__device__ __forceinline__ float mac(float acc, float a, float b) {
    if (STRICT) return __fadd_rn(acc, __fmul_rn(a, b));
    return fmaf(a, b, acc);
}

// streaming 128-bit global access (data touched once: keep it out of L1)
__device__ __forceinline__ uint4 ldg_stream(const uint4 *p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream(uint4 *p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// ---- TMA 1-D bulk copies + mbarrier (sm_90+; SASS: UBLKCP / SYNCS) ----------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared, completion counted on `bar`.  dst/src 16-byte aligned, bytes % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global, tracked by the bulk async-group of the issuing thread.
__device__ __forceinline__ void bulk_s2g(void *gdst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" :: "n"(N) : "memory");
}
// make generic-proxy shared-memory writes visible to the async proxy (before bulk_s2g)
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#endif  // __CUDACC__

}  // namespace b200
