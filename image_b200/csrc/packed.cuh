// packed.cuh -- device helpers shared by the resampler kernels: packed f32x2 arithmetic,
// integer-subpixel -> float conversion, cp.async.
#pragma once

#include "common.cuh"

namespace b200 {

// ---- packed f32x2 arithmetic (sm_100+: FFMA2 / FADD2).  Same FLOP rate as the scalar forms
// but half the issue slots, which is what these issue-bound passes are short of. ----
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// 32-bit word of subpixels -> floats.  Integer subpixels are converted with the 2^23 trick
// (PRMT builds the float 8388608 + v, one FADD removes the bias: exact) because I2F issues on
// the quarter-rate XU pipe (measured 15 lanes/clk/SM vs > 34 for PRMT + FADD), which made the
// vertical pass XU-bound in the first profile.
template <typename T> struct Word;
template <> struct Word<uint8_t> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void biased(uint32_t w, float (&f)[4]) {  // 2^23 + v
        f[0] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540));
        f[1] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7541));
        f[2] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7542));
        f[3] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7543));
    }
};
template <> struct Word<uint16_t> {
    static constexpr int N = 2;
    static __device__ __forceinline__ void biased(uint32_t w, float (&f)[2]) {
        f[0] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7510));
        f[1] = __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7532));
    }
};
template <> struct Word<float> {
    static constexpr int N = 1;
    static __device__ __forceinline__ void biased(uint32_t w, float (&f)[1]) { f[0] = __uint_as_float(w); }
};
template <typename T> __device__ __forceinline__ float unbias(float m) { return m - 8388608.0f; }
template <> __device__ __forceinline__ float unbias<float>(float m) { return m; }

// (v + 0.5) as u8 of four values (NearestFrom<f32> for u8, src/primitive_sealed.rs:110-120) packed into one word, on the FMA
// pipe only: clamp, then an addition to 2^23 that rounds toward zero leaves floor(v + 0.5) in the low mantissa byte.
__device__ __forceinline__ uint32_t pack_u8x4(float r, float g, float b, float al) {
    auto q = [](float v) -> uint32_t {
        const float t = fminf(fmaxf(__fadd_rn(v, 0.5f), 0.0f), 255.5f);
        return __float_as_uint(__fadd_rz(t, 8388608.0f));
    };
    const uint32_t rg = __byte_perm(q(r), q(g), 0x0040), ba = __byte_perm(q(b), q(al), 0x0040);
    return __byte_perm(rg, ba, 0x5410);
}

// ... and (v + 0.5) as u16 of two values packed into one word, the same way (low 16 mantissa bits)
__device__ __forceinline__ uint32_t pack_u16x2(float a, float b) {
    auto q = [](float v) -> uint32_t {
        const float t = fminf(fmaxf(__fadd_rn(v, 0.5f), 0.0f), 65535.5f);
        return __float_as_uint(__fadd_rz(t, 8388608.0f));
    };
    return __byte_perm(q(a), q(b), 0x5410);
}

__device__ __forceinline__ void cp_async(void *smem_dst, const void *gsrc, int bytes) {
    const uint32_t d = smem_u32(smem_dst);
    if (bytes == 16) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gsrc) : "memory");
    else if (bytes == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(d), "l"(gsrc) : "memory");
    else asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

}  // namespace b200
