/*
 * rust_sem.h -- the handful of Rust numeric semantics the oracle must reproduce
 * in C++ (TEST INFRASTRUCTURE ONLY; see oracle.h).
 *
 *   `x as u8/u16/u32/i64` from a float: truncate toward zero, saturate, NaN -> 0
 *   f32::min / f32::max: ignore a NaN operand (fminf / fmaxf do the same)
 *   f32::round: half away from zero (roundf)
 *   f32::powi(n): compiler-rt __powisf2 = square-and-multiply
 */
#ifndef B200IMG_ORACLE_RUST_SEM_H
#define B200IMG_ORACLE_RUST_SEM_H

#include <cmath>
#include <cstdint>
#include <limits>

namespace rs {

template <typename I, typename F>
static inline I sat_cast(F v) {
    if (v != v) return 0;  // NaN -> 0
    const F lo = static_cast<F>(std::numeric_limits<I>::min());
    // max() may not be representable in F (e.g. i64 in f32): compare against 2^bits
    const F hi_excl = std::ldexp(static_cast<F>(1), std::numeric_limits<I>::digits);
    if (v <= lo) return std::numeric_limits<I>::min();
    if (v >= hi_excl) return std::numeric_limits<I>::max();
    return static_cast<I>(v);  // in range: C truncation == Rust truncation
}

static inline float powi(float a, int b) {
    // compiler-rt lib/builtins/powisf2.c
    const bool recip = b < 0;
    float r = 1.0f;
    while (true) {
        if (b & 1) r *= a;
        b /= 2;
        if (b == 0) break;
        a *= a;
    }
    return recip ? 1.0f / r : r;
}

template <typename N>
static inline N clamp(N a, N mn, N mx) {
    // src/utils/mod.rs:96-107
    if (a < mn) return mn;
    if (a > mx) return mx;
    return a;
}

// NearestFrom<f32>::clamp_nearest_from, src/primitive_sealed.rs:107-147
template <typename T> static inline T clamp_nearest_from(float v);
template <> inline uint8_t clamp_nearest_from<uint8_t>(float v) { return sat_cast<uint8_t>(v + 0.5f); }
template <> inline uint16_t clamp_nearest_from<uint16_t>(float v) { return sat_cast<uint16_t>(v + 0.5f); }
template <> inline float clamp_nearest_from<float>(float v) { return fminf(fmaxf(v, 0.0f), 1.0f); }

// NearestFrom<f32>::nearest_from, src/primitive_sealed.rs:110-133 (f32: identity, no clamp)
template <typename T> static inline T nearest_from(float v);
template <> inline uint8_t nearest_from<uint8_t>(float v) { return sat_cast<uint8_t>(v + 0.5f); }
template <> inline uint16_t nearest_from<uint16_t>(float v) { return sat_cast<uint16_t>(v + 0.5f); }
template <> inline float nearest_from<float>(float v) { return v; }

}  // namespace rs

extern int g_orc_fused_mac;

// multiply_accumulate, src/math/utils.rs:7-47
static inline float orc_mac(float acc, float a, float b) {
    return g_orc_fused_mac ? fmaf(a, b, acc) : acc + a * b;
}
static inline uint32_t orc_mac(uint32_t acc, uint32_t a, uint32_t b) { return acc + a * b; }

#endif
