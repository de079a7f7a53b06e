// Microbenchmark: issue rate of FFMA (scalar) vs FFMA2 (fma.rn.f32x2, sm_100+) and of I2F.U8 vs PRMT+FADD.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
__global__ void k_ffma(float *out, float a, float b) {
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = threadIdx.x + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_ffma2(float *out, float a, float b) {
    unsigned long long acc[8], aa, bb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(aa) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll
    for (int i = 0; i < 8; ++i) { float x = threadIdx.x + i, y = x + 0.5f; asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(x), "f"(y)); }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(aa), "l"(bb));
    }
    float s = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float x, y; asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(acc[i])); s += x + y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_i2f(float *out, const unsigned *in) {
    unsigned w = in[threadIdx.x & 31];
    float s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int it = 0; it < ITERS; ++it) {
        s0 += (float)(w & 0xff); s1 += (float)((w >> 8) & 0xff); s2 += (float)((w >> 16) & 0xff); s3 += (float)(w >> 24);
        w = w * 1664525u + 1013904223u;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
}
__global__ void k_prmt(float *out, const unsigned *in) {
    unsigned w = in[threadIdx.x & 31];
    float s0 = 0, s1 = 0, s2 = 0, s3 = 0;
    for (int it = 0; it < ITERS; ++it) {
        s0 += __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7540)) - 8388608.0f;
        s1 += __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7541)) - 8388608.0f;
        s2 += __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7542)) - 8388608.0f;
        s3 += __uint_as_float(__byte_perm(w, 0x4B000000u, 0x7543)) - 8388608.0f;
        w = w * 1664525u + 1013904223u;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s0 + s1 + s2 + s3;
}
template <typename F> float timeit(F f) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(a); for (int i = 0; i < 5; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); return ms / 5;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 8, threads = 256;
    float *out; cudaMalloc(&out, blocks * threads * 4);
    unsigned *in; cudaMalloc(&in, 128); cudaMemset(in, 0x5a, 128);
    float t1 = timeit([&] { k_ffma<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    float t2 = timeit([&] { k_ffma2<<<blocks, threads>>>(out, 1.0001f, 0.5f); });
    double fmas = (double)blocks * threads * ITERS * 16;
    printf("FFMA : %.3f ms  %.1f TFMA/s (%.1f lanes/clk/SM @1.965GHz)\n", t1, fmas / t1 / 1e9, fmas / (t1 * 1e-3) / sms / 1.965e9);
    printf("FFMA2: %.3f ms  %.1f TFMA/s (%.1f lanes/clk/SM @1.965GHz)\n", t2, fmas / t2 / 1e9, fmas / (t2 * 1e-3) / sms / 1.965e9);
    float t3 = timeit([&] { k_i2f<<<blocks, threads>>>(out, in); });
    float t4 = timeit([&] { k_prmt<<<blocks, threads>>>(out, in); });
    double cv = (double)blocks * threads * ITERS * 4;
    printf("I2F.U8 loop   : %.3f ms  %.1f Gcvt/s\nPRMT+FADD loop: %.3f ms  %.1f Gcvt/s\n", t3, cv / t3 / 1e6, t4, cv / t4 / 1e6);
    return 0;
}
