// Microbenchmark: the fused Gaussian's column-pass and row-pass inner loops (conv_blocks on a 16-entry
// register window, FFMA2 with weights from the kernel-parameter bank) run back to back on shared
// memory with NO barriers and NO producers, at 2 / 3 / 4 / 6 warps per SM sub-partition: what the
// FMA pipe can be driven to by this instruction mix alone.   nvcc -arch=sm_100a -O3 conv_micro.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long f32x2;
struct W { f32x2 w[72]; };
__device__ __forceinline__ f32x2 pack2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &a, float &b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ void taps8(const f32x2 (&lo)[8], const f32x2 (&hi)[8], const f32x2 *w, f32x2 (&acc)[8]) {
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
        const f32x2 wk = w[kk];
#pragma unroll
        for (int r = 0; r < 8; ++r) acc[r] = fma2(r + kk < 8 ? lo[r + kk] : hi[r + kk - 8], wk, acc[r]);
    }
}
constexpr int RP = 194;
template <bool ROW>
__global__ void __launch_bounds__(768, 1) k(float *out, int iters, int nfull, const __grid_constant__ W tw) {
    extern __shared__ float sm[];
    for (int i = threadIdx.x; i < 128 * RP; i += blockDim.x) sm[i] = (float)(i & 255);
    __syncthreads();
    const int tid = threadIdx.x, lane = tid & 31;
    f32x2 acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) acc[r] = 0ull;
    for (int it = 0; it < iters; ++it) {
        f32x2 A[8], B[8];
        if (ROW) {
            const float *pa = sm + lane * 385 % (100 * RP) + (it & 7) * 3, *pb = pa + 96;
#pragma unroll
            for (int i = 0; i < 8; ++i) A[i] = pack2(pa[i * 3], pb[i * 3]);
#pragma unroll 1
            for (int ch = 0; ch + 2 <= nfull; ch += 2) {
                pa += 24; pb += 24;
#pragma unroll
                for (int i = 0; i < 8; ++i) B[i] = pack2(pa[i * 3], pb[i * 3]);
                taps8(A, B, tw.w + ch * 8, acc);
                pa += 24; pb += 24;
#pragma unroll
                for (int i = 0; i < 8; ++i) A[i] = pack2(pa[i * 3], pb[i * 3]);
                taps8(B, A, tw.w + ch * 8 + 8, acc);
            }
        } else {
            const float *p = sm + 2 * (tid % 96) + ((it & 3) * 8) * RP;
#pragma unroll
            for (int i = 0; i < 8; ++i) A[i] = *reinterpret_cast<const f32x2 *>(p + i * RP);
#pragma unroll 1
            for (int ch = 0; ch + 2 <= nfull; ch += 2) {
                p += 8 * RP;
#pragma unroll
                for (int i = 0; i < 8; ++i) B[i] = *reinterpret_cast<const f32x2 *>(p + i * RP);
                taps8(A, B, tw.w + ch * 8, acc);
                p += 8 * RP;
#pragma unroll
                for (int i = 0; i < 8; ++i) A[i] = *reinterpret_cast<const f32x2 *>(p + i * RP);
                taps8(B, A, tw.w + ch * 8 + 8, acc);
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int r = 0; r < 8; ++r) { float x, y; unpack2(acc[r], x, y); s += x + y; }
    out[blockIdx.x * blockDim.x + tid] = s;
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float *out; cudaMalloc(&out, sms * 768 * 4);
    W tw; for (int i = 0; i < 72; ++i) { float f = 1.0f / (i + 3); unsigned b; memcpy(&b, &f, 4); tw.w[i] = ((f32x2)b << 32) | b; }
    const size_t smem = 128 * RP * 4;
    cudaFuncSetAttribute(k<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(k<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000, nfull = 8;
    for (int row = 0; row < 2; ++row)
        for (int warps : {4, 8, 12, 16, 24}) {
            cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
            auto go = [&] { if (row) k<true><<<sms, warps * 32, smem>>>(out, iters, nfull, tw); else k<false><<<sms, warps * 32, smem>>>(out, iters, nfull, tw); };
            go(); cudaDeviceSynchronize();
            cudaEventRecord(a); go(); cudaEventRecord(b); cudaEventSynchronize(b);
            float ms; cudaEventElapsedTime(&ms, a, b);
            const double fma = (double)sms * warps * 32 * iters * nfull * 64 * 2;
            printf("%s pass, %2d warps/SM: %.3f ms  %.1f lanes/clk/SM @1.965GHz (%s)\n", row ? "row" : "col", warps, ms,
                   fma / (ms * 1e-3) / sms / 1.965e9, cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
