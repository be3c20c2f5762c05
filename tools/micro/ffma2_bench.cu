// Microbenchmark: warp-instruction throughput of scalar FFMA / FMUL / FADD vs packed FFMA2 on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float2 upk(u64 v) { float2 r; asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, int iters) {
    float x[8];
    u64 X[8];
    for (int i = 0; i < 8; i++) { x[i] = threadIdx.x * 0.001f + i; X[i] = pk(x[i], x[i] + 0.5f); }
    const u64 A = pk(a, a), B = pk(b, b);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) x[i] = __fmaf_rn(x[i], a, b);                                  // FFMA 3-reg
            if (MODE == 1) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(X[i]) : "l"(A), "l"(B));  // FFMA2
            if (MODE == 2) { x[i] = __fmaf_rn(x[i], a, b); x[i] = __int_as_float(__float_as_int(x[i]) ^ it); }  // FFMA + LOP3 (alu pipe)
            if (MODE == 4) x[i] = __fmaf_rn(x[i], x[(i + 1) & 7], x[(i + 3) & 7]);           // FFMA, three register operands
            if (MODE == 5) { x[i] = __fmaf_rn(x[i], x[(i + 1) & 7], x[(i + 3) & 7]); X[i] ^= (u64)it; }  // + 64-bit LOP (2 alu)
            if (MODE == 6) x[i] = __fmul_rn(x[i], x[(i + 1) & 7]);                           // FMUL reg,reg
            if (MODE == 7) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(X[i]) : "l"(X[(i + 1) & 7]), "l"(X[(i + 3) & 7]));
            if (MODE == 3) { asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(X[i]) : "l"(A), "l"(B)); x[i] = __int_as_float(__float_as_int(x[i]) ^ it); }
        }
    }
    float s = 0;
    for (int i = 0; i < 8; i++) { float2 u = upk(X[i]); s += x[i] + u.x + u.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, int per_iter_instr) {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    k<MODE><<<148 * 8, 256>>>(d, 1.0001f, 0.5f, 100);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(d, 1.0001f, 0.5f, iters);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double winstr = 148.0 * 8 * 8 /*warps*/ * (double)iters * per_iter_instr;
    double per_clk_smsp = winstr / (ms * 1e-3) / (148.0 * 4) / 1.965e9;
    printf("%-28s %8.3f ms  %.3f warp-instr/clk/SMSP (at 1965 MHz)\n", name, ms, per_clk_smsp);
    cudaFree(d);
}
int main() {
    run<0>("FFMA (3-reg)", 8);
    run<1>("FFMA2 (packed)", 8);
    run<2>("FFMA + LOP3 interleaved", 16);
    run<3>("FFMA2 + LOP3 interleaved", 16);
    run<4>("FFMA (3 distinct regs)", 8);
    run<5>("FFMA 3-reg + 2 LOP3", 24);
    run<6>("FMUL reg,reg", 8);
    run<7>("FFMA2 (3 distinct reg pairs)", 8);
    return 0;
}
