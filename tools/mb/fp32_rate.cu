// Micro-benchmark: FP32 FFMA / FADD / FMUL issue rate per SM with independent chains (register only),
// alone and mixed with MUFU.EX2, to size the polynomial-exp2 share of the NT-Xent epilogues.
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define N 16
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int MODE>
__global__ void k(float* out, long long* cycles, float a, float b) {
    float v[N];
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = 0.001f * (threadIdx.x + i);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (MODE == 0) v[i] = fmaf(v[i], a, b);                 // FFMA, 2 distinct source regs + dest
            if (MODE == 1) v[i] = v[i] + b;                         // FADD
            if (MODE == 2) v[i] = fmaf(v[i], v[(i + 1) % N], b);    // FFMA 3 distinct regs
            if (MODE == 3) { v[i] = fmaf(v[i], a, b); if ((i & 3) == 0) v[i] = ex2f(v[i]); }   // 4 FFMA : 1 MUFU
            if (MODE == 4) { v[i] = fmaf(v[i], a, b); if ((i & 1) == 0) v[i] = ex2f(v[i]); }   // 2 FFMA : 1 MUFU
            if (MODE == 5) { v[i] = ex2f(fmaf(v[i], a, b)) + b; }                              // FFMA MUFU FADD
            if (MODE == 6) { v[i] = __int_as_float(__float_as_int(v[i]) + (__float_as_int(b) << 23)); }   // LEA/int
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < N; ++i) s += v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int ops_per_elem) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&cyc, 148 * sizeof(long long));
    for (int warps : {4, 8, 16, 32}) {
        k<MODE><<<148, warps * 32>>>(out, cyc, 0.999f, 0.001f);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        double elems = (double)warps * 32 * ITERS * N;
        printf("%-26s warps/SM %2d : %7.2f elem/clk/SM (%6.1f thread-instr/clk/SM)\n", name, warps, elems / avg,
               elems * ops_per_elem / avg);
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("FFMA r,r,c,c", 1);
    run<1>("FADD", 1);
    run<2>("FFMA 3 regs", 1);
    run<3>("4 FFMA : 1 EX2", 1);
    run<4>("2 FFMA : 1 EX2", 1);
    run<5>("FFMA EX2 FADD", 3);
    run<6>("SHL+IADD (int)", 1);
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
