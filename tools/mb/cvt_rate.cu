// Micro-benchmark: which pipe does the fp32 -> packed bf16 conversion (F2FP) use, and does it
// contend with MUFU.EX2?  Modes time each instruction alone and mixed in the epilogues' ratio
// (two exponentials per packed conversion), plus an integer-only packing (IADD, IADD, PRMT).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define ITERS 4096
#define N 16
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t cvt_pack(float lo, float hi) {
    uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo)); return r;
}
__device__ __forceinline__ uint32_t int_pack(float lo, float hi) {     // round half up, then take the high halves
    const uint32_t a = __float_as_uint(lo) + 0x8000u, b = __float_as_uint(hi) + 0x8000u;
    return __byte_perm(a, b, 0x7632);
}
__device__ __forceinline__ uint32_t trunc_pack(float lo, float hi) {
    return __byte_perm(__float_as_uint(lo), __float_as_uint(hi), 0x7632);
}

template <int MODE>
__global__ void k(float* out, long long* cycles, float a, float b) {
    float v[N];
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = 0.001f * (threadIdx.x + i);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < N; i += 2) {
            if (MODE == 0) { v[i] = __uint_as_float(cvt_pack(v[i], v[i + 1])); }
            if (MODE == 1) { v[i] = ex2f(v[i]); v[i + 1] = ex2f(v[i + 1]); }
            if (MODE == 2) { v[i] = ex2f(v[i]); v[i + 1] = ex2f(v[i + 1]); v[i] = __uint_as_float(cvt_pack(v[i], v[i + 1])); }
            if (MODE == 3) { v[i] = __uint_as_float(int_pack(v[i], v[i + 1])); }
            if (MODE == 4) { v[i] = ex2f(v[i]); v[i + 1] = ex2f(v[i + 1]); v[i] = __uint_as_float(int_pack(v[i], v[i + 1])); }
            if (MODE == 5) { v[i] = ex2f(v[i]); v[i + 1] = ex2f(v[i + 1]); v[i] = __uint_as_float(trunc_pack(v[i], v[i + 1])); }
            if (MODE == 6) { v[i] = fmaf(v[i], a, b); v[i + 1] = fmaf(v[i + 1], a, b); v[i] = __uint_as_float(cvt_pack(v[i], v[i + 1])); }
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
void run(const char* name) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&cyc, 148 * sizeof(long long));
    for (int warps : {4, 8, 16}) {
        k<MODE><<<148, warps * 32>>>(out, cyc, 0.999f, 0.001f);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        double pairs = (double)warps * 32 * ITERS * (N / 2);
        printf("%-34s warps/SM %2d : %7.2f pairs/clk/SM  (%6.2f clk per warp-pair per SM sub-partition)\n", name, warps,
               pairs / avg, avg / ((double)warps / 4 * ITERS * (N / 2)));
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("F2FP (cvt.rn.bf16x2.f32)");
    run<1>("2 EX2");
    run<2>("2 EX2 + F2FP");
    run<3>("IADD IADD PRMT");
    run<4>("2 EX2 + IADD IADD PRMT");
    run<5>("2 EX2 + PRMT (truncate)");
    run<6>("2 FFMA + F2FP");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
