// Micro-benchmark: packed fp32x2 FMA/ADD/MUL (Blackwell) versus scalar FFMA, alone and in the
// forward-epilogue mix (per PAIR of logits: one fma.f32x2, two ex2, one add.f32x2).
#include <cstdio>
#include <cuda_runtime.h>
#define ITERS 4096
#define N 16
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ unsigned long long pack(float a, float b) {
    unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpack(unsigned long long v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
    unsigned long long d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
    unsigned long long d; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

template <int MODE>
__global__ void k(float* out, long long* cycles, float a, float b) {
    unsigned long long v[N], A = pack(a, a), B = pack(b, b), acc0 = pack(0.f, 0.f), acc1 = acc0;
    float s[2 * N];
#pragma unroll
    for (int i = 0; i < N; ++i) { v[i] = pack(0.001f * (threadIdx.x + i), 0.002f * i); s[2 * i] = 0.001f * i; s[2 * i + 1] = 0.002f * i; }
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (MODE == 0) v[i] = fma2(v[i], A, B);                               // packed FMA chain
            if (MODE == 1) { s[2 * i] = fmaf(s[2 * i], a, b); s[2 * i + 1] = fmaf(s[2 * i + 1], a, b); }   // 2 scalar FFMA
            if (MODE == 2) {                                                      // packed fwd mix
                unsigned long long x = fma2(v[i], A, B);
                float x0, x1; unpack(x, x0, x1);
                unsigned long long e = pack(ex2f(x0), ex2f(x1));
                if (i & 1) acc1 = add2(acc1, e); else acc0 = add2(acc0, e);
            }
            if (MODE == 3) {                                                      // scalar fwd mix
                float x0, x1; unpack(v[i], x0, x1);
                s[0] += ex2f(fmaf(x0, a, b)); s[1] += ex2f(fmaf(x1, a, b));
            }
        }
    }
    long long t1 = clock64();
    float r = 0, p, q;
#pragma unroll
    for (int i = 0; i < N; ++i) { unpack(v[i], p, q); r += p + q + s[2 * i] + s[2 * i + 1]; }
    unpack(acc0, p, q); r += p + q; unpack(acc1, p, q); r += p + q;
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&cyc, 148 * sizeof(long long));
    for (int warps : {8, 16, 32}) {
        k<MODE><<<148, warps * 32>>>(out, cyc, 0.999f, -0.001f);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        double elems = (double)warps * 32 * ITERS * N * 2;     // scalar elements
        printf("%-28s warps/SM %2d : %7.2f fp32 elements/clk/SM\n", name, warps, elems / avg);
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("fma.f32x2");
    run<1>("2 x FFMA");
    run<2>("fma2 + 2 ex2 + add2 (pair)");
    run<3>("2 x (FFMA ex2 FADD)");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
