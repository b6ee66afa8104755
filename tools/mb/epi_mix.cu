// Micro-benchmark: throughput of the epilogue instruction mixes per SM, inputs re-read from
// shared memory every iteration (stands in for tcgen05.ld) so nothing is hoisted.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

#define ITERS 512

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int MODE>
__global__ void k(float* out, long long* cycles, float c1, float alpha) {
    __shared__ __align__(16) float in[32 * 33];
    __shared__ __align__(16) float beta[128];
    for (int i = threadIdx.x; i < 32 * 33; i += blockDim.x) in[i] = -0.01f * (i % 97);
    for (int i = threadIdx.x; i < 128; i += blockDim.x) beta[i] = 0.001f * i;
    __syncthreads();
    const float nc1 = -c1;
    float acc[4] = {0, 0, 0, 0};
    unsigned pkacc = 0;
    const float4* src = reinterpret_cast<const float4*>(in) + (threadIdx.x & 31);
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; ++it) {
        float v[32];
#pragma unroll
        for (int q = 0; q < 8; ++q) {   // 8 x LDS.128 = 32 values
            float4 t;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                         : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                         : "r"((unsigned)__cvta_generic_to_shared(src + ((q + it) & 7) * 32)));
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            float a = fmaf(v[j], c1, nc1);
            if (MODE == 0) { acc[j & 3] += a; }                                 // FFMA + FADD only
            if (MODE == 1) { acc[j & 3] += ex2f(a); }                           // fwd mix: FFMA MUFU FADD
            if (MODE >= 2) {
                float b = MODE >= 4 ? beta[(j + it) & 127] : 0.003f * j;
                float e = ex2f(a) * (alpha + b);                                // + FADD FMUL
                if (MODE == 2) acc[j & 3] += e;
                if (MODE >= 3) {                                                // + pack to bf16x2
                    v[j] = e;
                }
            }
        }
        if (MODE >= 3) {
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
                __nv_bfloat162 h = __floats2bfloat162_rn(v[j], v[j + 1]);
                pkacc ^= *reinterpret_cast<unsigned*>(&h);
            }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc[0] + acc[1] + acc[2] + acc[3] + __uint_as_float(pkacc);
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name) {
    float* out; long long* cyc;
    cudaMalloc(&out, 148 * 1024 * sizeof(float));
    cudaMalloc(&cyc, 148 * sizeof(long long));
    for (int warps : {4, 8, 12, 16}) {
        k<MODE><<<148, warps * 32>>>(out, cyc, 2.885f, 0.01f);
        cudaDeviceSynchronize();
        long long h[148];
        cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
        double elems = (double)warps * 32 * ITERS * 32;
        printf("%-34s warps/SM %2d : %6.2f elem/clk/SM = %6.0f cycles per 128x128 tile\n", name, warps, elems / avg,
               16384.0 / (elems / avg));
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("LDS + FFMA + FADD");
    run<1>("LDS + FFMA + EX2 + FADD (fwd)");
    run<2>("LDS + FFMA EX2 FADD FMUL + FADD");
    run<3>("LDS + FFMA EX2 FADD FMUL + F2FP");
    run<4>("... + beta from smem (bwd)");
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
