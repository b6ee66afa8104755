// Micro-benchmark: how fast can ONE SM run the backward kernel's tensor-core instruction mix?
//   GEMM1-like: 8 x tcgen05.mma SS  M128 N128 K16 (A and B K-major from shared memory)
//   GEMM2-like: 8 x tcgen05.mma TS  M128 N128 K16 (A from TMEM, B MN-major from shared memory)
//   copies    : one 32 KB bulk copy global -> shared per tile (the key tile the TMA producer brings in)
// Each part alone and together, two issuing warps like the kernel, no epilogue at all.  The SS form
// reads 8 KB of shared memory per MMA = 128 B/clk at the nominal 64 clk, i.e. the whole bandwidth.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simclr_b200/csrc/ntx_umma.cuh"
using namespace umma;

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

struct Ctl { uint64_t done1, done2, full[4], empty[4]; uint32_t tmem_slot; };

// mode bits: 1 = SS stream, 2 = TS stream, 4 = copies, 8 = SS uses N=256 (two key tiles per MMA)
__device__ __forceinline__ uint32_t rnd_bf16x2(uint32_t i) {      // two bf16 values in (-1, 1) with random mantissas
    uint32_t h = i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    const uint32_t lo = (h & 0x807Fu) | (0x7Du + ((h >> 8) & 1u)) << 7;
    const uint32_t hi = ((h >> 16) & 0x807Fu) | (0x7Du + ((h >> 25) & 1u)) << 7;
    return lo | hi << 16;
}
// mode bit 16: random operand data instead of one repeated value
// mode bit 32: 16 more warps run an FFMA/MUFU mix (the epilogues' arithmetic) for the whole duration
__global__ void __launch_bounds__(672, 1) k(long long* out, const uint8_t* src, int tiles, int mode) {
    extern __shared__ __align__(1024) uint8_t smem[];          // Q 32 KB + 4 x 32 KB ring (+ 32 KB for N=256)
    __shared__ Ctl ctl;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(&ctl.done1, 1); mbar_init(&ctl.done2, 1);
        for (int s = 0; s < 4; ++s) { mbar_init(&ctl.full[s], 1); mbar_init(&ctl.empty[s], 1); }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&ctl.tmem_slot, 512);
    for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x)
        reinterpret_cast<uint32_t*>(smem)[i] = (mode & 16) ? rnd_bf16x2(i + blockIdx.x * 77777) : 0x3c003c00u;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp < 4) {                                             // the TS form's A operand lives in TMEM columns [128, 512)
        for (int c = 128; c < 512; c += 16) {
            uint32_t v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = (mode & 16) ? rnd_bf16x2((c + j) * 131 + threadIdx.x) : 0x3c003c00u;
            tmem_st_32x16(((uint32_t)(warp * 32) << 16) + c, v);
        }
        tmem_st_wait();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint8_t* q = smem;
    uint8_t* ring = smem + 32768;
    long long t0 = clock64();
    if (warp == 1 && (mode & 1)) {
        if (elect_one()) {
            for (int r = 0; r < tiles; ++r) {
                const uint8_t* kt = ring + (r & 3) * 32768;
                if (mode & 4) mbar_wait(&ctl.full[r & 3], (r >> 2) & 1);
                if (mode & 8) {
                    if (r & 1) { if (!(mode & 2)) mma_commit(&ctl.empty[r & 3]); continue; }
                    constexpr uint32_t idesc = instr_desc_bf16(128, 256, 0);
#pragma unroll
                    for (int b = 0; b < 2; ++b)
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4) {
                            const uint64_t a = smem_desc_sw128(smem_u32(q + b * 16384 + k4 * 32), 16, 1024);
                            const uint64_t bd = smem_desc_sw128(smem_u32(kt + b * 32768 + k4 * 32), 16, 1024);
                            mma_bf16_ss(128, a, bd, idesc, (b | k4) ? 1u : 0u);
                        }
                } else {
                    constexpr uint32_t idesc = instr_desc_bf16(128, 128, 0);
                    const uint32_t d = 128 + (r % 3) * 128;
#pragma unroll
                    for (int b = 0; b < 2; ++b)
#pragma unroll
                        for (int k4 = 0; k4 < 4; ++k4) {
                            const uint64_t a = smem_desc_sw128(smem_u32(q + b * 16384 + k4 * 32), 16, 1024);
                            const uint64_t bd = smem_desc_sw128(smem_u32(kt + b * 16384 + k4 * 32), 16, 1024);
                            mma_bf16_ss(d, a, bd, idesc, (b | k4) ? 1u : 0u);
                        }
                }
                if (!(mode & 2)) mma_commit(&ctl.empty[r & 3]);
            }
            mma_commit(&ctl.done1);
        }
        __syncwarp();
        mbar_wait(&ctl.done1, 0);
    } else if (warp == 2 && (mode & 2)) {
        if (elect_one()) {
            constexpr uint32_t idesc2 = instr_desc_bf16(128, 128, 1);
            for (int r = 0; r < tiles; ++r) {
                const uint8_t* kt = ring + (r & 3) * 32768;
                const uint32_t c = 128 + (r % 3) * 128;
                if (mode & 4) mbar_wait(&ctl.full[r & 3], (r >> 2) & 1);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint64_t bd = smem_desc_sw128(smem_u32(kt + kk * 2048), 16384, 1024);
                    mma_bf16_ts(0, c + kk * 8, bd, idesc2, 1u);
                }
                mma_commit(&ctl.empty[r & 3]);
            }
            mma_commit(&ctl.done2);
        }
        __syncwarp();
        mbar_wait(&ctl.done2, 0);
    } else if (warp == 3 && (mode & 4)) {
        if (elect_one()) {
            // a real ring: stage s is refilled once the MMAs that read it have completed
            for (int r = 0; r < tiles; ++r) {
                const int s = r & 3;
                if (mode & 3) { if (r >= 4) mbar_wait(&ctl.empty[s], ((r >> 2) & 1) ^ 1); }
                else if (r >= 4) mbar_wait(&ctl.full[s], ((r >> 2) & 1) ^ 1);          // unpaced: wait for the old copy
                mbar_expect_tx(&ctl.full[s], 32768);
                const uint8_t* from = src + (size_t)((blockIdx.x * 7 + r) & 255) * 32768;
                bulk_g2s(ring + s * 32768, from, 16384, &ctl.full[s]);
                bulk_g2s(ring + s * 32768 + 16384, from + 16384, 16384, &ctl.full[s]);
            }
            for (int r = tiles - 4; r < tiles; ++r) mbar_wait(&ctl.full[r & 3], (r >> 2) & 1);
        }
        __syncwarp();
    }
    else if (warp >= 5 && (mode & 32)) {
        float a[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) a[j] = 0.001f * (threadIdx.x + j);
        const long long budget = (long long)tiles * 1100;
        while (clock64() - t0 < budget) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    a[j] = fmaf(a[j], 0.999f, 0.001f);
                    if ((j & 1) == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[j]));
                    a[j] = fmaf(a[j], a[(j + 1) & 15], 0.5f);
                }
        }
        float sum = 0;
#pragma unroll
        for (int j = 0; j < 16; ++j) sum += a[j];
        if (sum == 123.456f) out[0] = 1;
    }
    long long t1 = clock64();
    if ((threadIdx.x & 31) == 0 && warp < 8) out[blockIdx.x * 8 + warp] = t1 - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(0, 512);
}

int main() {
    long long* d; cudaMalloc(&d, 148 * 8 * sizeof(long long));
    uint8_t* src; cudaMalloc(&src, 256 * 32768); cudaMemset(src, 0x3c, 256 * 32768);
    { uint32_t* h = (uint32_t*)malloc(256 * 32768); for (int i = 0; i < 256 * 32768 / 4; ++i) { uint32_t x = i * 2654435761u; x ^= x >> 15; x *= 2246822519u; x ^= x >> 13; h[i] = (x & 0x807F807Fu) | 0x3E803E80u; } cudaMemcpy(src, h, 256 * 32768, cudaMemcpyHostToDevice); free(h); }
    const int smem = 192 * 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int tiles = 2048;
    const struct { int mode; const char* name; } runs[] = {
        {3, "SS + TS"}, {16 + 7, "random: SS + TS + copies"}, {32 + 16 + 7, "random: SS + TS + copies + ALU warps"},
        {32 + 7, "const: SS + TS + copies + ALU warps"}, {32 + 16 + 7, "random: SS + TS + copies + ALU warps"},
        {16 + 7, "random: SS + TS + copies"}};
    for (auto& r : runs) {
        cudaMemset(d, 0, 148 * 8 * sizeof(long long));
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k<<<148, (r.mode & 32) ? 672 : 160, smem>>>(d, src, tiles, r.mode);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h[148 * 8]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
        double w[4] = {0, 0, 0, 0};
        for (int b = 0; b < 148; ++b) for (int x = 1; x < 4; ++x) w[x] += (double)h[b * 8 + x] / 148;
        printf("%-26s: SS warp %7.1f  TS warp %7.1f  copy warp %7.1f clk/tile   kernel %.1f us (%.2f GHz eff)  [%s]\n",
               r.name, w[1] / tiles, w[2] / tiles, w[3] / tiles, ms * 1e3,
               (w[1] > w[2] ? (w[1] > w[3] ? w[1] : w[3]) : (w[2] > w[3] ? w[2] : w[3])) / (ms * 1e6), cudaGetErrorString(e));
    }
    return 0;
}
