// Micro-benchmark: tcgen05.mma issue/execute rate for M=128, N in {64,128,256}, K=16 (bf16, SS and TS),
// the cost of tcgen05.commit, and the latency of mbarrier.try_wait on an already-completed barrier.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../simclr_b200/csrc/ntx_umma.cuh"
using namespace umma;

template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) k(long long* out, int reps, int commit_every) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc(&tmem_slot, 512);
    for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) {
        constexpr uint32_t idesc = instr_desc_bf16(128, N, 0);
        long long t0 = 0, t1 = 0;
        uint32_t commits = 0;
        if (elect_one()) {
            t0 = clock64();
            for (int r = 0; r < reps; ++r) {
#pragma unroll
                for (int k4 = 0; k4 < 8; ++k4) {
                    const uint64_t a = smem_desc_sw128(smem_u32(smem + (k4 & 3) * 32), 16, 1024);
                    const uint64_t b = smem_desc_sw128(smem_u32(smem + 32768 + (k4 & 3) * 32), 16, 1024);
                    if (TS) mma_bf16_ts(256, 0 + k4 * 8, b, idesc, 1);
                    else mma_bf16_ss(256, a, b, idesc, 1);
                }
                if (commit_every && (r % commit_every) == 0) { mma_commit(&bar); ++commits; }
            }
            t1 = clock64();
            mma_commit(&bar);
        }
        __syncwarp();
        // wait for everything, measure drain
        if (threadIdx.x == 32) {
            // barrier phases completed so far: commits + 1 ; wait for the last
            uint32_t total = commits + 1;
            for (uint32_t ph = 0; ph < total; ++ph) mbar_wait(&bar, ph & 1);
            long long t2 = clock64();
            // latency of a try_wait on a completed phase
            long long t3 = clock64();
            for (int r = 0; r < 64; ++r) mbar_wait(&bar, (total - 1) & 1);
            long long t4 = clock64();
            out[0] = t1 - t0; out[1] = t2 - t0; out[2] = (t4 - t3) / 64;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(0, 512);
}

template <int N, bool TS>
void run(const char* name, int commit_every) {
    long long* d; cudaMalloc(&d, 64);
    cudaFuncSetAttribute(k<N, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int reps = 256;
    k<N, TS><<<1, 128, 64 * 1024>>>(d, reps, commit_every);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[3]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-10s commit_every=%d: issue %.1f cyc/MMA, issue+drain %.1f cyc/MMA (nominal %d), try_wait(done) %lld cyc  [%s]\n",
           name, commit_every, (double)h[0] / (reps * 8), (double)h[1] / (reps * 8), 128 * N / 256, h[2], cudaGetErrorString(e));
    cudaFree(d);
}

int main() {
    run<128, false>("SS N=128", 0);
    run<128, false>("SS N=128", 1);
    run<256, false>("SS N=256", 0);
    run<256, false>("SS N=256", 1);
    run<64, false>("SS N=64", 0);
    run<128, true>("TS N=128", 0);
    run<128, true>("TS N=128", 1);
    run<256, true>("TS N=256", 0);
    return 0;
}
