// Micro-benchmark: the backward kernel's pipeline without its arithmetic.
//   producer warp : one 32 KB bulk copy per tile into a ring of NS stages (waits k_empty)
//   GEMM1 warp    : waits k_full + buf_free, 8 x tcgen05.mma SS into S buffer i % NB, commits s_full
//   "epilogue"    : NE warps take tiles round robin: wait s_full, spin for E cycles, arrive c_full
//   GEMM2 warp    : waits c_full, 8 x tcgen05.mma TS (A = the buffer in TMEM), commits k_empty + buf_free
// Reports cycles per tile for several E and NE: how much of the real kernel's ~1400 clk/tile is the
// pipeline's own latency, and how it reacts to the epilogue time.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#include "../../simclr_b200/csrc/ntx_umma.cuh"
using namespace umma;

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

#ifndef NB_
#define NB_ 3
#endif
#ifndef KEYS_
#define KEYS_ 128       // keys per tile: 128 (3 buffers of 128 TMEM columns) or 96 (4 buffers of 96)
#endif
constexpr int NS = 5, NB = NB_, KEYS = KEYS_, MAXE = 8;
struct Ctl { uint64_t k_full[NS], k_empty[NS], s_full[NB], c_full[NB], c_half[NB], buf_free[NB]; uint32_t tmem_slot; };

// variant bits: 1 = GEMM1 and GEMM2 issued by ONE warp in the order G1(i), G2(i - lag)
// variant bit 2: four warps per epilogue group (all wait and arrive); bit 4: they also read the S tile from TMEM
// and write 64 columns back (the real epilogue's TMEM traffic)
__global__ void __launch_bounds__(640, 1) k(long long* out, const uint8_t* src, int tiles, int E, int NE, int variant,
                                            int lag) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ Ctl ctl;
    __shared__ int pipe_lock;                                    // variant bit 128: a GEMM's MMAs are issued under this lock
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        pipe_lock = 0;
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl.k_full[s], 1); mbar_init(&ctl.k_empty[s], 1); }
        for (int b = 0; b < NB; ++b) { mbar_init(&ctl.s_full[b], 1); mbar_init(&ctl.c_full[b], (variant & 2) ? 4 : 1); mbar_init(&ctl.buf_free[b], 1); mbar_init(&ctl.c_half[b], (variant & 2) ? 4 : 1); }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&ctl.tmem_slot, 512);
    for (int i = threadIdx.x; i < 192 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint8_t* q = smem;
    uint8_t* ring = smem + 32768;
    constexpr uint32_t idesc1 = instr_desc_bf16(128, KEYS, 0), idesc2 = instr_desc_bf16(128, 128, 1);
    auto lock = [&]() { if (variant & 128) while (atomicCAS(&pipe_lock, 0, 1) != 0) {} };
    auto unlock = [&]() { if (variant & 128) atomicExch(&pipe_lock, 0); };
    auto gemm1 = [&](int i) {
        lock();
        const uint8_t* kt = ring + (i % NS) * 32768;
        const uint32_t d = 128 + (i % NB) * KEYS;
#pragma unroll
        for (int b = 0; b < 2; ++b)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
                mma_bf16_ss(d, smem_desc_sw128(smem_u32(q + b * 16384 + k4 * 32), 16, 1024),
                            smem_desc_sw128(smem_u32(kt + b * 16384 + k4 * 32), 16, 1024), idesc1, (b | k4) ? 1u : 0u);
        mma_commit(&ctl.s_full[i % NB]);
        unlock();
    };
    auto gemm2_half = [&](int i, int hf) {
        const uint8_t* kt = ring + (i % NS) * 32768;
        const uint32_t c = 128 + (i % NB) * 128;
#pragma unroll
        for (int kk = 0; kk < 4; ++kk)
            mma_bf16_ts(0, c + (hf * 4 + kk) * 8, smem_desc_sw128(smem_u32(kt + (hf * 4 + kk) * 2048), 16384, 1024), idesc2, 1u);
        if (hf) { mma_commit(&ctl.k_empty[i % NS]); mma_commit(&ctl.buf_free[i % NB]); }
    };
    auto gemm2 = [&](int i) {
        lock();
        const uint8_t* kt = ring + (i % NS) * 32768;
        const uint32_t c = 128 + (i % NB) * KEYS;
#pragma unroll
        for (int kk = 0; kk < KEYS / 16; ++kk)
            mma_bf16_ts(0, c + kk * 8, smem_desc_sw128(smem_u32(kt + kk * 2048), 16384, 1024), idesc2, 1u);
        mma_commit(&ctl.k_empty[i % NS]);
        mma_commit(&ctl.buf_free[i % NB]);
        unlock();
    };
    long long t0 = clock64();
    if (warp == 1) {                                             // producer
        if (lane == 0)
            for (int i = 0; i < tiles; ++i) {
                const int s = i % NS;
                mbar_wait(&ctl.k_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&ctl.k_full[s], 32768);
                const uint8_t* from = src + (size_t)((blockIdx.x * 7 + i) & 255) * 32768;
                bulk_g2s(ring + s * 32768, from, 16384, &ctl.k_full[s]);
                bulk_g2s(ring + s * 32768 + 16384, from + 16384, 16384, &ctl.k_full[s]);
            }
    } else if (warp == 2) {                                      // GEMM1 (or both)
        if (variant & 1) {
            for (int i = 0; i < tiles + lag; ++i) {
                if (i < tiles) {
                    mbar_wait(&ctl.k_full[i % NS], (i / NS) & 1);
                    if (!(variant & 8)) mbar_wait(&ctl.buf_free[i % NB], ((i / NB) & 1) ^ 1);   // bit 8: rely on in-order MMAs
                    tc_fence_after();
                    if (elect_one()) gemm1(i);
                    __syncwarp();
                }
                const int j = i - lag;
                if (j >= 0) {
                    mbar_wait(&ctl.c_full[j % NB], (j / NB) & 1);
                    tc_fence_after();
                    if (elect_one()) gemm2(j);
                    __syncwarp();
                }
            }
        } else {
            for (int i = 0; i < tiles; i += (variant & 64) ? 2 : 1) {    // bit 64: a second GEMM1 warp takes the odd tiles
                mbar_wait(&ctl.k_full[i % NS], (i / NS) & 1);
                mbar_wait(&ctl.buf_free[i % NB], ((i / NB) & 1) ^ 1);
                tc_fence_after();
                if (elect_one()) gemm1(i);
                __syncwarp();
            }
        }
    } else if (warp == 19 && (variant & 64)) {
        for (int i = 1; i < tiles; i += 2) {
            mbar_wait(&ctl.k_full[i % NS], (i / NS) & 1);
            mbar_wait(&ctl.buf_free[i % NB], ((i / NB) & 1) ^ 1);
            tc_fence_after();
            if (elect_one()) gemm1(i);
            __syncwarp();
        }
    } else if (warp == 3) {                                      // GEMM2
        if (!(variant & 1))
            for (int i = 0; i < tiles; ++i) {
                if (variant & 16) {
                    mbar_wait(&ctl.c_half[i % NB], (i / NB) & 1);
                    tc_fence_after();
                    if (elect_one()) gemm2_half(i, 0);
                    __syncwarp();
                    mbar_wait(&ctl.c_full[i % NB], (i / NB) & 1);
                    tc_fence_after();
                    if (elect_one()) gemm2_half(i, 1);
                    __syncwarp();
                    continue;
                }
                mbar_wait(&ctl.c_full[i % NB], (i / NB) & 1);
                tc_fence_after();
                if (elect_one()) gemm2(i);
                __syncwarp();
            }
    } else if (warp >= 4 && warp < 4 + NE * ((variant & 2) ? 4 : 1)) {      // stand-ins for the epilogue groups
        const int e = (variant & 2) ? (warp - 4) >> 2 : warp - 4;
        const uint32_t t_lane = (uint32_t)((warp & 3) * 32) << 16;
        for (int i = e; i < tiles; i += NE) {
            mbar_wait(&ctl.s_full[i % NB], (i / NB) & 1);
            tc_fence_after();
            const long long t = clock64();
            if (variant & 4) {
                const uint32_t base = t_lane + 128 + (i % NB) * 128;
#pragma unroll 1
                for (int ch = 0; ch < 4; ++ch) {
                    uint32_t v[32], pk[16];
                    tmem_ld_32x32(base + ch * 32, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) pk[j] = v[j] + v[j + 16];
                    while (clock64() - t < (long long)E * (ch + 1) / 4) {}
                    tmem_st_32x16(base + ch * 16, pk);
                }
                tmem_st_wait();
            } else if (variant & 16) {
                while (clock64() - t < E / 2) {}
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl.c_half[i % NB]);
                while (clock64() - t < E) {}
            } else if (variant & 32) {                           // the delay is spent on an FFMA / MUFU mix
                float a[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) a[j] = 0.001f * (threadIdx.x + j);
                while (clock64() - t < E) {
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
            } else {
                while (clock64() - t < E) {}
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl.c_full[i % NB]);
        }
    }
    // everyone waits for the last GEMM2
    {
        const int last = tiles - 1;
        if (warp == 2 || warp == 3) mbar_wait(&ctl.buf_free[last % NB], (last / NB) & 1);
    }
    long long t1 = clock64();
    if (lane == 0 && warp < 8) out[blockIdx.x * 8 + warp] = t1 - t0;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(0, 512);
}

int main() {
    long long* d; cudaMalloc(&d, 148 * 8 * sizeof(long long));
    uint8_t* src; cudaMalloc(&src, 256 * 32768); cudaMemset(src, 0x3c, 256 * 32768);
    const int smem = 192 * 1024;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    const int tiles = 600;
    const int Es[] = {0, 500, 1000, 1500, 2000, 2500, 3000};
    const struct { int variant, NE, lag; const char* name; } runs[] = {
        {2, 3, 1, "two issuers, 3 groups x 4 warps "}, {2, 4, 1, "two issuers, 4 groups x 4 warps "}};
    printf("NB = %d buffers, %d keys per tile (E is per tile of that size)\n", NB, KEYS);
    for (auto& r : runs) {
        printf("%s:", r.name);
        for (int E : Es) {
            cudaMemset(d, 0, 148 * 8 * sizeof(long long));
            k<<<148, 640, smem>>>(d, src, tiles, E, r.NE, r.variant, r.lag);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf(" [%s]", cudaGetErrorString(e)); break; }
            long long h[148 * 8]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
            double w = 0;
            for (int b = 0; b < 148; ++b) w += (double)h[b * 8 + 2] / 148;
            printf("  E=%d: %5.0f", E, w / tiles);
        }
        printf("   clk/tile\n");
    }
    return 0;
}
