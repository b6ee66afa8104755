// bf16 mode of the fused NT-Xent kernels on the 5th-gen tensor cores (tcgen05).
//
// One persistent CTA per SM, 10 warps, warp specialised:
//   warps 0-3 / 4-7 : two epilogue warpgroups; tile i of the CTA goes to group i % 2.
//                     Thread = one row of the 128 x 128 similarity tile (TMEM lane).
//   warp 8          : TMA producer (cp.async.bulk.tensor, 128B swizzle, mbarrier tx)
//   warp 9          : TMEM allocator + the one thread that issues tcgen05.mma
//
// Forward  (loss.py:42-61): S = Q . K^T lands in TMEM (4 buffers of 128 fp32 columns);
//   the epilogue reads it with tcgen05.ld, forms e = exp2(s*c1 - c1) (fixed shift
//   1/tau, |cos| <= 1), zeroes the self column / columns past the last key in edge
//   tiles, and keeps per-row sums in registers.  The 2N x 2N logits never leave the SM.
// Backward (autograd of the same lines): recompute S, C = e * (coef_q[r] + coef_k[c])
//   as bf16 into a 128B-swizzled smem tile, second MMA  dZ += C . K  with the SAME key
//   tile read MN-major, accumulator resident in TMEM across the CTA's key tiles.
//
// Work partition and scratch slots: see NtxPlan / ntx_common.cuh.
#include <stdlib.h>

#include "ntx_common.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace umma;

constexpr int kThreads = 320;
constexpr int kProducerWarp = 8;
constexpr int kMmaWarp = 9;
constexpr uint32_t kBlkBytes = 128 * 128;   // one [128 rows x 64 bf16] swizzled block
constexpr uint32_t kTmemCols = 512;

template <int NDB> struct FwdCfg { static constexpr int NS = NDB == 1 ? 6 : (NDB == 2 ? 4 : 2); };
template <int NDB> struct BwdCfg {
    static constexpr int NS = NDB == 1 ? 4 : (NDB == 2 ? 3 : 2);
    static constexpr int NP = NDB == 4 ? 1 : 2;
};

template <int NS> struct FwdCtl {
    uint64_t q_full, q_empty, k_full[NS], k_empty[NS], s_full[4], s_empty[4];
    uint32_t tmem_base;
    alignas(16) float xch[2][128];
};
template <int NS> struct BwdCtl {
    uint64_t q_full, q_empty, k_full[NS], k_empty[NS], s_full[2], s_empty[2], p_full[2], p_empty[2],
        dz_full, dz_empty;
    uint32_t tmem_base;
    alignas(16) float beta[2][2][128];
};

// dynamic smem is declared __align__(1024) (128B-swizzle atoms are 1024 B); refuse to run otherwise
__device__ __forceinline__ uint8_t* checked_1024(uint8_t* p) {
    if (smem_u32(p) & 1023u) {
        if (threadIdx.x == 0) printf("ntxent: dynamic shared memory is not 1024-byte aligned\n");
        __trap();
    }
    return p;
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void named_bar_sync(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
// tcgen05.wait::ld that also names the destination registers, so the compiler cannot
// schedule a use of them above the wait
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&r)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]),
                   "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]),
                   "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]), "+r"(r[17]), "+r"(r[18]),
                   "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                   "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]),
                   "+r"(r[31])
                 :
                 : "memory");
}

// does key tile kt hold a self column of row block rb, or columns past the last key?
__device__ __forceinline__ bool edge_tile(int rb, int kt, const NtxDims& dm) {
    const int col0 = kt * 128;
    const int self_lo = ntx_row_offset(dm.n_local, dm.rank) + rb * 128;
    return (col0 + 128 > dm.k_rows) || (self_lo <= col0 + 127 && self_lo + 127 >= col0);
}

// S = Q . K^T for one key tile: NDB blocks of 64 features, 4 MMAs (K = 16) per block
template <int NDB>
__device__ __forceinline__ void issue_gemm1(uint32_t d_tmem, const uint8_t* q_smem, const uint8_t* k_tile) {
    constexpr uint32_t idesc = instr_desc_bf16(128, 128, 0);
#pragma unroll
    for (int b = 0; b < NDB; ++b)
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
            const uint64_t a = smem_desc_sw128(smem_u32(q_smem + b * kBlkBytes + k4 * 32), 16, 1024);
            const uint64_t bd = smem_desc_sw128(smem_u32(k_tile + b * kBlkBytes + k4 * 32), 16, 1024);
            mma_bf16_ss(d_tmem, a, bd, idesc, (b | k4) ? 1u : 0u);
        }
}

// ============================================================================ forward
template <int NDB>
__global__ void __launch_bounds__(kThreads, 1)
tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              NtxDims dm, int key_tiles, long long T, int rows_pad, float* __restrict__ partial) {
    constexpr int NS = FwdCfg<NDB>::NS;
    constexpr uint32_t kTileBytes = NDB * kBlkBytes;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = checked_1024(smem_raw);
    uint8_t* q_smem = smem;
    uint8_t* k_smem = smem + kTileBytes;
    FwdCtl<NS>* ctl = reinterpret_cast<FwdCtl<NS>*>(k_smem + NS * kTileBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    const long long t_begin = ntx_chunk_begin(T, G, cta);
    const int n_tiles = (int)(ntx_chunk_begin(T, G, cta + 1) - t_begin);

    if (threadIdx.x == 0) {
        mbar_init(&ctl->q_full, 1);
        mbar_init(&ctl->q_empty, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl->k_full[s], 1); mbar_init(&ctl->k_empty[s], 1); }
        for (int b = 0; b < 4; ++b) { mbar_init(&ctl->s_full[b], 1); mbar_init(&ctl->s_empty[b], 4); }
        fence_barrier_init();
    }
    if (warp == kMmaWarp) tmem_alloc(&ctl->tmem_base, kTmemCols);
    if (warp == kProducerWarp && lane == 0) { tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == kProducerWarp) {
        if (lane == 0) {
            int cur_rb = -1, seg = 0;
            for (int i = 0; i < n_tiles; ++i) {
                const long long t = t_begin + i;
                const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
                if (rb != cur_rb) {
                    if (cur_rb >= 0) mbar_wait(&ctl->q_empty, (seg - 1) & 1);
                    mbar_expect_tx(&ctl->q_full, kTileBytes);
#pragma unroll
                    for (int b = 0; b < NDB; ++b)
                        tma_load_2d(q_smem + b * kBlkBytes, &tmap_q, b * 64, rb * 128, &ctl->q_full);
                    cur_rb = rb;
                    ++seg;
                }
                const int s = i % NS;
                mbar_wait(&ctl->k_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&ctl->k_full[s], kTileBytes);
#pragma unroll
                for (int b = 0; b < NDB; ++b)
                    tma_load_2d(k_smem + s * kTileBytes + b * kBlkBytes, &tmap_k, b * 64, kt * 128,
                                &ctl->k_full[s]);
            }
        }
    } else if (warp == kMmaWarp) {
        int cur_rb = -1, seg = -1;
        for (int i = 0; i < n_tiles; ++i) {
            const long long t = t_begin + i;
            const int rb = (int)(t / key_tiles);
            if (rb != cur_rb) {
                ++seg;
                mbar_wait(&ctl->q_full, seg & 1);
                cur_rb = rb;
            }
            const int s = i % NS, buf = i & 3;
            mbar_wait(&ctl->k_full[s], (i / NS) & 1);
            mbar_wait(&ctl->s_empty[buf], ((i >> 2) & 1) ^ 1);
            tc_fence_after();
            if (lane == 0) {
                issue_gemm1<NDB>(tmem_base + buf * 128, q_smem, k_smem + s * kTileBytes);
                mma_commit(&ctl->k_empty[s]);
                mma_commit(&ctl->s_full[buf]);
                const bool last_of_seg = (i + 1 == n_tiles) || ((int)((t + 1) / key_tiles) != rb);
                if (last_of_seg) mma_commit(&ctl->q_empty);
            }
            __syncwarp();
        }
    } else {
        // ---------------------------------------------------------------- epilogue groups
        const int g = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const float c1 = dm.c1, nc1 = -dm.c1;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        int cur_rb = -1, flushes = 0;

        auto flush = [&](int rb) {
            float rs = (acc[0] + acc[1]) + (acc[2] + acc[3]);
            float* x = ctl->xch[flushes & 1];
            if (g == 1) x[row_in_tile] = rs;
            named_bar_sync(1, 256);
            if (g == 0) {
                rs += x[row_in_tile];
                const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
                partial[(size_t)slot * rows_pad + rb * 128 + row_in_tile] = rs;
            }
            ++flushes;
            acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
        };

        for (int i = 0; i < n_tiles; ++i) {
            const long long t = t_begin + i;
            const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
            if (rb != cur_rb) {
                if (cur_rb >= 0) flush(cur_rb);
                cur_rb = rb;
            }
            if ((i & 1) != g) continue;
            const int buf = i & 3;
            mbar_wait(&ctl->s_full[buf], (i >> 2) & 1);
            tc_fence_after();
            const uint32_t taddr = t_lane + buf * 128;
            const bool edge = edge_tile(rb, kt, dm);
            const int self = ntx_self_col(rb * 128 + row_in_tile, dm.n_local, dm.rank);
            uint32_t v[2][32];
            tmem_ld_32x32(taddr, v[0]);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                tmem_ld_wait_dep(v[ch & 1]);
                if (ch < 3) {
                    tmem_ld_32x32(taddr + (ch + 1) * 32, v[(ch + 1) & 1]);
                } else {   // the whole tile is in registers: hand the TMEM buffer back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ctl->s_empty[buf]);
                }
                if (!edge) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        acc[j & 3] += ex2(fmaf(__uint_as_float(v[ch & 1][j]), c1, nc1));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int col = kt * 128 + ch * 32 + j;
                        float e = ex2(fmaf(__uint_as_float(v[ch & 1][j]), c1, nc1));
                        if (col == self || col >= dm.k_rows) e = 0.f;   // mask of loss.py:38,46-47
                        acc[j & 3] += e;
                    }
                }
            }
        }
        if (cur_rb >= 0) flush(cur_rb);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// ============================================================================ backward
template <int NDB>
__global__ void __launch_bounds__(kThreads, 1)
tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              NtxDims dm, int key_tiles, long long T, int rows_pad, const float* __restrict__ coef_q,
              const float* __restrict__ coef_k, float* __restrict__ partial, uint32_t b2_lbo,
              uint32_t b2_sbo) {
    constexpr int NS = BwdCfg<NDB>::NS, NP = BwdCfg<NDB>::NP;
    constexpr int DP = NDB * 64;                      // dZ columns (fp32) held in TMEM [0, DP)
    constexpr uint32_t kSCol = 256;                   // S buffers at TMEM columns 256 and 384
    constexpr uint32_t kTileBytes = NDB * kBlkBytes;
    constexpr uint32_t kPBytes = 2 * kBlkBytes;       // [128 rows x 128 keys] bf16
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = checked_1024(smem_raw);
    uint8_t* q_smem = smem;
    uint8_t* k_smem = smem + kTileBytes;
    uint8_t* p_smem = k_smem + NS * kTileBytes;
    BwdCtl<NS>* ctl = reinterpret_cast<BwdCtl<NS>*>(p_smem + NP * kPBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    const long long t_begin = ntx_chunk_begin(T, G, cta);
    const int n_tiles = (int)(ntx_chunk_begin(T, G, cta + 1) - t_begin);

    if (threadIdx.x == 0) {
        mbar_init(&ctl->q_full, 1);
        mbar_init(&ctl->q_empty, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl->k_full[s], 1); mbar_init(&ctl->k_empty[s], 1); }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&ctl->s_full[b], 1);
            mbar_init(&ctl->s_empty[b], 4);
            mbar_init(&ctl->p_full[b], 4);
            mbar_init(&ctl->p_empty[b], 1);
        }
        mbar_init(&ctl->dz_full, 1);
        mbar_init(&ctl->dz_empty, 8);
        fence_barrier_init();
    }
    if (warp == kMmaWarp) tmem_alloc(&ctl->tmem_base, kTmemCols);
    if (warp == kProducerWarp && lane == 0) { tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;

    if (warp == kProducerWarp) {
        if (lane == 0) {
            int cur_rb = -1, seg = 0;
            for (int i = 0; i < n_tiles; ++i) {
                const long long t = t_begin + i;
                const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
                if (rb != cur_rb) {
                    if (cur_rb >= 0) mbar_wait(&ctl->q_empty, (seg - 1) & 1);
                    mbar_expect_tx(&ctl->q_full, kTileBytes);
#pragma unroll
                    for (int b = 0; b < NDB; ++b)
                        tma_load_2d(q_smem + b * kBlkBytes, &tmap_q, b * 64, rb * 128, &ctl->q_full);
                    cur_rb = rb;
                    ++seg;
                }
                const int s = i % NS;
                mbar_wait(&ctl->k_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&ctl->k_full[s], kTileBytes);
#pragma unroll
                for (int b = 0; b < NDB; ++b)
                    tma_load_2d(k_smem + s * kTileBytes + b * kBlkBytes, &tmap_k, b * 64, kt * 128,
                                &ctl->k_full[s]);
            }
        }
    } else if (warp == kMmaWarp) {
        constexpr uint32_t idesc2 = instr_desc_bf16(128, DP, 1);
        int seg2 = -1;
        // dZ += C(tile j) . K(tile j): 8 MMAs of K = 16 keys, N = DP features
        auto gemm2 = [&](int j) {
            const long long tj = t_begin + j;
            const int rbj = (int)(tj / key_tiles);
            const bool first_of_seg = (j == 0) || ((int)((tj - 1) / key_tiles) != rbj);
            const bool last_of_seg = (j + 1 == n_tiles) || ((int)((tj + 1) / key_tiles) != rbj);
            if (first_of_seg) {
                ++seg2;
                mbar_wait(&ctl->dz_empty, (seg2 & 1) ^ 1);   // previous row block drained
            }
            const int pb = j % NP, sj = j % NS;
            mbar_wait(&ctl->p_full[pb], (j / NP) & 1);
            tc_fence_after();
            if (lane == 0) {
                const uint8_t* p_tile = p_smem + pb * kPBytes;
                const uint8_t* k_tile = k_smem + sj * kTileBytes;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint64_t a = smem_desc_sw128(
                        smem_u32(p_tile + (kk >> 2) * kBlkBytes + (kk & 3) * 32), 16, 1024);
                    const uint64_t bd = smem_desc_sw128(smem_u32(k_tile + kk * 2048), b2_lbo, b2_sbo);
                    mma_bf16_ss(tmem_base, a, bd, idesc2, (first_of_seg && kk == 0) ? 0u : 1u);
                }
                mma_commit(&ctl->k_empty[sj]);
                mma_commit(&ctl->p_empty[pb]);
                if (last_of_seg) mma_commit(&ctl->dz_full);
            }
            __syncwarp();
        };

        int cur_rb = -1, seg = -1;
        for (int i = 0; i < n_tiles; ++i) {
            const long long t = t_begin + i;
            const int rb = (int)(t / key_tiles);
            if (rb != cur_rb) {
                ++seg;
                mbar_wait(&ctl->q_full, seg & 1);
                cur_rb = rb;
            }
            const int s = i % NS, sb = i & 1;
            mbar_wait(&ctl->k_full[s], (i / NS) & 1);
            mbar_wait(&ctl->s_empty[sb], ((i >> 1) & 1) ^ 1);
            tc_fence_after();
            if (lane == 0) {
                issue_gemm1<NDB>(tmem_base + kSCol + sb * 128, q_smem, k_smem + s * kTileBytes);
                mma_commit(&ctl->s_full[sb]);
                const bool last_of_seg = (i + 1 == n_tiles) || ((int)((t + 1) / key_tiles) != rb);
                if (last_of_seg) mma_commit(&ctl->q_empty);
            }
            __syncwarp();
            if (i > 0) gemm2(i - 1);
        }
        if (n_tiles > 0) gemm2(n_tiles - 1);
    } else {
        // ---------------------------------------------------------------- epilogue groups
        const int g = warp >> 2;
        const int tid_g = threadIdx.x & 127;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const float c1 = dm.c1, nc1 = -dm.c1;
        const uint32_t sw = (uint32_t)(row_in_tile & 7);
        int cur_rb = -1, seg = -1, mine = 0;
        float alpha = 0.f;
        bool row_ok = false;

        // move this group's half of the dZ accumulator to its scratch slot
        auto drain = [&](int rb, int sg) {
            mbar_wait(&ctl->dz_full, sg & 1);
            tc_fence_after();
            const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
            float* dst = partial + ((size_t)slot * rows_pad + rb * 128 + row_in_tile) * DP + g * (DP / 2);
#pragma unroll
            for (int c = 0; c < DP / 2; c += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t_lane + g * (DP / 2) + c, v);
                tmem_ld_wait_dep(v);
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<uint4*>(dst + c + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->dz_empty);
        };

        for (int i = 0; i < n_tiles; ++i) {
            const long long t = t_begin + i;
            const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
            if (rb != cur_rb) {
                if (cur_rb >= 0) drain(cur_rb, seg);
                cur_rb = rb;
                ++seg;
                const int row = rb * 128 + row_in_tile;
                row_ok = row < dm.q_rows;
                alpha = row_ok ? __ldg(coef_q + row) : 0.f;
            }
            if ((i & 1) != g) continue;

            float* bt = ctl->beta[g][mine & 1];
            {
                const int c = kt * 128 + tid_g;
                bt[tid_g] = c < dm.k_rows ? __ldg(coef_k + c) : 0.f;
            }
            ++mine;
            named_bar_sync(2 + g, 128);

            const int sb = i & 1, pb = i % NP;
            mbar_wait(&ctl->s_full[sb], (i >> 1) & 1);
            tc_fence_after();
            mbar_wait(&ctl->p_empty[pb], ((i / NP) & 1) ^ 1);   // GEMM2 that read this P buffer is done

            const uint32_t taddr = t_lane + kSCol + sb * 128;
            const bool edge = edge_tile(rb, kt, dm) || (rb * 128 + 128 > dm.q_rows);
            const int self = ntx_self_col(rb * 128 + row_in_tile, dm.n_local, dm.rank);
            uint8_t* p_row = p_smem + pb * kPBytes + row_in_tile * 128;
            uint32_t v[2][32];
            tmem_ld_32x32(taddr, v[0]);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                tmem_ld_wait_dep(v[ch & 1]);
                if (ch < 3) {
                    tmem_ld_32x32(taddr + (ch + 1) * 32, v[(ch + 1) & 1]);
                } else {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ctl->s_empty[sb]);
                }
#pragma unroll
                for (int j8 = 0; j8 < 4; ++j8) {           // 8 keys -> one 16-byte swizzled chunk
                    uint32_t packed[4];
                    const float4 b0 = *reinterpret_cast<const float4*>(bt + ch * 32 + j8 * 8);
                    const float4 b1 = *reinterpret_cast<const float4*>(bt + ch * 32 + j8 * 8 + 4);
                    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                    float cv[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const float e = ex2(fmaf(__uint_as_float(v[ch & 1][j8 * 8 + j]), c1, nc1));
                        cv[j] = e * (alpha + bb[j]);
                    }
                    if (edge) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const int col = kt * 128 + ch * 32 + j8 * 8 + j;
                            if (col == self || col >= dm.k_rows || !row_ok) cv[j] = 0.f;
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        __nv_bfloat162 h = __floats2bfloat162_rn(cv[2 * j], cv[2 * j + 1]);
                        packed[j] = *reinterpret_cast<uint32_t*>(&h);
                    }
                    const int cg = ch * 4 + j8;             // 16-byte chunk index along the 128 keys
                    uint8_t* dst = p_row + (cg >> 3) * kBlkBytes + ((((uint32_t)cg & 7u) ^ sw) << 4);
                    *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                }
            }
            fence_proxy_async_smem();      // generic-proxy stores -> visible to tcgen05.mma
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->p_full[pb]);
        }
        if (cur_rb >= 0) drain(cur_rb, seg);
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// ============================================================================ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

int get_encode_fn(EncodeTiledFn* out) {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        NTX_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        if (!p || qres != cudaDriverEntryPointSuccess) {
            ntx_set_error("cuTensorMapEncodeTiled not available from the driver");
            return NTX_E_CUDA;
        }
        fn = (EncodeTiledFn)p;
    }
    *out = fn;
    return NTX_OK;
}

// [rows x ld] bf16 row buffer -> tiles of 128 rows x 64 features, 128B swizzle, zero OOB fill
int make_rows_tmap(CUtensorMap* m, const __nv_bfloat16* base, int rows, int ld) {
    EncodeTiledFn enc;
    int rc = get_encode_fn(&enc);
    if (rc) return rc;
    const cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(__nv_bfloat16)};
    const cuuint32_t box[2] = {64, 128};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        ntx_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d ld=%d)", (int)r, rows, ld);
        return NTX_E_CUDA;
    }
    return NTX_OK;
}

template <int NDB> size_t fwd_smem_bytes() {
    return (size_t)(1 + FwdCfg<NDB>::NS) * NDB * kBlkBytes + sizeof(FwdCtl<FwdCfg<NDB>::NS>);
}
template <int NDB> size_t bwd_smem_bytes() {
    return (size_t)(1 + BwdCfg<NDB>::NS) * NDB * kBlkBytes + (size_t)BwdCfg<NDB>::NP * 2 * kBlkBytes +
           sizeof(BwdCtl<BwdCfg<NDB>::NS>);
}

template <int NDB>
int launch_fwd(const CUtensorMap& mq, const CUtensorMap& mk, const NtxDims& dm, const NtxPlan& plan,
               float* partial, cudaStream_t st) {
    const size_t smem = fwd_smem_bytes<NDB>();
    NTX_CUDA(cudaFuncSetAttribute(tc_fwd_kernel<NDB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    tc_fwd_kernel<NDB><<<plan.grid, kThreads, smem, st>>>(mq, mk, dm, plan.key_tiles, T,
                                                           plan.row_blocks * plan.tile_rows, partial);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

template <int NDB>
int launch_bwd(const CUtensorMap& mq, const CUtensorMap& mk, const NtxDims& dm, const NtxPlan& plan,
               const float* coef_q, const float* coef_k, float* partial, cudaStream_t st) {
    const size_t smem = bwd_smem_bytes<NDB>();
    NTX_CUDA(cudaFuncSetAttribute(tc_bwd_kernel<NDB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    // MN-major key-tile descriptor: next 64 features one 16 KB block away, 8-key groups 1 KB apart.
    uint32_t lbo = kBlkBytes, sbo = 1024;
    if (const char* e = getenv("NTX_DEBUG_B2_LBO")) lbo = (uint32_t)atoi(e);   // bring-up only
    if (const char* e = getenv("NTX_DEBUG_B2_SBO")) sbo = (uint32_t)atoi(e);
    tc_bwd_kernel<NDB><<<plan.grid, kThreads, smem, st>>>(mq, mk, dm, plan.key_tiles, T,
                                                           plan.row_blocks * plan.tile_rows, coef_q,
                                                           coef_k, partial, lbo, sbo);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

}  // namespace

int ntx_tc_fwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, float* partial, cudaStream_t st) {
    CUtensorMap mq, mk;
    int rc;
    if ((rc = make_rows_tmap(&mq, q, dm.q_rows, dm.ld)) || (rc = make_rows_tmap(&mk, k, dm.k_rows, dm.ld)))
        return rc;
    switch (plan.d_pad) {
        case 64: return launch_fwd<1>(mq, mk, dm, plan, partial, st);
        case 128: return launch_fwd<2>(mq, mk, dm, plan, partial, st);
        case 256: return launch_fwd<4>(mq, mk, dm, plan, partial, st);
    }
    ntx_set_error("tensor-core forward: padded feature dim %d not built", plan.d_pad);
    return NTX_E_UNSUPPORTED;
}

int ntx_tc_bwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, const float* coef_q, const float* coef_k, float* partial,
                      cudaStream_t st) {
    CUtensorMap mq, mk;
    int rc;
    if ((rc = make_rows_tmap(&mq, q, dm.q_rows, dm.ld)) || (rc = make_rows_tmap(&mk, k, dm.k_rows, dm.ld)))
        return rc;
    switch (plan.d_pad) {
        case 64: return launch_bwd<1>(mq, mk, dm, plan, coef_q, coef_k, partial, st);
        case 128: return launch_bwd<2>(mq, mk, dm, plan, coef_q, coef_k, partial, st);
        case 256: return launch_bwd<4>(mq, mk, dm, plan, coef_q, coef_k, partial, st);
    }
    ntx_set_error("tensor-core backward: padded feature dim %d not built", plan.d_pad);
    return NTX_E_UNSUPPORTED;
}
