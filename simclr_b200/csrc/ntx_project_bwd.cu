// Backward of the projection head's last layer: both GEMMs of
//   z = [h0 ; h1] . W^T        (Linear(h -> d, bias=False), model.py:69)
// in one tcgen05 kernel, given dz = d loss / d z from ntx_bwd_rows:
//   dh = dz . W          (2n x hdim)   -- per 128-row tile, fresh accumulator
//   dW = dz^T . [h0;h1]  (d x hdim)    -- split over the row tiles, accumulated in TMEM across a CTA's tiles
// The reference gets both from autograd as two cuBLAS GEMMs plus a transpose; here one dz tile in shared memory
// feeds both MMAs: read K-major as the A operand of dh, and MN-major (transposed by the descriptor, no data
// movement) as the A operand of dW.  W and h tiles are B operands read MN-major straight from their row-major
// layouts.  CTA (j, nt) owns hdim columns [128 nt, 128 nt + 128) and a contiguous range of row tiles; its W tile
// stays resident.  dW partials go to one scratch slot per CTA and are added in CTA order by a second small kernel
// (deterministic).  d <= 128 (one M tile); HBM-bound: h (2n x hdim) is read once, dh written once.
#include "ntx_common.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace umma;

constexpr int kThreads = 192;                 // 4 epilogue warps + TMA warp + MMA warp
constexpr uint32_t kBlk = 128 * 128;          // [128 rows x 64 bf16], 128B swizzle
constexpr int kStages = 2;
constexpr uint32_t kStageBytes = 4 * kBlk;    // dz tile (2 blocks: 128 features) + h tile (2 blocks: 128 columns)
constexpr uint32_t kColDh = 0, kColDw = 256;  // TMEM: two dh accumulators of 128 columns, then the dW accumulator

struct PbCtl {
    uint64_t w_full, full[kStages], empty[kStages], dh_full[2], dh_empty[2], dw_full;
    uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
project_bwd_kernel(const __grid_constant__ CUtensorMap tmap_dz0, const __grid_constant__ CUtensorMap tmap_dz1,
                   const __grid_constant__ CUtensorMap tmap_h0, const __grid_constant__ CUtensorMap tmap_h1,
                   const __grid_constant__ CUtensorMap tmap_w, int n, int hdim, int tiles_per_view,
                   __nv_bfloat16* __restrict__ dh0, __nv_bfloat16* __restrict__ dh1, float* __restrict__ partial) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    if (smem_u32(smem) & 1023u) __trap();
    uint8_t* w_smem = smem;                               // [128 d-rows x 128 h-columns]: 2 blocks
    uint8_t* ring = smem + 2 * kBlk;
    PbCtl* ctl = reinterpret_cast<PbCtl*>(ring + kStages * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, j = blockIdx.x, nt = blockIdx.y;
    const int RT = 2 * tiles_per_view;
    const int t_begin = (int)((long long)j * RT / G), t_end = (int)((long long)(j + 1) * RT / G);
    const int n_tiles = t_end - t_begin;

    if (threadIdx.x == 0) {
        mbar_init(&ctl->w_full, 1);
        for (int s = 0; s < kStages; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&ctl->dh_full[b], 1); mbar_init(&ctl->dh_empty[b], 4); }
        mbar_init(&ctl->dw_full, 1);
        fence_barrier_init();
    }
    if (warp == 5) tmem_alloc(&ctl->tmem_base, 512);
    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmap_dz0); tma_prefetch_desc(&tmap_dz1);
        tma_prefetch_desc(&tmap_h0); tma_prefetch_desc(&tmap_h1);
        tma_prefetch_desc(&tmap_w);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    ntx_dep_trigger();
    ntx_dep_wait();

    if (warp == 4) {
        if (elect_one()) {
            mbar_expect_tx(&ctl->w_full, 2 * kBlk);
            for (int b = 0; b < 2; ++b) tma_load_2d(w_smem + b * kBlk, &tmap_w, nt * 128 + b * 64, 0, &ctl->w_full);
            for (int i = 0; i < n_tiles; ++i) {
                const int t = t_begin + i, view = t / tiles_per_view, tile = t % tiles_per_view;
                const int s = i % kStages;
                mbar_wait(&ctl->empty[s], ((i / kStages) & 1) ^ 1);
                mbar_expect_tx(&ctl->full[s], kStageBytes);
                uint8_t* st = ring + s * kStageBytes;
                const CUtensorMap* mz = view == 0 ? &tmap_dz0 : &tmap_dz1;
                const CUtensorMap* mh = view == 0 ? &tmap_h0 : &tmap_h1;
                for (int b = 0; b < 2; ++b) tma_load_2d(st + b * kBlk, mz, b * 64, tile * 128, &ctl->full[s]);
                for (int b = 0; b < 2; ++b)
                    tma_load_2d(st + (2 + b) * kBlk, mh, nt * 128 + b * 64, tile * 128, &ctl->full[s]);
            }
        }
    } else if (warp == 5) {
        constexpr uint32_t idesc_dh = instr_desc_bf16(128, 128, 1);          // A K-major, B MN-major
        constexpr uint32_t idesc_dw = instr_desc_bf16_ex(128, 128, 1, 1);    // A and B MN-major
        mbar_wait(&ctl->w_full, 0);
        for (int i = 0; i < n_tiles; ++i) {
            const int s = i % kStages, b = i & 1;
            mbar_wait(&ctl->full[s], (i / kStages) & 1);
            mbar_wait(&ctl->dh_empty[b], ((i >> 1) & 1) ^ 1);
            tc_fence_after();
            if (elect_one()) {
                const uint8_t* dz_t = ring + s * kStageBytes;
                const uint8_t* h_t = dz_t + 2 * kBlk;
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    // dh tile = dz (128 rows x 128 features, K-major) . W (features x 128 columns, MN-major)
                    const uint64_t a = smem_desc_sw128(smem_u32(dz_t + (kk >> 2) * kBlk + (kk & 3) * 32), 16, 1024);
                    const uint64_t bd = smem_desc_sw128(smem_u32(w_smem + kk * 2048), kBlk, 1024);
                    mma_bf16_ss(tmem_base + kColDh + b * 128, a, bd, idesc_dh, kk ? 1u : 0u);
                }
                mma_commit(&ctl->dh_full[b]);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    // dW tile += dz^T (128 features x 128 rows: the same tile, MN-major) . h (rows x 128 columns, MN-major)
                    const uint64_t a = smem_desc_sw128(smem_u32(dz_t + kk * 2048), kBlk, 1024);
                    const uint64_t bd = smem_desc_sw128(smem_u32(h_t + kk * 2048), kBlk, 1024);
                    mma_bf16_ss(tmem_base + kColDw, a, bd, idesc_dw, (i | kk) ? 1u : 0u);
                }
                mma_commit(&ctl->empty[s]);
                if (i + 1 == n_tiles) mma_commit(&ctl->dw_full);
            }
            __syncwarp();
        }
    } else {
        // epilogue: thread = accumulator lane.  dh: lane = row of the tile; dW: lane = feature index.
        const int lane_row = warp * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
        for (int i = 0; i < n_tiles; ++i) {
            const int t = t_begin + i, view = t / tiles_per_view, tile = t % tiles_per_view;
            const int b = i & 1;
            mbar_wait(&ctl->dh_full[b], (i >> 1) & 1);
            tc_fence_after();
            const int r = tile * 128 + lane_row;
            __nv_bfloat16* dst = (view == 0 ? dh0 : dh1) + (size_t)r * hdim + nt * 128;
#pragma unroll 1
            for (int c = 0; c < 128; c += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t_lane + kColDh + b * 128 + c, v);
                tmem_ld_wait();
                if (r < n) {
#pragma unroll
                    for (int j8 = 0; j8 < 4; ++j8) {
                        uint32_t pk[4];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            __nv_bfloat162 hh = __floats2bfloat162_rn(__uint_as_float(v[j8 * 8 + q * 2]),
                                                                      __uint_as_float(v[j8 * 8 + q * 2 + 1]));
                            pk[q] = *reinterpret_cast<uint32_t*>(&hh);
                        }
                        if (nt * 128 + c + j8 * 8 < hdim)          // hdim is a multiple of 64: whole 16-byte groups
                            *reinterpret_cast<uint4*>(dst + c + j8 * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->dh_empty[b]);
        }
        // this CTA's share of dW: [128 features x 128 columns] fp32 into its slot
        float* slot = partial + ((size_t)(nt * G + j) * 128 + lane_row) * 128;
        if (n_tiles > 0) {
            mbar_wait(&ctl->dw_full, 0);
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 128; c += 32) {
                uint32_t v[32];
                tmem_ld_32x32(t_lane + kColDw + c, v);
                tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 32; q += 4)
                    *reinterpret_cast<uint4*>(slot + c + q) = make_uint4(v[q], v[q + 1], v[q + 2], v[q + 3]);
            }
        } else {
            for (int c = 0; c < 128; c += 4) *reinterpret_cast<float4*>(slot + c) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(tmem_base, 512);
}

// dW[f][c] = sum over the CTAs j of a column tile, in j order, of partial[(nt, j)][f][c]  -> bf16
__global__ void __launch_bounds__(256)
project_bwd_reduce_kernel(const float* __restrict__ partial, int G, int d, int hdim, __nv_bfloat16* __restrict__ dw) {
    ntx_dep_trigger();
    ntx_dep_wait();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;      // one thread per 4 columns
    const int c4 = (idx % (hdim / 4)) * 4, f = idx / (hdim / 4);
    if (f >= d) return;
    const int nt = c4 / 128, c = c4 % 128;
    const float* src = partial + ((size_t)nt * G * 128 + f) * 128 + c;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j0 = 0; j0 < G; j0 += 4) {
        float4 t[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            t[u] = j0 + u < G ? __ldg(reinterpret_cast<const float4*>(src + (size_t)(j0 + u) * 128 * 128))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int u = 0; u < 4; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
    }
    __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t*>(&lo);
    o.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(dw + (size_t)f * hdim + c4) = o;
}

int grid_rows(int n, int hdim, int sm_count) {
    const int RT = 2 * ((n + 127) / 128), NT = (hdim + 127) / 128;
    int g = sm_count / NT;
    if (g < 1) g = 1;
    if (g > RT) g = RT;
    return g;
}

int sm_count_now(int* out) {
    static int counts[64];
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) dev = 0;
    if (!counts[dev]) NTX_CUDA(cudaDeviceGetAttribute(&counts[dev], cudaDevAttrMultiProcessorCount, dev));
    *out = counts[dev];
    return NTX_OK;
}

}  // namespace

extern "C" {

size_t ntx_project_bwd_workspace_bytes(int n, int hdim, int sm_count) {
    if (n < 1 || hdim < 64) return 0;
    if (sm_count <= 0 && sm_count_now(&sm_count)) return 0;
    const int NT = (hdim + 127) / 128;
    return (size_t)NT * grid_rows(n, hdim, sm_count) * 128 * 128 * sizeof(float);
}

int ntx_project_bwd(const void* dz, const void* h0, const void* h1, const void* weight, int n, int hdim, int d,
                    void* dh0, void* dh1, void* dweight, void* workspace, size_t workspace_bytes, void* stream) {
    NTX_CHECK_ARG(n >= 1 && d >= 1 && hdim >= 64, "bad sizes n=%d hdim=%d d=%d", n, hdim, d);
    NTX_CHECK_ARG(hdim % 64 == 0, "hidden size %d must be a multiple of 64 (one 128-byte TMA row of bf16)", hdim);
    NTX_CHECK_ARG(d <= 128 && d % 8 == 0, "ntx_project_bwd is built for d <= 128, a multiple of 8 (got %d)", d);
    NTX_CHECK_ARG(dz && h0 && h1 && weight && dh0 && dh1 && dweight && workspace, "NULL buffer");
    for (const void* p : {dz, h0, h1, weight, (const void*)dh0, (const void*)dh1, (const void*)dweight,
                          (const void*)workspace})
        NTX_CHECK_ARG(reinterpret_cast<uintptr_t>(p) % 16 == 0, "buffers must be 16-byte aligned");
    NTX_CHECK_ARG(((size_t)n * d * 2) % 16 == 0, "n * d must be a multiple of 8 (the second view's dz block starts there)");
    int sms = 0, rc;
    if ((rc = sm_count_now(&sms))) return rc;
    const size_t need = ntx_project_bwd_workspace_bytes(n, hdim, sms);
    if (workspace_bytes < need) {
        ntx_set_error("ntx_project_bwd needs %llu bytes of workspace, got %llu", (unsigned long long)need,
                      (unsigned long long)workspace_bytes);
        return NTX_E_WORKSPACE;
    }
    const __nv_bfloat16* dzb = static_cast<const __nv_bfloat16*>(dz);
    CUtensorMap mz0, mz1, mh0, mh1, mw;
    if ((rc = ntx_make_rows_tmap(&mz0, dzb, n, d, 128)) || (rc = ntx_make_rows_tmap(&mz1, dzb + (size_t)n * d, n, d, 128)) ||
        (rc = ntx_make_rows_tmap(&mh0, h0, n, hdim, 128)) || (rc = ntx_make_rows_tmap(&mh1, h1, n, hdim, 128)) ||
        (rc = ntx_make_rows_tmap(&mw, weight, d, hdim, 128)))
        return rc;
    const int tiles_per_view = (n + 127) / 128, NT = (hdim + 127) / 128;
    const int G = grid_rows(n, hdim, sms);
    const size_t smem = 2 * kBlk + kStages * kStageBytes + sizeof(PbCtl);
    NTX_SMEM_OPT_IN(project_bwd_kernel, smem);
    cudaStream_t st = (cudaStream_t)stream;
    NTX_CUDA(ntx_launch_dep(project_bwd_kernel, dim3(G, NT), kThreads, smem, st, mz0, mz1, mh0, mh1, mw, n, hdim,
                            tiles_per_view, (__nv_bfloat16*)dh0, (__nv_bfloat16*)dh1, (float*)workspace));
    const int threads = d * (hdim / 4);
    NTX_CUDA(ntx_launch_dep(project_bwd_reduce_kernel, (threads + 255) / 256, 256, 0, st, (const float*)workspace, G, d,
                            hdim, (__nv_bfloat16*)dweight));
    return NTX_OK;
}

}  // extern "C"
