// Projection head's last layer fused with the normalise:
//   rows = normalize([h0 ; h1] . W^T)        h: (n x hdim) bf16, W: (d x hdim) bf16 (nn.Linear layout, bias=False)
// replaces the reference's Linear(h -> d, bias=False) at model.py:69 followed by F.normalize at loss.py:33-34;
// the (2n x d) projections z are never written to memory, only their normalised rows and 1/||z||.
//
// One CTA per 128-row tile of one view: TMA streams 64-feature blocks of the h tile and of W through a
// 4- to 6-stage ring (128B swizzle), tcgen05.mma (M128 x N = d_pad x K16, bf16 -> fp32) accumulates z in TMEM
// over hdim / 16 steps, and the four epilogue warps (one thread per row) read the accumulator twice:
// sum of squares, then z / max(||z||, 1e-12) packed to bf16 and stored.  HBM-bound: the h tile (128 x hdim
// bf16) is read once, W (d x hdim) comes from L2.
#include "ntx_common.cuh"
#include "ntx_umma.cuh"

namespace {

using namespace umma;

// ring depth: 6 x (16 KB of h + 16 KB of W) = 192 KB at d_pad <= 128, 4 x 48 KB at d_pad = 256
template <int NP> struct ProjStages { static constexpr int N = NP == 256 ? 4 : 6; };
constexpr int kMaxStages = 6;
constexpr int kThreads = 192;                 // 4 epilogue warps + TMA warp + MMA warp
constexpr uint32_t kABytes = 128 * 128;       // [128 rows x 64 bf16]

struct ProjCtl {
    uint64_t full[kMaxStages], empty[kMaxStages], acc_full;
    uint32_t tmem_base;
};

template <int NP>   // NP = padded output width (64, 128 or 256) = MMA N = TMEM columns
__global__ void __launch_bounds__(kThreads, 1)
project_l2norm_kernel(const __grid_constant__ CUtensorMap tmap_h0, const __grid_constant__ CUtensorMap tmap_h1,
                      const __grid_constant__ CUtensorMap tmap_w, int n, int hdim, int d, int tiles_per_view,
                      __nv_bfloat16* __restrict__ rows, int ld, float* __restrict__ inv_norm, float row_scale) {
    constexpr uint32_t kBBytes = NP * 128;    // [NP rows x 64 bf16]
    constexpr uint32_t kStageBytes = kABytes + kBBytes;
    constexpr int kStages = ProjStages<NP>::N;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = smem_raw;
    if (smem_u32(smem) & 1023u) __trap();
    ProjCtl* ctl = reinterpret_cast<ProjCtl*>(smem + kStages * kStageBytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int view = blockIdx.x / tiles_per_view, tile = blockIdx.x % tiles_per_view;
    const int kblocks = hdim / 64;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(&ctl->full[s], 1); mbar_init(&ctl->empty[s], 1); }
        mbar_init(&ctl->acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 5) tmem_alloc(&ctl->tmem_base, NP < 32 ? 32 : NP);
    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmap_h0);
        tma_prefetch_desc(&tmap_h1);
        tma_prefetch_desc(&tmap_w);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = ctl->tmem_base;
    // our TMEM is allocated: only now may the next kernel on the stream be staged (its CTAs allocate TMEM too)
    ntx_dep_trigger();
    ntx_dep_wait();

    if (warp == 4) {
        if (elect_one()) {
            const CUtensorMap* th = view == 0 ? &tmap_h0 : &tmap_h1;
            for (int kb = 0; kb < kblocks; ++kb) {
                const int s = kb % kStages;
                mbar_wait(&ctl->empty[s], ((kb / kStages) & 1) ^ 1);
                mbar_expect_tx(&ctl->full[s], kStageBytes);
                tma_load_2d(smem + s * kStageBytes, th, kb * 64, tile * 128, &ctl->full[s]);
                tma_load_2d(smem + s * kStageBytes + kABytes, &tmap_w, kb * 64, 0, &ctl->full[s]);
            }
        }
    } else if (warp == 5) {
        constexpr uint32_t idesc = instr_desc_bf16(128, NP, 0);
        for (int kb = 0; kb < kblocks; ++kb) {
            const int s = kb % kStages;
            mbar_wait(&ctl->full[s], (kb / kStages) & 1);
            tc_fence_after();
            if (elect_one()) {
                const uint8_t* a_tile = smem + s * kStageBytes;
                const uint8_t* b_tile = a_tile + kABytes;
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4) {
                    const uint64_t a = smem_desc_sw128(smem_u32(a_tile + k4 * 32), 16, 1024);
                    const uint64_t b = smem_desc_sw128(smem_u32(b_tile + k4 * 32), 16, 1024);
                    mma_bf16_ss(tmem_base, a, b, idesc, (kb | k4) ? 1u : 0u);
                }
                mma_commit(&ctl->empty[s]);
                if (kb + 1 == kblocks) mma_commit(&ctl->acc_full);
            }
            __syncwarp();
        }
    } else {
        // one thread per row: lane i of warp w owns accumulator lane 32 w + i
        const int r_in_tile = warp * 32 + lane;
        const int r_view = tile * 128 + r_in_tile;
        const uint32_t t_lane = tmem_base + ((uint32_t)(warp * 32) << 16);
        mbar_wait(&ctl->acc_full, 0);
        tc_fence_after();
        float ss = 0.f;
#pragma unroll 1
        for (int c = 0; c < NP; c += 32) {
            uint32_t v[32];
            tmem_ld_32x32(t_lane + c, v);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float x = (c + j < d) ? __uint_as_float(v[j]) : 0.f;
                ss = fmaf(x, x, ss);
            }
        }
        const float denom = fmaxf(sqrtf(ss), 1e-12f);          // clamp_min(eps) of F.normalize (loss.py:33-34)
        const bool live = r_view < n;
        const size_t r_out = (size_t)view * n + r_view;
        if (live) inv_norm[r_out] = 1.0f / denom;
        __nv_bfloat16* dst = rows + r_out * ld;
#pragma unroll 1
        for (int c = 0; c < NP && c < ld; c += 32) {
            uint32_t v[32];
            tmem_ld_32x32(t_lane + c, v);
            tmem_ld_wait();
            if (live) {
#pragma unroll
                for (int j8 = 0; j8 < 4; ++j8) {
                    uint32_t pk[4];
#pragma unroll
                    for (int h = 0; h < 4; ++h) {
                        const int j = j8 * 8 + h * 2;
                        const float x0 = (c + j < d) ? __uint_as_float(v[j]) / denom * row_scale : 0.f;
                        const float x1 = (c + j + 1 < d) ? __uint_as_float(v[j + 1]) / denom * row_scale : 0.f;
                        __nv_bfloat162 hh = __floats2bfloat162_rn(x0, x1);
                        pk[h] = *reinterpret_cast<uint32_t*>(&hh);
                    }
                    if (c + j8 * 8 < ld) *reinterpret_cast<uint4*>(dst + c + j8 * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(tmem_base, NP < 32 ? 32 : NP);
}

template <int NP>
int launch(const CUtensorMap& m0, const CUtensorMap& m1, const CUtensorMap& mw, int n, int hdim, int d, void* rows,
           int ld, float* inv_norm, float row_scale, cudaStream_t st) {
    const size_t smem = (size_t)ProjStages<NP>::N * (kABytes + NP * 128) + sizeof(ProjCtl);
    NTX_SMEM_OPT_IN(project_l2norm_kernel<NP>, smem);
    const int tiles_per_view = (n + 127) / 128;
    NTX_CUDA(ntx_launch_dep(project_l2norm_kernel<NP>, 2 * tiles_per_view, kThreads, smem, st, m0, m1, mw, n, hdim, d,
                            tiles_per_view, (__nv_bfloat16*)rows, ld, inv_norm, row_scale));
    return NTX_OK;
}

}  // namespace

int ntx_project_launch(const void* h0, const void* h1, const void* weight, int n, int hdim, int d, void* rows, int ld,
                       float* inv_norm, float row_scale, cudaStream_t st) {
    CUtensorMap m0, m1, mw;
    int rc;
    if ((rc = ntx_make_rows_tmap(&m0, h0, n, hdim, 128)) || (rc = ntx_make_rows_tmap(&m1, h1, n, hdim, 128)) ||
        (rc = ntx_make_rows_tmap(&mw, weight, d, hdim, ld)))
        return rc;
    switch (ld) {
        case 64: return launch<64>(m0, m1, mw, n, hdim, d, rows, ld, inv_norm, row_scale, st);
        case 128: return launch<128>(m0, m1, mw, n, hdim, d, rows, ld, inv_norm, row_scale, st);
        case 256: return launch<256>(m0, m1, mw, n, hdim, d, rows, ld, inv_norm, row_scale, st);
    }
    ntx_set_error("ntx_project_l2norm: row pitch %d not built (64, 128, 256)", ld);
    return NTX_E_UNSUPPORTED;
}
