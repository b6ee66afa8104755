// fp32 mode of the fused NT-Xent kernels on the CUDA cores (FFMA).
//
// The fp32 contract (loss and gradients within 1e-5 of the reference, which runs
// IEEE fp32 GEMMs: loss.py:42,43,49) cannot be met by tcgen05's tf32/bf16 inputs
// without multi-term splitting, so fp32 rows take this path; bf16 rows take the
// tcgen05 path in ntx_tc.cu.  Same structure as the tensor-core kernels:
//   forward : per (row block, key tile) similarity tile in registers ->
//             e = exp2(dot*c1 - c1), self / out-of-range columns zeroed ->
//             row sums; the logits never leave the SM        (loss.py:42-61)
//   backward: recompute the tile -> C = e * (w_r/R_r + w_c/R_c) ->
//             dZ[row block] += C . K[key tile]               (autograd of the above)
// Work is cut by the NtxPlan partition; every (CTA, row block) segment writes its
// partial into its slot of the scratch buffer and the finalize kernels add the
// slots in a fixed order.
#include "ntx_common.cuh"

namespace {

constexpr int BM = 64;      // rows per tile
constexpr int BN = 64;      // keys per tile
constexpr int BK = 32;      // feature chunk staged per step
constexpr int NT = 256;     // threads: 16 x 16, each owns a 4 x 4 micro tile
constexpr int PITCH = BM + 4;
constexpr int BD = 128;     // features of dZ one CTA accumulates (backward)
constexpr int KC = 16;      // keys staged per step of the second GEMM

// [64 rows x 32 features] of a row buffer -> smem transposed [32][PITCH], zero filled
__device__ __forceinline__ void stage_chunk_t(float (*dst)[PITCH], const float* __restrict__ base,
                                              int row0, int rows, int ld, int d0) {
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        const int idx = threadIdx.x + p * NT;
        const int r = idx >> 3, c4 = (idx & 7) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < rows && d0 + c4 < ld)
            v = __ldg(reinterpret_cast<const float4*>(base + (size_t)(row0 + r) * ld + d0 + c4));
        dst[c4 + 0][r] = v.x;
        dst[c4 + 1][r] = v.y;
        dst[c4 + 2][r] = v.z;
        dst[c4 + 3][r] = v.w;
    }
}

// acc[i][j] = <row (ty*4+i), key (tx*4+j)> over all features
__device__ __forceinline__ void similarity_tile(float (&acc)[4][4], float (*As)[PITCH],
                                                float (*Bs)[PITCH], const float* __restrict__ q,
                                                const float* __restrict__ k, int row0, int key0,
                                                const NtxDims& dm, int tx, int ty) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    for (int d0 = 0; d0 < dm.ld; d0 += BK) {
        __syncthreads();
        stage_chunk_t(As, q, row0, dm.q_rows, dm.ld, d0);
        stage_chunk_t(Bs, k, key0, dm.k_rows, dm.ld, d0);
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
    }
}

// Peer exchange: one thread waits until every rank's data of this step has landed in this rank's block (the fp32
// path is not the fast one: no overlap is attempted), then all threads pick the key buffer of the step.
__device__ __forceinline__ size_t simt_gate(const NtxGate& gate, bool backward, size_t key_buf_elems) {
    if (!gate.step) return 0;
    __shared__ uint32_t step_sh;
    if (threadIdx.x == 0) {
        const uint32_t step = *gate.step;
        for (int p = 0; p < gate.world; ++p) {
            ntx_gate_wait(gate.ctl, NTX_CTL_KFLAG + p, step);
            if (backward) {
                ntx_gate_wait(gate.ctl, NTX_CTL_RFLAG + p, step);
                ntx_gate_wait(gate.ctl, NTX_CTL_WFLAG + p, step);
            }
        }
        step_sh = step;
    }
    __syncthreads();
    return (step_sh & 1u) ? key_buf_elems : 0;
}

// kStable (temperature below NTX_MIN_TEMPERATURE): the fixed shift 1/tau would push every term of a row out of fp32
// range, so each thread keeps a running maximum with its sum (log2 domain) and a slot holds the pair (max, sum);
// the finalize kernel merges the pairs into lse_r.  Same tiles, same masks; one more max / rescale per tile.
template <bool kStable>
__global__ void __launch_bounds__(NT)
simt_fwd_kernel(const float* __restrict__ q, const float* __restrict__ k, NtxDims dm, int key_tiles,
                long long T, int rows_pad, float* __restrict__ partial,
                unsigned int* __restrict__ done_counter, NtxGate gate, size_t key_buf_elems) {
    __shared__ __align__(16) float As[BK][PITCH];
    __shared__ __align__(16) float Bs[BK][PITCH];
    {
        const size_t off = simt_gate(gate, false, key_buf_elems);
        q += off;
        k += off;
    }
    const int G = gridDim.x, cta = blockIdx.x;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const long long t_begin = ntx_chunk_begin(T, G, cta), t_end = ntx_chunk_begin(T, G, cta + 1);
    if (cta == 0 && threadIdx.x == 0)                       // the finalize kernel's 64-bit arrival word
        *reinterpret_cast<unsigned long long*>(done_counter) = 0ull;

    int cur_rb = -1;
    float rowsum[4] = {0.f, 0.f, 0.f, 0.f};
    float rowmax[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};      // kStable only
    auto flush = [&](int rb) {
        const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float v = rowsum[i];
            if constexpr (kStable) {
                float m = rowmax[i];
#pragma unroll
                for (int o = 1; o < 16; o <<= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                v = rowmax[i] == -INFINITY ? 0.f : v * exp2f(rowmax[i] - m);
#pragma unroll
                for (int o = 1; o < 16; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (tx == 0)
                    reinterpret_cast<float2*>(partial)[(size_t)slot * rows_pad + rb * BM + ty * 4 + i] = make_float2(m, v);
                rowmax[i] = -INFINITY;
            } else {
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                if (tx == 0) partial[(size_t)slot * rows_pad + rb * BM + ty * 4 + i] = v;
            }
            rowsum[i] = 0.f;
        }
    };

    for (long long t = t_begin; t < t_end; ++t) {
        const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
        if (rb != cur_rb) {
            if (cur_rb >= 0) flush(cur_rb);
            cur_rb = rb;
        }
        float acc[4][4];
        similarity_tile(acc, As, Bs, q, k, rb * BM, kt * BN, dm, tx, ty);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int self = ntx_self_col(rb * BM + ty * 4 + i, dm.n_local, dm.rank);
            if constexpr (kStable) {
                float x[4], tmax = -INFINITY;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = kt * BN + tx * 4 + j;
                    x[j] = (col == self || col >= dm.k_rows) ? -INFINITY : acc[i][j] * dm.c1;   // mask of loss.py:38,46-47
                    tmax = fmaxf(tmax, x[j]);
                }
                if (tmax > -INFINITY) {
                    const float mnew = fmaxf(rowmax[i], tmax);
                    float sum = rowsum[i] * exp2f(rowmax[i] - mnew);          // exp2(-inf) = 0 on the first tile
#pragma unroll
                    for (int j = 0; j < 4; ++j) sum += exp2f(x[j] - mnew);
                    rowsum[i] = sum;
                    rowmax[i] = mnew;
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const int col = kt * BN + tx * 4 + j;
                    float e = exp2f(fmaf(acc[i][j], dm.c1, -dm.c1));
                    if (col == self || col >= dm.k_rows) e = 0.f;   // mask of loss.py:38,46-47
                    rowsum[i] += e;
                }
            }
        }
    }
    if (cur_rb >= 0) flush(cur_rb);
}

// kStable: the row sums arrive as lse_r (see simt_fwd_kernel): C_rc = w_r exp(s_rc - lse_r) + w_c exp(s_rc - lse_c),
// two exponentials per logit instead of one, neither of which can overflow (s_rc <= lse_r and s_rc <= lse_c)
template <bool kStable>
__global__ void __launch_bounds__(NT)
simt_bwd_kernel(const float* __restrict__ q, const float* __restrict__ k, NtxDims dm, int key_tiles,
                long long T, int rows_pad, int d_pad, NtxWeights wt, float* __restrict__ partial, NtxGate gate,
                size_t key_buf_elems) {
    __shared__ __align__(16) float As[BK][PITCH];
    __shared__ __align__(16) float Bs[BK][PITCH];
    {
        const size_t off = simt_gate(gate, true, key_buf_elems);
        q += off;
        k += off;
    }
    __shared__ __align__(16) float Ct[BN][PITCH];     // coefficient tile, [key][row]
    __shared__ __align__(16) float Ks[KC][BD];        // keys x features of this CTA's dZ slab
    const int G = gridDim.x, cta = blockIdx.x;
    const int dc0 = blockIdx.y * BD;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const long long t_begin = ntx_chunk_begin(T, G, cta), t_end = ntx_chunk_begin(T, G, cta + 1);

    float dz[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) dz[i][j] = 0.f;

    int cur_rb = -1;
    auto flush = [&](int rb) {
        const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float* dst = partial + ((size_t)slot * rows_pad + rb * BM + ty * 4 + i) * d_pad + dc0;
            if (dc0 + tx * 4 < d_pad)
                *reinterpret_cast<float4*>(dst + tx * 4) = make_float4(dz[i][0], dz[i][1], dz[i][2], dz[i][3]);
            if (dc0 + 64 + tx * 4 < d_pad)
                *reinterpret_cast<float4*>(dst + 64 + tx * 4) = make_float4(dz[i][4], dz[i][5], dz[i][6], dz[i][7]);
#pragma unroll
            for (int j = 0; j < 8; ++j) dz[i][j] = 0.f;
        }
    };

    for (long long t = t_begin; t < t_end; ++t) {
        const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
        if (rb != cur_rb) {
            if (cur_rb >= 0) flush(cur_rb);
            cur_rb = rb;
        }
        float acc[4][4];
        similarity_tile(acc, As, Bs, q, k, rb * BM, kt * BN, dm, tx, ty);

        // coefficient tile  C_rc = e_rc * (w_r/R_r + w_c/R_c), zero on self / out-of-range
        float alpha[4], beta[4];
        float alpha_l[4] = {0.f, 0.f, 0.f, 0.f}, beta_l[4] = {0.f, 0.f, 0.f, 0.f};   // kStable: lse of rows / keys, log2 units
        int self[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int r = rb * BM + ty * 4 + i;
            if constexpr (kStable) {
                alpha[i] = r < dm.q_rows ? ntx_ld_weight(wt.w_q, (size_t)r * wt.w_stride, wt.w_bf16) * wt.w_scale : 0.f;
                alpha_l[i] = r < dm.q_rows ? __ldg(wt.rowsum_q + r) * NTX_LOG2E : 0.f;
            } else {
                alpha[i] = r < dm.q_rows
                               ? ntx_ld_weight(wt.w_q, (size_t)r * wt.w_stride, wt.w_bf16) * wt.w_scale / __ldg(wt.rowsum_q + r)
                               : 0.f;
            }
            self[i] = ntx_self_col(r, dm.n_local, dm.rank);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = kt * BN + tx * 4 + j;
            const size_t wi = ntx_key_weight_index(c, wt.w_stride, wt.w_chunk);
            if (c >= dm.k_rows) {
                beta[j] = 0.f;
            } else {
                // with a peer exchange both values were written by other ranks during the step: skip L1
                const float wv = gate.step ? ntx_ld_weight(wt.w_k, wi, wt.w_bf16, true) : ntx_ld_weight(wt.w_k, wi, wt.w_bf16);
                const float rv = gate.step ? ntx_ld_l2(wt.rowsum_k + c) : __ldg(wt.rowsum_k + c);
                if constexpr (kStable) {
                    beta[j] = wv * wt.w_scale;
                    beta_l[j] = rv * NTX_LOG2E;
                } else {
                    beta[j] = wv * wt.w_scale / rv;
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = kt * BN + tx * 4 + j;
            float cv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool masked = col == self[i] || col >= dm.k_rows || rb * BM + ty * 4 + i >= dm.q_rows;
                if constexpr (kStable) {
                    const float x = acc[i][j] * dm.c1;
                    cv[i] = masked ? 0.f : alpha[i] * exp2f(x - alpha_l[i]) + beta[j] * exp2f(x - beta_l[j]);
                } else {
                    float e = exp2f(fmaf(acc[i][j], dm.c1, -dm.c1));
                    if (masked) e = 0.f;
                    cv[i] = e * (alpha[i] + beta[j]);
                }
            }
            *reinterpret_cast<float4*>(&Ct[tx * 4 + j][ty * 4]) = make_float4(cv[0], cv[1], cv[2], cv[3]);
        }

        // dZ[rows, dc0:dc0+128] += C . K[key tile, dc0:dc0+128]
        for (int kc = 0; kc < BN; kc += KC) {
            __syncthreads();   // Ct visible (first pass) / previous Ks consumed
#pragma unroll
            for (int p = 0; p < 2; ++p) {
                const int idx = threadIdx.x + p * NT;
                const int kr = idx >> 5, c4 = (idx & 31) * 4;
                const int key = kt * BN + kc + kr;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (key < dm.k_rows && dc0 + c4 < dm.ld)
                    v = __ldg(reinterpret_cast<const float4*>(k + (size_t)key * dm.ld + dc0 + c4));
                *reinterpret_cast<float4*>(&Ks[kr][c4]) = v;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < KC; ++kk) {
                const float4 a = *reinterpret_cast<const float4*>(&Ct[kc + kk][ty * 4]);
                const float4 b0 = *reinterpret_cast<const float4*>(&Ks[kk][tx * 4]);
                const float4 b1 = *reinterpret_cast<const float4*>(&Ks[kk][64 + tx * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w};
                const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 8; ++j) dz[i][j] = fmaf(av[i], bv[j], dz[i][j]);
            }
        }
        // the next tile's similarity_tile starts with __syncthreads(), which also
        // orders this tile's last Ct/Ks reads before they are overwritten
    }
    if (cur_rb >= 0) flush(cur_rb);
}

}  // namespace

int ntx_simt_fwd_launch(const float* q, const float* k, const NtxDims& dm, const NtxPlan& plan,
                        float* partial, unsigned int* counter, const NtxGate& gate, size_t key_buf_elems,
                        cudaStream_t st) {
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    ntx_note_launch();
    if (dm.stable)
        simt_fwd_kernel<true><<<plan.grid, NT, 0, st>>>(q, k, dm, plan.key_tiles, T, rows_pad, partial, counter, gate,
                                                        key_buf_elems);
    else
        simt_fwd_kernel<false><<<plan.grid, NT, 0, st>>>(q, k, dm, plan.key_tiles, T, rows_pad, partial, counter, gate,
                                                         key_buf_elems);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

int ntx_simt_bwd_launch(const float* q, const float* k, const NtxDims& dm, const NtxPlan& plan,
                        const NtxWeights& wt, float* partial, const NtxGate& gate, size_t key_buf_elems,
                        cudaStream_t st) {
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    dim3 grid(plan.grid, (plan.d_pad + BD - 1) / BD);
    ntx_note_launch();
    if (dm.stable)
        simt_bwd_kernel<true><<<grid, NT, 0, st>>>(q, k, dm, plan.key_tiles, T, rows_pad, plan.d_pad, wt, partial, gate,
                                                   key_buf_elems);
    else
        simt_bwd_kernel<false><<<grid, NT, 0, st>>>(q, k, dm, plan.key_tiles, T, rows_pad, plan.d_pad, wt, partial, gate,
                                                    key_buf_elems);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}
