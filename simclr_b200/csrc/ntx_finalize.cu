// Finalize kernels shared by the fp32 and the tensor-core paths.
//
//  fwd_finalize : add the row-sum partials of a row's slots in slot order, take the
//                 positive logit as a direct dot product with the partner row
//                 (targets = arange(N) of loss.py:37 -> ntx_pos_row), and form
//                 row_loss = log(rowsum) + 1/tau - pos  (= logsumexp - positive,
//                 the per-row cross entropy of loss.py:56-61).
//  loss_reduce  : fixed-order sum of the row losses (loss.py:60-61).
//  bwd_prep     : row weights from the upstream gradient and coef = w / rowsum.
//  bwd_finalize : add the dZ partials, subtract the positive-pair term, scale by
//                 1/tau and apply the Jacobian of x / max(||x||, eps)
//                 (loss.py:33-34 backward), write dview0 / dview1.
// All are HBM/L2-bound O(rows * d) passes; one warp per row.
#include "ntx_common.cuh"

namespace {

constexpr int kWarps = 8;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float ldf(const float* p, size_t i) { return __ldg(p + i); }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p, size_t i) {
    return __bfloat162float(p[i]);
}
__device__ __forceinline__ void stf(float* p, size_t i, float v) { p[i] = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, size_t i, float v) {
    p[i] = __float2bfloat16_rn(v);
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32)
fwd_finalize_kernel(const T* __restrict__ q, NtxDims dm, int tile_rows, int key_tiles, long long Tt,
                    int G, int rows_pad, const float* __restrict__ partial,
                    float* __restrict__ rowsum, float* __restrict__ pos_logit,
                    float* __restrict__ row_loss) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (r >= dm.q_rows) return;
    const int rb = r / tile_rows;
    const int ns = ntx_num_slots(Tt, G, key_tiles, rb);
    float R = 0.f;
    for (int s = 0; s < ns; ++s) R += __ldg(partial + (size_t)s * rows_pad + r);

    const int p = ntx_pos_row(r, dm.n_local);
    const T* a = q + (size_t)r * dm.ld;
    const T* b = q + (size_t)p * dm.ld;
    float dot = 0.f;
    for (int c = lane; c < dm.ld; c += 32) dot = fmaf(ldf(a, c), ldf(b, c), dot);
    dot = warp_sum(dot);
    if (lane == 0) {
        const float pos = dot * dm.inv_tau;
        rowsum[r] = R;
        pos_logit[r] = pos;
        row_loss[r] = logf(R) + dm.inv_tau - pos;
    }
}

__global__ void __launch_bounds__(1024)
loss_reduce_kernel(const float* __restrict__ row_loss, int rows, float* __restrict__ loss_sum) {
    __shared__ double part[32];
    double acc = 0.0;
    for (int i = threadIdx.x; i < rows; i += 1024) acc += (double)row_loss[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 32) {
        double v = part[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) {
            loss_sum[0] = (float)v;                    // reduction="sum"
            loss_sum[1] = (float)(v / (double)rows);   // reduction="mean": sum / N * 0.5, rows = 2N
        }
    }
}

__global__ void bwd_prep_kernel(const float* __restrict__ rowsum, const float* __restrict__ grad_out,
                                int reduction, int q_rows, float* __restrict__ w,
                                float* __restrict__ coef) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= q_rows) return;
    float wr;
    if (reduction == NTX_REDUCE_NONE) wr = grad_out[r];
    else if (reduction == NTX_REDUCE_MEAN) wr = grad_out[0] / (float)q_rows;   // loss / N * 0.5
    else wr = grad_out[0];
    w[r] = wr;
    coef[r] = wr / rowsum[r];
}

template <typename TIn>
__global__ void __launch_bounds__(kWarps * 32)
bwd_finalize_kernel(NtxDims dm, int tile_rows, int key_tiles, long long Tt, int G, int rows_pad,
                    int d_pad, const float* __restrict__ partial, const float* __restrict__ w_q,
                    const TIn* __restrict__ view0, const TIn* __restrict__ view1,
                    const float* __restrict__ inv_norm, TIn* __restrict__ dview0,
                    TIn* __restrict__ dview1) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (r >= dm.q_rows) return;
    const int n = dm.n_local, d = dm.d;
    const int p = ntx_pos_row(r, n);
    const int rb = r / tile_rows;
    const int ns = ntx_num_slots(Tt, G, key_tiles, rb);
    const TIn* zr = r < n ? view0 + (size_t)r * d : view1 + (size_t)(r - n) * d;
    const TIn* zp = p < n ? view0 + (size_t)p * d : view1 + (size_t)(p - n) * d;
    TIn* out = r < n ? dview0 + (size_t)r * d : dview1 + (size_t)(r - n) * d;
    const float inr = __ldg(inv_norm + r), inp = __ldg(inv_norm + p);
    const float wpair = __ldg(w_q + r) + __ldg(w_q + p);   // (P - Y) and its transpose both hit the pair
    const float* part = partial + (size_t)r * d_pad;
    const size_t slot_stride = (size_t)rows_pad * d_pad;

    auto dzhat = [&](int c) {
        float acc = 0.f;
        for (int s = 0; s < ns; ++s) acc += __ldg(part + s * slot_stride + c);
        return (acc - wpair * (ldf(zp, c) * inp)) * dm.inv_tau;
    };

    float proj = 0.f;
    for (int c = lane; c < d; c += 32) proj = fmaf(ldf(zr, c) * inr, dzhat(c), proj);
    proj = warp_sum(proj);
    const bool clamped = inr >= 1e12f;    // ||z|| <= eps: denominator was the constant eps
    for (int c = lane; c < d; c += 32) {
        const float g = dzhat(c);
        const float v = clamped ? g * inr : (g - ldf(zr, c) * inr * proj) * inr;
        stf(out, c, v);
    }
}

__global__ void debug_indices_kernel(int n_local, int rank, int32_t* __restrict__ self_col,
                                     int32_t* __restrict__ pos_col) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= 2 * n_local) return;
    self_col[r] = ntx_self_col(r, n_local, rank);
    pos_col[r] = ntx_pos_col(r, n_local, rank);
}

}  // namespace

int ntx_debug_indices_launch(int n_local, int rank, int32_t* self_col, int32_t* pos_col,
                             cudaStream_t st) {
    debug_indices_kernel<<<(2 * n_local + 255) / 256, 256, 0, st>>>(n_local, rank, self_col, pos_col);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

int ntx_fwd_finalize_launch(const void* q, int dtype, const NtxDims& dm, const NtxPlan& plan,
                            const float* partial, float* rowsum, float* pos_logit, float* row_loss,
                            float* loss_sum, cudaStream_t st) {
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    const int grid = (dm.q_rows + kWarps - 1) / kWarps;
    if (dtype == NTX_F32)
        fwd_finalize_kernel<float><<<grid, kWarps * 32, 0, st>>>(
            (const float*)q, dm, plan.tile_rows, plan.key_tiles, T, plan.grid, rows_pad, partial,
            rowsum, pos_logit, row_loss);
    else
        fwd_finalize_kernel<__nv_bfloat16><<<grid, kWarps * 32, 0, st>>>(
            (const __nv_bfloat16*)q, dm, plan.tile_rows, plan.key_tiles, T, plan.grid, rows_pad,
            partial, rowsum, pos_logit, row_loss);
    NTX_CUDA(cudaGetLastError());
    loss_reduce_kernel<<<1, 1024, 0, st>>>(row_loss, dm.q_rows, loss_sum);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

int ntx_bwd_prep_launch(const float* rowsum, const float* grad_out, int reduction, int q_rows,
                        float* w, float* coef, cudaStream_t st) {
    bwd_prep_kernel<<<(q_rows + 255) / 256, 256, 0, st>>>(rowsum, grad_out, reduction, q_rows, w, coef);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

int ntx_bwd_finalize_launch(const NtxDims& dm, const NtxPlan& plan, const float* partial,
                            const float* w_q, const void* view0, const void* view1, int in_dtype,
                            const float* inv_norm, void* dview0, void* dview1, cudaStream_t st) {
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    const int grid = (dm.q_rows + kWarps - 1) / kWarps;
    if (in_dtype == NTX_F32)
        bwd_finalize_kernel<float><<<grid, kWarps * 32, 0, st>>>(
            dm, plan.tile_rows, plan.key_tiles, T, plan.grid, rows_pad, plan.d_pad, partial, w_q,
            (const float*)view0, (const float*)view1, inv_norm, (float*)dview0, (float*)dview1);
    else
        bwd_finalize_kernel<__nv_bfloat16><<<grid, kWarps * 32, 0, st>>>(
            dm, plan.tile_rows, plan.key_tiles, T, plan.grid, rows_pad, plan.d_pad, partial, w_q,
            (const __nv_bfloat16*)view0, (const __nv_bfloat16*)view1, inv_norm,
            (__nv_bfloat16*)dview0, (__nv_bfloat16*)dview1);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}
