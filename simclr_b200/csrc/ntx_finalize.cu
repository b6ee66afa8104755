// Finalize kernels shared by the fp32 and the tensor-core paths.
//
//  fwd_finalize : add the row-sum partials of a row's slots in slot order, take the
//                 positive logit as a direct dot product with the partner row
//                 (targets = arange(N) of loss.py:37 -> ntx_pos_row), and form
//                 row_loss = log(rowsum) + 1/tau - pos  (= logsumexp - positive,
//                 the per-row cross entropy of loss.py:56-61).  The last block to
//                 finish adds all row losses in a fixed order (loss.py:60-65).
//  bwd_finalize : add the dZ partials, subtract the positive-pair term, scale by
//                 1/tau and apply the Jacobian of x / max(||x||, eps)
//                 (loss.py:33-34 backward), write dview0 / dview1.
// Both are HBM/L2-bound O(rows * d) passes with 128-bit accesses.
#include "ntx_common.cuh"

namespace {

// block sizes measured on B200 at the benchmark size (one box, alternating runs): backward finalize 512 -> 256
// threads -0.9 us per step, forward finalize 1024 -> 256 -0.3 us (more, smaller blocks hide the load latency better;
// the forward's arrival word takes any number of blocks)
constexpr int kThreads = 256;        // backward finalize / generic kernels
constexpr int kFwdThreads = 256;     // forward finalize

__device__ __forceinline__ float ldf(const float* p, size_t i) { return __ldg(p + i); }
__device__ __forceinline__ float ldf(const __nv_bfloat16* p, size_t i) { return __bfloat162float(p[i]); }
__device__ __forceinline__ void stf(float* p, size_t i, float v) { p[i] = v; }
__device__ __forceinline__ void stf(__nv_bfloat16* p, size_t i, float v) { p[i] = __float2bfloat16_rn(v); }

// 4 consecutive elements as floats (16-byte fp32 / 8-byte bf16 accesses)
__device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(p));
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 t;
    t.x = *reinterpret_cast<uint32_t*>(&a);
    t.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = t;
}

// sum over the `width` consecutive lanes that share a row (width = power of two <= 32)
__device__ __forceinline__ float group_sum(float v, int width) {
    for (int o = width >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------ forward finalize
// `lanes` threads per row (power of two), each covers 4-element groups sub, sub+lanes, ...
template <typename T>
__global__ void __launch_bounds__(kFwdThreads)
fwd_finalize_kernel(const T* __restrict__ q, NtxDims dm, int lanes, int tile_rows, int key_tiles,
                    long long Tt, int G, int rows_pad, const float* __restrict__ partial,
                    const float* __restrict__ colpart, unsigned int* __restrict__ counter,
                    float* __restrict__ rowsum,
                    float* __restrict__ pos_logit, float* __restrict__ row_loss,
                    float* __restrict__ loss_sum, NtxPeerOut peers, size_t rowsum_off, NtxGate gate,
                    size_t key_buf_elems) {
    ntx_dep_trigger();
    ntx_dep_wait();
    // peer exchange: the rows sit in the key buffer of this step; the row sums also go to every rank's block
    const uint32_t step = gate.step ? *gate.step : 0u;      // written by the normalise kernel, earlier on the stream
    if (step & 1u) q += key_buf_elems;
    // rows of one block share a row block (tile_rows is a multiple of rows_per_block): one slot count.
    // colpart != NULL: the symmetric forward's upper-triangle partition (Tt, G describe that one).
    __shared__ int ns_sh;       // (computed once and broadcast: ~0.5 us faster than the 64-bit divisions in every thread)
    if (threadIdx.x == 0) {
        const int rb0 = (int)((long long)blockIdx.x * ((int)blockDim.x / lanes) / tile_rows);
        ns_sh = colpart ? ntx_tri_num_slots(Tt, G, key_tiles, rb0) : ntx_num_slots(Tt, G, key_tiles, rb0);
    }
    __syncthreads();
    double my_loss = 0.0;
    const int nthreads = (int)blockDim.x;                    // kFwdThreads
    const int rows_per_block = nthreads / lanes;
    // The positive logit's operands first: their loads are in flight while the partial sums below are gathered
    // (these kernels are chains of dependent memory round trips; every one that overlaps another is ~0.5 us).
    const int sub = threadIdx.x % lanes;
    const int r_raw = blockIdx.x * rows_per_block + threadIdx.x / lanes;
    const bool live = r_raw < dm.q_rows;
    const int r = live ? r_raw : dm.q_rows - 1;              // keep every lane in the shuffles
    float dot = 0.f;
    {
        const int p = ntx_pos_row(r, dm.n_local);
        const T* a = q + (size_t)r * dm.ld;
        const T* b = q + (size_t)p * dm.ld;
        if (dm.split) {                                          // two-term rows: x = hi + lo, exact in fp32
            for (int c = sub * 4; c < dm.dh; c += lanes * 4) {
                const float4 xh = ld4(a + c), xl = ld4(a + dm.dh + c), yh = ld4(b + c), yl = ld4(b + dm.dh + c);
                const float4 x = make_float4(xh.x + xl.x, xh.y + xl.y, xh.z + xl.z, xh.w + xl.w);
                const float4 y = make_float4(yh.x + yl.x, yh.y + yl.y, yh.z + yl.z, yh.w + yl.w);
                dot = fmaf(x.x, y.x, fmaf(x.y, y.y, fmaf(x.z, y.z, fmaf(x.w, y.w, dot))));
            }
        } else {
            for (int c = sub * 4; c < dm.ld; c += lanes * 4) {  // ld is a multiple of 4
                const float4 x = ld4(a + c), y = ld4(b + c);
                dot = fmaf(x.x, y.x, fmaf(x.y, y.y, fmaf(x.z, y.z, fmaf(x.w, y.w, dot))));
            }
        }
    }
    // Row sums first, with the block laid out rows-fastest so that a warp reads 32 consecutive rows of one
    // slot (or of one row block's column partials): thread t adds the terms part, part + lanes, ... of row
    // t % rows_per_block; the `lanes` parts of a row are then added in part order.
    __shared__ float rs_sh[kFwdThreads];
    {
        const int row_local = threadIdx.x % rows_per_block, part = threadIdx.x / rows_per_block;
        const int r2 = min(blockIdx.x * rows_per_block + row_local, dm.q_rows - 1);
        // terms part, part + lanes, ... are added in that order; the loads of eight terms are issued together
        // (these kernels are latency-bound: a load per loop iteration costs an L2 round trip each)
        auto strided_sum = [&](const float* base, int count) {
            float acc = 0.f;
            for (int i0 = part; i0 < count; i0 += 8 * lanes) {
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int i = i0 + u * lanes;
                    v[u] = i < count ? __ldg(base + (size_t)i * rows_pad + r2) : 0.f;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) acc += v[u];
            }
            return acc;
        };
        float c = 0.f;
        if (dm.stable) {
            // max-shifted form (small temperatures, fp32 rows): a slot holds (max, sum of exp2(x - max)); merged by
            // one thread per row into lse in log2 units, the other parts add nothing
            if (part == 0) {
                const float2* p2 = reinterpret_cast<const float2*>(partial);
                float M = -INFINITY;
                for (int sl = 0; sl < ns_sh; ++sl) M = fmaxf(M, p2[(size_t)sl * rows_pad + r2].x);
                float S = 0.f;
                for (int sl = 0; sl < ns_sh; ++sl) {
                    const float2 t = p2[(size_t)sl * rows_pad + r2];
                    if (t.x > -INFINITY) S += t.y * exp2f(t.x - M);
                }
                c = M + log2f(S);
            }
        } else {
            if (colpart) c = strided_sum(colpart, r2 / tile_rows);   // symmetric forward: what the tiles ABOVE this row's block saw in this row's column
            c += strided_sum(partial, ns_sh);
        }
        rs_sh[part * rows_per_block + row_local] = c;
    }
    __syncthreads();
    {
        dot = group_sum(dot, lanes);
        if (sub == 0 && live) {
            float R = 0.f;
            for (int part = 0; part < lanes; ++part) R += rs_sh[part * rows_per_block + threadIdx.x / lanes];
            const float pos = dot * dm.pos_scale;            // 1/tau, or ln 2 for rows that carry sqrt(log2e/tau)
            R *= dm.rowsum_scale;                            // sums of e * 2^c1 -> sums of exp(s - 1/tau) (scaled rows)
            float l;
            if (dm.stable) {
                R *= 0.6931471805599453f;                    // lse_r in natural-log units: what "rowsum" holds in this mode
                l = R - pos;
            } else {
                l = logf(R) + dm.inv_tau - pos;
            }
            rowsum[r] = R;
            pos_logit[r] = pos;
            row_loss[r] = l;
            my_loss = (double)l;
            if (peers.world > 0)
                ntx_peer_st_f32(peers, rowsum_off + ((size_t)peers.rank * dm.q_rows + r) * sizeof(float), R);
        }
    }
    // The block's loss sum, warps added in order, goes into ONE 64-bit word as fixed point (16 fractional bits, upper
    // 44 bits) together with a count of arrived blocks (lower 20 bits): integer addition does not depend on the
    // order of arrival, so the total is the same in every run, and the block that brings the count to the grid size
    // holds the total in the value its own atomicAdd returned -- no second pass over per-block sums (loss.py:60-65).
    __shared__ double part[kFwdThreads / 32];
    double acc = my_loss;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = 0.0;
        for (int w = 0; w < nthreads / 32; ++w) v += part[w];
        const long long fixed = __double2ll_rn(v * 65536.0);
        const unsigned long long add = ((unsigned long long)fixed << 20) + 1ull;
        if (peers.world > 0) __threadfence();                // the block's peer stores (barrier above) before its arrival
        unsigned long long* word = reinterpret_cast<unsigned long long*>(counter);
        const unsigned long long old = atomicAdd(word, add);
        if ((old & 0xFFFFFull) == (unsigned long long)gridDim.x - 1ull) {
            const long long total = (long long)(old + add) >> 20;             // arithmetic shift keeps the sign
            ntx_store_loss(loss_sum, (double)total / 65536.0, dm.q_rows, dm.loss_bf16);   // "sum" and "mean" (sum / N * 0.5)
            *word = 0ull;
            if (peers.world > 0) ntx_raise_flags(peers, NTX_CTL_RFLAG, step);   // every block's row sums are out
        }
    }
}

// ------------------------------------------------------------------ backward finalize
struct BwdFinArgs {
    NtxDims dm;
    int tile_rows, key_tiles, G, rows_pad, d_pad;
    long long Tt;
    const float* partial;
    NtxWeights w;
    const float* inv_norm;
    int in_ld;          // elements between consecutive input rows (d for the caller's views)
    int normalised;     // 1: view0/view1 hold the NORMALISED rows (ntx_bwd_rows), 0: the raw inputs
};

__device__ __forceinline__ float pair_weight(const NtxWeights& w, int r, int p) {
    // (P - Y) and its transpose both hit the positive pair: weight w_r + w_p
    return (ntx_ld_weight(w.w_q, (size_t)r * w.w_stride, w.w_bf16) + ntx_ld_weight(w.w_q, (size_t)p * w.w_stride, w.w_bf16)) *
           w.w_scale;
}

// fast path: d % 4 == 0, one group of `lanes` threads per row, up to kMaxV float4 per thread
template <typename TIn, int kMaxV>
__global__ void __launch_bounds__(kThreads)
bwd_finalize_vec_kernel(BwdFinArgs A, int lanes, const TIn* __restrict__ view0,
                        const TIn* __restrict__ view1, TIn* __restrict__ dview0, TIn* __restrict__ dview1) {
    ntx_dep_trigger();
    ntx_dep_wait();
    const NtxDims& dm = A.dm;
    const int rows_per_block = (int)blockDim.x / lanes;      // kThreads, or fewer when that leaves SMs idle
    // rows of one block share a row block (tile_rows is a multiple of rows_per_block): one slot count
    __shared__ int ns_sh;       // (computed once and broadcast: 0.6 us faster than the 64-bit divisions in every thread)
    if (threadIdx.x == 0)
        ns_sh = ntx_num_slots(A.Tt, A.G, A.key_tiles, (int)((long long)blockIdx.x * rows_per_block / A.tile_rows));
    __syncthreads();
    const int sub = threadIdx.x % lanes;
    const int r_raw = blockIdx.x * rows_per_block + threadIdx.x / lanes;
    const bool live = r_raw < dm.q_rows;
    const int r = live ? r_raw : dm.q_rows - 1;              // keep every lane in the shuffles
    const int n = dm.n_local, d = dm.d;
    const int p = ntx_pos_row(r, n);
    const int ns = ns_sh;
    const TIn* zr = r < n ? view0 + (size_t)r * A.in_ld : view1 + (size_t)(r - n) * A.in_ld;
    const TIn* zp = p < n ? view0 + (size_t)p * A.in_ld : view1 + (size_t)(p - n) * A.in_ld;
    TIn* out = r < n ? dview0 + (size_t)r * d : dview1 + (size_t)(r - n) * d;
    const float inr = __ldg(A.inv_norm + r), inp = __ldg(A.inv_norm + p);
    // input rows -> unit rows: 1/||z|| for the caller's views, 1/sqrt(c1) for normalised rows that carry the scale
    const float sr = A.normalised ? dm.inv_rs : inr, sp = A.normalised ? dm.inv_rs : inp;
    const float wpair = pair_weight(A.w, r, p);
    const float* part = A.partial + (size_t)r * A.d_pad;
    const size_t slot_stride = (size_t)A.rows_pad * A.d_pad;

    float4 g[kMaxV], zh[kMaxV];
    float proj = 0.f;
#pragma unroll
    for (int v = 0; v < kMaxV; ++v) {
        const int c = (sub + v * lanes) * 4;
        if (c < d) {
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int s0 = 0; s0 < ns; s0 += 4) {           // slot order, four loads in flight
                float4 t[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    t[u] = s0 + u < ns ? __ldg(reinterpret_cast<const float4*>(part + (s0 + u) * slot_stride + c))
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 4; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
            }
            const float4 zpv = ld4(zp + c), zrv = ld4(zr + c);
            const float k = wpair * sp, it = dm.inv_tau, ir = dm.inv_rs;       // dZ' = C . K' carries sqrt(c1)
            g[v] = make_float4((acc.x * ir - k * zpv.x) * it, (acc.y * ir - k * zpv.y) * it, (acc.z * ir - k * zpv.z) * it,
                               (acc.w * ir - k * zpv.w) * it);
            zh[v] = make_float4(zrv.x * sr, zrv.y * sr, zrv.z * sr, zrv.w * sr);
            proj = fmaf(zh[v].x, g[v].x, fmaf(zh[v].y, g[v].y, fmaf(zh[v].z, g[v].z, fmaf(zh[v].w, g[v].w, proj))));
        }
    }
    proj = group_sum(proj, lanes);
    const bool clamped = inr >= 1e12f;    // ||z|| <= eps: denominator was the constant eps
#pragma unroll
    for (int v = 0; v < kMaxV; ++v) {
        const int c = (sub + v * lanes) * 4;
        if (c < d && live) {
            float4 o;
            if (clamped) {
                o = make_float4(g[v].x * inr, g[v].y * inr, g[v].z * inr, g[v].w * inr);
            } else {
                o = make_float4((g[v].x - zh[v].x * proj) * inr, (g[v].y - zh[v].y * proj) * inr,
                                (g[v].z - zh[v].z * proj) * inr, (g[v].w - zh[v].w * proj) * inr);
            }
            st4(out + c, o);
        }
    }
}

// generic path: any d, one warp per row, two passes
template <typename TIn>
__global__ void __launch_bounds__(kThreads)
bwd_finalize_kernel(BwdFinArgs A, const TIn* __restrict__ view0, const TIn* __restrict__ view1,
                    TIn* __restrict__ dview0, TIn* __restrict__ dview1) {
    const NtxDims& dm = A.dm;
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
    if (r >= dm.q_rows) return;                              // whole warps leave together
    const int n = dm.n_local, d = dm.d;
    const int p = ntx_pos_row(r, n);
    const int ns = ntx_num_slots(A.Tt, A.G, A.key_tiles, r / A.tile_rows);
    const TIn* zr = r < n ? view0 + (size_t)r * A.in_ld : view1 + (size_t)(r - n) * A.in_ld;
    const TIn* zp = p < n ? view0 + (size_t)p * A.in_ld : view1 + (size_t)(p - n) * A.in_ld;
    TIn* out = r < n ? dview0 + (size_t)r * d : dview1 + (size_t)(r - n) * d;
    const float inr = __ldg(A.inv_norm + r), inp = __ldg(A.inv_norm + p);
    const float sr = A.normalised ? dm.inv_rs : inr, sp = A.normalised ? dm.inv_rs : inp;
    const float wpair = pair_weight(A.w, r, p);
    const float* part = A.partial + (size_t)r * A.d_pad;
    const size_t slot_stride = (size_t)A.rows_pad * A.d_pad;

    auto dzhat = [&](int c) {
        float acc = 0.f;
        for (int s = 0; s < ns; ++s) acc += __ldg(part + s * slot_stride + c);
        return (acc * dm.inv_rs - wpair * (ldf(zp, c) * sp)) * dm.inv_tau;
    };
    float proj = 0.f;
    for (int c = lane; c < d; c += 32) proj = fmaf(ldf(zr, c) * sr, dzhat(c), proj);
    proj = group_sum(proj, 32);
    const bool clamped = inr >= 1e12f;
    for (int c = lane; c < d; c += 32) {
        const float g = dzhat(c);
        stf(out, c, clamped ? g * inr : (g - ldf(zr, c) * sr * proj) * inr);
    }
}

__global__ void debug_indices_kernel(int n_local, int rank, int32_t* __restrict__ self_col,
                                     int32_t* __restrict__ pos_col) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= 2 * n_local) return;
    self_col[r] = ntx_self_col(r, n_local, rank);
    pos_col[r] = ntx_pos_col(r, n_local, rank);
}

// (16 or more, so that the rows of one block, <= 1024 / 16, never straddle a 64-row tile: one slot count per block)
// threads per row: 16 up to 256 features (64 four-element groups), so that a block covers enough rows for the
// whole grid to be resident at once (these kernels are latency-bound: one wave, not two), 32 beyond
// (8 and 32 lanes at 128 features measured on one box: +4.6 / +5 us per step)
int lanes_for(int groups4) { return groups4 <= 64 ? 16 : 32; }

template <typename TIn>
int launch_bwd_finalize(const BwdFinArgs& A, const void* view0, const void* view1, void* dview0,
                        void* dview1, cudaStream_t st) {
    const int d = A.dm.d;
    const bool vec = d % 4 == 0 && d <= 1024 && A.in_ld % 4 == 0 &&
                     reinterpret_cast<uintptr_t>(view0) % 16 == 0 && reinterpret_cast<uintptr_t>(view1) % 16 == 0 &&
                     reinterpret_cast<uintptr_t>(dview0) % 16 == 0 && reinterpret_cast<uintptr_t>(dview1) % 16 == 0;
    if (vec) {
        const int lanes = lanes_for(d / 4);
        // latency-bound: with few rows (one of several ranks) smaller blocks keep every SM busy
        int threads = kThreads;
        while (threads > 128 && (A.dm.q_rows + threads / lanes - 1) / (threads / lanes) < 296) threads >>= 1;
        const int grid = (A.dm.q_rows + threads / lanes - 1) / (threads / lanes);
        const int per_thread = (d / 4 + lanes - 1) / lanes;          // float4 groups each thread covers
#define NTX_BWD_FIN(V)                                                                                        \
    NTX_CUDA(ntx_launch_dep(bwd_finalize_vec_kernel<TIn, V>, grid, threads, 0, st, A, lanes, (const TIn*)view0, \
                            (const TIn*)view1, (TIn*)dview0, (TIn*)dview1))
        if (per_thread <= 1) NTX_BWD_FIN(1);
        else if (per_thread <= 2) NTX_BWD_FIN(2);
        else if (per_thread <= 4) NTX_BWD_FIN(4);
        else NTX_BWD_FIN(8);
#undef NTX_BWD_FIN
    } else {
        const int grid = (A.dm.q_rows + kThreads / 32 - 1) / (kThreads / 32);
        ntx_note_launch();
        bwd_finalize_kernel<TIn><<<grid, kThreads, 0, st>>>(A, (const TIn*)view0, (const TIn*)view1, (TIn*)dview0,
                                                            (TIn*)dview1);
    }
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

}  // namespace

int ntx_debug_indices_launch(int n_local, int rank, int32_t* self_col, int32_t* pos_col,
                             cudaStream_t st) {
    ntx_note_launch();
    debug_indices_kernel<<<(2 * n_local + 255) / 256, 256, 0, st>>>(n_local, rank, self_col, pos_col);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

static int fwd_finalize_threads(int, int) { return kFwdThreads; }

int ntx_fwd_finalize_blocks(int q_rows, int ld) {
    const int lanes = lanes_for(ld / 4);
    const int threads = fwd_finalize_threads(q_rows, lanes);
    return (q_rows + threads / lanes - 1) / (threads / lanes);
}

int ntx_fwd_finalize_launch(const void* q, int dtype, const NtxDims& dm, const NtxPlan& plan,
                            const float* partial, const float* colpart, unsigned int* counter,
                            float* rowsum, float* pos_logit, float* row_loss,
                            float* loss_sum, const NtxPeerOut& peers, size_t rowsum_off, const NtxGate& gate,
                            size_t key_buf_elems, cudaStream_t st) {
    const long long T = colpart ? (long long)plan.row_blocks * (plan.row_blocks + 1) / 2
                                : (long long)plan.row_blocks * plan.key_tiles;
    const int G = colpart ? plan.fwd_grid : plan.grid;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    const int lanes = lanes_for(dm.ld / 4);
    const int grid = ntx_fwd_finalize_blocks(dm.q_rows, dm.ld);
    const int threads = fwd_finalize_threads(dm.q_rows, lanes);
    if (dtype == NTX_F32)
        NTX_CUDA(ntx_launch_dep(fwd_finalize_kernel<float>, grid, threads, 0, st, (const float*)q, dm, lanes,
                                plan.tile_rows, plan.key_tiles, T, G, rows_pad, partial, colpart, counter,
                                rowsum, pos_logit, row_loss, loss_sum, peers, rowsum_off, gate,
                                key_buf_elems));
    else
        NTX_CUDA(ntx_launch_dep(fwd_finalize_kernel<__nv_bfloat16>, grid, threads, 0, st,
                                (const __nv_bfloat16*)q, dm, lanes, plan.tile_rows, plan.key_tiles, T, G, rows_pad,
                                partial, colpart, counter, rowsum, pos_logit, row_loss, loss_sum, peers,
                                rowsum_off, gate, key_buf_elems));
    return NTX_OK;
}

int ntx_bwd_finalize_launch(const NtxDims& dm, const NtxPlan& plan, const float* partial,
                            const NtxWeights& w, const void* view0, const void* view1, int in_dtype,
                            const float* inv_norm, void* dview0, void* dview1, cudaStream_t st, int in_ld,
                            int normalised) {
    BwdFinArgs A;
    A.dm = dm;
    A.tile_rows = plan.tile_rows;
    A.key_tiles = plan.key_tiles;
    A.G = plan.grid;
    A.rows_pad = plan.row_blocks * plan.tile_rows;
    A.d_pad = plan.d_pad;
    A.Tt = (long long)plan.row_blocks * plan.key_tiles;
    A.partial = partial;
    A.w = w;
    A.inv_norm = inv_norm;
    A.in_ld = in_ld > 0 ? in_ld : dm.d;
    A.normalised = normalised;
    if (in_dtype == NTX_F32) return launch_bwd_finalize<float>(A, view0, view1, dview0, dview1, st);
    return launch_bwd_finalize<__nv_bfloat16>(A, view0, view1, dview0, dview1, st);
}
