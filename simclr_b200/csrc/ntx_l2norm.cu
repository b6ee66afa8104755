// K1: row-wise L2 normalisation of both views into one row buffer.
// Replaces F.normalize(view, p=2, dim=1) at loss.py:33-34: x / max(||x||_2, 1e-12).
//
// HBM-bound: reads 2n*d input elements once (second pass hits L1), writes 2n*ld
// output elements + 2n floats.  A group of lanes per row, 128-bit accesses when the row
// length allows, fp32 accumulation.  Output rows are zero padded from d to ld so
// the tensor-core kernels can TMA whole 64-element swizzle atoms.
#include "ntx_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 8;      // (4 and 16 measured: +0.2 / +0.5 us per step)

template <typename T> struct Vec;   // 16-byte vector of T
template <> struct Vec<float> {
    static constexpr int N = 4;
    static __device__ __forceinline__ void load(const float* p, float (&v)[4]) {
        float4 t = __ldg(reinterpret_cast<const float4*>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
};
template <> struct Vec<__nv_bfloat16> {
    static constexpr int N = 8;
    static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
        uint4 t = __ldg(reinterpret_cast<const uint4*>(p));
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            float2 f = __bfloat1622float2(h[i]);
            v[2 * i] = f.x; v[2 * i + 1] = f.y;
        }
    }
};

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
__device__ __forceinline__ void store1(float* p, float x) { *p = x; }
__device__ __forceinline__ void store1(__nv_bfloat16* p, float x) { *p = __float2bfloat16_rn(x); }

template <int N>
__device__ __forceinline__ void storev(float* p, const float (&v)[N]) {
#pragma unroll
    for (int i = 0; i < N; i += 4)
        *reinterpret_cast<float4*>(p + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
}
template <int N>
__device__ __forceinline__ void storev(__nv_bfloat16* p, const float (&v)[N]) {
    static_assert(N % 2 == 0, "");
    __nv_bfloat162 h[N / 2];
#pragma unroll
    for (int i = 0; i < N / 2; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    if constexpr (N == 8) {
        *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(h);
    } else {
        *reinterpret_cast<uint2*>(p) = *reinterpret_cast<const uint2*>(h);
    }
}

// `lanes` consecutive threads (power of two <= 32) share one row, so short rows
// (d = 128 bf16 is only 16 sixteen-byte loads) still keep every lane busy.
//
// Peer mode (peers.world > 0): the rows go into the key buffer of the step in EVERY rank's exchange block
// (rows [rank*2n, (rank+1)*2n) there), over NVLink for the other ranks.  The step number is this rank's own
// count of normalise calls (control word 0 of its block): it picks the key buffer (step & 1), is handed to the
// later kernels of the step through *step_out, and is what the last CTA to finish stores -- after a system-scope
// fence -- into kflag[rank] of every block, which is what the readers on the other ranks wait for.
template <typename TIn, typename TOut, bool kVector>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
l2norm_kernel(const TIn* __restrict__ view0, const TIn* __restrict__ view1, int n, int d,
              TOut* __restrict__ rows_local, NtxPeerOut peers, size_t keys_off, size_t key_buf_bytes,
              uint32_t* __restrict__ step_out, int ld, float* __restrict__ inv_norm, int lanes, float row_scale) {
    ntx_dep_trigger();
    ntx_dep_wait();
    const int sub = threadIdx.x % lanes;
    const int rows_per_block = kWarpsPerBlock * 32 / lanes;
    const int r_raw = blockIdx.x * rows_per_block + threadIdx.x / lanes;
    const bool live = r_raw < 2 * n;
    const int r = live ? r_raw : 2 * n - 1;                   // keep every lane in the shuffles
    const TIn* src = (r < n) ? view0 + (size_t)r * d : view1 + (size_t)(r - n) * d;
    constexpr int V = Vec<TIn>::N;
    const int ndst = peers.world > 0 ? peers.world : 1;
    uint32_t* ctl = nullptr;
    uint32_t step = 0;
    size_t dst_off = (size_t)r * ld;                          // elements, same place in every destination
    size_t dst_bytes = 0;
    if (peers.world > 0) {
        ctl = reinterpret_cast<uint32_t*>(peers.base[peers.rank]);
        step = ctl[NTX_CTL_STEP] + 1;     // plain cached load (one word for the whole grid: an L2-only load from every
                                          // thread serialises on it); only the last CTA, once all have read it, bumps it
        dst_off += (size_t)peers.rank * 2 * n * ld;
        dst_bytes = keys_off + (step & 1u) * key_buf_bytes;
    }
    auto dst_rows = [&](int p) -> TOut* {
        return peers.world > 0 ? reinterpret_cast<TOut*>(peers.base[p] + dst_bytes) : rows_local;
    };

    float ss = 0.f;
    if constexpr (kVector) {
        for (int c = sub * V; c < d; c += lanes * V) {
            float v[V];
            Vec<TIn>::load(src + c, v);
#pragma unroll
            for (int i = 0; i < V; ++i) ss = fmaf(v[i], v[i], ss);
        }
    } else {
        for (int c = sub; c < d; c += lanes) {
            float v = to_f32(src[c]);
            ss = fmaf(v, v, ss);
        }
    }
    for (int o = lanes >> 1; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float denom = fmaxf(sqrtf(ss), 1e-12f);   // clamp_min(eps) of F.normalize
    if (live) {
        if (sub == 0) inv_norm[r] = 1.0f / denom;
        if constexpr (kVector) {
            for (int c = sub * V; c < ld; c += lanes * V) {   // ld is a multiple of V on this path
                float v[V];
                if (c < d) {
                    Vec<TIn>::load(src + c, v);
#pragma unroll
                    for (int i = 0; i < V; ++i) v[i] = v[i] / denom * row_scale;     // row_scale == 1: exact
                } else {
#pragma unroll
                    for (int i = 0; i < V; ++i) v[i] = 0.f;
                }
                if (peers.world > 0 && peers.mc) {            // multicast stores: 16 bytes each (8 for 4 bf16 values)
                    __align__(16) TOut tmp[V];
                    storev<V>(tmp, v);
                    constexpr int kBytes = V * (int)sizeof(TOut);
                    const size_t at = dst_bytes + (dst_off + c) * sizeof(TOut);
                    if constexpr (kBytes == 8) {
                        ntx_mc_st_8B(peers.mc, at, *reinterpret_cast<const uint2*>(tmp));
                    } else {
#pragma unroll
                        for (int q = 0; q < kBytes / 16; ++q)
                            ntx_mc_st_16B(peers.mc, at + q * 16, reinterpret_cast<const uint4*>(tmp)[q]);
                    }
                } else {
                    for (int p = 0; p < ndst; ++p) storev<V>(dst_rows(p) + dst_off + c, v);
                }
            }
        } else {
            for (int c = sub; c < ld; c += lanes) {
                const float x = c < d ? to_f32(src[c]) / denom * row_scale : 0.f;
                for (int p = 0; p < ndst; ++p) store1(dst_rows(p) + dst_off + c, x);
            }
        }
    }
    if (peers.world > 0) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();                                  // the CTA's stores (barrier above) before its ticket
            if (atomicAdd(ctl + NTX_CTL_TICKET, 1u) == gridDim.x - 1) {
                __threadfence();
                ctl[NTX_CTL_TICKET] = 0;
                ctl[NTX_CTL_STEP] = step;
                *step_out = step;
                ntx_raise_flags(peers, NTX_CTL_KFLAG, step);
            }
        }
    }
}

template <typename TIn, typename TOut>
int launch(const void* v0, const void* v1, int n, int d, void* rows, const NtxPeerOut& peers, size_t keys_off,
           size_t key_buf_bytes, uint32_t* step_out, int ld, float* inv_norm, float row_scale, cudaStream_t st) {
    constexpr int V = Vec<TIn>::N;
    const bool vec = (d % V == 0) && (ld % V == 0) &&
                     (reinterpret_cast<uintptr_t>(v0) % 16 == 0) &&
                     (reinterpret_cast<uintptr_t>(v1) % 16 == 0) &&
                     [&] {
                         if (peers.world <= 0) return reinterpret_cast<uintptr_t>(rows) % 16 == 0;
                         for (int p = 0; p < peers.world; ++p)
                             if (reinterpret_cast<uintptr_t>(peers.base[p]) % 16 != 0) return false;
                         return true;
                     }();
    const int work = vec ? (ld + V - 1) / V : ld;
    int lanes = 1;
    while (lanes < work && lanes < 32) lanes <<= 1;
    const int rows_per_block = kWarpsPerBlock * 32 / lanes;
    const int grid = (2 * n + rows_per_block - 1) / rows_per_block;
    if (vec)
        NTX_CUDA(ntx_launch_dep(l2norm_kernel<TIn, TOut, true>, grid, kWarpsPerBlock * 32, 0, st, (const TIn*)v0,
                                (const TIn*)v1, n, d, (TOut*)rows, peers, keys_off, key_buf_bytes, step_out, ld,
                                inv_norm, lanes, row_scale));
    else
        NTX_CUDA(ntx_launch_dep(l2norm_kernel<TIn, TOut, false>, grid, kWarpsPerBlock * 32, 0, st, (const TIn*)v0,
                                (const TIn*)v1, n, d, (TOut*)rows, peers, keys_off, key_buf_bytes, step_out, ld,
                                inv_norm, lanes, row_scale));
    return NTX_OK;
}

// Row weights of the backward (the upstream gradient of this rank's rows) into the weight area of every rank's
// block, then wflag[rank] = step there.  w_stride == 0 (one upstream gradient for the whole rank): ONE value, at
// weight[rank]; w_stride == 1: q_rows values at weight[rank * q_rows ..].  The last CTA raises the flags.
__global__ void __launch_bounds__(1024)
post_weights_kernel(const float* __restrict__ w, int w_stride, int w_bf16, int q_rows, NtxPeerOut peers,
                    size_t weight_off, const uint32_t* __restrict__ step_ptr) {
    ntx_dep_trigger();
    ntx_dep_wait();
    if (w_stride == 0) {
        if (threadIdx.x == 0)
            ntx_peer_st_f32(peers, weight_off + (size_t)peers.rank * sizeof(float), ntx_ld_weight(w, 0, w_bf16));
    } else {
        const int r = blockIdx.x * blockDim.x + threadIdx.x;
        if (r < q_rows)
            ntx_peer_st_f32(peers, weight_off + ((size_t)peers.rank * q_rows + r) * sizeof(float),
                            ntx_ld_weight(w, r, w_bf16));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t* ctl = reinterpret_cast<uint32_t*>(peers.base[peers.rank]);
        __threadfence();
        if (gridDim.x == 1 || atomicAdd(ctl + NTX_CTL_POST_TICKET, 1u) == gridDim.x - 1) {
            __threadfence();
            ctl[NTX_CTL_POST_TICKET] = 0;
            ntx_raise_flags(peers, NTX_CTL_WFLAG, *step_ptr);
        }
    }
}


// NTX_BF16X2 rows: the fp32 normalised (and scaled) row x as two bf16 terms, hi = bf16(x) in columns [0, dh) and
// lo = bf16(x - hi) in [dh, 2 dh), zero padded from d to dh.  One warp per row, four columns (8 bytes of each term)
// per lane and pass.  Peer mode as in l2norm_kernel: the row goes into the step's key buffer of every rank's block.
template <typename TIn>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
l2norm_split_kernel(const TIn* __restrict__ view0, const TIn* __restrict__ view1, int n, int d,
                    __nv_bfloat16* __restrict__ rows_local, NtxPeerOut peers, size_t keys_off, size_t key_buf_bytes,
                    uint32_t* __restrict__ step_out, int dh, float* __restrict__ inv_norm, float row_scale) {
    ntx_dep_trigger();
    ntx_dep_wait();
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    const int ndst = peers.world > 0 ? peers.world : 1;
    uint32_t* ctl = nullptr;
    uint32_t step = 0;
    size_t dst_off = (size_t)r * 2 * dh;                      // elements, same place in every destination
    size_t dst_bytes = 0;
    if (peers.world > 0) {
        ctl = reinterpret_cast<uint32_t*>(peers.base[peers.rank]);
        step = ctl[NTX_CTL_STEP] + 1;                         // see l2norm_kernel
        dst_off += (size_t)peers.rank * 2 * n * 2 * dh;
        dst_bytes = keys_off + (step & 1u) * key_buf_bytes;
    }
    if (r < 2 * n) {                                          // whole warps share a row
        const TIn* src = (r < n) ? view0 + (size_t)r * d : view1 + (size_t)(r - n) * d;
        float ss = 0.f;
        for (int c = lane; c < d; c += 32) {
            const float v = to_f32(src[c]);
            ss = fmaf(v, v, ss);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
        const float denom = fmaxf(sqrtf(ss), 1e-12f);         // clamp_min(eps) of F.normalize
        if (lane == 0) inv_norm[r] = 1.0f / denom;
        for (int c = 4 * lane; c < dh; c += 128) {
            float x[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = c + j < d ? to_f32(src[c + j]) / denom * row_scale : 0.f;
            __nv_bfloat162 hi[2], lo[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                hi[j] = __floats2bfloat162_rn(x[2 * j], x[2 * j + 1]);
                const float2 hf = __bfloat1622float2(hi[j]);
                lo[j] = __floats2bfloat162_rn(x[2 * j] - hf.x, x[2 * j + 1] - hf.y);   // the differences are exact in fp32
            }
            const uint2 hv = *reinterpret_cast<const uint2*>(hi), lv = *reinterpret_cast<const uint2*>(lo);
            if (peers.world > 0 && peers.mc) {                // one multicast store per 8 bytes instead of one per rank
                const size_t at = dst_bytes + (dst_off + c) * sizeof(__nv_bfloat16);
                ntx_mc_st_8B(peers.mc, at, hv);
                ntx_mc_st_8B(peers.mc, at + (size_t)dh * sizeof(__nv_bfloat16), lv);
            } else {
                for (int p = 0; p < ndst; ++p) {
                    __nv_bfloat16* dst = (peers.world > 0 ? reinterpret_cast<__nv_bfloat16*>(peers.base[p] + dst_bytes)
                                                          : rows_local) + dst_off;
                    *reinterpret_cast<uint2*>(dst + c) = hv;
                    *reinterpret_cast<uint2*>(dst + dh + c) = lv;
                }
            }
        }
    }
    if (peers.world > 0) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();                                  // the CTA's stores (barrier above) before its ticket
            if (atomicAdd(ctl + NTX_CTL_TICKET, 1u) == gridDim.x - 1) {
                __threadfence();
                ctl[NTX_CTL_TICKET] = 0;
                ctl[NTX_CTL_STEP] = step;
                *step_out = step;
                ntx_raise_flags(peers, NTX_CTL_KFLAG, step);
            }
        }
    }
}

}  // namespace

int ntx_l2norm_split_launch(const void* v0, const void* v1, int n, int d, int in_dtype, void* rows,
                            const NtxPeerOut& peers, size_t keys_off, size_t key_buf_bytes, uint32_t* step_out, int ld,
                            float* inv_norm, float row_scale, cudaStream_t st) {
    const int grid = (2 * n + kWarpsPerBlock - 1) / kWarpsPerBlock;
    if (in_dtype == NTX_F32)
        NTX_CUDA(ntx_launch_dep(l2norm_split_kernel<float>, grid, kWarpsPerBlock * 32, 0, st, (const float*)v0,
                                (const float*)v1, n, d, (__nv_bfloat16*)rows, peers, keys_off, key_buf_bytes, step_out,
                                ld / 2, inv_norm, row_scale));
    else
        NTX_CUDA(ntx_launch_dep(l2norm_split_kernel<__nv_bfloat16>, grid, kWarpsPerBlock * 32, 0, st,
                                (const __nv_bfloat16*)v0, (const __nv_bfloat16*)v1, n, d, (__nv_bfloat16*)rows, peers,
                                keys_off, key_buf_bytes, step_out, ld / 2, inv_norm, row_scale));
    return NTX_OK;
}

int ntx_post_weights_launch(const void* w, int w_stride, int w_bf16, int q_rows, const NtxPeerOut& peers,
                            size_t weight_off, const uint32_t* step, cudaStream_t st) {
    const int threads = w_stride == 0 ? 32 : 256;
    const int grid = w_stride == 0 ? 1 : (q_rows + threads - 1) / threads;
    NTX_CUDA(ntx_launch_dep(post_weights_kernel, grid, threads, 0, st, (const float*)w, w_stride, w_bf16, q_rows, peers,
                            weight_off, step));
    return NTX_OK;
}

int ntx_l2norm_launch(const void* v0, const void* v1, int n, int d, int in_dtype, void* rows, const NtxPeerOut& peers,
                      size_t keys_off, size_t key_buf_bytes, uint32_t* step_out, int ld, int rows_dtype,
                      float* inv_norm, float row_scale, cudaStream_t st) {
    if (in_dtype == NTX_F32 && rows_dtype == NTX_F32)
        return launch<float, float>(v0, v1, n, d, rows, peers, keys_off, key_buf_bytes, step_out, ld, inv_norm, row_scale, st);
    if (in_dtype == NTX_F32 && rows_dtype == NTX_BF16)
        return launch<float, __nv_bfloat16>(v0, v1, n, d, rows, peers, keys_off, key_buf_bytes, step_out, ld, inv_norm, row_scale, st);
    if (in_dtype == NTX_BF16 && rows_dtype == NTX_BF16)
        return launch<__nv_bfloat16, __nv_bfloat16>(v0, v1, n, d, rows, peers, keys_off, key_buf_bytes, step_out, ld, inv_norm, row_scale, st);
    if (in_dtype == NTX_BF16 && rows_dtype == NTX_F32)
        return launch<__nv_bfloat16, float>(v0, v1, n, d, rows, peers, keys_off, key_buf_bytes, step_out, ld, inv_norm, row_scale, st);
    ntx_set_error("ntx_l2norm: unsupported dtype pair (%d -> %d)", in_dtype, rows_dtype);
    return NTX_E_ARG;
}
