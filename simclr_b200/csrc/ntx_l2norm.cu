// K1: row-wise L2 normalisation of both views into one row buffer.
// Replaces F.normalize(view, p=2, dim=1) at loss.py:33-34: x / max(||x||_2, 1e-12).
//
// HBM-bound: reads 2n*d input elements once (second pass hits L1), writes 2n*ld
// output elements + 2n floats.  One warp per row, 128-bit accesses when the row
// length allows, fp32 accumulation.  Output rows are zero padded from d to ld so
// the tensor-core kernels can TMA whole 64-element swizzle atoms.
#include "ntx_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 8;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

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

template <typename TIn, typename TOut, bool kVector>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
l2norm_kernel(const TIn* __restrict__ view0, const TIn* __restrict__ view1, int n, int d,
              TOut* __restrict__ rows, int ld, float* __restrict__ inv_norm) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= 2 * n) return;
    const TIn* src = (r < n) ? view0 + (size_t)r * d : view1 + (size_t)(r - n) * d;
    TOut* dst = rows + (size_t)r * ld;
    constexpr int V = Vec<TIn>::N;

    float ss = 0.f;
    if constexpr (kVector) {
        for (int c = lane * V; c < d; c += 32 * V) {
            float v[V];
            Vec<TIn>::load(src + c, v);
#pragma unroll
            for (int i = 0; i < V; ++i) ss = fmaf(v[i], v[i], ss);
        }
    } else {
        for (int c = lane; c < d; c += 32) {
            float v = to_f32(src[c]);
            ss = fmaf(v, v, ss);
        }
    }
    ss = warp_sum(ss);
    const float denom = fmaxf(sqrtf(ss), 1e-12f);   // clamp_min(eps) of F.normalize
    if (lane == 0) inv_norm[r] = 1.0f / denom;

    if constexpr (kVector) {
        for (int c = lane * V; c < ld; c += 32 * V) {   // ld is a multiple of V on this path
            float v[V];
            if (c < d) {
                Vec<TIn>::load(src + c, v);
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = v[i] / denom;
            } else {
#pragma unroll
                for (int i = 0; i < V; ++i) v[i] = 0.f;
            }
            storev<V>(dst + c, v);
        }
    } else {
        for (int c = lane; c < ld; c += 32) store1(dst + c, c < d ? to_f32(src[c]) / denom : 0.f);
    }
}

template <typename TIn, typename TOut>
int launch(const void* v0, const void* v1, int n, int d, void* rows, int ld, float* inv_norm,
           cudaStream_t st) {
    constexpr int V = Vec<TIn>::N;
    const bool vec = (d % V == 0) && (ld % V == 0) &&
                     (reinterpret_cast<uintptr_t>(v0) % 16 == 0) &&
                     (reinterpret_cast<uintptr_t>(v1) % 16 == 0) &&
                     (reinterpret_cast<uintptr_t>(rows) % 16 == 0);
    const int grid = (2 * n + kWarpsPerBlock - 1) / kWarpsPerBlock;
    if (vec)
        l2norm_kernel<TIn, TOut, true><<<grid, kWarpsPerBlock * 32, 0, st>>>(
            (const TIn*)v0, (const TIn*)v1, n, d, (TOut*)rows, ld, inv_norm);
    else
        l2norm_kernel<TIn, TOut, false><<<grid, kWarpsPerBlock * 32, 0, st>>>(
            (const TIn*)v0, (const TIn*)v1, n, d, (TOut*)rows, ld, inv_norm);
    NTX_CUDA(cudaGetLastError());
    return NTX_OK;
}

}  // namespace

int ntx_l2norm_launch(const void* v0, const void* v1, int n, int d, int in_dtype, void* rows, int ld,
                      int rows_dtype, float* inv_norm, cudaStream_t st) {
    if (in_dtype == NTX_F32 && rows_dtype == NTX_F32)
        return launch<float, float>(v0, v1, n, d, rows, ld, inv_norm, st);
    if (in_dtype == NTX_F32 && rows_dtype == NTX_BF16)
        return launch<float, __nv_bfloat16>(v0, v1, n, d, rows, ld, inv_norm, st);
    if (in_dtype == NTX_BF16 && rows_dtype == NTX_BF16)
        return launch<__nv_bfloat16, __nv_bfloat16>(v0, v1, n, d, rows, ld, inv_norm, st);
    if (in_dtype == NTX_BF16 && rows_dtype == NTX_F32)
        return launch<__nv_bfloat16, float>(v0, v1, n, d, rows, ld, inv_norm, st);
    ntx_set_error("ntx_l2norm: unsupported dtype pair (%d -> %d)", in_dtype, rows_dtype);
    return NTX_E_ARG;
}
