// Fused LARC + SGD-momentum step over all parameter tensors in ONE launch.
//
// Replaces, for fp32 parameters, what the reference does after loss.backward() (main.py:85-94,116):
//   apex.parallel.LARC(torch.optim.SGD(lr, momentum, nesterov=False, weight_decay=0 + per-group decay),
//                      trust_coefficient=0.001, clip=False).step()
// i.e. per parameter tensor p with gradient g, momentum buffer m and weight decay wd (apex LARC.step, then
// torch.optim.SGD's update):
//   if ||p|| != 0 and ||g|| != 0:  a = trust * ||p|| / (||g|| + wd * ||p|| + eps)      [clip: a = min(a / lr, 1)]
//                                   g' = (g + wd * p) * a
//   else:                           g' = g                                              (no decay either)
//   m = momentum * m + g';  p = p - lr * m          (a zero-initialised m reproduces SGD's first step: m = g')
// The reference launches ~10 small kernels per tensor (two norms with a host sync each through `!= 0`, the in-place
// gradient edits, the SGD update); here a cooperative grid walks a chunk list: sums of squares per chunk, a grid
// barrier, per-tensor totals added in a fixed order (deterministic), a second barrier, the update.  HBM-bound:
// 3 reads + 2 writes of 4 bytes per element in the update, the norm pass re-reads p and g (mostly from L2).
#include "ntx_common.cuh"

namespace {

constexpr int kThreads = 256;

struct Partial {
    double pp, gg;
};

__device__ __forceinline__ float block_sum(float v, float* sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
    if (threadIdx.x < kThreads / 32) t = sh[threadIdx.x];
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    }
    __syncthreads();
    return t;   // valid in thread 0
}

// all CTAs are co-resident (cooperative launch): arrive, then spin until every CTA has
__device__ __forceinline__ void grid_barrier(unsigned int* word, unsigned int target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(word, 1u);
        while (*reinterpret_cast<volatile unsigned int*>(word) < target) __nanosleep(64);
        __threadfence();
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kThreads)
larc_sgd_kernel(const NtxLarcParam* __restrict__ table, int n_tensors, const int2* __restrict__ chunks, int n_chunks,
                int chunk_elems, float lr, float momentum, float trust, float eps, int clip,
                Partial* __restrict__ partial, Partial* __restrict__ totals, unsigned int* __restrict__ bar) {
    __shared__ float sh[2][kThreads / 32];
    const int G = gridDim.x, cta = blockIdx.x;
    const int i_begin = (int)((long long)cta * n_chunks / G), i_end = (int)((long long)(cta + 1) * n_chunks / G);

    // ---- pass 1: sum of squares of p and g per chunk
    for (int i = i_begin; i < i_end; ++i) {
        const int2 ch = chunks[i];                                   // (tensor, chunk within the tensor)
        const NtxLarcParam P = table[ch.x];
        const long long base = (long long)ch.y * chunk_elems;
        const int len = (int)min((long long)chunk_elems, P.numel - base);
        const float* p = P.param + base;
        const float* g = P.grad + base;
        float sp = 0.f, sg = 0.f;
        const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g)) & 15) == 0;
        const int nv = vec ? len / 4 : 0;
        for (int v = threadIdx.x; v < nv; v += kThreads) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(p) + v);
            const float4 b = __ldg(reinterpret_cast<const float4*>(g) + v);
            sp = fmaf(a.x, a.x, fmaf(a.y, a.y, fmaf(a.z, a.z, fmaf(a.w, a.w, sp))));
            sg = fmaf(b.x, b.x, fmaf(b.y, b.y, fmaf(b.z, b.z, fmaf(b.w, b.w, sg))));
        }
        for (int e = nv * 4 + threadIdx.x; e < len; e += kThreads) {
            const float a = __ldg(p + e), b = __ldg(g + e);
            sp = fmaf(a, a, sp);
            sg = fmaf(b, b, sg);
        }
        const float tp = block_sum(sp, sh[0]);
        const float tg = block_sum(sg, sh[1]);
        if (threadIdx.x == 0) partial[i] = Partial{(double)tp, (double)tg};
    }
    grid_barrier(bar, (unsigned)G);

    // ---- pass 1.5: per-tensor totals by one warp per tensor: lane l adds chunks l, l + 32, ... in order, then a
    // fixed shuffle tree adds the 32 lane sums -- the same order in every run
    {
        const int lane = threadIdx.x & 31;
        const int warps = G * (kThreads / 32);
        for (int t = cta * (kThreads / 32) + (threadIdx.x >> 5); t < n_tensors; t += warps) {
            const NtxLarcParam P = table[t];
            double pp = 0.0, gg = 0.0;
            for (int c = lane; c < P.n_chunks; c += 32) {
                const Partial q = partial[P.first_chunk + c];
                pp += q.pp;
                gg += q.gg;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                pp += __shfl_xor_sync(0xffffffffu, pp, o);
                gg += __shfl_xor_sync(0xffffffffu, gg, o);
            }
            if (lane == 0) totals[t] = Partial{pp, gg};
        }
    }
    grid_barrier(bar, 2u * (unsigned)G);

    // ---- pass 2: the update, walking the chunks backwards (the last chunks of pass 1 are the likeliest L2 hits)
    for (int i = i_end - 1; i >= i_begin; --i) {
        const int2 ch = chunks[i];
        const NtxLarcParam P = table[ch.x];
        const Partial tot = totals[ch.x];
        const float pn = (float)sqrt(tot.pp), gn = (float)sqrt(tot.gg);
        const float wd = P.weight_decay;
        float a = 1.f, decay = 0.f;
        if (pn != 0.f && gn != 0.f) {                                // apex LARC.step: both norms non-zero
            a = trust * pn / (gn + pn * wd + eps);
            if (clip) a = fminf(a / lr, 1.f);
            decay = wd;
        }
        const long long base = (long long)ch.y * chunk_elems;
        const int len = (int)min((long long)chunk_elems, P.numel - base);
        float* p = P.param + base;
        const float* g = P.grad + base;
        float* m = P.momentum + base;
        const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
                           reinterpret_cast<uintptr_t>(m)) & 15) == 0;
        const int nv = vec ? len / 4 : 0;
        auto upd = [&](float pv, float gv, float mv, float& p_out, float& m_out) {
            const float g2 = (gv + decay * pv) * a;
            m_out = momentum * mv + g2;
            p_out = pv - lr * m_out;
        };
        for (int v = threadIdx.x; v < nv; v += kThreads) {
            float4 pv = reinterpret_cast<const float4*>(p)[v];
            const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + v);
            float4 mv = reinterpret_cast<const float4*>(m)[v];
            upd(pv.x, gv.x, mv.x, pv.x, mv.x);
            upd(pv.y, gv.y, mv.y, pv.y, mv.y);
            upd(pv.z, gv.z, mv.z, pv.z, mv.z);
            upd(pv.w, gv.w, mv.w, pv.w, mv.w);
            reinterpret_cast<float4*>(p)[v] = pv;
            reinterpret_cast<float4*>(m)[v] = mv;
        }
        for (int e = nv * 4 + threadIdx.x; e < len; e += kThreads) upd(p[e], __ldg(g + e), m[e], p[e], m[e]);
    }

    // ---- leave the barrier word at zero for the next step: the last CTA out resets it
    __syncthreads();
    if (threadIdx.x == 0) {
        if (atomicAdd(bar + 1, 1u) == (unsigned)G - 1) {
            bar[0] = 0;
            bar[1] = 0;
        }
    }
}

}  // namespace

extern "C" {

size_t ntx_larc_workspace_bytes(int n_tensors, int n_chunks) {
    if (n_tensors < 0 || n_chunks < 0) return 0;
    return 256 + sizeof(Partial) * ((size_t)n_chunks + (size_t)n_tensors);
}

int ntx_larc_sgd_step(const NtxLarcParam* table, int n_tensors, const int32_t* chunks, int n_chunks, int chunk_elems,
                      float lr, float momentum, float trust_coefficient, float eps, int clip, void* workspace,
                      size_t workspace_bytes, void* stream) {
    NTX_CHECK_ARG(n_tensors >= 0 && n_chunks >= 0, "bad counts n_tensors=%d n_chunks=%d", n_tensors, n_chunks);
    if (n_tensors == 0 || n_chunks == 0) return NTX_OK;
    NTX_CHECK_ARG(table && chunks && workspace, "NULL buffer");
    NTX_CHECK_ARG(chunk_elems >= 4 && chunk_elems % 4 == 0, "chunk_elems=%d must be a positive multiple of 4", chunk_elems);
    NTX_CHECK_ARG(reinterpret_cast<uintptr_t>(workspace) % 16 == 0 && reinterpret_cast<uintptr_t>(table) % 8 == 0 &&
                      reinterpret_cast<uintptr_t>(chunks) % 8 == 0,
                  "workspace must be 16-byte aligned, the tables 8-byte aligned");
    NTX_CHECK_ARG(lr >= 0.f && momentum >= 0.f && trust_coefficient >= 0.f && eps >= 0.f, "negative hyper-parameter");
    const size_t need = ntx_larc_workspace_bytes(n_tensors, n_chunks);
    if (workspace_bytes < need) {
        ntx_set_error("ntx_larc_sgd_step needs %llu bytes of workspace, got %llu", (unsigned long long)need,
                      (unsigned long long)workspace_bytes);
        return NTX_E_WORKSPACE;
    }
    static int per_sm[64];
    static int sms[64];
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) dev = 0;
    if (!per_sm[dev]) {
        int n = 0, s = 0;
        NTX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, larc_sgd_kernel, kThreads, 0));
        NTX_CUDA(cudaDeviceGetAttribute(&s, cudaDevAttrMultiProcessorCount, dev));
        per_sm[dev] = n > 0 ? n : 1;
        sms[dev] = s;
    }
    int grid = per_sm[dev] * sms[dev];
    if (grid > n_chunks) grid = n_chunks;
    unsigned int* bar = static_cast<unsigned int*>(workspace);
    Partial* partial = reinterpret_cast<Partial*>(static_cast<char*>(workspace) + 256);
    Partial* totals = partial + n_chunks;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;      // every CTA resident at once: the grid barriers cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ntx_note_launch();
    NTX_CUDA(cudaLaunchKernelEx(&cfg, larc_sgd_kernel, table, n_tensors, reinterpret_cast<const int2*>(chunks), n_chunks,
                                chunk_elems, lr, momentum, trust_coefficient, eps, clip, partial, totals, bar));
    return NTX_OK;
}

}  // extern "C"
