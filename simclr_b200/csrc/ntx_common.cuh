// Shared host/device helpers of the NT-Xent library: error plumbing, the index
// arithmetic of the self / positive columns, and the tile partition that both
// main kernels and their finalize kernels follow.
#pragma once

#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ntxent_b200.h"

#define NTX_HD __host__ __device__ __forceinline__

// ---------------------------------------------------------------------------
// errors (thread-local message; no exception crosses the C boundary)
// ---------------------------------------------------------------------------
void ntx_set_error(const char* fmt, ...);

#define NTX_CHECK_ARG(cond, ...)            \
    do {                                    \
        if (!(cond)) {                      \
            ntx_set_error(__VA_ARGS__);     \
            return NTX_E_ARG;               \
        }                                   \
    } while (0)

#define NTX_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            ntx_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),  \
                          __FILE__, __LINE__);                                      \
            return NTX_E_CUDA;                                                      \
        }                                                                           \
    } while (0)

// ---------------------------------------------------------------------------
// index arithmetic -- the bit-exact part of the contract (loss.py:37-38,46-52)
// ---------------------------------------------------------------------------
// Row r of this rank (0 <= r < 2*n_local; view0 rows first) among rank-major keys.
NTX_HD int ntx_row_offset(int n_local, int rank) { return rank * 2 * n_local; }
NTX_HD int ntx_self_col(int r, int n_local, int rank) { return ntx_row_offset(n_local, rank) + r; }
NTX_HD int ntx_pos_row(int r, int n_local) { return r < n_local ? r + n_local : r - n_local; }
NTX_HD int ntx_pos_col(int r, int n_local, int rank) {
    return ntx_row_offset(n_local, rank) + ntx_pos_row(r, n_local);
}

// ---------------------------------------------------------------------------
// tile partition
// ---------------------------------------------------------------------------
// Tiles are numbered row-block-major: t = rb * key_tiles + kt.  CTA i owns the
// contiguous range [i*T/G, (i+1)*T/G).  A row block can be shared by several
// consecutive CTAs; each writes its partial into slot (i - first owner of the
// block) and the finalize kernel adds the slots in order (deterministic).
NTX_HD long long ntx_chunk_begin(long long T, int G, int i) { return (long long)i * T / G; }
NTX_HD int ntx_tile_owner(long long T, int G, long long t) { return (int)(((t + 1) * G - 1) / T); }
NTX_HD int ntx_first_owner(long long T, int G, int key_tiles, int rb) {
    return ntx_tile_owner(T, G, (long long)rb * key_tiles);
}
NTX_HD int ntx_num_slots(long long T, int G, int key_tiles, int rb) {
    return ntx_tile_owner(T, G, (long long)rb * key_tiles + key_tiles - 1) -
           ntx_tile_owner(T, G, (long long)rb * key_tiles) + 1;
}

// Upper-triangle enumeration used by the symmetric forward (keys == rows, S = S^T): only tiles
// (rb, kt >= rb) are computed.  Row block rb has key_tiles - rb of them, so the long rows are paired with
// the short ones: the order is 0, n-1, 1, n-2, ... and every pair holds n + 1 tiles.  An equal share of
// the tile list then also holds an equal share of row-block changes and diagonal tiles for every CTA
// (numbered 0, 1, 2, ... the last CTA got all the short rows: ~10 row-block changes against ~1).
// Row block rb starts at tile ntx_tri_first(rb); its tiles are contiguous.
NTX_HD long long ntx_tri_first(int key_tiles, int rb) {
    const int partner = key_tiles - 1 - rb;
    return rb <= partner ? (long long)rb * (key_tiles + 1)                                   // first of its pair
                         : (long long)partner * (key_tiles + 1) + (key_tiles - partner);     // after its partner
}
NTX_HD int ntx_tri_first_owner(long long T, int G, int key_tiles, int rb) {
    return ntx_tile_owner(T, G, ntx_tri_first(key_tiles, rb));
}
NTX_HD int ntx_tri_num_slots(long long T, int G, int key_tiles, int rb) {
    const long long f = ntx_tri_first(key_tiles, rb);
    return ntx_tile_owner(T, G, f + (key_tiles - rb) - 1) - ntx_tile_owner(T, G, f) + 1;
}

// Opt a kernel in to more than 48 KB of dynamic shared memory once per device instead of on every launch
// (the attribute call costs a few microseconds of host time; a repeated set from two threads is harmless).
#define NTX_SMEM_OPT_IN(kernel, bytes)                                                                     \
    do {                                                                                                   \
        static unsigned done_mask__ = 0;                                                                   \
        int dev__ = 0;                                                                                     \
        NTX_CUDA(cudaGetDevice(&dev__));                                                                   \
        if (!((done_mask__ >> (dev__ & 31)) & 1u)) {                                                       \
            NTX_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            done_mask__ |= 1u << (dev__ & 31);                                                             \
        }                                                                                                  \
    } while (0)

// Programmatic dependent launch: a kernel launched through ntx_launch_dep may start while the previous kernel
// on the stream is still running; it does its set-up (barriers, TMEM, descriptor prefetch), then ntx_dep_wait()
// returns once that previous kernel has completed and its writes are visible.  Every kernel launched this way
// calls ntx_dep_wait() before it touches global memory, and ntx_dep_trigger() right at its start so that the
// kernel after it can be staged early.  NTXENT_NO_PDL=1 in the environment turns the launch attribute off.
#ifdef __CUDACC__
__device__ __forceinline__ void ntx_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void ntx_dep_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
bool ntx_pdl_enabled();
void ntx_note_launch();          // every kernel launch of the library is counted (ntx_launch_count)
template <typename... KArgs, typename... Args>
inline cudaError_t ntx_launch_dep(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                  Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = ntx_pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    ntx_note_launch();
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// log2(e) / tau and friends live in one place so every kernel shifts identically:
// e_rc = exp(s_rc - 1/tau) = exp2(dot_rc * c1 - c1),  c1 = log2(e) / tau.
#define NTX_LOG2E 1.4426950408889634f
// exp2 argument range is [-2*c1, 0]: NTX_MIN_TEMPERATURE (public header) keeps it inside fp32 normal range.

struct NtxDims {
    int q_rows, k_rows, d, ld;
    int n_local, rank;
    float c1;        // log2e / tau
    float inv_tau;
    int stable;      // temperature below NTX_MIN_TEMPERATURE (fp32 rows only): max-shifted logsumexp, and the
                     // "rowsum" vectors hold lse_r = log sum_c exp(s_rc) instead of sum_c exp(s_rc - 1/tau)
    // Rows stored as zhat * sqrt(log2e / tau) (NtxLayout.row_scale, bf16 rows): the similarity MMA then delivers the
    // exponent log2e/tau * <q, k> itself and the tensor-core epilogues use e' = exp2(exponent) = e * 2^c1 without the
    // scale-and-shift FFMA2.  The constant factors are taken out again where the sums are used:
    int scaled;          // 1: rows carry the factor
    float rowsum_scale;  // 2^-c1 : sums of e' -> sums of exp(s - 1/tau)                    (1 when not scaled)
    float pos_scale;     // <row, partner row> -> positive logit s: ln 2                     (1/tau when not scaled)
    float inv_rs;        // 1 / sqrt(c1) : dZ' = C . K' and the rows themselves -> unit rows (1 when not scaled)
    int loss_bf16;       // loss_sum receives two bf16 values instead of two floats (NtxLayout.loss_dtype)
    int split;           // NTX_BF16X2 rows: [hi | lo] bf16 terms, each dh = ld / 2 elements wide
    int dh;
};
#ifdef __CUDACC__
// loss.py:60-65: the reduced loss, sum and mean (= sum / N * 0.5 over the 2N rows), in the dtype the caller asked for
__device__ __forceinline__ void ntx_store_loss(float* loss_sum, double total, int q_rows, int bf16) {
    const float sum = (float)total, mean = (float)(total / (double)q_rows);
    if (bf16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(loss_sum);
        o[0] = __float2bfloat16_rn(sum);
        o[1] = __float2bfloat16_rn(mean);
    } else {
        loss_sum[0] = sum;
        loss_sum[1] = mean;
    }
}
#endif

// Backward weights: alpha_r = w_q[r*stride]*scale / rowsum_q[r],  beta_c = w_k[c*stride]*scale / rowsum_k[c]
struct NtxWeights {
    const float* rowsum_q;
    const float* rowsum_k;
    const float* w_q;
    const float* w_k;
    int w_stride;      // 0: one upstream gradient for every row (mean / sum), 1: one per row (none)
    float w_scale;     // 1 / q_rows for mean, 1 otherwise
    int w_chunk;       // > 0 (peer exchange, one weight per RANK): the weight of key c is w_k[c / w_chunk]
    int w_bf16;        // 1: w_q / w_k point at bf16 values (the upstream gradient of a bf16 loss), 0: fp32
    float coef_scale;  // scale of the weights inside the coefficients alpha, beta: w_scale, times 2^-c1 for scaled rows
};
#ifdef __CUDACC__
// one weight, fp32 or bf16 (read-only for the kernel's lifetime unless `l2`, which skips L1: data written by a peer)
__device__ __forceinline__ float ntx_ld_weight(const float* base, size_t i, int bf16, bool l2 = false) {
    if (bf16) {
        const unsigned short* p = reinterpret_cast<const unsigned short*>(base) + i;
        unsigned short v;
        if (l2) asm volatile("ld.relaxed.gpu.global.u16 %0, [%1];" : "=h"(v) : "l"(p) : "memory");
        else v = __ldg(p);
        return __uint_as_float((unsigned)v << 16);
    }
    if (l2) {
        float v;
        asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(base + i) : "memory");
        return v;
    }
    return __ldg(base + i);
}
#endif
NTX_HD size_t ntx_key_weight_index(int c, int w_stride, int w_chunk) {
    return w_chunk > 0 ? (size_t)(c / w_chunk) : (size_t)c * w_stride;
}

// ---------------------------------------------------------------------------
// peer exchange (one node, NVLink peer memory; include/ntxent_b200.h "Multi-rank exchange")
// ---------------------------------------------------------------------------
// Every rank owns one exchange block, mapped into every process:
//   [0, 256)            control words (uint32): [0] steps taken, [1] / [2] tickets of the normalise / weight-post kernel,
//                       [8 + p] kflag, [16 + p] rflag, [24 + p] wflag: rank p's keys / row sums / row weights
//                       of step >= the stored value have landed in THIS block
//   [256, ...)          float rowsum[world * q_rows], float weight[world * q_rows]  (rank-major, like the keys; with one
//                       upstream gradient per rank only weight[0 .. world) is used: one value per rank)
//   [keys_off, ...)     two key buffers [world * q_rows x ld], used by alternate steps
#define NTX_CTL_BYTES 256
#define NTX_CTL_STEP 0
#define NTX_CTL_TICKET 1
#define NTX_CTL_POST_TICKET 2
#define NTX_CTL_KFLAG 8
#define NTX_CTL_RFLAG 16
#define NTX_CTL_WFLAG 24

struct NtxPeerGeom {           // byte offsets inside an exchange block
    size_t rowsum_off, weight_off, keys_off, key_buf_bytes, total;
};
NTX_HD NtxPeerGeom ntx_peer_geom(int q_rows, int ld, int esz, int world) {
    NtxPeerGeom g;
    const size_t vec = (size_t)world * q_rows * sizeof(float);
    g.rowsum_off = NTX_CTL_BYTES;
    g.weight_off = g.rowsum_off + vec;
    g.keys_off = (g.weight_off + vec + 1023) & ~(size_t)1023;
    g.key_buf_bytes = (((size_t)world * q_rows * ld * esz) + 1023) & ~(size_t)1023;
    g.total = g.keys_off + 2 * g.key_buf_bytes;
    return g;
}

// What a kernel needs to store into every rank's block and to raise the flags there.
struct NtxPeerOut {
    uint8_t* base[NTX_MAX_PEERS];   // block of rank p as mapped into this process; base[rank] is this rank's own
    uint8_t* mc;                    // multicast mapping of the blocks (one multimem.st lands in EVERY rank's block:
                                    // the NVSwitch replicates it), or NULL: one store per rank instead
    int world, rank;                // world == 0: no exchange
};
// What a kernel needs to wait for data other ranks store into this rank's block.
struct NtxGate {
    const uint32_t* step;           // number of the step being computed (device memory; NULL = no exchange)
    const uint32_t* ctl;            // this rank's control words
    int chunk_rows;                 // keys per rank
    int world;
};

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t ntx_ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Raising the flags of a step.  A system-scope fence after stores to peer memory waits for their round trip over
// NVLink (~2 us measured), and st.release.sys carries one each: a fence per thread, per CTA or per flag multiplies
// that.  So: every CTA orders its stores with a GPU-scope fence and takes a ticket; the last CTA alone issues ONE
// system-scope fence -- cumulative over everything the tickets made visible to it -- and then writes the flag
// word of every rank with relaxed stores, back to back.
__device__ __forceinline__ void ntx_fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void ntx_st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// stores that land at the same offset of every rank's block: one multimem.st through the multicast mapping
__device__ __forceinline__ void ntx_mc_st_u32(uint8_t* mc, size_t off, uint32_t v) {
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(mc + off), "r"(v) : "memory");
}
__device__ __forceinline__ void ntx_mc_st_f32(uint8_t* mc, size_t off, float v) {
    asm volatile("multimem.st.weak.global.f32 [%0], %1;" ::"l"(mc + off), "f"(v) : "memory");
}
__device__ __forceinline__ void ntx_mc_st_16B(uint8_t* mc, size_t off, uint4 v) {
    asm volatile("multimem.st.weak.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc + off), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
                 : "memory");
}
__device__ __forceinline__ void ntx_mc_st_8B(uint8_t* mc, size_t off, uint2 v) {
    asm volatile("multimem.st.weak.global.v2.f32 [%0], {%1, %2};" ::"l"(mc + off), "f"(__uint_as_float(v.x)),
                 "f"(__uint_as_float(v.y))
                 : "memory");
}
// one float to the same place of every rank's block
__device__ __forceinline__ void ntx_peer_st_f32(const NtxPeerOut& peers, size_t off, float v) {
    if (peers.mc) {
        ntx_mc_st_f32(peers.mc, off, v);
    } else {
        for (int p = 0; p < peers.world; ++p) *reinterpret_cast<float*>(peers.base[p] + off) = v;
    }
}
__device__ __forceinline__ void ntx_raise_flags(const NtxPeerOut& peers, int word, uint32_t step) {
    ntx_fence_sys();
    const size_t off = (size_t)(word + peers.rank) * sizeof(uint32_t);
    if (peers.mc) {
        ntx_mc_st_u32(peers.mc, off, step);
    } else {
        for (int p = 0; p < peers.world; ++p)
            ntx_st_relaxed_sys(reinterpret_cast<uint32_t*>(peers.base[p] + off), step);
    }
}
__device__ __forceinline__ uint32_t ntx_ld_relaxed(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ntx_ld_l2(const float* p) {            // skips L1: the line may have been written by a peer
    float v;
    asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
// spin until flag word `which + p` of this rank's block has reached `step` (wrap-safe)
__device__ __forceinline__ void ntx_gate_wait(const uint32_t* ctl, int word, uint32_t step) {
    while ((int)(ntx_ld_acquire_sys(ctl + word) - step) < 0) __nanosleep(32);
}
#endif

// ---------------------------------------------------------------------------
// finalize work folded into the tensor-core kernels
// ---------------------------------------------------------------------------
// The CTAs that share a row block each write a partial into their slot, announce it on the row block's arrival
// counter, and -- once they have run out of tiles -- wait until the row block is complete and finish an equal
// share of its rows (slot s of ns takes rows [128 s / ns, 128 (s + 1) / ns)): slots are added in slot order, so
// the result does not depend on who finishes when.  No second kernel, no launch gap, and the wait is short
// because the CTAs of a row block own equally long tile ranges that they started together.
// The counters live at the head of the caller's scratch, which arrives uninitialised: word 0 is a tag.  At kernel
// start every CTA looks at it; unless it holds NTX_CTR_MAGIC, CTA 0 zeroes the counters and then writes the tag
// while the others wait for it (all CTAs of these kernels are co-resident, and the first arrival comes a whole
// tile later).  After that the counters are plain atomicAdd words (a compare-and-swap counter serialises ~150
// simultaneous arrivals at one L2 round trip each), and the last CTA through a counter stores 0 again, so the next
// launch -- a replayed CUDA graph included -- starts from clean words and skips the initialisation.
#define NTX_CTR_MAGIC 0x4E585443u
struct NtxFoldScratch {
    unsigned int* tag;                // NTX_CTR_MAGIC once the counters below have been zeroed
    unsigned int* done_total;         // segments (CTA, row block) finished in this launch
    unsigned int* arrive;             // [row_blocks]
    double* seg_loss;                 // forward: loss of each finished share, indexed first_owner(rb) + slot + rb
    int n_segments;                   // sum over row blocks of their slot counts
    int row_blocks;
    int enabled;                      // 0: the kernel only writes its slots (NTXENT_NO_FOLD=1: separate finalize kernel)
    unsigned int* legacy_ticket;      // the finalize kernel's ticket, zeroed by the main kernel when !enabled
};
NTX_HD size_t ntx_fold_ctl_bytes(int row_blocks) { return (64 + (size_t)row_blocks * 4 + 255) & ~(size_t)255; }
NTX_HD size_t ntx_fold_seg_bytes(int row_blocks, int grid) {
    return ((size_t)(row_blocks + grid) * 8 + 255) & ~(size_t)255;
}
#ifdef __CUDACC__
__device__ __forceinline__ unsigned int ntx_ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// called by ONE thread of every CTA before the CTA's first arrival
__device__ __forceinline__ void ntx_fold_init(unsigned int* tag, unsigned int* done_total, unsigned int* arrive,
                                              int row_blocks, int cta) {
    // the tag also names how many counters were zeroed: a launch with another number of row blocks starts over
    const unsigned int want = NTX_CTR_MAGIC + (unsigned int)row_blocks;
    if (ntx_ld_acquire_gpu(tag) == want) return;
    if (cta == 0) {
        *done_total = 0u;
        for (int i = 0; i < row_blocks; ++i) arrive[i] = 0u;
        __threadfence();
        asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(tag), "r"(want) : "memory");
    } else {
        while (ntx_ld_acquire_gpu(tag) != want) __nanosleep(64);
    }
}
#endif

// launch helpers implemented per translation unit
int ntx_l2norm_launch(const void* v0, const void* v1, int n, int d, int in_dtype, void* rows, const NtxPeerOut& peers,
                      size_t keys_off, size_t key_buf_bytes, uint32_t* step_out, int ld, int rows_dtype,
                      float* inv_norm, float row_scale, cudaStream_t st);
// NTX_BF16X2 rows: [hi | lo] bf16 terms of the fp32 normalised row (ntx_l2norm.cu)
int ntx_l2norm_split_launch(const void* v0, const void* v1, int n, int d, int in_dtype, void* rows,
                            const NtxPeerOut& peers, size_t keys_off, size_t key_buf_bytes, uint32_t* step_out, int ld,
                            float* inv_norm, float row_scale, cudaStream_t st);
int ntx_post_weights_launch(const void* w, int w_stride, int w_bf16, int q_rows, const NtxPeerOut& peers, size_t weight_off,
                            const uint32_t* step, cudaStream_t st);
int ntx_make_rows_tmap(void* map, const void* base, int rows, int ld, int box_rows);      // map = CUtensorMap*
int ntx_project_launch(const void* h0, const void* h1, const void* weight, int n, int hdim, int d, void* rows, int ld,
                       float* inv_norm, float row_scale, cudaStream_t st);
// q / k: the rows; with an exchange (gate.step != NULL) they point at key buffer 0 and buffer 1 lies key_buf_elems
// elements further on -- the kernels pick the buffer of the step
int ntx_simt_fwd_launch(const float* q, const float* k, const NtxDims& dm, const NtxPlan& plan,
                        float* partial, unsigned int* counter, const NtxGate& gate, size_t key_buf_elems,
                        cudaStream_t st);
int ntx_simt_bwd_launch(const float* q, const float* k, const NtxDims& dm, const NtxPlan& plan,
                        const NtxWeights& w, float* partial, const NtxGate& gate, size_t key_buf_elems,
                        cudaStream_t st);
// the full-tile forward finishes its rows itself (no finalize kernel): outputs and fold scratch go in
struct NtxFwdOut {
    float* rowsum;
    float* pos_logit;
    float* row_loss;
    float* loss_sum;
    NtxPeerOut peers;      // world > 0: row sums also into every rank's block, rflag raised at the end
    size_t rowsum_off;
};
int ntx_tc_fwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, float* partial, const NtxFoldScratch& fold, const NtxFwdOut& out,
                      const NtxGate& gate, size_t key_buf_elems, cudaStream_t st);
int ntx_tc_fwd_sym_launch(const __nv_bfloat16* rows, const NtxDims& dm, const NtxPlan& plan, float* partial,
                          float* colpart, unsigned int* counter, cudaStream_t st);
// enabled: the backward kernel finishes its rows itself (dview0 / dview1 written, no finalize kernel)
struct NtxBwdOut {
    const void* view0;
    const void* view1;
    void* dview0;
    void* dview1;
    const float* inv_norm;
    unsigned int* tag;              // fold counters at the head of the scratch (NtxFoldScratch)
    unsigned int* done_total;
    unsigned int* arrive;
    int row_blocks;
    int in_ld, normalised, in_bf16, enabled;
};
int ntx_tc_bwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, const NtxWeights& w, float* partial, const NtxGate& gate,
                      size_t key_buf_elems, const NtxBwdOut& out, cudaStream_t st);
int ntx_fwd_finalize_blocks(int q_rows, int ld);   // grid of the forward finalize
// colpart != NULL: symmetric forward (upper-triangle partition, column partials added per row)
// peers.world > 0: the row sums also go to every rank's block (rowsum_off) and the last block raises rflag
int ntx_fwd_finalize_launch(const void* q, int dtype, const NtxDims& dm, const NtxPlan& plan,
                            const float* partial, const float* colpart, unsigned int* counter,
                            float* rowsum, float* pos_logit, float* row_loss,
                            float* loss_sum, const NtxPeerOut& peers, size_t rowsum_off, const NtxGate& gate,
                            size_t key_buf_elems, cudaStream_t st);
int ntx_debug_indices_launch(int n_local, int rank, int32_t* self_col, int32_t* pos_col,
                             cudaStream_t st);
int ntx_bwd_finalize_launch(const NtxDims& dm, const NtxPlan& plan, const float* partial,
                            const NtxWeights& w, const void* view0, const void* view1, int in_dtype,
                            const float* inv_norm, void* dview0, void* dview1, cudaStream_t st, int in_ld = 0,
                            int normalised = 0);
