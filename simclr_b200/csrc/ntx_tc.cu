// The fused NT-Xent kernels on the 5th-gen tensor cores (tcgen05): bf16 rows, and fp32 inputs held as two bf16
// terms per value (NTX_BF16X2 rows, the kSplit instantiations: three partial products per similarity, the
// backward's coefficient tile split the same way -- fp32-grade results at 3x the MMA count).
//
// One persistent CTA per SM, warp specialised:
//   NWG epilogue warpgroups (4 warps each); tile i of the CTA goes to group i % NWG.
//        Thread = one row of the 128 x 128 similarity tile (= one TMEM lane).
//   1 TMA producer warp (cp.async.bulk.tensor, 128B swizzle, mbarrier transaction bytes)
//   1 warp that allocates TMEM and whose lane 0 issues every tcgen05.mma
//
// Forward  (loss.py:42-61): S = Q . K^T lands in TMEM (4 buffers of 128 fp32 columns,
//   one per epilogue group); the group reads it with tcgen05.ld, forms
//   e = exp2(s*c1 - c1) (fixed shift 1/tau, |cos| <= 1), zeroes the self column and
//   columns past the last key in edge tiles, and keeps per-row sums in registers.
//   The 2N x 2N logits never leave the SM.
// Backward (autograd of the same lines): recompute S, turn it IN PLACE in TMEM into the
//   bf16 coefficient tile C = e * (w_r/R_r + w_c/R_c) (tcgen05.st over the fp32 columns
//   already consumed), second MMA  dZ += C . K  with A read from TMEM and the SAME key
//   tile read MN-major from shared memory; the dZ accumulator stays in TMEM across the
//   CTA's key tiles.
//
// Work partition and scratch slots: see NtxPlan / ntx_common.cuh.
#include "ntx_common.cuh"
#include "ntx_umma.cuh"

#ifdef NTX_TIMING
// bring-up only: per-CTA cycle counters (fwd: slots 0-7, bwd: slots 8-15), read by ntx_debug_timers()
__device__ long long g_ntx_timers[148][24];   // 16..23: NTX_TIMELINE points (cycles from kernel entry)
#define NTX_T(var) long long var = 0
#define NTX_TIC(var) const long long var##_t0 = clock64()
#define NTX_TOC(var) var += clock64() - var##_t0
#define NTX_TSAVE(slot, var) g_ntx_timers[blockIdx.x % 148][slot] = (var)
#else
#define NTX_T(var)
#define NTX_TIC(var)
#define NTX_TOC(var)
#define NTX_TSAVE(slot, var)
#endif

namespace {

using namespace umma;

constexpr uint32_t kBlkBytes = 128 * 128;   // one [128 rows x 64 bf16] swizzled block
constexpr uint32_t kTmemCols = 512;

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
// exp2 on the FMA / ALU pipes for a share of the elements: MUFU.EX2 (16 per clock per SM) is the
// unit that bounds these kernels, the FP32 pipes have slack.  x = n + f, n = round(x) taken from
// the low mantissa bits of x + 1.5 * 2^23, 2^f by a degree-3 minimax polynomial on [-0.5, 0.5]
// (max relative error 7.5e-5, mean 5e-6: far inside the bf16 mode's 1e-2), 2^n by adding n to
// the exponent field.  Valid for x in (-126, 0], which the temperature floor guarantees.
__device__ __forceinline__ float ex2_poly(float x) {
    const float t = x + 12582912.f;
    const float f = x - (t - 12582912.f);
    float p = fmaf(f, 0.0551716685f, 0.242611125f);
    p = fmaf(p, f, 0.693260968f);
    p = fmaf(p, f, 0.999928057f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
// Packed fp32x2 arithmetic (Blackwell FFMA2 / FADD2 / FMUL2): two logits per instruction.  The
// epilogues are bound by instruction issue and MUFU, not by the FP32 datapath, so halving the
// number of FP32 instructions is what buys time.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
// exp2 of a pair: on MUFU, or by the packed form of ex2_poly (3 FADD2-class + 3 FFMA2 + 2 integer ops)
__device__ __forceinline__ f32x2 ex2_pair_mufu(f32x2 x) {
    float lo, hi;
    unpack2(x, lo, hi);
    return pack2(ex2(lo), ex2(hi));
}
__device__ __forceinline__ f32x2 ex2_pair_poly(f32x2 x) {
    const f32x2 kMagic = pack2(12582912.f, 12582912.f), kNegMagic = pack2(-12582912.f, -12582912.f);
    const f32x2 kNegOne = pack2(-1.f, -1.f);
    const f32x2 k3 = pack2(0.0551716685f, 0.0551716685f), k2 = pack2(0.242611125f, 0.242611125f);
    const f32x2 k1 = pack2(0.693260968f, 0.693260968f), k0 = pack2(0.999928057f, 0.999928057f);
    const f32x2 t = add2(x, kMagic);
    const f32x2 f = fma2(add2(t, kNegMagic), kNegOne, x);       // x - round(x)
    f32x2 p = fma2(f, k3, k2);
    p = fma2(p, f, k1);
    p = fma2(p, f, k0);
    float p0, p1, t0, t1;
    unpack2(p, p0, p1);
    unpack2(t, t0, t1);
    return pack2(__int_as_float(__float_as_int(p0) + (__float_as_int(t0) << 23)),
                 __int_as_float(__float_as_int(p1) + (__float_as_int(t1) << 23)));
}
#ifndef NTX_FWD_POLY
#define NTX_FWD_POLY 1      // of every 4 PAIRS of exponentials in the forward epilogue, how many use the polynomial
                            // (B200, N=8192 D=128: 1 -> 75.4 us, 2 -> 79.2 us, 3 -> 90.5 us; scalar code without any: 86.6 us)
#endif
#ifndef NTX_SYM_POLY
#define NTX_SYM_POLY 0      // same for the symmetric forward, whose FMA pipe also packs the bf16 tile (0 -> 67.9 us, 1 -> 72.4 us, 2 -> 82.6 us)
#endif
#ifndef NTX_BWD_POLY
#define NTX_BWD_POLY 1      // same (pairs of every 4) for the backward epilogue (0 -> 115.2 us, 1 -> 112.7 us)
#endif
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

// does key tile kt hold the positive column of some row of row block rb?  The positives of rows [r0, r1] are the
// columns ntx_pos_col(r): consecutive, except where the block straddles the two views (r0 < n_local <= r1).
__device__ __forceinline__ bool positive_tile(int rb, int kt, const NtxDims& dm) {
    const int r0 = rb * 128, r1 = min(rb * 128 + 127, dm.q_rows - 1);
    auto hit = [&](int a, int b) {      // rows a..b on one side of the view boundary
        return ntx_pos_col(a, dm.n_local, dm.rank) / 128 <= kt && kt <= ntx_pos_col(b, dm.n_local, dm.rank) / 128;
    };
    if (r0 < dm.n_local && r1 >= dm.n_local) return hit(r0, dm.n_local - 1) || hit(dm.n_local, r1);
    return hit(r0, r1);
}

// S = Q . K^T for one key tile: NDB blocks of 64 features, 4 MMAs (K = 16) per block.
// kSplit (NTX_BF16X2 rows): the tile holds NDB / 2 blocks of hi terms followed by NDB / 2 blocks of lo terms and
// S = hi.lo + lo.hi + hi.hi, accumulated in fp32 in TMEM.  The lo.lo products (2^-18 of a term, below the 2^-17 to
// which two bf16 terms hold a value anyway) are left out: with them the error against the fp64 oracle was the same.
template <int NDB, bool kSplit = false>
__device__ __forceinline__ void issue_gemm1(uint32_t d_tmem, const uint8_t* q_smem, const uint8_t* k_tile) {
    constexpr uint32_t idesc = instr_desc_bf16(128, 128, 0);
    if constexpr (kSplit) {
        constexpr int NH = NDB / 2;
#pragma unroll
        for (int t = 0; t < 3; ++t)                // (q term, k term): (hi, lo), (lo, hi), (hi, hi)
#pragma unroll
            for (int b = 0; b < NH; ++b)
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4) {
                    const int qb = t == 1 ? NH + b : b, kb = t == 0 ? NH + b : b;
                    const uint64_t a = smem_desc_sw128(smem_u32(q_smem + qb * kBlkBytes + k4 * 32), 16, 1024);
                    const uint64_t bd = smem_desc_sw128(smem_u32(k_tile + kb * kBlkBytes + k4 * 32), 16, 1024);
                    mma_bf16_ss(d_tmem, a, bd, idesc, (t | b | k4) ? 1u : 0u);
                }
    } else {
#pragma unroll
        for (int b = 0; b < NDB; ++b)
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4) {
                const uint64_t a = smem_desc_sw128(smem_u32(q_smem + b * kBlkBytes + k4 * 32), 16, 1024);
                const uint64_t bd = smem_desc_sw128(smem_u32(k_tile + b * kBlkBytes + k4 * 32), 16, 1024);
                mma_bf16_ss(d_tmem, a, bd, idesc, (b | k4) ? 1u : 0u);
            }
    }
}

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Cross-rank gate of the TMA producer: before the first key tile that touches rank c's chunk of the key buffer
// is loaded, wait until that rank's rows of this step have landed (kflag[c] >= step, written by its normalise
// kernel after a system-scope fence), then order the async proxy (TMA) behind the acquire.
struct KeyGate {
    NtxGate g;
    uint32_t step, ready;
    int k_rows;
    __device__ __forceinline__ void init(const NtxGate& gate, int k_rows_) {
        g = gate;
        k_rows = k_rows_;
        ready = 0;
        step = gate.step ? *gate.step : 0u;                  // written by the normalise kernel, earlier on the stream
    }
    __device__ __forceinline__ void wait_tile(int kt) {
        if (!g.step) return;
        const int c_lo = kt * 128 / g.chunk_rows, c_hi = min(kt * 128 + 127, k_rows - 1) / g.chunk_rows;
        bool fresh = false;
        for (int c = c_lo; c <= c_hi; ++c)
            if (!((ready >> c) & 1u)) {
                ntx_gate_wait(g.ctl, NTX_CTL_KFLAG + c, step);
                ready |= 1u << c;
                fresh = true;
            }
        if (fresh) fence_proxy_async_all();
    }
};

// TMA producer shared by both kernels: the Q tile of each row block, then its key tiles
template <int NDB, int NS, class Ctl>
__device__ __forceinline__ void producer_loop(Ctl* ctl, uint8_t* q_smem, uint8_t* k_smem,
                                              const CUtensorMap* tmap_q, const CUtensorMap* tmap_k,
                                              long long t_begin, int n_tiles, int key_tiles, KeyGate& gate) {
    constexpr uint32_t kTileBytes = NDB * kBlkBytes;
    int cur_rb = -1, seg = 0;
    for (int i = 0; i < n_tiles; ++i) {
        const long long t = t_begin + i;
        const int rb = (int)(t / key_tiles), kt = (int)(t % key_tiles);
        if (rb != cur_rb) {
            if (cur_rb >= 0) mbar_wait(&ctl->q_empty, (seg - 1) & 1);   // GEMM1s of the old block done
            mbar_expect_tx(&ctl->q_full, kTileBytes);
#pragma unroll
            for (int b = 0; b < NDB; ++b)
                tma_load_2d(q_smem + b * kBlkBytes, tmap_q, b * 64, rb * 128, &ctl->q_full);
            cur_rb = rb;
            ++seg;
        }
        const int s = i % NS;
        gate.wait_tile(kt);
        mbar_wait(&ctl->k_empty[s], ((i / NS) & 1) ^ 1);
        mbar_expect_tx(&ctl->k_full[s], kTileBytes);
#pragma unroll
        for (int b = 0; b < NDB; ++b)
            tma_load_2d(k_smem + s * kTileBytes + b * kBlkBytes, tmap_k, b * 64, kt * 128, &ctl->k_full[s]);
    }
}

// ============================================================================ forward
constexpr int kFwdWG = 4;                                  // epilogue groups = S buffers
constexpr int kFwdThreads = kFwdWG * 128 + 96;             // + producer warp + two MMA-issuing warps
// NS divides kFwdWG so the MMA loop can be unrolled by kFwdWG with compile-time stage / buffer indices
template <int NDB> struct FwdCfg { static constexpr int NS = NDB == 4 ? 2 : 4; };
template <int NS> struct FwdCtl {
    uint64_t q_full, q_empty, k_full[NS], k_empty[NS], s_full[kFwdWG], s_empty[kFwdWG];
    uint32_t tmem_base;
    uint32_t fin_last;                 // finish phase: this CTA took the launch's last segment
    alignas(16) float xch[2][kFwdWG - 1][128];
    alignas(16) double fin_sum[4];     // finish phase: per-warp loss sums of group 0
};

// kScaled: the rows carry sqrt(log2e / tau) (NtxDims.scaled), so a similarity value IS the exponent: no FFMA2, and
// the row sums are sums of e' = e * 2^c1 (NtxDims.rowsum_scale takes the factor out where they are finished)
// kSplit: NTX_BF16X2 rows (issue_gemm1 above); every exponential on MUFU, the polynomial's 7.5e-5 is too coarse here
template <int NDB, bool kScaled, bool kSplit = false>
__global__ void __launch_bounds__(kFwdThreads, 1)
tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              const __grid_constant__ CUtensorMap tmap_q1, const __grid_constant__ CUtensorMap tmap_k1,
              NtxDims dm, int key_tiles, long long T, int rows_pad, float* __restrict__ partial,
              NtxFoldScratch fold, NtxFwdOut out, NtxGate gate) {
#ifdef NTX_TIMELINE
    const long long t_entry = clock64();
#endif
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
    constexpr int kProducerWarp = kFwdWG * 4, kMmaWarp = kFwdWG * 4 + 1;   // MMA warps: kMmaWarp, kMmaWarp + 1

    if (threadIdx.x == 0) {
        mbar_init(&ctl->q_full, 1);
        mbar_init(&ctl->q_empty, 2);                       // both MMA warps release the Q tile
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl->k_full[s], 1); mbar_init(&ctl->k_empty[s], 1); }
        for (int b = 0; b < kFwdWG; ++b) { mbar_init(&ctl->s_full[b], 1); mbar_init(&ctl->s_empty[b], 4); }
        fence_barrier_init();
    }
    if (warp == kMmaWarp) tmem_alloc(&ctl->tmem_base, kTmemCols);
    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k);
        tma_prefetch_desc(&tmap_q1); tma_prefetch_desc(&tmap_k1);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // our TMEM is allocated: only now may the next kernel on the stream be staged (its CTAs allocate TMEM too)
    ntx_dep_trigger();
    // the whole TMEM (512 columns) is ours, so the allocation starts at column 0, lane 0;
    // using the literal keeps every TMEM address a compile-time constant
    if (ctl->tmem_base != 0) {
        if (threadIdx.x == 0) printf("ntxent: unexpected TMEM base %u\n", ctl->tmem_base);
        __trap();
    }
    constexpr uint32_t tmem_base = 0;
    // everything above overlapped the tail of the previous kernel on the stream; from here on its results are read
    ntx_dep_wait();
#ifdef NTX_TIMELINE
    if (threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][20] = clock64() - t_entry;
#endif

    if (warp == kProducerWarp) {
        if (elect_one()) {
            KeyGate kg;
            kg.init(gate, dm.k_rows);
            const bool alt = kg.step & 1u;                   // key buffer of this step (peer exchange)
            producer_loop<NDB, NS>(ctl, q_smem, k_smem, alt ? &tmap_q1 : &tmap_q, alt ? &tmap_k1 : &tmap_k, t_begin,
                                   n_tiles, key_tiles, kg);
        }
    } else if (warp >= kMmaWarp) {
        // Two issuing warps take alternate tiles: tcgen05.mma issue blocks while the tensor pipe's
        // short queue is full, so one warp's barrier waits (~150-200 cycles each, even when the
        // phase is already complete) run while the other warp's MMAs execute.
        // Unrolled by kFwdWG: stage and buffer indices are compile-time constants, so every
        // descriptor is a uniform value and tcgen05.mma issues back to back.
        const int m = warp - kMmaWarp;
        int rb = (int)(t_begin / key_tiles), kt = (int)(t_begin % key_tiles), seg = 0;
        bool new_seg = true;
        NTX_T(tm_k); NTX_T(tm_s); NTX_T(tm_issue); NTX_T(tm_total);
        NTX_TIC(tm_total);
        for (int i0 = 0; i0 < n_tiles; i0 += kFwdWG) {
            const uint32_t ph_k = (uint32_t)(i0 / NS), ph_s = (uint32_t)(i0 / kFwdWG);
#pragma unroll
            for (int j = 0; j < kFwdWG; ++j) {
                const int i = i0 + j;
                if (i >= n_tiles) break;
                const int s = j % NS, buf = j;
                const bool last_of_seg = (i + 1 == n_tiles) || (kt + 1 == key_tiles);
                if ((j & 1) == m) {
                    if (new_seg) mbar_wait(&ctl->q_full, seg & 1);
                    new_seg = false;
                    { NTX_TIC(tm_k); mbar_wait(&ctl->k_full[s], (ph_k + j / NS) & 1); NTX_TOC(tm_k); }
                    { NTX_TIC(tm_s); mbar_wait(&ctl->s_empty[buf], (ph_s & 1) ^ 1); NTX_TOC(tm_s); }
                    tc_fence_after();
                    NTX_TIC(tm_issue);
                    if (elect_one()) {
                        issue_gemm1<NDB, kSplit>(tmem_base + buf * 128, q_smem, k_smem + s * kTileBytes);
                        mma_commit(&ctl->k_empty[s]);
                        mma_commit(&ctl->s_full[buf]);
                    }
                    __syncwarp();
                    NTX_TOC(tm_issue);
                }
                if (last_of_seg) {
                    // both warps release Q: every GEMM1 this warp issued for the row block is done
                    if (elect_one()) mma_commit(&ctl->q_empty);
                    __syncwarp();
                    ++seg;
                    new_seg = true;
                }
                if (++kt == key_tiles) { kt = 0; ++rb; }
            }
        }
        NTX_TOC(tm_total);
        if (lane == 0 && m == 0) { NTX_TSAVE(0, tm_total); NTX_TSAVE(1, tm_k); NTX_TSAVE(2, tm_s); NTX_TSAVE(3, tm_issue); }
        (void)rb;
    } else {
        // ---------------------------------------------------------------- epilogue groups
        const int g = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + g * 128;
        const float c1 = dm.c1, nc1 = -dm.c1;
        const f32x2 c1x2 = pack2(c1, c1), nc1x2 = pack2(nc1, nc1);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};          // edge tiles (scalar, masked)
        f32x2 pacc[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};   // interior tiles (packed pairs)
        int flushes = 0;
        NTX_T(te_wait); NTX_T(te_total); NTX_T(te_ld);
        NTX_TIC(te_total);

        // row sums of the groups are added in group order: deterministic
        auto flush = [&](int rb) {
            float p0, p1, p2, p3;
            unpack2(pacc[0], p0, p1);
            unpack2(pacc[1], p2, p3);
            float rs = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((p0 + p1) + (p2 + p3));
            pacc[0] = pacc[1] = pack2(0.f, 0.f);
            float (*x)[128] = ctl->xch[flushes & 1];
            if (g > 0) x[g - 1][row_in_tile] = rs;
            named_bar_sync(1, kFwdWG * 128);
            if (g == 0) {
#pragma unroll
                for (int o = 0; o < kFwdWG - 1; ++o) rs += x[o][row_in_tile];
                const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
                partial[(size_t)slot * rows_pad + rb * 128 + row_in_tile] = rs;
                // this CTA's part of the row block is out (its positive logits too: the barrier above orders every
                // group's stores before this thread): announce it
                if (fold.enabled) {
                    named_bar_sync(3, 128);
                    if (threadIdx.x == 0) {
                        __threadfence();
                        atomicAdd(fold.arrive + rb, 1u);
                    }
                }
            }
            ++flushes;
            acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
        };

        if (threadIdx.x == 0) {
            if (fold.enabled) ntx_fold_init(fold.tag, fold.done_total, fold.arrive, fold.row_blocks, cta);
            else if (cta == 0) *reinterpret_cast<unsigned long long*>(fold.legacy_ticket) = 0ull;   // the finalize kernel's arrival word
        }
        int rb = (int)(t_begin / key_tiles), kt = (int)(t_begin % key_tiles);
        int own = 0, turn = 0;
        for (int i = 0; i < n_tiles; ++i) {
            const bool last_of_seg = (i + 1 == n_tiles) || (kt + 1 == key_tiles);
            if (turn == g) {
                { NTX_TIC(te_wait); mbar_wait(&ctl->s_full[g], own & 1); NTX_TOC(te_wait); }
#ifdef NTX_TIMELINE
                if (own == 0 && threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][21] = clock64() - t_entry;
#endif
                ++own;
                tc_fence_after();
                // folded finalize only: tiles that hold a positive column go through the masked path too, the positive
                // logit of a row is taken from the similarity tile itself (the same MMA value its exponential uses).
                // Otherwise the finalize kernel computes it, and the slow path is kept to the self / ragged tiles
                // (one of 8 ranks: a second masked tile among a CTA's 14 made the slowest CTA 47 % longer than the mean)
                const bool has_pos = fold.enabled && positive_tile(rb, kt, dm);
                const bool edge = edge_tile(rb, kt, dm) || has_pos;
                const int self = ntx_self_col(rb * 128 + row_in_tile, dm.n_local, dm.rank);
                const int my_row = rb * 128 + row_in_tile;
                const int my_pos = my_row < dm.q_rows ? ntx_pos_col(my_row, dm.n_local, dm.rank) : -1;
                float pos_val = 0.f;
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
                        if (lane == 0) mbar_arrive(&ctl->s_empty[g]);
                    }
                    // the masked form only for the 32-key chunks that hold a self (or, folded, a positive) column
                    // of this warp's rows or run past the last key: warp-uniform
                    bool masked = edge;
                    if (edge) {
                        const int c0 = kt * 128 + ch * 32;
                        const bool mine = (self >= c0 && self < c0 + 32) || (has_pos && my_pos >= c0 && my_pos < c0 + 32);
                        masked = c0 + 32 > dm.k_rows || __any_sync(0xffffffffu, mine);
                    }
                    if (!masked) {
                        // two logits per instruction: x = s * c1 - c1 (FFMA2), exp2 on MUFU or by the
                        // polynomial for NTX_FWD_POLY of every 4 pairs, row sums by FADD2
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const f32x2 raw = pack2(__uint_as_float(v[ch & 1][2 * j]), __uint_as_float(v[ch & 1][2 * j + 1]));
                            const f32x2 x = kScaled ? raw : fma2(raw, c1x2, nc1x2);
                            const bool poly = !kSplit && NTX_FWD_POLY > 0 && ((j * NTX_FWD_POLY) & 3) < NTX_FWD_POLY;
                            pacc[j & 1] = add2(pacc[j & 1], poly ? ex2_pair_poly(x) : ex2_pair_mufu(x));
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const int col = kt * 128 + ch * 32 + j;
                            float e = ex2(kScaled ? __uint_as_float(v[ch & 1][j]) : fmaf(__uint_as_float(v[ch & 1][j]), c1, nc1));
                            if (col == self || col >= dm.k_rows) e = 0.f;   // mask of loss.py:38,46-47
                            if (col == my_pos) pos_val = __uint_as_float(v[ch & 1][j]);
                            acc[j & 3] += e;
                        }
                    }
                }
                if (has_pos && my_pos >= 0 && my_pos / 128 == kt) out.pos_logit[my_row] = pos_val * dm.pos_scale;
            }
            if (last_of_seg) flush(rb);
            if (++kt == key_tiles) { kt = 0; ++rb; }
            if (++turn == kFwdWG) turn = 0;
        }
        NTX_TOC(te_total);
        if (threadIdx.x == 0) { NTX_TSAVE(4, te_total); NTX_TSAVE(5, te_wait); }
#ifdef NTX_TIMELINE
        if (threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][22] = clock64() - t_entry;
#endif

        // ---------------------------------------------------------------- finish: this CTA's share of its row blocks
        // (loss.py:54-65: logsumexp - positive per row, then the reduction).  Group 0, one thread per row.
        if (fold.enabled && g == 0 && n_tiles > 0) {
            const int rb_first = (int)(t_begin / key_tiles), rb_last = (int)((t_begin + n_tiles - 1) / key_tiles);
            const int tid = threadIdx.x;                       // 0..127
            for (int frb = rb_first; frb <= rb_last; ++frb) {
                const int first = ntx_first_owner(T, G, key_tiles, frb);
                const int ns = ntx_num_slots(T, G, key_tiles, frb), slot = cta - first;
                if (tid == 0) {
                    while (ntx_ld_acquire_gpu(fold.arrive + frb) < (unsigned)ns) __nanosleep(32);
                }
                named_bar_sync(3, 128);
                const int lo = slot * 128 / ns, hi = (slot + 1) * 128 / ns;
                const int row = frb * 128 + lo + tid;
                double my_loss = 0.0;
                if (lo + tid < hi && row < dm.q_rows) {
                    float R = 0.f;
                    for (int s0 = 0; s0 < ns; s0 += 8) {       // slot order, eight loads in flight
                        float t[8];
#pragma unroll
                        for (int u = 0; u < 8; ++u)
                            t[u] = s0 + u < ns ? __ldcg(partial + (size_t)(s0 + u) * rows_pad + row) : 0.f;
#pragma unroll
                        for (int u = 0; u < 8; ++u) R += t[u];
                    }
                    R *= dm.rowsum_scale;
                    const float pos = __ldcg(out.pos_logit + row);
                    const float l = logf(R) + dm.inv_tau - pos;
                    out.rowsum[row] = R;
                    out.row_loss[row] = l;
                    if (out.peers.world > 0)
                        ntx_peer_st_f32(out.peers, out.rowsum_off + ((size_t)out.peers.rank * dm.q_rows + row) * sizeof(float), R);
                    my_loss = (double)l;
                }
                // the share's loss: lanes, then the four warps, in a fixed order
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) my_loss += __shfl_xor_sync(0xffffffffu, my_loss, o);
                if (lane == 0) ctl->fin_sum[warp] = my_loss;
                named_bar_sync(3, 128);
                if (tid == 0) {
                    fold.seg_loss[first + slot + frb] = (ctl->fin_sum[0] + ctl->fin_sum[1]) + (ctl->fin_sum[2] + ctl->fin_sum[3]);
                    __threadfence();
                    // second pass over the row block's counter: the last CTA out leaves it clean for the next launch
                    if (atomicAdd(fold.arrive + frb, 1u) == 2u * (unsigned)ns - 1u) fold.arrive[frb] = 0u;
                    ctl->fin_last = atomicAdd(fold.done_total, 1u) == (unsigned)fold.n_segments - 1u ? 1u : 0u;
                }
                named_bar_sync(3, 128);
                if (ctl->fin_last) {
                    // every share of every row block is done: add the share losses in (row block, slot) order
                    __threadfence();
                    const int row_blocks = (dm.q_rows + 127) / 128;
                    double part = 0.0;
                    for (int b = tid; b < row_blocks; b += 128) {
                        const int f = ntx_first_owner(T, G, key_tiles, b), n2 = ntx_num_slots(T, G, key_tiles, b);
                        for (int s2 = 0; s2 < n2; ++s2) part += __ldcg(fold.seg_loss + f + s2 + b);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
                    if (lane == 0) ctl->fin_sum[warp] = part;
                    named_bar_sync(3, 128);
                    if (tid == 0) {
                        const double v = (ctl->fin_sum[0] + ctl->fin_sum[1]) + (ctl->fin_sum[2] + ctl->fin_sum[3]);
                        ntx_store_loss(out.loss_sum, v, dm.q_rows, dm.loss_bf16);
                        *fold.done_total = 0u;
                        if (out.peers.world > 0)                                 // every rank's row sums are out
                            ntx_raise_flags(out.peers, NTX_CTL_RFLAG, gate.step ? *gate.step : 0u);
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// ============================================================================ symmetric forward
// One rank, keys == rows: S = S^T, so only the tiles (rb, kt >= rb) are computed.  An off-diagonal
// tile feeds the row sums of its 128 rows (registers, as in tc_fwd_kernel) AND the row sums of its
// 128 columns: the epilogue group writes e as a bf16 tile to shared memory and a second, tiny MMA
// (M = 128 columns x N = 16, A = E^T read MN-major from that tile, B = ones) reduces it over the
// rows on the tensor core; lane c of the result is the column sum of key c, which the group stores
// to colpart[rb][c].  Half the exponentials and half the similarity MMAs of the full-tile forward.
constexpr int kSymWG = 3;                                   // epilogue groups = S buffers = column-sum slots
constexpr int kSymThreads = kSymWG * 128 + 96;              // + producer, GEMM1 warp, column-sum warp
constexpr int kSymNS = 2;                                   // key-tile ring; MMA loops unroll by 6 = lcm(ring, groups)
constexpr uint32_t kSymCsCol = 384;                         // column-sum results: TMEM columns 384 + 32 * g
struct SymCtl {
    uint64_t q_full, q_empty, k_full[kSymNS], k_empty[kSymNS], s_full[kSymWG], s_free[kSymWG],
        e_full[kSymWG], e_free[kSymWG], cs_full[kSymWG], cs_free[kSymWG];
    uint32_t tmem_base;
    alignas(16) float xch[2][kSymWG - 1][128];
};

// walks the upper-triangle tiles in the paired order of ntx_tri_first
struct TriIter {
    int rb, kt, key_tiles;
    __device__ __forceinline__ void init(long long t, int kts) {
        key_tiles = kts;
        const int v = (int)(t / (kts + 1)), off = (int)(t - (long long)v * (kts + 1));   // pair, offset in the pair
        if (off < kts - v) { rb = v; kt = v + off; }
        else { rb = kts - 1 - v; kt = rb + off - (kts - v); }
    }
    __device__ __forceinline__ bool last_in_row() const { return kt + 1 == key_tiles; }
    __device__ __forceinline__ void next() {
        if (++kt == key_tiles) {            // next row block in the order 0, n-1, 1, n-2, ...
            const int partner = key_tiles - 1 - rb;
            rb = rb < partner ? partner : partner + 1;
            kt = rb;
        }
    }
};

template <int NDB, bool kScaled>
__global__ void __launch_bounds__(kSymThreads, 1)
tc_fwd_sym_kernel(const __grid_constant__ CUtensorMap tmap, NtxDims dm, int key_tiles, long long T,
                  int rows_pad, float* __restrict__ partial, float* __restrict__ colpart,
                  unsigned int* __restrict__ done_counter) {
    constexpr int NS = kSymNS;
    constexpr uint32_t kTileBytes = NDB * kBlkBytes;
    constexpr uint32_t kEBytes = 2 * kBlkBytes;             // [128 rows x 128 keys] bf16
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = checked_1024(smem_raw);
    uint8_t* q_smem = smem;
    uint8_t* k_smem = smem + kTileBytes;
    uint8_t* e_smem = k_smem + NS * kTileBytes;
    uint8_t* ones_smem = e_smem + kSymWG * kEBytes;         // [16 x 64] bf16 ones: one 2 KB block serves every K step
    SymCtl* ctl = reinterpret_cast<SymCtl*>(ones_smem + 2048);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    const long long t_begin = ntx_chunk_begin(T, G, cta);
    const int n_tiles = (int)(ntx_chunk_begin(T, G, cta + 1) - t_begin);
    constexpr int kProducerWarp = kSymWG * 4, kMmaWarp = kSymWG * 4 + 1, kCsWarp = kSymWG * 4 + 2;

    if (threadIdx.x == 0) {
        mbar_init(&ctl->q_full, 1);
        mbar_init(&ctl->q_empty, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl->k_full[s], 1); mbar_init(&ctl->k_empty[s], 1); }
        for (int b = 0; b < kSymWG; ++b) {
            mbar_init(&ctl->s_full[b], 1);
            mbar_init(&ctl->s_free[b], 4);
            mbar_init(&ctl->cs_full[b], 1);
            mbar_init(&ctl->cs_free[b], 4);
        }
        for (int b = 0; b < kSymWG; ++b) { mbar_init(&ctl->e_full[b], 4); mbar_init(&ctl->e_free[b], 1); }
        fence_barrier_init();
    }
    for (int i = threadIdx.x; i < 2048 / 4; i += kSymThreads)
        reinterpret_cast<uint32_t*>(ones_smem)[i] = 0x3F803F80u;       // bf16 1.0 pairs
    fence_proxy_async_smem();
    if (warp == kMmaWarp) tmem_alloc(&ctl->tmem_base, kTmemCols);
    if (warp == kProducerWarp && lane == 0) tma_prefetch_desc(&tmap);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // our TMEM is allocated: only now may the next kernel on the stream be staged (its CTAs allocate TMEM too)
    ntx_dep_trigger();
    if (ctl->tmem_base != 0) {
        if (threadIdx.x == 0) printf("ntxent: unexpected TMEM base %u\n", ctl->tmem_base);
        __trap();
    }
    constexpr uint32_t tmem_base = 0;
    // everything above overlapped the tail of the previous kernel on the stream; from here on its results are read
    ntx_dep_wait();
    if (threadIdx.x == 0 && cta == 0)                        // the finalize kernel's 64-bit arrival word
        *reinterpret_cast<unsigned long long*>(done_counter) = 0ull;

    if (warp == kProducerWarp) {
        if (elect_one()) {
            TriIter it;
            it.init(t_begin, key_tiles);
            int seg = 0;
            bool new_seg = true;
            for (int i = 0; i < n_tiles; ++i) {
                if (new_seg) {
                    if (seg > 0) mbar_wait(&ctl->q_empty, (seg - 1) & 1);
                    mbar_expect_tx(&ctl->q_full, kTileBytes);
#pragma unroll
                    for (int b = 0; b < NDB; ++b)
                        tma_load_2d(q_smem + b * kBlkBytes, &tmap, b * 64, it.rb * 128, &ctl->q_full);
                    new_seg = false;
                }
                const int s = i % NS;
                mbar_wait(&ctl->k_empty[s], ((i / NS) & 1) ^ 1);
                mbar_expect_tx(&ctl->k_full[s], kTileBytes);
#pragma unroll
                for (int b = 0; b < NDB; ++b)
                    tma_load_2d(k_smem + s * kTileBytes + b * kBlkBytes, &tmap, b * 64, it.kt * 128, &ctl->k_full[s]);
                if (it.last_in_row()) { ++seg; new_seg = true; }
                it.next();
            }
        }
    } else if (warp == kMmaWarp) {
        // GEMM1 issuer; unrolled by 6 so stage (i % 2) and buffer (i % 3) indices are compile-time
        TriIter it;
        it.init(t_begin, key_tiles);
        int seg = 0;
        bool new_seg = true;
        NTX_T(tm_k); NTX_T(tm_s); NTX_T(tm_issue); NTX_T(tm_total);
        NTX_TIC(tm_total);
        for (int i0 = 0; i0 < n_tiles; i0 += 6) {
#pragma unroll
            for (int jj = 0; jj < 6; ++jj) {
                const int i = i0 + jj;
                if (i >= n_tiles) break;
                const int st = jj % NS, j = jj % kSymWG;
                const bool last_of_seg = (i + 1 == n_tiles) || it.last_in_row();
                if (new_seg) mbar_wait(&ctl->q_full, seg & 1);
                new_seg = false;
                { NTX_TIC(tm_k); mbar_wait(&ctl->k_full[st], (i / NS) & 1); NTX_TOC(tm_k); }
                { NTX_TIC(tm_s); mbar_wait(&ctl->s_free[j], ((i / kSymWG) & 1) ^ 1); NTX_TOC(tm_s); }
                tc_fence_after();
                NTX_TIC(tm_issue);
                if (elect_one()) {
                    issue_gemm1<NDB>(tmem_base + j * 128, q_smem, k_smem + st * kTileBytes);
                    mma_commit(&ctl->k_empty[st]);
                    mma_commit(&ctl->s_full[j]);
                    if (last_of_seg) mma_commit(&ctl->q_empty);
                }
                __syncwarp();
                NTX_TOC(tm_issue);
                if (it.last_in_row()) { ++seg; new_seg = true; }
                it.next();
            }
        }
        NTX_TOC(tm_total);
        if (lane == 0) { NTX_TSAVE(0, tm_total); NTX_TSAVE(1, tm_k); NTX_TSAVE(2, tm_s); NTX_TSAVE(3, tm_issue); }
    } else if (warp == kCsWarp) {
        // column sums of every off-diagonal tile: D[c][0..15] = sum_r E[r][c] * 1
        constexpr uint32_t idesc = instr_desc_bf16_ex(128, 16, 1, 0);
        TriIter it;
        it.init(t_begin, key_tiles);
        int uses[kSymWG] = {0, 0, 0};                       // off-diagonal tiles of each group so far
        for (int i0 = 0; i0 < n_tiles; i0 += kSymWG) {
#pragma unroll
            for (int j = 0; j < kSymWG; ++j) {
                const int i = i0 + j;
                if (i >= n_tiles) break;
                if (it.kt != it.rb) {
                    mbar_wait(&ctl->e_full[j], uses[j] & 1);
                    mbar_wait(&ctl->cs_free[j], (uses[j] & 1) ^ 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint8_t* e_tile = e_smem + j * kEBytes;
#pragma unroll
                        for (int kk = 0; kk < 8; ++kk) {
                            // A = E^T, MN-major: 64 keys per 128 B row, rows 128 B apart, next 64 keys 16 KB away
                            const uint64_t a = smem_desc_sw128(smem_u32(e_tile + kk * 2048), kBlkBytes, 1024);
                            const uint64_t b = smem_desc_sw128(smem_u32(ones_smem + (kk & 3) * 32), 16, 1024);
                            mma_bf16_ss(tmem_base + kSymCsCol + j * 32, a, b, idesc, kk ? 1u : 0u);
                        }
                        mma_commit(&ctl->e_free[j]);
                        mma_commit(&ctl->cs_full[j]);
                    }
                    __syncwarp();
                    ++uses[j];
                }
                it.next();
            }
        }
    } else {
        // ---------------------------------------------------------------- epilogue groups
        const int g = warp >> 2;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t taddr = t_lane + g * 128;
        const float c1 = dm.c1, nc1 = -dm.c1;
        const f32x2 c1x2 = pack2(c1, c1), nc1x2 = pack2(nc1, nc1);
        const uint32_t sw = (uint32_t)(row_in_tile & 7);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        f32x2 pacc[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};
        int flushes = 0, own = 0, e_uses = 0, cs_uses = 0;
        bool pending = false;                               // a column-sum result of this group not yet stored
        int pend_rb = 0, pend_kt = 0;
        NTX_T(te_total); NTX_T(te_wait); NTX_T(te_ef); NTX_T(te_loop); NTX_T(te_cs);
        NTX_TIC(te_total);

        auto flush = [&](int rb) {
            float p0, p1, p2, p3;
            unpack2(pacc[0], p0, p1);
            unpack2(pacc[1], p2, p3);
            float rs = ((acc[0] + acc[1]) + (acc[2] + acc[3])) + ((p0 + p1) + (p2 + p3));
            pacc[0] = pacc[1] = pack2(0.f, 0.f);
            float (*x)[128] = ctl->xch[flushes & 1];
            if (g > 0) x[g - 1][row_in_tile] = rs;
            named_bar_sync(1, kSymWG * 128);
            if (g == 0) {
#pragma unroll
                for (int o = 0; o < kSymWG - 1; ++o) rs += x[o][row_in_tile];
                const int slot = cta - ntx_tri_first_owner(T, G, key_tiles, rb);
                partial[(size_t)slot * rows_pad + rb * 128 + row_in_tile] = rs;
            }
            ++flushes;
            acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
        };
        // lane c of the column-sum MMA result = sum over this tile's rows of e[., c]
        auto store_colsum = [&]() {
            mbar_wait(&ctl->cs_full[g], cs_uses & 1);
            ++cs_uses;
            tc_fence_after();
            const uint32_t v = tmem_ld_32x1(t_lane + kSymCsCol + g * 32);
            tmem_ld_wait();
            colpart[(size_t)pend_rb * rows_pad + pend_kt * 128 + row_in_tile] = __uint_as_float(v);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->cs_free[g]);
            pending = false;
        };

        TriIter it;
        it.init(t_begin, key_tiles);
        int turn = 0;
        for (int i = 0; i < n_tiles; ++i) {
            const int rb = it.rb, kt = it.kt;
            const bool last_of_seg = (i + 1 == n_tiles) || it.last_in_row();
            const bool offdiag = kt != rb;
            if (turn == g) {
                { NTX_TIC(te_wait); mbar_wait(&ctl->s_full[g], own & 1); NTX_TOC(te_wait); }
                ++own;
                tc_fence_after();
                uint32_t v[2][32];
                tmem_ld_32x32(taddr, v[0]);
                // the previous tile's column sums are read out behind the first S load: one TMEM round trip for both
                { NTX_TIC(te_cs); if (pending) store_colsum(); NTX_TOC(te_cs); }
                const bool row_ok = rb * 128 + row_in_tile < dm.q_rows;
                const bool edge = !offdiag || (kt * 128 + 128 > dm.k_rows) || (rb * 128 + 128 > dm.q_rows);
                const int self = rb * 128 + row_in_tile;
                uint8_t* e_row = e_smem + g * kEBytes + row_in_tile * 128;         // this group's own E tile
                { NTX_TIC(te_ef); if (offdiag) mbar_wait(&ctl->e_free[g], (e_uses & 1) ^ 1); NTX_TOC(te_ef); }   // column-sum MMA that read it is done
                NTX_TIC(te_loop);
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    tmem_ld_wait_dep(v[ch & 1]);
                    if (ch < 3) {
                        tmem_ld_32x32(taddr + (ch + 1) * 32, v[(ch + 1) & 1]);
                    } else {   // the whole tile is in registers: hand the S buffer back
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&ctl->s_free[g]);
                    }
                    if (!edge) {
                        // interior off-diagonal tile: packed arithmetic, no masks, E tile always written
#pragma unroll
                        for (int j8 = 0; j8 < 4; ++j8) {       // 8 keys -> one 16-byte chunk of the E tile
                            uint32_t packed[4];
#pragma unroll
                            for (int h = 0; h < 4; ++h) {
                                const int j = j8 * 8 + h * 2;
                                const f32x2 raw = pack2(__uint_as_float(v[ch & 1][j]), __uint_as_float(v[ch & 1][j + 1]));
                                const f32x2 x = kScaled ? raw : fma2(raw, c1x2, nc1x2);
                                const bool poly = NTX_SYM_POLY > 0 && (((j8 * 4 + h) * NTX_SYM_POLY) & 3) < NTX_SYM_POLY;
                                const f32x2 e = poly ? ex2_pair_poly(x) : ex2_pair_mufu(x);
                                pacc[h & 1] = add2(pacc[h & 1], e);
                                float e0, e1;
                                unpack2(e, e0, e1);
                                __nv_bfloat162 hh = __floats2bfloat162_rn(e0, e1);
                                packed[h] = *reinterpret_cast<uint32_t*>(&hh);
                            }
                            const int cg = ch * 4 + j8;         // 16-byte chunk index along the 128 keys
                            uint8_t* dst = e_row + (cg >> 3) * kBlkBytes + ((((uint32_t)cg & 7u) ^ sw) << 4);
                            *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                        }
                    } else {
                        // diagonal tile (self column, no E tile) or a tile with rows / keys past the end
#pragma unroll
                        for (int j8 = 0; j8 < 4; ++j8) {
                            uint32_t packed[4];
#pragma unroll
                            for (int h = 0; h < 4; ++h) {
                                const int j = j8 * 8 + h * 2;
                                const int col = kt * 128 + ch * 32 + j;
                                float e0 = ex2(kScaled ? __uint_as_float(v[ch & 1][j]) : fmaf(__uint_as_float(v[ch & 1][j]), c1, nc1));
                                float e1 = ex2(kScaled ? __uint_as_float(v[ch & 1][j + 1])
                                                       : fmaf(__uint_as_float(v[ch & 1][j + 1]), c1, nc1));
                                if (col == self || col >= dm.k_rows || !row_ok) e0 = 0.f;   // mask of loss.py:38,46-47
                                if (col + 1 == self || col + 1 >= dm.k_rows || !row_ok) e1 = 0.f;
                                acc[h] += e0 + e1;
                                __nv_bfloat162 hh = __floats2bfloat162_rn(e0, e1);
                                packed[h] = *reinterpret_cast<uint32_t*>(&hh);
                            }
                            if (offdiag) {
                                const int cg = ch * 4 + j8;
                                uint8_t* dst = e_row + (cg >> 3) * kBlkBytes + ((((uint32_t)cg & 7u) ^ sw) << 4);
                                *reinterpret_cast<uint4*>(dst) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
                            }
                        }
                    }
                }
                NTX_TOC(te_loop);
                if (offdiag) {
                    fence_proxy_async_smem();       // generic-proxy stores -> visible to tcgen05.mma
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&ctl->e_full[g]);
                    ++e_uses;
                    pending = true;
                    pend_rb = rb;
                    pend_kt = kt;
                }
            }
            if (last_of_seg) flush(rb);
            it.next();
            if (++turn == kSymWG) turn = 0;
        }
        if (pending) store_colsum();
        NTX_TOC(te_total);
        if (threadIdx.x == 0) {
            NTX_TSAVE(4, te_total); NTX_TSAVE(5, te_wait); NTX_TSAVE(6, te_ef); NTX_TSAVE(7, te_loop);
            NTX_TSAVE(8, te_cs);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) tmem_dealloc(tmem_base, kTmemCols);
}

// ============================================================================ backward
// NDB = 1, 2 (d_pad 64, 128): dZ in TMEM columns [0, 128), three S/C buffers, three groups.
// NDB = 4       (d_pad 256) : dZ in TMEM columns [0, 256), two S/C buffers, two groups.
// kSplit (NTX_BF16X2 rows, NDB = 2 or 4 blocks of [hi | lo] terms, d_pad = NDB * 32): dZ in columns [0, 128), two
//   buffers of 192 columns: S (128 fp32 columns), over which the hi term of the coefficient tile is written in place
//   (64 columns), then 64 columns for its lo term; two groups.
template <int NDB, bool kSplit = false> struct BwdCfg {
    static constexpr int NWG = (NDB == 4 || kSplit) ? 2 : 3;   // epilogue groups = S/C buffers
    static constexpr int NS = NDB == 4 ? 2 : (kSplit ? 4 : 6); // multiple of NWG: the MMA loop unrolls by NS
    static constexpr int DP = kSplit ? NDB * 32 : NDB * 64;    // dZ columns
    static constexpr uint32_t kSCol = (NDB == 4 && !kSplit) ? 256 : 128;
    static constexpr uint32_t kBufCols = kSplit ? 192 : 128;   // TMEM columns per S/C buffer
    static constexpr int kThreads = NWG * 128 + 96;            // + producer warp + GEMM1 warp + GEMM2 warp
};
template <int NS, int NWG> struct BwdCtl {
    uint64_t q_full, q_empty, k_full[NS], k_empty[NS], s_full[NWG], c_full[NWG], buf_free[NWG], dz_full,
        dz_empty;
    uint32_t tmem_base;
    alignas(16) float beta[NWG][128];
};

struct BwdWeights {
    const float* rowsum_q;
    const float* rowsum_k;
    const float* w_q;
    const float* w_k;
    int w_stride;       // 0: one weight for every row, 1: one per row
    float w_scale;
    int w_chunk;        // > 0: one weight per rank, key c takes w_k[c / w_chunk]
    int w_bf16;         // weights are bf16 values
    float coef_scale;   // scale of the weights in alpha / beta (w_scale, times 2^-c1 for scaled rows)
};

// The rows of a finished row block, turned into dview0 / dview1 inside the backward kernel (see ntx_common.cuh,
// "finalize work folded into the tensor-core kernels"): what bwd_finalize_vec_kernel does, with the input dtype a
// run-time flag.  enabled == 0: the kernel only writes its slots and the separate finalize kernel runs.
struct BwdFuse {
    const void* view0;
    const void* view1;
    void* dview0;
    void* dview1;
    const float* inv_norm;
    unsigned int* tag;              // fold counters at the head of the scratch (NtxFoldScratch)
    unsigned int* done_total;
    unsigned int* arrive;           // [row_blocks]
    int row_blocks;
    int in_ld, normalised, in_bf16, d, enabled;
};
__device__ __forceinline__ float4 ld4_any(const void* base, size_t i, int bf16) {     // 4 elements from index i
    if (bf16) {
        const uint2 t = __ldg(reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(base) + i));
        const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.x));
        const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&t.y));
        return make_float4(a.x, a.y, b.x, b.y);
    }
    return __ldg(reinterpret_cast<const float4*>(static_cast<const float*>(base) + i));
}
__device__ __forceinline__ void st4_any(void* base, size_t i, float4 v, int bf16) {
    if (bf16) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        uint2 t;
        t.x = *reinterpret_cast<uint32_t*>(&a);
        t.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(base) + i) = t;
    } else {
        *reinterpret_cast<float4*>(static_cast<float*>(base) + i) = v;
    }
}

// 32 similarity values of one row -> 32 coefficients, packed bf16x2.  kSplit: the coefficients as two bf16 terms,
// pk = bf16(c), pk_lo = bf16(c - pk), and every exponential on MUFU.
template <bool kEdge, bool kScaled, bool kSplit = false>
__device__ __forceinline__ void coef_chunk(const uint32_t (&v)[32], uint32_t (&pk)[16], uint32_t (&pk_lo)[16],
                                           const float* bt, float alpha, float c1, float nc1, int col0, int self,
                                           int k_rows, bool row_ok) {
    const f32x2 c1x2 = pack2(c1, c1), nc1x2 = pack2(nc1, nc1), alpha2 = pack2(alpha, alpha);
#pragma unroll
    for (int j4 = 0; j4 < 8; ++j4) {
        const float4 b = *reinterpret_cast<const float4*>(bt + j4 * 4);
        const f32x2 beta2[2] = {pack2(b.x, b.y), pack2(b.z, b.w)};
#pragma unroll
        for (int h = 0; h < 2; ++h) {        // two keys per instruction (FFMA2, FADD2, FMUL2)
            const int j = j4 * 4 + h * 2;
            const f32x2 raw = pack2(__uint_as_float(v[j]), __uint_as_float(v[j + 1]));
            const f32x2 x = kScaled ? raw : fma2(raw, c1x2, nc1x2);
            const bool poly = !kEdge && !kSplit && NTX_BWD_POLY > 0 && (((j4 * 2 + h) * NTX_BWD_POLY) & 3) < NTX_BWD_POLY;
            const f32x2 e = poly ? ex2_pair_poly(x) : ex2_pair_mufu(x);
            float c0, c1v;
            unpack2(mul2(e, add2(alpha2, beta2[h])), c0, c1v);
            if (kEdge) {
                const int col = col0 + j;
                if (col == self || col >= k_rows || !row_ok) c0 = 0.f;
                if (col + 1 == self || col + 1 >= k_rows || !row_ok) c1v = 0.f;
            }
            __nv_bfloat162 hh = __floats2bfloat162_rn(c0, c1v);
            pk[j4 * 2 + h] = *reinterpret_cast<uint32_t*>(&hh);
            if constexpr (kSplit) {
                const float2 hf = __bfloat1622float2(hh);
                __nv_bfloat162 ll = __floats2bfloat162_rn(c0 - hf.x, c1v - hf.y);
                pk_lo[j4 * 2 + h] = *reinterpret_cast<uint32_t*>(&ll);
            }
        }
    }
}

// kGate: peer exchange (the row sums and weights of other ranks' keys arrive during the kernel); a separate
// instantiation so the single-GPU kernel keeps exactly its own instruction stream
template <int NDB, bool kGate, bool kScaled, bool kSplit = false>
__global__ void __launch_bounds__(BwdCfg<NDB, kSplit>::kThreads, 1)
tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_k,
              const __grid_constant__ CUtensorMap tmap_q1, const __grid_constant__ CUtensorMap tmap_k1,
              NtxDims dm, int key_tiles, long long T, int rows_pad, BwdWeights wt,
              float* __restrict__ partial, NtxGate gate, BwdFuse fuse) {
    using Cfg = BwdCfg<NDB, kSplit>;
#ifdef NTX_TIMELINE
    const long long t_entry = clock64();
#endif
    constexpr int NS = Cfg::NS, NWG = Cfg::NWG, DP = Cfg::DP;
    constexpr uint32_t kSCol = Cfg::kSCol, kBufCols = Cfg::kBufCols;
    constexpr uint32_t kTileBytes = NDB * kBlkBytes;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = checked_1024(smem_raw);
    uint8_t* q_smem = smem;
    uint8_t* k_smem = smem + kTileBytes;
    BwdCtl<NS, NWG>* ctl = reinterpret_cast<BwdCtl<NS, NWG>*>(k_smem + NS * kTileBytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x, cta = blockIdx.x;
    const long long t_begin = ntx_chunk_begin(T, G, cta);
    const int n_tiles = (int)(ntx_chunk_begin(T, G, cta + 1) - t_begin);
    constexpr int kProducerWarp = NWG * 4, kMmaWarp = NWG * 4 + 1;

    if (threadIdx.x == 0) {
        mbar_init(&ctl->q_full, 1);
        mbar_init(&ctl->q_empty, 1);
        for (int s = 0; s < NS; ++s) { mbar_init(&ctl->k_full[s], 1); mbar_init(&ctl->k_empty[s], 1); }
        for (int b = 0; b < NWG; ++b) {
            mbar_init(&ctl->s_full[b], 1);
            mbar_init(&ctl->c_full[b], 4);
            mbar_init(&ctl->buf_free[b], 1);
        }
        mbar_init(&ctl->dz_full, 1);
        mbar_init(&ctl->dz_empty, NWG * 4);
        fence_barrier_init();
    }
    if (warp == kMmaWarp) tmem_alloc(&ctl->tmem_base, kTmemCols);
    if (warp == kProducerWarp && lane == 0) {
        tma_prefetch_desc(&tmap_q); tma_prefetch_desc(&tmap_k);
        tma_prefetch_desc(&tmap_q1); tma_prefetch_desc(&tmap_k1);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // our TMEM is allocated: only now may the next kernel on the stream be staged (its CTAs allocate TMEM too)
    ntx_dep_trigger();
    // the whole TMEM (512 columns) is ours, so the allocation starts at column 0, lane 0;
    // using the literal keeps every TMEM address a compile-time constant
    if (ctl->tmem_base != 0) {
        if (threadIdx.x == 0) printf("ntxent: unexpected TMEM base %u\n", ctl->tmem_base);
        __trap();
    }
    constexpr uint32_t tmem_base = 0;
    // everything above overlapped the tail of the previous kernel on the stream; from here on its results are read.
    // (Measured and dropped: with a peer exchange the key tiles are ordered by the exchange's flag words, so the
    // producer and the MMA warps could fill the pipeline before this wait; on two GPUs the step did not get shorter.)
    ntx_dep_wait();
#ifdef NTX_TIMELINE
    if (threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][16] = clock64() - t_entry;
#endif

    if (warp == kProducerWarp) {
        if (elect_one()) {
            KeyGate kg;
            kg.init(gate, dm.k_rows);
            const bool alt = kg.step & 1u;                   // key buffer the forward of this step used
            producer_loop<NDB, NS>(ctl, q_smem, k_smem, alt ? &tmap_q1 : &tmap_q, alt ? &tmap_k1 : &tmap_k, t_begin,
                                   n_tiles, key_tiles, kg);
        }
    } else if (warp == kMmaWarp) {
        // GEMM1 issuer: S(i) = Q . K(i)^T as soon as the key tile has landed and S/C buffer i % NWG
        // is free again.  Unrolled by NS (a multiple of NWG): stage / buffer indices are compile-time
        // constants, so descriptors and TMEM addresses are uniform values.
        int kt = (int)(t_begin % key_tiles), seg = 0;
        bool new_seg = true;
        NTX_T(tm_k); NTX_T(tm_free); NTX_T(tm_issue); NTX_T(tm_total);
        NTX_TIC(tm_total);
        for (int i0 = 0; i0 < n_tiles; i0 += NS) {
#pragma unroll
            for (int jj = 0; jj < NS; ++jj) {
                const int i = i0 + jj;
                if (i >= n_tiles) break;
                const int s = jj, sb = jj % NWG;
                const bool last_of_seg = (i + 1 == n_tiles) || (kt + 1 == key_tiles);
                if (new_seg) mbar_wait(&ctl->q_full, seg & 1);
                new_seg = false;
                { NTX_TIC(tm_k); mbar_wait(&ctl->k_full[s], (i0 / NS) & 1); NTX_TOC(tm_k); }
                { NTX_TIC(tm_free); mbar_wait(&ctl->buf_free[sb], ((i / NWG) & 1) ^ 1); NTX_TOC(tm_free); }
                tc_fence_after();
                NTX_TIC(tm_issue);
                if (elect_one()) {
                    issue_gemm1<NDB, kSplit>(tmem_base + kSCol + sb * kBufCols, q_smem, k_smem + s * kTileBytes);
                    mma_commit(&ctl->s_full[sb]);
                    if (last_of_seg) mma_commit(&ctl->q_empty);
                }
                __syncwarp();
                NTX_TOC(tm_issue);
                if (last_of_seg) { ++seg; new_seg = true; }
                if (++kt == key_tiles) kt = 0;
            }
        }
        NTX_TOC(tm_total);
        if (lane == 0) { NTX_TSAVE(8, tm_total); NTX_TSAVE(9, tm_k); NTX_TSAVE(10, tm_free); NTX_TSAVE(12, tm_issue); }
    } else if (warp == kMmaWarp + 1) {
        // GEMM2 issuer: dZ += C(j) . K(j) once the epilogue group has written C(j) into TMEM.
        // C is read from TMEM (bf16, 8 columns per 16 keys), K MN-major from the same smem tile.
        // A second issuing warp keeps the tensor pipe fed while the GEMM1 warp sits in its waits.
        constexpr uint32_t idesc2 = instr_desc_bf16(128, DP, 1);
        int kt = (int)(t_begin % key_tiles), seg = 0;
        bool first_of_seg = true;
        NTX_T(tm_c); NTX_T(tm_issue);
        for (int j0 = 0; j0 < n_tiles; j0 += NS) {
#pragma unroll
            for (int jj = 0; jj < NS; ++jj) {
                const int j = j0 + jj;
                if (j >= n_tiles) break;
                const int sj = jj, cb = jj % NWG;
                const bool last_of_seg = (j + 1 == n_tiles) || (kt + 1 == key_tiles);
                if (first_of_seg) mbar_wait(&ctl->dz_empty, (seg & 1) ^ 1);   // previous row block drained
                { NTX_TIC(tm_c); mbar_wait(&ctl->c_full[cb], (j / NWG) & 1); NTX_TOC(tm_c); }
                tc_fence_after();
                NTX_TIC(tm_issue);
                if (elect_one()) {
                    const uint32_t c_tmem = tmem_base + kSCol + cb * kBufCols;
                    const uint8_t* k_tile = k_smem + sj * kTileBytes;
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        // MN-major key tile: next 64 features one 16 KB block away, 8-key groups 1 KB apart
                        const uint64_t bd = smem_desc_sw128(smem_u32(k_tile + kk * 2048), kBlkBytes, 1024);
                        if constexpr (kSplit) {
                            // dZ += C_lo . K_hi + C_hi . K_lo + C_hi . K_hi; the lo blocks of the tile follow its hi blocks
                            const uint64_t bd_lo =
                                smem_desc_sw128(smem_u32(k_tile + (NDB / 2) * kBlkBytes + kk * 2048), kBlkBytes, 1024);
                            mma_bf16_ts(tmem_base, c_tmem + 128 + kk * 8, bd, idesc2, (first_of_seg && kk == 0) ? 0u : 1u);
                            mma_bf16_ts(tmem_base, c_tmem + kk * 8, bd_lo, idesc2, 1u);
                            mma_bf16_ts(tmem_base, c_tmem + kk * 8, bd, idesc2, 1u);
                        } else {
                            mma_bf16_ts(tmem_base, c_tmem + kk * 8, bd, idesc2, (first_of_seg && kk == 0) ? 0u : 1u);
                        }
                    }
                    mma_commit(&ctl->k_empty[sj]);
                    mma_commit(&ctl->buf_free[cb]);
                    if (last_of_seg) mma_commit(&ctl->dz_full);
                }
                __syncwarp();
                NTX_TOC(tm_issue);
                first_of_seg = last_of_seg;
                if (last_of_seg) ++seg;
                if (++kt == key_tiles) kt = 0;
            }
        }
        if (lane == 0) { NTX_TSAVE(11, tm_c); NTX_TSAVE(15, tm_issue); }
    } else {
        // ---------------------------------------------------------------- epilogue groups
        const int g = warp >> 2;
        const int tid_g = threadIdx.x & 127;
        const int row_in_tile = (warp & 3) * 32 + lane;
        const uint32_t t_lane = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const uint32_t taddr = t_lane + kSCol + g * kBufCols;
        const float c1 = dm.c1, nc1 = -dm.c1;
        float alpha = 0.f;
        bool row_ok = false;
        NTX_T(te_wait); NTX_T(te_total); NTX_T(te_drain);
        NTX_TIC(te_total);

        // move this group's 32-column chunks of the dZ accumulator to the scratch slot
        auto drain = [&](int rb, int sg) {
            mbar_wait(&ctl->dz_full, sg & 1);
            tc_fence_after();
            const int slot = cta - ntx_first_owner(T, G, key_tiles, rb);
            float* dst = partial + ((size_t)slot * rows_pad + rb * 128 + row_in_tile) * DP;
#pragma unroll 1
            for (int c = g * 32; c < DP; c += NWG * 32) {
                uint32_t v[32];
                tmem_ld_32x32(t_lane + c, v);
                tmem_ld_wait_dep(v);
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<uint4*>(dst + c + j) = make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&ctl->dz_empty);
            if (fuse.enabled) {          // every group's part of the slot is stored: announce it on the row block
                named_bar_sync(6, NWG * 128);
                if (threadIdx.x == 0) {
                    __threadfence();
                    atomicAdd(fuse.arrive + rb, 1u);
                }
            }
        };

        // w_c and R_c of one key are fetched a whole tile ahead (raw: the division waits until the
        // values are consumed, so the load latency stays off the critical path)
        // With a peer exchange the values of rank c's keys are stored into this rank's block by rank c: its forward
        // finalize (row sums, rflag) and the start of its backward (weights, wflag) -- wait for both flags once.
        float w_next = 0.f, r_next = 1.f;
        const uint32_t gate_step = kGate ? *gate.step : 0u;
        uint32_t gate_ready = 0;
        auto fetch_beta = [&](int key_tile) {
            const int c = key_tile * 128 + tid_g;
            const bool ok = c < dm.k_rows;
            if constexpr (kGate) {
                // one poller per warp (every thread polling the same word serialises on its L2 slice); the warp
                // barrier passes the acquire on to the other lanes
                const int c0 = key_tile * 128 + (warp & 3) * 32;
                if (c0 < dm.k_rows) {
                    const int ch_lo = c0 / gate.chunk_rows, ch_hi = min(c0 + 31, dm.k_rows - 1) / gate.chunk_rows;
                    for (int ch = ch_lo; ch <= ch_hi; ++ch)
                        if (!((gate_ready >> ch) & 1u)) {
                            if (lane == 0) {
                                ntx_gate_wait(gate.ctl, NTX_CTL_RFLAG + ch, gate_step);
                                ntx_gate_wait(gate.ctl, NTX_CTL_WFLAG + ch, gate_step);
                            }
                            gate_ready |= 1u << ch;
                            __syncwarp();
                        }
                }
                w_next = ok ? ntx_ld_weight(wt.w_k, ntx_key_weight_index(c, wt.w_stride, wt.w_chunk), wt.w_bf16, true) : 0.f;
                r_next = ok ? ntx_ld_l2(wt.rowsum_k + c) : 1.f;
            } else {
                w_next = ok ? ntx_ld_weight(wt.w_k, ntx_key_weight_index(c, wt.w_stride, wt.w_chunk), wt.w_bf16) : 0.f;
                r_next = ok ? __ldg(wt.rowsum_k + c) : 1.f;
            }
        };
        int rb = (int)(t_begin / key_tiles), kt = (int)(t_begin % key_tiles);
        if (fuse.enabled && threadIdx.x == 0) ntx_fold_init(fuse.tag, fuse.done_total, fuse.arrive, fuse.row_blocks, cta);
        int seg = 0, own = 0, turn = 0;
        bool new_seg = true;
        if (g < n_tiles) fetch_beta((kt + g) % key_tiles);

        for (int i = 0; i < n_tiles; ++i) {
            if (new_seg) {
                const int row = rb * 128 + row_in_tile;
                row_ok = row < dm.q_rows;
                alpha = row_ok ? ntx_ld_weight(wt.w_q, (size_t)row * wt.w_stride, wt.w_bf16) * wt.coef_scale / __ldg(wt.rowsum_q + row)
                               : 0.f;
                new_seg = false;
            }
            const bool last_of_seg = (i + 1 == n_tiles) || (kt + 1 == key_tiles);
            if (turn == g) {
                float* bt = ctl->beta[g];
                const float beta = w_next * wt.coef_scale / r_next;
                named_bar_sync(2 + g, 128);          // everyone is done with the previous tile's values
                bt[tid_g] = beta;
                named_bar_sync(2 + g, 128);
                if (i + NWG < n_tiles) fetch_beta((kt + NWG) % key_tiles);

                { NTX_TIC(te_wait); mbar_wait(&ctl->s_full[g], own & 1); NTX_TOC(te_wait); }
#ifdef NTX_TIMELINE
                if (own == 0 && threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][17] = clock64() - t_entry;
#endif
                ++own;
                tc_fence_after();
                const bool ragged = rb * 128 + 128 > dm.q_rows;
                const bool edge_any = edge_tile(rb, kt, dm) || ragged;
                const int self = ntx_self_col(rb * 128 + row_in_tile, dm.n_local, dm.rank);
#pragma unroll 1
                for (int ch = 0; ch < 4; ++ch) {
                    uint32_t v[32], pk[16], pk_lo[16];                    // pk_lo: two-term rows only
                    tmem_ld_32x32(taddr + ch * 32, v);
                    // the masked form only for the 32-key chunks that hold a self column of this warp's rows or run
                    // past the last key (warp-uniform), and for every chunk of a ragged last row block
                    bool edge = edge_any;
                    if (edge_any && !ragged) {
                        const int c0 = kt * 128 + ch * 32;
                        edge = c0 + 32 > dm.k_rows || __any_sync(0xffffffffu, self >= c0 && self < c0 + 32);
                    }
                    tmem_ld_wait_dep(v);
                    if (!edge)
                        coef_chunk<false, kScaled, kSplit>(v, pk, pk_lo, bt + ch * 32, alpha, c1, nc1, 0, 0, 0, true);
                    else
                        coef_chunk<true, kScaled, kSplit>(v, pk, pk_lo, bt + ch * 32, alpha, c1, nc1, kt * 128 + ch * 32,
                                                          self, dm.k_rows, row_ok);
                    if constexpr (kSplit) tmem_st_32x16(taddr + 128 + ch * 16, pk_lo);   // lo terms behind the S columns
                    // bf16 pairs over fp32 columns already read: keys [32ch, 32ch+32) -> columns [16ch, +16)
                    tmem_st_32x16(taddr + ch * 16, pk);
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&ctl->c_full[g]);
            }
            if (last_of_seg) {
#ifdef NTX_TIMELINE
                if (i + 1 == n_tiles && threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][18] = clock64() - t_entry;
#endif
                { NTX_TIC(te_drain); drain(rb, seg); NTX_TOC(te_drain); }
                ++seg;
                new_seg = true;
            }
            if (++kt == key_tiles) { kt = 0; ++rb; }
            if (++turn == NWG) turn = 0;
        }
        NTX_TOC(te_total);
        if (threadIdx.x == 0) { NTX_TSAVE(13, te_total); NTX_TSAVE(14, te_wait); NTX_TSAVE(7, te_drain); }
#ifdef NTX_TIMELINE
        if (threadIdx.x == 0) g_ntx_timers[blockIdx.x % 148][19] = clock64() - t_entry;
#endif

        // ---------------------------------------------------------------- finish: this CTA's share of its row blocks
        // add the dZ slots in slot order, subtract the positive-pair term, 1/tau, Jacobian of x / max(||x||, eps)
        // (loss.py:33-34 backward), write dview0 / dview1.  16 threads per row, up to DP / 64 float4 each.
        if (fuse.enabled && n_tiles > 0) {
            constexpr int kLanes = 16, kRowsPerPass = NWG * 128 / kLanes, kV = DP / 64;
            const int tid = threadIdx.x, sub = tid % kLanes;
            const int n = dm.n_local, d = fuse.d;
            const size_t slot_stride = (size_t)rows_pad * DP;
            const int rb_first = (int)(t_begin / key_tiles), rb_last = (int)((t_begin + n_tiles - 1) / key_tiles);
            for (int frb = rb_first; frb <= rb_last; ++frb) {
                const int first = ntx_first_owner(T, G, key_tiles, frb);
                const int ns = ntx_num_slots(T, G, key_tiles, frb), slot = cta - first;
                if (tid == 0) {
                    while (ntx_ld_acquire_gpu(fuse.arrive + frb) < (unsigned)ns) __nanosleep(32);
                }
                named_bar_sync(6, NWG * 128);
                const int lo = slot * 128 / ns, hi = (slot + 1) * 128 / ns;
                for (int r0 = lo; r0 < hi; r0 += kRowsPerPass) {
                    const int rl = r0 + tid / kLanes;
                    const bool live = rl < hi && frb * 128 + rl < dm.q_rows;
                    const int r = min(frb * 128 + min(rl, hi - 1), dm.q_rows - 1);   // keep every lane in the shuffles
                    const int p = ntx_pos_row(r, n);
                    const void* zr_base = r < n ? fuse.view0 : fuse.view1;
                    const void* zp_base = p < n ? fuse.view0 : fuse.view1;
                    const size_t zr_off = (size_t)(r < n ? r : r - n) * fuse.in_ld, zp_off = (size_t)(p < n ? p : p - n) * fuse.in_ld;
                    const float inr = __ldg(fuse.inv_norm + r), inp = __ldg(fuse.inv_norm + p);
                    // rows -> unit rows: 1/||z|| for the caller's inputs, 1/sqrt(c1) for normalised rows that carry the scale
                    const float sr = fuse.normalised ? dm.inv_rs : inr, sp = fuse.normalised ? dm.inv_rs : inp;
                    const float wpair = (ntx_ld_weight(wt.w_q, (size_t)r * wt.w_stride, wt.w_bf16) +
                                         ntx_ld_weight(wt.w_q, (size_t)p * wt.w_stride, wt.w_bf16)) * wt.w_scale;
                    const float* part = partial + (size_t)r * DP;
                    float4 gq[kV], zh[kV];
                    float proj = 0.f;
#pragma unroll
                    for (int v = 0; v < kV; ++v) {
                        const int c = (sub + v * kLanes) * 4;
                        if (c < d) {
                            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                            for (int s0 = 0; s0 < ns; s0 += 4) {           // slot order, four loads in flight
                                float4 t[4];
#pragma unroll
                                for (int u = 0; u < 4; ++u)
                                    t[u] = s0 + u < ns ? __ldcg(reinterpret_cast<const float4*>(part + (s0 + u) * slot_stride + c))
                                                       : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                                for (int u = 0; u < 4; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
                            }
                            const float4 zpv = ld4_any(zp_base, zp_off + c, fuse.in_bf16);
                            const float4 zrv = ld4_any(zr_base, zr_off + c, fuse.in_bf16);
                            const float k = wpair * sp, it = dm.inv_tau, ir = dm.inv_rs;     // dZ' = C . K' carries sqrt(c1)
                            gq[v] = make_float4((acc.x * ir - k * zpv.x) * it, (acc.y * ir - k * zpv.y) * it,
                                                (acc.z * ir - k * zpv.z) * it, (acc.w * ir - k * zpv.w) * it);
                            zh[v] = make_float4(zrv.x * sr, zrv.y * sr, zrv.z * sr, zrv.w * sr);
                            proj = fmaf(zh[v].x, gq[v].x, fmaf(zh[v].y, gq[v].y, fmaf(zh[v].z, gq[v].z, fmaf(zh[v].w, gq[v].w, proj))));
                        }
                    }
#pragma unroll
                    for (int o = kLanes >> 1; o > 0; o >>= 1) proj += __shfl_xor_sync(0xffffffffu, proj, o);
                    const bool clamped = inr >= 1e12f;    // ||z|| <= eps: denominator was the constant eps
                    void* out_base = r < n ? fuse.dview0 : fuse.dview1;
                    const size_t out_off = (size_t)(r < n ? r : r - n) * d;
#pragma unroll
                    for (int v = 0; v < kV; ++v) {
                        const int c = (sub + v * kLanes) * 4;
                        if (c < d && live) {
                            float4 o4;
                            if (clamped) {
                                o4 = make_float4(gq[v].x * inr, gq[v].y * inr, gq[v].z * inr, gq[v].w * inr);
                            } else {
                                o4 = make_float4((gq[v].x - zh[v].x * proj) * inr, (gq[v].y - zh[v].y * proj) * inr,
                                                 (gq[v].z - zh[v].z * proj) * inr, (gq[v].w - zh[v].w * proj) * inr);
                            }
                            st4_any(out_base, out_off + c, o4, fuse.in_bf16);
                        }
                    }
                }
                // second pass over the row block's counter: the last CTA out leaves it clean for the next launch
                named_bar_sync(6, NWG * 128);
                if (tid == 0 && atomicAdd(fuse.arrive + frb, 1u) == 2u * (unsigned)ns - 1u) fuse.arrive[frb] = 0u;
            }
        }
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
// cuTensorMapEncodeTiled is a DRIVER call: it needs a context bound to the calling thread.  Runtime calls bind the
// primary context lazily, but a thread whose first CUDA work is this library (PyTorch's autograd thread, when the
// backward is its first CUDA call) has none yet: bind it once per (thread, device).
int bind_context() {
    thread_local int bound_dev = -1;
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev != bound_dev) {
        NTX_CUDA(cudaFree(nullptr));
        bound_dev = dev;
    }
    return NTX_OK;
}

int make_rows_tmap(CUtensorMap* m, const __nv_bfloat16* base, int rows, int ld, int box_rows = 128) {
    EncodeTiledFn enc;
    int rc = get_encode_fn(&enc);
    if (rc) return rc;
    if ((rc = bind_context())) return rc;
    const cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(__nv_bfloat16)};
    const cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
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
template <int NDB, bool kSplit = false> size_t bwd_smem_bytes() {
    using Cfg = BwdCfg<NDB, kSplit>;
    return (size_t)(1 + Cfg::NS) * NDB * kBlkBytes + sizeof(BwdCtl<Cfg::NS, Cfg::NWG>);
}

// the maps of a launch: rows and keys in key buffer 0 and, with a peer exchange, in key buffer 1
struct RowMaps {
    CUtensorMap q[2], k[2];
};
int make_row_maps(RowMaps* m, const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm, const NtxGate& gate,
                  size_t key_buf_elems) {
    int rc;
    if ((rc = make_rows_tmap(&m->q[0], q, dm.q_rows, dm.ld)) || (rc = make_rows_tmap(&m->k[0], k, dm.k_rows, dm.ld)))
        return rc;
    if (gate.step) {
        if ((rc = make_rows_tmap(&m->q[1], q + key_buf_elems, dm.q_rows, dm.ld)) ||
            (rc = make_rows_tmap(&m->k[1], k + key_buf_elems, dm.k_rows, dm.ld)))
            return rc;
    } else {
        m->q[1] = m->q[0];
        m->k[1] = m->k[0];
    }
    return NTX_OK;
}

template <int NDB, bool kScaled, bool kSplit = false>
int launch_fwd_s(const RowMaps& m, const NtxDims& dm, const NtxPlan& plan, float* partial, const NtxFoldScratch& fold,
                 const NtxFwdOut& out, const NtxGate& gate, cudaStream_t st) {
    const size_t smem = fwd_smem_bytes<NDB>();
    NTX_SMEM_OPT_IN((tc_fwd_kernel<NDB, kScaled, kSplit>), smem);
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    NTX_CUDA(ntx_launch_dep(tc_fwd_kernel<NDB, kScaled, kSplit>, plan.grid, kFwdThreads, smem, st, m.q[0], m.k[0], m.q[1],
                            m.k[1], dm, plan.key_tiles, T, plan.row_blocks * plan.tile_rows, partial, fold, out, gate));
    return NTX_OK;
}
template <int NDB, bool kSplit = false>
int launch_fwd(const RowMaps& m, const NtxDims& dm, const NtxPlan& plan, float* partial, const NtxFoldScratch& fold,
               const NtxFwdOut& out, const NtxGate& gate, cudaStream_t st) {
    return dm.scaled ? launch_fwd_s<NDB, true, kSplit>(m, dm, plan, partial, fold, out, gate, st)
                     : launch_fwd_s<NDB, false, kSplit>(m, dm, plan, partial, fold, out, gate, st);
}

template <int NDB, bool kGate, bool kScaled, bool kSplit = false>
int launch_bwd_gs(const RowMaps& m, const NtxDims& dm, const NtxPlan& plan, const BwdWeights& wt, float* partial,
                  const NtxGate& gate, const BwdFuse& fuse, cudaStream_t st) {
    const size_t smem = bwd_smem_bytes<NDB, kSplit>();
    NTX_SMEM_OPT_IN((tc_bwd_kernel<NDB, kGate, kScaled, kSplit>), smem);
    const long long T = (long long)plan.row_blocks * plan.key_tiles;
    NTX_CUDA(ntx_launch_dep(tc_bwd_kernel<NDB, kGate, kScaled, kSplit>, plan.grid, BwdCfg<NDB, kSplit>::kThreads, smem, st,
                            m.q[0], m.k[0], m.q[1], m.k[1], dm, plan.key_tiles, T, plan.row_blocks * plan.tile_rows, wt,
                            partial, gate, fuse));
    return NTX_OK;
}
template <int NDB>
int launch_bwd(const RowMaps& m, const NtxDims& dm, const NtxPlan& plan, const BwdWeights& wt, float* partial,
               const NtxGate& gate, const BwdFuse& fuse, cudaStream_t st) {
    if (dm.scaled)
        return gate.step ? launch_bwd_gs<NDB, true, true>(m, dm, plan, wt, partial, gate, fuse, st)
                         : launch_bwd_gs<NDB, false, true>(m, dm, plan, wt, partial, gate, fuse, st);
    return gate.step ? launch_bwd_gs<NDB, true, false>(m, dm, plan, wt, partial, gate, fuse, st)
                     : launch_bwd_gs<NDB, false, false>(m, dm, plan, wt, partial, gate, fuse, st);
}
// two-term rows (NTX_BF16X2)
template <int NDB>
int launch_bwd_split(const RowMaps& m, const NtxDims& dm, const NtxPlan& plan, const BwdWeights& wt, float* partial,
                     const NtxGate& gate, const BwdFuse& fuse, cudaStream_t st) {
    if (dm.scaled)
        return gate.step ? launch_bwd_gs<NDB, true, true, true>(m, dm, plan, wt, partial, gate, fuse, st)
                         : launch_bwd_gs<NDB, false, true, true>(m, dm, plan, wt, partial, gate, fuse, st);
    return gate.step ? launch_bwd_gs<NDB, true, false, true>(m, dm, plan, wt, partial, gate, fuse, st)
                     : launch_bwd_gs<NDB, false, false, true>(m, dm, plan, wt, partial, gate, fuse, st);
}

}  // namespace

// [rows x ld] bf16 row-major -> TMA map with boxes of box_rows rows x 64 columns, 128B swizzle (for ntx_project.cu)
int ntx_make_rows_tmap(void* map, const void* base, int rows, int ld, int box_rows) {
    return make_rows_tmap(static_cast<CUtensorMap*>(map), static_cast<const __nv_bfloat16*>(base), rows, ld, box_rows);
}

#ifdef NTX_TIMING
extern "C" int ntx_debug_timers(long long* host_out) {   // 148 x 24 cycle counters
    return cudaMemcpyFromSymbol(host_out, g_ntx_timers, sizeof(long long) * 148 * 24) == cudaSuccess ? 0 : -3;
}
#endif

int ntx_tc_fwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, float* partial, const NtxFoldScratch& fold, const NtxFwdOut& out,
                      const NtxGate& gate, size_t key_buf_elems, cudaStream_t st) {
    RowMaps m;
    int rc = make_row_maps(&m, q, k, dm, gate, key_buf_elems);
    if (rc) return rc;
    if (dm.split) {                     // tiles of [hi | lo] terms: plan.d_pad is the width of one term
        if (plan.d_pad == 64) return launch_fwd<2, true>(m, dm, plan, partial, fold, out, gate, st);
        if (plan.d_pad == 128) return launch_fwd<4, true>(m, dm, plan, partial, fold, out, gate, st);
        ntx_set_error("tensor-core forward on two-term rows: padded feature dim %d not built", plan.d_pad);
        return NTX_E_UNSUPPORTED;
    }
    switch (plan.d_pad) {
        case 64: return launch_fwd<1>(m, dm, plan, partial, fold, out, gate, st);
        case 128: return launch_fwd<2>(m, dm, plan, partial, fold, out, gate, st);
        case 256: return launch_fwd<4>(m, dm, plan, partial, fold, out, gate, st);
    }
    ntx_set_error("tensor-core forward: padded feature dim %d not built", plan.d_pad);
    return NTX_E_UNSUPPORTED;
}

namespace {
template <int NDB, bool kScaled>
int launch_sym(const CUtensorMap& m, const NtxDims& dm, const NtxPlan& plan, float* partial, float* colpart,
               unsigned int* counter, cudaStream_t st) {
    const long long T = (long long)plan.row_blocks * (plan.row_blocks + 1) / 2;
    const int rows_pad = plan.row_blocks * plan.tile_rows;
    const size_t smem = (size_t)(1 + kSymNS) * NDB * kBlkBytes + kSymWG * 2 * kBlkBytes + 2048 + sizeof(SymCtl);
    NTX_SMEM_OPT_IN((tc_fwd_sym_kernel<NDB, kScaled>), smem);
    NTX_CUDA(ntx_launch_dep(tc_fwd_sym_kernel<NDB, kScaled>, plan.fwd_grid, kSymThreads, smem, st, m, dm, plan.row_blocks, T,
                            rows_pad, partial, colpart, counter));
    return NTX_OK;
}
}  // namespace

int ntx_tc_fwd_sym_launch(const __nv_bfloat16* rows, const NtxDims& dm, const NtxPlan& plan, float* partial,
                          float* colpart, unsigned int* counter, cudaStream_t st) {
    CUtensorMap m;
    int rc = make_rows_tmap(&m, rows, dm.q_rows, dm.ld);
    if (rc) return rc;
    if (plan.d_pad == 64)
        return dm.scaled ? launch_sym<1, true>(m, dm, plan, partial, colpart, counter, st)
                         : launch_sym<1, false>(m, dm, plan, partial, colpart, counter, st);
    if (plan.d_pad == 128)
        return dm.scaled ? launch_sym<2, true>(m, dm, plan, partial, colpart, counter, st)
                         : launch_sym<2, false>(m, dm, plan, partial, colpart, counter, st);
    ntx_set_error("symmetric forward: padded feature dim %d not built", plan.d_pad);
    return NTX_E_UNSUPPORTED;
}

int ntx_tc_bwd_launch(const __nv_bfloat16* q, const __nv_bfloat16* k, const NtxDims& dm,
                      const NtxPlan& plan, const NtxWeights& w, float* partial, const NtxGate& gate,
                      size_t key_buf_elems, const NtxBwdOut& o, cudaStream_t st) {
    RowMaps m;
    int rc = make_row_maps(&m, q, k, dm, gate, key_buf_elems);
    if (rc) return rc;
    BwdWeights wt{w.rowsum_q, w.rowsum_k, w.w_q, w.w_k, w.w_stride, w.w_scale, w.w_chunk, w.w_bf16, w.coef_scale};
    const BwdFuse fuse{o.view0, o.view1, o.dview0, o.dview1, o.inv_norm, o.tag, o.done_total, o.arrive, o.row_blocks,
                       o.in_ld, o.normalised, o.in_bf16, dm.d, o.enabled};
    if (dm.split) {
        if (plan.d_pad == 64) return launch_bwd_split<2>(m, dm, plan, wt, partial, gate, fuse, st);
        if (plan.d_pad == 128) return launch_bwd_split<4>(m, dm, plan, wt, partial, gate, fuse, st);
        ntx_set_error("tensor-core backward on two-term rows: padded feature dim %d not built", plan.d_pad);
        return NTX_E_UNSUPPORTED;
    }
    switch (plan.d_pad) {
        case 64: return launch_bwd<1>(m, dm, plan, wt, partial, gate, fuse, st);
        case 128: return launch_bwd<2>(m, dm, plan, wt, partial, gate, fuse, st);
        case 256: return launch_bwd<4>(m, dm, plan, wt, partial, gate, fuse, st);
    }
    ntx_set_error("tensor-core backward: padded feature dim %d not built", plan.d_pad);
    return NTX_E_UNSUPPORTED;
}
