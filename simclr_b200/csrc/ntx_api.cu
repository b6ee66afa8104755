// extern "C" surface of libntxent_b200.so (see include/ntxent_b200.h).
// Validates arguments, derives the work partition and enqueues kernels on the
// caller's stream.  Allocates nothing, never synchronises.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "ntx_common.cuh"

static thread_local char g_err[512] = "";

bool ntx_pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("NTXENT_NO_PDL");
        return !(e && e[0] && e[0] != '0');
    }();
    return on;
}

static std::atomic<long long> g_launches{0};
void ntx_note_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

void ntx_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

namespace {

// armed by ntx_profile_next(); consumed by the next matching call on this thread
thread_local cudaEvent_t g_prof_start[2] = {nullptr, nullptr};
thread_local cudaEvent_t g_prof_stop[2] = {nullptr, nullptr};

struct ProfScope {
    int which;
    cudaStream_t st;
    ProfScope(int w, cudaStream_t s) : which(w), st(s) {
        if (g_prof_start[which]) cudaEventRecord(g_prof_start[which], st);
    }
    ~ProfScope() {
        if (g_prof_stop[which]) cudaEventRecord(g_prof_stop[which], st);
        g_prof_start[which] = g_prof_stop[which] = nullptr;
    }
};

constexpr int kTcMaxD = 256;   // widest padded feature dim the tcgen05 kernels hold in smem

// bytes per stored element of a rows layout (NTX_BF16X2 rows are bf16 elements, twice as many)
inline size_t esz_of(int dtype) { return dtype == NTX_F32 ? 4 : 2; }
inline bool rows_dtype_ok(int dtype) { return dtype == NTX_F32 || dtype == NTX_BF16 || dtype == NTX_BF16X2; }

int cached_sm_count(int* out) {
    static int counts[64];
    static bool have[64];
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) dev = 0;
    if (!have[dev]) {
        int n = 0;
        NTX_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        counts[dev] = n;
        have[dev] = true;
    }
    *out = counts[dev];
    return NTX_OK;
}

int check_layout(const NtxLayout* L, int q_rows, int k_rows) {
    NTX_CHECK_ARG(L != nullptr, "layout is NULL");
    NTX_CHECK_ARG(L->n_local >= 1 && L->world >= 1 && L->rank >= 0 && L->rank < L->world,
                  "bad layout (n_local=%d world=%d rank=%d)", L->n_local, L->world, L->rank);
    NTX_CHECK_ARG(q_rows == 2 * L->n_local, "q_rows=%d must equal 2*n_local=%d", q_rows,
                  2 * L->n_local);
    NTX_CHECK_ARG((long long)k_rows == (long long)q_rows * L->world,
                  "k_rows=%d must equal q_rows*world=%lld", k_rows, (long long)q_rows * L->world);
    return NTX_OK;
}

int make_dims(int q_rows, int k_rows, int d, int ld, int dtype, const NtxLayout* L, float temperature,
              NtxDims* dm) {
    NTX_CHECK_ARG(dtype == NTX_F32 || dtype == NTX_BF16 || dtype == NTX_BF16X2, "rows dtype %d not supported", dtype);
    NTX_CHECK_ARG(d >= 1, "feature dim d=%d must be positive", d);
    if (dtype == NTX_BF16X2 && d > 128) {
        ntx_set_error("two-term rows (NTX_BF16X2) support d <= 128 (got %d)", d);
        return NTX_E_UNSUPPORTED;
    }
    NTX_CHECK_ARG(ld == ntx_row_pitch(d, dtype), "ld=%d must be ntx_row_pitch(d=%d)=%d", ld, d,
                  ntx_row_pitch(d, dtype));
    NTX_CHECK_ARG(temperature > 0.f, "temperature must be positive (loss.py:16)");
    if (temperature < NTX_MIN_TEMPERATURE && dtype != NTX_F32) {
        ntx_set_error("temperature %g is below %g: the tensor-core kernels' fixed-shift logsumexp would leave fp32 "
                      "range; use fp32 rows, whose kernels switch to a max-shifted form",
                      (double)temperature, (double)NTX_MIN_TEMPERATURE);
        return NTX_E_UNSUPPORTED;
    }
    int rc = check_layout(L, q_rows, k_rows);
    if (rc) return rc;
    dm->q_rows = q_rows; dm->k_rows = k_rows; dm->d = d; dm->ld = ld;
    dm->n_local = L->n_local; dm->rank = L->rank;
    dm->inv_tau = 1.0f / temperature;
    dm->c1 = NTX_LOG2E / temperature;
    dm->stable = temperature < NTX_MIN_TEMPERATURE ? 1 : 0;
    NTX_CHECK_ARG(L->loss_dtype == NTX_F32 || L->loss_dtype == NTX_BF16, "layout.loss_dtype %d not supported", L->loss_dtype);
    dm->loss_bf16 = L->loss_dtype == NTX_BF16 ? 1 : 0;
    dm->split = dtype == NTX_BF16X2 ? 1 : 0;
    dm->dh = dm->split ? ld / 2 : ld;
    // rows that carry sqrt(log2e / tau): only the value ntx_row_scale() hands out, and only bf16 rows
    const float rs = L->row_scale > 0.f ? L->row_scale : 1.0f;
    dm->scaled = 0;
    dm->rowsum_scale = 1.0f;
    dm->pos_scale = dm->inv_tau;
    dm->inv_rs = 1.0f;
    if (rs != 1.0f) {
        const float want = ntx_row_scale(temperature, dtype);
        NTX_CHECK_ARG(dtype != NTX_F32 && want != 1.0f && fabsf(rs - want) <= 1e-6f * want,
                      "layout.row_scale=%g: rows must be unit rows (1) or, for bf16 rows, scaled by ntx_row_scale(%g)=%g",
                      (double)rs, (double)temperature, (double)want);
        dm->scaled = 1;
        dm->rowsum_scale = exp2f(-dm->c1);
        dm->pos_scale = 0.6931471805599453f * (dm->c1 / (rs * rs));     // ln 2 up to the rounding of rs
        dm->inv_rs = 1.0f / rs;
    }
    return NTX_OK;
}

// Do the tensor-core kernels finish their rows themselves (see fwd_impl)?  Opt-in, NTXENT_FOLD=1: the folded form
// makes the CTAs of one launch wait for each other, which is only safe while the whole grid is resident (it is on
// a GPU of its own; an MPS share or green context with fewer SMs than the device reports would stall it), and it
// pays only where many CTAs share a row block.
bool fold_wanted(const NtxPlan& plan) {
    static const bool on = [] {
        const char* e = getenv("NTXENT_FOLD");
        return e && e[0] && e[0] != '0';
    }();
    return on && plan.max_slots >= 2;
}

int aligned16(const void* p, const char* what) {
    NTX_CHECK_ARG(p != nullptr, "%s is NULL", what);
    NTX_CHECK_ARG(reinterpret_cast<uintptr_t>(p) % 16 == 0, "%s must be 16-byte aligned", what);
    return NTX_OK;
}

}  // namespace

extern "C" {

int ntx_version(void) { return NTX_VERSION; }
long long ntx_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
const char* ntx_last_error(void) { return g_err; }

float ntx_row_scale(float temperature, int dtype) {
    if ((dtype != NTX_BF16 && dtype != NTX_BF16X2) || !(temperature >= NTX_MIN_TEMPERATURE)) return 1.0f;
    return sqrtf(NTX_LOG2E / temperature);
}

int ntx_row_pitch(int d, int dtype) {
    if (d < 1) return 0;
    const int g = dtype == NTX_F32 ? 4 : 64;
    return (d + g - 1) / g * g * (dtype == NTX_BF16X2 ? 2 : 1);
}

int ntx_plan(int q_rows, int k_rows, int d, int dtype, int sm_count, NtxPlan* plan) {
    NTX_CHECK_ARG(plan != nullptr, "plan is NULL");
    NTX_CHECK_ARG(q_rows >= 1 && k_rows >= q_rows && d >= 1, "bad sizes q_rows=%d k_rows=%d d=%d",
                  q_rows, k_rows, d);
    NTX_CHECK_ARG(dtype == NTX_F32 || dtype == NTX_BF16 || dtype == NTX_BF16X2, "rows dtype %d not supported", dtype);
    if (sm_count <= 0) {
        int rc = cached_sm_count(&sm_count);
        if (rc) return rc;
    }
    memset(plan, 0, sizeof(*plan));
    plan->d_pad = ntx_row_pitch(d, dtype == NTX_BF16X2 ? NTX_BF16 : dtype);
    plan->tensor_cores = dtype == NTX_BF16 ? 1 : dtype == NTX_BF16X2 ? 2 : 0;
    if (dtype == NTX_BF16X2 && plan->d_pad > 128) {
        ntx_set_error("two-term rows (NTX_BF16X2) support d <= 128 (got %d)", d);
        return NTX_E_UNSUPPORTED;
    }
    if (plan->tensor_cores) {
        if (plan->d_pad > kTcMaxD) {
            ntx_set_error("bf16 rows: d=%d pads to %d > %d features per tile", d, plan->d_pad, kTcMaxD);
            return NTX_E_UNSUPPORTED;
        }
        if (plan->d_pad == 192) plan->d_pad = 256;   // kernels are built for 64 / 128 / 256
        plan->tile_rows = 128;
        plan->tile_keys = 128;
    } else {
        plan->tile_rows = 64;
        plan->tile_keys = 64;
    }
    plan->row_blocks = (q_rows + plan->tile_rows - 1) / plan->tile_rows;
    plan->key_tiles = (k_rows + plan->tile_keys - 1) / plan->tile_keys;
    const long long T = (long long)plan->row_blocks * plan->key_tiles;
    const long long ctas = (long long)sm_count * (plan->tensor_cores ? 1 : 2);
    plan->grid = (int)(T < ctas ? T : ctas);
    // A CTA whose tile range crosses a row-block boundary pays a pipeline drain and refill there (the backward's dZ
    // accumulator: ~4 us), and the kernel ends with its slowest CTA.  With a grid that is a multiple of the number
    // of row blocks the ranges i*T/G line up with the row blocks and no CTA crosses one; taken when it costs at
    // most 5 % of the CTAs (one of 8 ranks at config 4: 16 row blocks -> 144 of 148 CTAs).
    if (plan->tensor_cores && plan->row_blocks >= 2 && plan->row_blocks < ctas && T >= ctas) {
        const long long aligned = ctas / plan->row_blocks * plan->row_blocks;
        if (aligned * 100 >= ctas * 95) plan->grid = (int)aligned;
    }
    int max_slots = 1, segments = 0;
    for (int rb = 0; rb < plan->row_blocks; ++rb) {
        const int s = ntx_num_slots(T, plan->grid, plan->key_tiles, rb);
        if (s > max_slots) max_slots = s;
        segments += s;
    }
    plan->max_slots = max_slots;
    plan->segments = segments;
    const uint64_t rows_pad = (uint64_t)plan->row_blocks * plan->tile_rows;
    // forward scratch: row-sum partials + either the fold counters and share losses of the tensor-core kernel that
    // finishes its rows itself, or the finalize kernel's ticket and one fp64 block sum per finalize block
    const uint64_t fin_bytes = 16 + (((uint64_t)ntx_fwd_finalize_blocks(q_rows, ntx_row_pitch(d, dtype)) * 8 + 15) & ~15ull);
    const uint64_t fold_bytes = ntx_fold_ctl_bytes(plan->row_blocks) + ntx_fold_seg_bytes(plan->row_blocks, plan->grid);
    // fp32 rows: a slot entry may be a (max, sum) pair (temperatures below NTX_MIN_TEMPERATURE)
    plan->fwd_ws_bytes_min = (uint64_t)max_slots * rows_pad * (plan->tensor_cores ? sizeof(float) : 2 * sizeof(float)) +
                             fin_bytes + (plan->tensor_cores ? fold_bytes : 0);
    plan->fwd_ws_bytes = plan->fwd_ws_bytes_min;
    plan->bwd_ws_bytes = (uint64_t)max_slots * rows_pad * plan->d_pad * sizeof(float) +
                         (plan->tensor_cores ? ntx_fold_ctl_bytes(plan->row_blocks) : 0);

    // Symmetric forward: the keys are the rows themselves (one rank), so S = S^T and only the tiles
    // (rb, kt >= rb) need computing.  Needs one column partial per (row block, key) besides the slots.
    plan->fwd_grid = plan->grid;
    plan->fwd_max_slots = max_slots;
    if (plan->tensor_cores == 1 && q_rows == k_rows && plan->d_pad <= 128 && plan->row_blocks >= 2 &&
        getenv("NTXENT_NO_SYMMETRIC_FWD") == nullptr) {
        const int RB = plan->row_blocks;
        const long long Tt = (long long)RB * (RB + 1) / 2;
        plan->fwd_symmetric = 1;
        plan->fwd_grid = (int)(Tt < ctas ? Tt : ctas);
        int ms = 1;
        for (int rb = 0; rb < RB; ++rb) {
            const int sl = ntx_tri_num_slots(Tt, plan->fwd_grid, RB, rb);
            if (sl > ms) ms = sl;
        }
        plan->fwd_max_slots = ms;
        plan->fwd_ws_bytes = (uint64_t)ms * rows_pad * sizeof(float) + fin_bytes +
                             (uint64_t)RB * rows_pad * sizeof(float);
        if (plan->fwd_ws_bytes < plan->fwd_ws_bytes_min) plan->fwd_ws_bytes = plan->fwd_ws_bytes_min;
    }
    return NTX_OK;
}

int ntx_l2norm(const void* view0, const void* view1, int n, int d, int in_dtype, void* rows, int ld,
               int rows_dtype, float* inv_norm, float row_scale, void* stream) {
    NTX_CHECK_ARG(n >= 1 && d >= 1, "bad sizes n=%d d=%d", n, d);
    NTX_CHECK_ARG(view0 && view1 && rows && inv_norm, "NULL buffer");
    NTX_CHECK_ARG(in_dtype == NTX_F32 || in_dtype == NTX_BF16, "view dtype %d not supported", in_dtype);
    NTX_CHECK_ARG(rows_dtype == NTX_F32 || rows_dtype == NTX_BF16 || rows_dtype == NTX_BF16X2,
                  "rows dtype %d not supported", rows_dtype);
    NTX_CHECK_ARG(row_scale >= 0.f, "row_scale must not be negative");
    if (rows_dtype == NTX_BF16X2) {
        // two-term rows: [hi | lo], each half ld / 2 = d rounded up to 64 wide
        NTX_CHECK_ARG(d <= 128 && ld == ntx_row_pitch(d, NTX_BF16X2),
                      "two-term rows: d=%d must be <= 128 and ld=%d equal to ntx_row_pitch(d, NTX_BF16X2)=%d", d, ld,
                      ntx_row_pitch(d, NTX_BF16X2));
        NTX_CHECK_ARG(reinterpret_cast<uintptr_t>(rows) % 16 == 0, "rows must be 16-byte aligned");
        return ntx_l2norm_split_launch(view0, view1, n, d, in_dtype, rows, NtxPeerOut{}, 0, 0, nullptr, ld, inv_norm,
                                       row_scale > 0.f ? row_scale : 1.0f, (cudaStream_t)stream);
    }
    NTX_CHECK_ARG(ld >= ntx_row_pitch(d, rows_dtype) && ld % (rows_dtype == NTX_BF16 ? 64 : 4) == 0,
                  "ld=%d too small or misaligned for d=%d", ld, d);
    return ntx_l2norm_launch(view0, view1, n, d, in_dtype, rows, NtxPeerOut{}, 0, 0, nullptr, ld, rows_dtype, inv_norm,
                             row_scale > 0.f ? row_scale : 1.0f, (cudaStream_t)stream);
}

int ntx_project_l2norm(const void* h0, const void* h1, const void* weight, int n, int hdim, int d, void* rows,
                       int ld, float* inv_norm, float row_scale, void* stream) {
    NTX_CHECK_ARG(n >= 1 && d >= 1 && hdim >= 64, "bad sizes n=%d hdim=%d d=%d", n, hdim, d);
    NTX_CHECK_ARG(hdim % 64 == 0, "hidden size %d must be a multiple of 64 (one 128-byte TMA row of bf16)", hdim);
    NTX_CHECK_ARG(h0 && h1 && weight && rows && inv_norm, "NULL buffer");
    NTX_CHECK_ARG(d <= 256, "bf16 rows support d <= 256 (got %d)", d);
    NTX_CHECK_ARG(ld == ntx_row_pitch(d, NTX_BF16), "ld=%d must be ntx_row_pitch(d=%d, bf16)=%d", ld, d,
                  ntx_row_pitch(d, NTX_BF16));
    NTX_CHECK_ARG(reinterpret_cast<uintptr_t>(h0) % 16 == 0 && reinterpret_cast<uintptr_t>(h1) % 16 == 0 &&
                      reinterpret_cast<uintptr_t>(weight) % 16 == 0 && reinterpret_cast<uintptr_t>(rows) % 16 == 0,
                  "buffers must be 16-byte aligned");
    NTX_CHECK_ARG(row_scale >= 0.f, "row_scale must not be negative");
    return ntx_project_launch(h0, h1, weight, n, hdim, d, rows, ld, inv_norm, row_scale > 0.f ? row_scale : 1.0f,
                              (cudaStream_t)stream);
}

}  // extern "C"

namespace {

// the exchange of one call: where its results go on every rank, what it waits for, where key buffer 1 lies
struct Exchange {
    NtxPeerOut out{};
    NtxGate gate{};
    NtxPeerGeom geom{};
    size_t key_buf_elems = 0;
};

int fwd_impl(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
             const NtxLayout* layout, float temperature, float* rowsum, float* pos_logit,
             float* row_loss, float* loss_sum, void* workspace, size_t workspace_bytes, const Exchange& ex,
             void* stream) {
    NtxDims dm;
    int rc = make_dims(q_rows, k_rows, d, ld, dtype, layout, temperature, &dm);
    if (rc) return rc;
    if ((rc = aligned16(q, "q rows")) || (rc = aligned16(k, "k rows"))) return rc;
    NTX_CHECK_ARG(rowsum && pos_logit && row_loss && loss_sum, "NULL output buffer");
    NtxPlan plan;
    if ((rc = ntx_plan(q_rows, k_rows, d, dtype, 0, &plan))) return rc;
    if (workspace == nullptr || workspace_bytes < plan.fwd_ws_bytes_min) {
        ntx_set_error("ntx_fwd needs at least %llu bytes of workspace (%llu for the symmetric forward), got %llu",
                      (unsigned long long)plan.fwd_ws_bytes_min, (unsigned long long)plan.fwd_ws_bytes,
                      (unsigned long long)workspace_bytes);
        return NTX_E_WORKSPACE;
    }
    if ((rc = aligned16(workspace, "workspace"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    // the symmetric kernel needs q and k to be the same buffer and the larger scratch
    const bool sym = plan.fwd_symmetric && q == k && layout->world == 1 && workspace_bytes >= plan.fwd_ws_bytes;
    const uint64_t rows_pad = (uint64_t)plan.row_blocks * plan.tile_rows;
    if (plan.tensor_cores && !sym) {
        // full-tile tensor-core forward.  With NTXENT_FOLD=1 the kernel finishes the rows itself, each CTA of a row
        // block a share of its rows (NtxFoldScratch at the head of the scratch).  Measured on B200, eager launches of
        // one rank's shapes at config 4: 82 against 88 us per step for 1 of 8 ranks (9 CTAs per row block, 14-row
        // shares), equal for 1 of 4, 168 against 164 us for 1 of 2 and 226 against 220 us on one GPU, where the
        // shares are large and the separate, fully occupied finalize kernel is faster; 83.3 against 85.0 us per
        // step in the 8-GPU bench.  Default: the separate finalize kernel.
        const bool no_fold = !fold_wanted(plan) || dm.split;
        char* ws = (char*)workspace;
        const uint64_t head = ntx_fold_ctl_bytes(plan.row_blocks) + ntx_fold_seg_bytes(plan.row_blocks, plan.grid);
        const uint64_t partial_bytes = (uint64_t)plan.max_slots * rows_pad * sizeof(float);
        float* partial = (float*)(ws + head);
        unsigned int* counter = (unsigned int*)(ws + head + partial_bytes);
        NtxFoldScratch fold;
        fold.tag = (unsigned int*)ws;
        fold.done_total = (unsigned int*)(ws + 4);
        fold.arrive = (unsigned int*)(ws + 64);
        fold.seg_loss = (double*)(ws + ntx_fold_ctl_bytes(plan.row_blocks));
        fold.n_segments = plan.segments;
        fold.row_blocks = plan.row_blocks;
        fold.enabled = no_fold ? 0 : 1;
        fold.legacy_ticket = counter;
        const NtxFwdOut out{rowsum, pos_logit, row_loss, loss_sum, ex.out, ex.geom.rowsum_off};
        {
            ProfScope prof(0, st);
            rc = ntx_tc_fwd_launch((const __nv_bfloat16*)q, (const __nv_bfloat16*)k, dm, plan, partial, fold, out,
                                   ex.gate, ex.key_buf_elems, st);
        }
        if (rc || !no_fold) return rc;
        return ntx_fwd_finalize_launch(q, dtype, dm, plan, partial, nullptr, counter, rowsum, pos_logit,
                                       row_loss, loss_sum, ex.out, ex.geom.rowsum_off, ex.gate, ex.key_buf_elems, st);
    }
    const uint64_t partial_bytes = (uint64_t)(sym ? plan.fwd_max_slots : plan.max_slots) * rows_pad *
                                   (plan.tensor_cores ? sizeof(float) : 2 * sizeof(float));
    const uint64_t fin_bytes = 16 + (((uint64_t)ntx_fwd_finalize_blocks(q_rows, ld) * 8 + 15) & ~15ull);
    float* partial = (float*)workspace;
    unsigned int* counter = (unsigned int*)((char*)workspace + partial_bytes);
    float* colpart = sym ? (float*)((char*)workspace + partial_bytes + fin_bytes) : nullptr;
    {
        ProfScope prof(0, st);
        if (sym)
            rc = ntx_tc_fwd_sym_launch((const __nv_bfloat16*)q, dm, plan, partial, colpart, counter, st);
        else
            rc = ntx_simt_fwd_launch((const float*)q, (const float*)k, dm, plan, partial, counter, ex.gate,
                                     ex.key_buf_elems, st);
    }
    if (rc) return rc;
    return ntx_fwd_finalize_launch(q, dtype, dm, plan, partial, colpart, counter, rowsum,
                                   pos_logit, row_loss, loss_sum, ex.out, ex.geom.rowsum_off, ex.gate,
                                   ex.key_buf_elems, st);
}

}  // namespace

extern "C" {

int ntx_fwd(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
            const NtxLayout* layout, float temperature, float* rowsum, float* pos_logit,
            float* row_loss, float* loss_sum, void* workspace, size_t workspace_bytes, void* stream) {
    return fwd_impl(q, k, q_rows, k_rows, d, ld, dtype, layout, temperature, rowsum, pos_logit, row_loss, loss_sum,
                    workspace, workspace_bytes, Exchange{}, stream);
}

}  // extern "C"

static int bwd_impl(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
                    const NtxLayout* layout, float temperature, const float* rowsum_q, const float* rowsum_k,
                    const void* w_q, const void* w_k, int w_stride, int w_dtype, float w_scale, const void* view0,
                    const void* view1, int in_dtype, int in_ld, int normalised, const float* inv_norm,
                    void* dview0, void* dview1, void* workspace, size_t workspace_bytes, void* stream,
                    const Exchange& ex = Exchange{}, int w_chunk = 0) {
    NtxDims dm;
    int rc = make_dims(q_rows, k_rows, d, ld, dtype, layout, temperature, &dm);
    if (rc) return rc;
    if ((rc = aligned16(q, "q rows")) || (rc = aligned16(k, "k rows"))) return rc;
    NTX_CHECK_ARG(rowsum_q && rowsum_k && w_q && w_k && view0 && view1 && inv_norm && dview0 && dview1,
                  "NULL buffer");
    NTX_CHECK_ARG(w_stride == 0 || w_stride == 1, "w_stride must be 0 (one weight) or 1 (one per row)");
    NTX_CHECK_ARG(w_dtype == NTX_F32 || w_dtype == NTX_BF16, "weight dtype %d not supported", w_dtype);
    NTX_CHECK_ARG(in_dtype == NTX_F32 || in_dtype == NTX_BF16, "view dtype %d not supported", in_dtype);
    NtxPlan plan;
    if ((rc = ntx_plan(q_rows, k_rows, d, dtype, 0, &plan))) return rc;
    if (workspace == nullptr || workspace_bytes < plan.bwd_ws_bytes) {
        ntx_set_error("ntx_bwd needs %llu bytes of workspace, got %llu",
                      (unsigned long long)plan.bwd_ws_bytes, (unsigned long long)workspace_bytes);
        return NTX_E_WORKSPACE;
    }
    if ((rc = aligned16(workspace, "workspace"))) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const NtxWeights wt{rowsum_q, rowsum_k, (const float*)w_q, (const float*)w_k, w_stride, w_scale, w_chunk,
                        w_dtype == NTX_BF16 ? 1 : 0, w_scale};
    const int row_pitch_in = in_ld > 0 ? in_ld : d;
    if (plan.tensor_cores) {
        // the tensor-core kernel also finishes the rows (fold counters at the head of the scratch) whenever the
        // vectorised finalize applies; odd shapes keep the separate, unaligned finalize kernel
        auto al16 = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
        const bool no_fold = !fold_wanted(plan) || dm.split;
        const bool fused = !no_fold && d % 4 == 0 && row_pitch_in % 4 == 0 && al16(view0) && al16(view1) && al16(dview0) &&
                           al16(dview1);
        char* ws = (char*)workspace;
        float* partial = (float*)(ws + ntx_fold_ctl_bytes(plan.row_blocks));
        const NtxBwdOut o{view0, view1, dview0, dview1, inv_norm, (unsigned int*)ws, (unsigned int*)(ws + 4),
                          (unsigned int*)(ws + 64), plan.row_blocks, row_pitch_in, normalised,
                          in_dtype == NTX_BF16 ? 1 : 0, fused ? 1 : 0};
        {
            // with scaled rows the kernel forms e' = e * 2^c1: the factor comes out of its coefficient weights; the
            // positive-pair weight of the finish keeps the caller's scale
            ProfScope prof(1, st);
            NtxWeights wt_tc = wt;
            wt_tc.coef_scale = w_scale * dm.rowsum_scale;
            rc = ntx_tc_bwd_launch((const __nv_bfloat16*)q, (const __nv_bfloat16*)k, dm, plan, wt_tc, partial, ex.gate,
                                   ex.key_buf_elems, o, st);
        }
        if (rc || fused) return rc;
        return ntx_bwd_finalize_launch(dm, plan, partial, wt, view0, view1, in_dtype, inv_norm, dview0, dview1, st, in_ld,
                                       normalised);
    }
    float* partial = (float*)workspace;
    {
        ProfScope prof(1, st);
        rc = ntx_simt_bwd_launch((const float*)q, (const float*)k, dm, plan, wt, partial, ex.gate, ex.key_buf_elems, st);
    }
    if (rc) return rc;
    return ntx_bwd_finalize_launch(dm, plan, partial, wt, view0, view1, in_dtype, inv_norm, dview0,
                                   dview1, st, in_ld, normalised);
}

namespace {

// NtxPeerBlocks -> what the kernels need; checks the table
int make_exchange(const NtxPeerBlocks* pb, int q_rows, int ld, int dtype, const uint32_t* step, Exchange* ex,
                  const char* who) {
    if (!pb) { ntx_set_error("%s: NULL peer blocks", who); return NTX_E_ARG; }
    if (pb->world < 1 || pb->world > NTX_MAX_PEERS || pb->rank < 0 || pb->rank >= pb->world) {
        ntx_set_error("%s: world=%d rank=%d outside [1, %d] (peer memory spans one node)", who, pb->world, pb->rank,
                      NTX_MAX_PEERS);
        return NTX_E_ARG;
    }
    for (int p = 0; p < pb->world; ++p) {
        if (!pb->base[p] || reinterpret_cast<uintptr_t>(pb->base[p]) % 1024 != 0) {
            ntx_set_error("%s: block of rank %d is NULL or not 1024-byte aligned", who, p);
            return NTX_E_ARG;
        }
        ex->out.base[p] = static_cast<uint8_t*>(pb->base[p]);
    }
    if (pb->multicast && reinterpret_cast<uintptr_t>(pb->multicast) % 1024 != 0) {
        ntx_set_error("%s: multicast mapping is not 1024-byte aligned", who);
        return NTX_E_ARG;
    }
    ex->out.mc = static_cast<uint8_t*>(pb->multicast);
    if (!step) { ntx_set_error("%s: NULL step word", who); return NTX_E_ARG; }
    const int esz = (int)esz_of(dtype);
    ex->out.world = pb->world;
    ex->out.rank = pb->rank;
    ex->geom = ntx_peer_geom(q_rows, ld, esz, pb->world);
    ex->key_buf_elems = ex->geom.key_buf_bytes / esz;
    ex->gate.step = step;
    ex->gate.ctl = reinterpret_cast<const uint32_t*>(pb->base[pb->rank]);
    ex->gate.chunk_rows = q_rows;
    ex->gate.world = pb->world;
    return NTX_OK;
}

}  // namespace

extern "C" {

size_t ntx_peer_block_bytes(int q_rows, int ld, int rows_dtype, int world) {
    if (q_rows < 1 || ld < 1 || world < 1 || world > NTX_MAX_PEERS) return 0;
    return ntx_peer_geom(q_rows, ld, (int)esz_of(rows_dtype), world).total;
}

int ntx_peer_l2norm(const void* view0, const void* view1, int n, int d, int in_dtype, const NtxPeerBlocks* blocks,
                    int ld, int rows_dtype, float* inv_norm, float row_scale, uint32_t* step, void* stream) {
    NTX_CHECK_ARG(n >= 1 && d >= 1, "bad sizes n=%d d=%d", n, d);
    NTX_CHECK_ARG(view0 && view1 && inv_norm, "NULL buffer");
    NTX_CHECK_ARG(in_dtype == NTX_F32 || in_dtype == NTX_BF16, "view dtype %d not supported", in_dtype);
    NTX_CHECK_ARG(rows_dtype_ok(rows_dtype), "rows dtype %d not supported", rows_dtype);
    NTX_CHECK_ARG(rows_dtype != NTX_BF16X2 || d <= 128, "two-term rows support d <= 128 (got %d)", d);
    NTX_CHECK_ARG(ld == ntx_row_pitch(d, rows_dtype), "ld=%d must be ntx_row_pitch(d=%d)=%d", ld, d,
                  ntx_row_pitch(d, rows_dtype));
    Exchange ex;
    int rc = make_exchange(blocks, 2 * n, ld, rows_dtype, step, &ex, "ntx_peer_l2norm");
    if (rc) return rc;
    NTX_CHECK_ARG(row_scale >= 0.f, "row_scale must not be negative");
    if (rows_dtype == NTX_BF16X2)
        return ntx_l2norm_split_launch(view0, view1, n, d, in_dtype, nullptr, ex.out, ex.geom.keys_off,
                                       ex.geom.key_buf_bytes, step, ld, inv_norm, row_scale > 0.f ? row_scale : 1.0f,
                                       (cudaStream_t)stream);
    return ntx_l2norm_launch(view0, view1, n, d, in_dtype, nullptr, ex.out, ex.geom.keys_off, ex.geom.key_buf_bytes,
                             step, ld, rows_dtype, inv_norm, row_scale > 0.f ? row_scale : 1.0f, (cudaStream_t)stream);
}

int ntx_peer_fwd(const NtxPeerBlocks* blocks, int n_local, int d, int ld, int dtype, float temperature,
                 float row_scale, int loss_dtype, const uint32_t* step, float* rowsum, float* pos_logit, float* row_loss, float* loss_sum,
                 void* workspace, size_t workspace_bytes, void* stream) {
    NTX_CHECK_ARG(n_local >= 1 && d >= 1, "bad sizes n_local=%d d=%d", n_local, d);
    NTX_CHECK_ARG(rows_dtype_ok(dtype), "rows dtype %d not supported", dtype);
    const int q_rows = 2 * n_local;
    Exchange ex;
    int rc = make_exchange(blocks, q_rows, ld, dtype, step, &ex, "ntx_peer_fwd");
    if (rc) return rc;
    const NtxLayout lay{n_local, blocks->world, blocks->rank, row_scale, loss_dtype};
    const size_t esz = esz_of(dtype);
    const uint8_t* keys = ex.out.base[blocks->rank] + ex.geom.keys_off;            // key buffer 0 of this rank's block
    const uint8_t* q = keys + (size_t)blocks->rank * q_rows * ld * esz;
    return fwd_impl(q, keys, q_rows, q_rows * blocks->world, d, ld, dtype, &lay, temperature, rowsum, pos_logit, row_loss,
                    loss_sum, workspace, workspace_bytes, ex, stream);
}

int ntx_peer_post_weights(const NtxPeerBlocks* blocks, int n_local, int ld, int dtype, const uint32_t* step,
                          const void* w, int w_stride, int w_dtype, void* stream) {
    NTX_CHECK_ARG(n_local >= 1, "bad size n_local=%d", n_local);
    NTX_CHECK_ARG(rows_dtype_ok(dtype), "rows dtype %d not supported", dtype);
    NTX_CHECK_ARG(w != nullptr, "NULL row weights");
    NTX_CHECK_ARG(w_stride == 0 || w_stride == 1, "w_stride must be 0 (one weight) or 1 (one per row)");
    NTX_CHECK_ARG(w_dtype == NTX_F32 || w_dtype == NTX_BF16, "weight dtype %d not supported", w_dtype);
    Exchange ex;
    int rc = make_exchange(blocks, 2 * n_local, ld, dtype, step, &ex, "ntx_peer_post_weights");
    if (rc) return rc;
    return ntx_post_weights_launch(w, w_stride, w_dtype == NTX_BF16 ? 1 : 0, 2 * n_local, ex.out, ex.geom.weight_off,
                                   step, (cudaStream_t)stream);
}

int ntx_peer_bwd(const NtxPeerBlocks* blocks, int n_local, int d, int ld, int dtype, float temperature,
                 float row_scale, const uint32_t* step, const float* rowsum_q, int w_stride, float w_scale, const void* view0,
                 const void* view1, int in_dtype, const float* inv_norm, void* dview0, void* dview1, void* workspace,
                 size_t workspace_bytes, void* stream) {
    NTX_CHECK_ARG(n_local >= 1 && d >= 1, "bad sizes n_local=%d d=%d", n_local, d);
    NTX_CHECK_ARG(w_stride == 0 || w_stride == 1, "w_stride must be 0 (one weight) or 1 (one per row)");
    NTX_CHECK_ARG(rows_dtype_ok(dtype), "rows dtype %d not supported", dtype);
    const int q_rows = 2 * n_local;
    Exchange ex;
    int rc = make_exchange(blocks, q_rows, ld, dtype, step, &ex, "ntx_peer_bwd");
    if (rc) return rc;
    const NtxLayout lay{n_local, blocks->world, blocks->rank, row_scale, NTX_F32};
    const size_t esz = esz_of(dtype);
    const uint8_t* own = ex.out.base[blocks->rank];
    const uint8_t* keys = own + ex.geom.keys_off;
    const uint8_t* q = keys + (size_t)blocks->rank * q_rows * ld * esz;
    const float* rowsum_k = reinterpret_cast<const float*>(own + ex.geom.rowsum_off);
    const float* w_k = reinterpret_cast<const float*>(own + ex.geom.weight_off);
    // one weight per rank (posted with w_stride 0): this rank's is w_k[rank], key c's is w_k[c / q_rows];
    // one per row: this rank's rows start at w_k[rank * q_rows], key c's is w_k[c]
    const float* w_q = w_stride == 0 ? w_k + blocks->rank : w_k + (size_t)blocks->rank * q_rows;
    return bwd_impl(q, keys, q_rows, q_rows * blocks->world, d, ld, dtype, &lay, temperature, rowsum_q, rowsum_k, w_q,
                    w_k, w_stride, NTX_F32, w_scale, view0, view1, in_dtype, 0, 0, inv_norm, dview0, dview1, workspace,
                    workspace_bytes, stream, ex, w_stride == 0 ? q_rows : 0);
}

int ntx_bwd(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
            const NtxLayout* layout, float temperature, const float* rowsum_q, const float* rowsum_k,
            const void* w_q, const void* w_k, int w_stride, int w_dtype, float w_scale, const void* view0,
            const void* view1, int in_dtype, const float* inv_norm, void* dview0, void* dview1,
            void* workspace, size_t workspace_bytes, void* stream) {
    return bwd_impl(q, k, q_rows, k_rows, d, ld, dtype, layout, temperature, rowsum_q, rowsum_k, w_q, w_k, w_stride,
                    w_dtype, w_scale, view0, view1, in_dtype, 0, 0, inv_norm, dview0, dview1, workspace, workspace_bytes,
                    stream);
}

int ntx_bwd_rows(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
                 const NtxLayout* layout, float temperature, const float* rowsum_q, const float* rowsum_k,
                 const void* w_q, const void* w_k, int w_stride, int w_dtype, float w_scale, const float* inv_norm,
                 void* dz, void* workspace, size_t workspace_bytes, void* stream) {
    NTX_CHECK_ARG(q != nullptr && dz != nullptr && layout != nullptr, "NULL buffer");
    NTX_CHECK_ARG(dtype == NTX_F32 || dtype == NTX_BF16, "rows dtype %d not supported", dtype);
    const size_t esz = esz_of(dtype);
    const int n = layout->n_local;
    // the normalised rows of this rank stand in for the views: rows [0, n) and [n, 2n), pitch ld
    const char* qb = static_cast<const char*>(q);
    char* out = static_cast<char*>(dz);
    return bwd_impl(q, k, q_rows, k_rows, d, ld, dtype, layout, temperature, rowsum_q, rowsum_k, w_q, w_k, w_stride,
                    w_dtype, w_scale, qb, qb + (size_t)n * ld * esz, dtype, ld, 1, inv_norm, out, out + (size_t)n * d * esz,
                    workspace, workspace_bytes, stream);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// ntx_fwd_static: the forward with its scratch in a fixed device array owned by the library.
// For callers that cannot hand us scratch memory (no device allocator of their own, or running
// where cudaMalloc is off limits).  One array per device, shared by every call on that device.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_ntx_static_scratch[NTX_STATIC_SCRATCH_BYTES];

namespace {

// `g_ntx_static_scratch` in host code is only a handle for the runtime, not a device address:
// the address each device gave the array is looked up once per device, with that device current.
int static_scratch_ptr(void** out) {
    static std::mutex mu;
    static void* ptrs[64] = {};
    int dev = 0;
    NTX_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) {
        ntx_set_error("ntx_fwd_static: device index %d out of range", dev);
        return NTX_E_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        NTX_CUDA(cudaGetSymbolAddress(&p, g_ntx_static_scratch));   // device `dev` is current
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return NTX_OK;
}

}  // namespace

extern "C" {

int ntx_fwd_static(const void* q, const void* k, int q_rows, int k_rows, int d, int ld, int dtype,
                   const NtxLayout* layout, float temperature, float* rowsum, float* pos_logit,
                   float* row_loss, float* loss_sum, void* stream) {
    NtxPlan plan;
    int rc = ntx_plan(q_rows, k_rows, d, dtype, 0, &plan);
    if (rc) return rc;
    if (plan.fwd_ws_bytes_min > NTX_STATIC_SCRATCH_BYTES) {
        ntx_set_error("ntx_fwd_static: this problem needs %llu bytes of scratch but the built-in array "
                      "holds %u; call ntx_fwd with a caller-provided workspace instead",
                      (unsigned long long)plan.fwd_ws_bytes_min, (unsigned)NTX_STATIC_SCRATCH_BYTES);
        return NTX_E_WORKSPACE;
    }
    void* scratch = nullptr;
    if ((rc = static_scratch_ptr(&scratch))) return rc;
    // Nothing in the array has to be cleared here: every partial slot that is read was written by
    // this call's main kernel, which also resets the finalize ticket before the finalize runs.
    return ntx_fwd(q, k, q_rows, k_rows, d, ld, dtype, layout, temperature, rowsum, pos_logit, row_loss,
                   loss_sum, scratch, NTX_STATIC_SCRATCH_BYTES, stream);
}

int ntx_profile_next(int which, void* ev_start, void* ev_stop) {
    NTX_CHECK_ARG(which == 0 || which == 1, "which must be 0 (forward) or 1 (backward)");
    g_prof_start[which] = (cudaEvent_t)ev_start;
    g_prof_stop[which] = (cudaEvent_t)ev_stop;
    return NTX_OK;
}

int ntx_debug_indices(const NtxLayout* layout, int32_t* self_col, int32_t* pos_col, void* stream) {
    NTX_CHECK_ARG(layout && self_col && pos_col, "NULL argument");
    int rc = check_layout(layout, 2 * layout->n_local, 2 * layout->n_local * layout->world);
    if (rc) return rc;
    return ntx_debug_indices_launch(layout->n_local, layout->rank, self_col, pos_col,
                                    (cudaStream_t)stream);
}

}  // extern "C"
