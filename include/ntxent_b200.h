/*
 * ntxent_b200.h -- C ABI of the B200-native NT-Xent library (libntxent_b200.so).
 *
 * The library replaces ONE path of nzw0301/SimCLR: the NT-Xent loss forward and
 * backward of `loss.py` (class NT_Xent, loss.py:4-65) plus the row-sharded
 * multi-GPU form of it.  The reference is pure Python and has no FFI of its
 * own; these entry points are what a ctypes binding inside the reference's
 * `loss.py` would bind (INTEGRATION.md shows that stub).  Each declaration
 * cites the reference lines it replaces.
 *
 * Conventions
 *  - every function returns 0 on success, a negative NTX_E_* code otherwise;
 *    ntx_last_error() returns a thread-local message for the last failure;
 *  - no C++ exception crosses the boundary, no torch type appears in a signature;
 *  - the caller owns every buffer (device memory, 16-byte aligned, row-major);
 *    the library allocates nothing, creates no stream and never synchronises
 *    the device -- all work is enqueued on `stream` (a cudaStream_t passed as
 *    void*; NULL = the legacy default stream);
 *  - callable from any host thread (PyTorch runs backward on its autograd
 *    thread, main.py:115); no thread-affine state is kept;
 *  - "views": the two (N, D) embedding matrices `view0`, `view1` of loss.py:25.
 *    "rows": the 2N L2-normalised embeddings ordered [view0 ; view1].
 *    "keys": the rows every query row is compared against -- the rank's own
 *    rows, or with W ranks the rank-major concatenation of all ranks' rows.
 */
#ifndef NTXENT_B200_H
#define NTXENT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NTX_VERSION 100 /* 0.1.0 */

/* element types of caller buffers */
#define NTX_F32 0
#define NTX_BF16 1
/* Normalised ROWS only (rows_dtype of ntx_l2norm, dtype of ntx_plan / ntx_fwd / ntx_bwd): fp32 rows held as two bf16
 * terms, x = hi + lo with hi = bf16(x), lo = bf16(x - hi) (x to 2^-17 relative).  A row is [hi (pitch/2) | lo
 * (pitch/2)] bf16 values, pitch = ntx_row_pitch(d, NTX_BF16X2) = 2 * (d rounded up to 64); d <= 128.  The tcgen05
 * kernels then form q.k = hi.lo + lo.hi + hi.hi with fp32 accumulation and split the backward's coefficient
 * tile the same way (C_hi.K_hi + C_hi.K_lo + C_lo.K_hi): fp32-grade results at tensor-core speed for callers whose
 * fp32 batches are large enough for the CUDA-core kernels to matter.  Error against the fp64 oracle, relative to
 * the largest gradient entry: 2e-7 at temperature 0.5, 1e-6 at 0.1, 1e-5 at 0.05 on random rows; 3e-6 at 0.5 and
 * 1.4e-5 at 0.1 on tightly clustered rows (cosines near 1).  It grows with 1/temperature; the module uses this layout
 * from temperature 0.35 upwards and the NTX_F32 kernels below.
 * Every entry point that takes a rows dtype accepts it (ntx_bwd_rows excepted), the peer exchange included. */
#define NTX_BF16X2 2

/* error codes */
#define NTX_OK 0
#define NTX_E_ARG (-1)       /* bad argument (shape, dtype, alignment, temperature) */
#define NTX_E_WORKSPACE (-2) /* scratch buffer missing or too small */
#define NTX_E_CUDA (-3)      /* a CUDA runtime / driver call failed */
#define NTX_E_UNSUPPORTED (-4)

/* Below this temperature the fixed-shift logsumexp of the fast kernels (exp(s - 1/tau), |cos| <= 1) would leave fp32
 * range.  fp32 rows then run a max-shifted form (two exponentials per logit in backward) and the "rowsum" vectors of
 * ntx_fwd / ntx_bwd hold lse_r = log sum_{c != r} exp(s_rc) instead of sum exp(s_rc - 1/tau); bf16 rows are refused
 * with NTX_E_UNSUPPORTED (the module switches to fp32 rows by itself).  The reference accepts any tau > 0 (loss.py:16). */
#define NTX_MIN_TEMPERATURE 0.0235f

/* Where this rank's rows sit among the keys.
 * n_local : pairs held by this rank (N of loss.py:36 when world == 1)
 * world   : number of ranks whose rows form the keys (1 = the reference's local loss)
 * rank    : this rank's chunk; its 2*n_local rows are keys [rank*2*n_local, (rank+1)*2*n_local)
 * Row r (0 <= r < 2*n_local) has self column rank*2*n_local + r (the diagonal the
 * mask at loss.py:38,46-47 removes) and positive column
 * rank*2*n_local + (r + n_local) mod 2*n_local (targets = arange(N), loss.py:37,51-52). */
typedef struct NtxLayout {
    int32_t n_local;
    int32_t world;
    int32_t rank;
    float row_scale; /* what the rows were multiplied by when they were normalised: 0 or 1 = unit rows; for bf16
                      * rows also ntx_row_scale(temperature, NTX_BF16) = sqrt(log2(e) / temperature), with which a
                      * similarity value is the exponent of its logit and the tensor-core kernels skip the scaling
                      * of every logit (about 2 % of the step; results agree to bf16 rounding) */
    int32_t loss_dtype; /* element type of the two values ntx_fwd writes to loss_sum: NTX_F32 (0, the default) or
                         * NTX_BF16 -- a bf16 loss comes out ready, without a conversion kernel behind the call */
} NtxLayout;

/* How a launch is cut into work: pure integer arithmetic, no device needed.
 * Filled by ntx_plan(); the same plan drives the main kernel and its finalize. */
typedef struct NtxPlan {
    int32_t tile_rows;     /* query rows per tile */
    int32_t tile_keys;     /* keys per tile */
    int32_t row_blocks;    /* ceil(q_rows / tile_rows) */
    int32_t key_tiles;     /* ceil(k_rows / tile_keys) */
    int32_t grid;          /* CTAs of the main kernel; CTA i owns tiles [i*T/grid, (i+1)*T/grid) */
    int32_t max_slots;     /* most CTAs that share one row block (partials per row) */
    int32_t d_pad;         /* feature dim padded to the kernel's granule (NTX_BF16X2: of one term, half the row pitch) */
    int32_t tensor_cores;  /* 1 = tcgen05 path (bf16 rows), 2 = tcgen05 path on two-term rows (NTX_BF16X2),
                            * 0 = fp32 CUDA-core path */
    uint64_t fwd_ws_bytes; /* scratch ntx_fwd wants (enables the symmetric forward when it applies) */
    uint64_t bwd_ws_bytes; /* scratch ntx_bwd needs */
    uint64_t fwd_ws_bytes_min; /* least scratch ntx_fwd accepts (full-tile forward) */
    int32_t fwd_symmetric; /* 1: keys == rows, bf16, d_pad <= 128 -> only tiles (rb, kt >= rb) are computed */
    int32_t fwd_grid;      /* CTAs of the symmetric forward kernel */
    int32_t fwd_max_slots; /* its row-sum slots per row block */
    int32_t segments;      /* (CTA, row block) pairs of the main kernels: sum over row blocks of their slot counts */
} NtxPlan;

int ntx_version(void);
const char* ntx_last_error(void);
/* Kernels this library has launched (or recorded into a CUDA graph under capture) in this process so far.
 * bench.py counts the launches of one step with it.  No counterpart in the reference. */
long long ntx_launch_count(void);

/* The factor NT_Xent's module stores into bf16 rows (see NtxLayout.row_scale): sqrt(log2(e) / temperature) for
 * NTX_BF16 / NTX_BF16X2 rows at temperature >= NTX_MIN_TEMPERATURE, 1 otherwise. */
float ntx_row_scale(float temperature, int dtype);

/* Leading dimension (elements per row) the library wants for normalised rows of
 * feature size d in dtype NTX_F32 / NTX_BF16 / NTX_BF16X2: d rounded up to 4 (fp32) or 64 (bf16), twice that for
 * two-term rows (bf16 elements). */
int ntx_row_pitch(int d, int dtype);

/* Integer planning shared by ntx_fwd/ntx_bwd and their finalize kernels.
 * sm_count <= 0 means "ask the current device". */
int ntx_plan(int q_rows, int k_rows, int d, int dtype, int sm_count, NtxPlan* plan);

/* L2-normalise both views into one [2n x ld] row buffer: rows [0,n) from view0,
 * rows [n,2n) from view1, zero padded from d to ld.  Replaces the two
 * F.normalize calls at loss.py:33-34 (x / max(||x||_2, 1e-12)).
 * inv_norm[r] = 1 / max(||row r||, 1e-12).  view pointers are (n x d), contiguous.
 * row_scale (0 = 1): the rows are stored multiplied by it; pass the same value in NtxLayout.row_scale later. */
int ntx_l2norm(const void* view0, const void* view1, int n, int d, int in_dtype,
               void* rows, int ld, int rows_dtype, float* inv_norm, float row_scale, void* stream);

/* ntx_l2norm with the projection head's last layer in front of it:
 *   rows = normalize([h0 ; h1] . weight^T),  h0/h1: (n x hdim) bf16, weight: (d x hdim) bf16 as nn.Linear stores it.
 * Replaces Linear(h -> d, bias=False) of the reference's head (model.py:69) followed by the two F.normalize calls
 * (loss.py:33-34); the (2n x d) projections are never written, only their normalised bf16 rows ([2n x ld],
 * ld = ntx_row_pitch(d, NTX_BF16)) and inv_norm.  hdim must be a multiple of 64, d <= 256.  One tcgen05 GEMM tile
 * per 128 rows with the normalisation as its epilogue. */
int ntx_project_l2norm(const void* h0, const void* h1, const void* weight, int n, int hdim, int d, void* rows,
                       int ld, float* inv_norm, float row_scale, void* stream);

/* Multi-rank exchange without a collective library call (one node, NVLink / NVSwitch peer memory).
 * Every rank owns one EXCHANGE BLOCK of ntx_peer_block_bytes() bytes that is mapped into every process
 * (e.g. torch.distributed._symmetric_memory): control words, the row sums and row weights of every key, and two
 * key buffers [world * 2n x ld] used by alternate steps.  NtxPeerBlocks.base[p] is rank p's block as THIS
 * process sees it (base[rank] = its own).  Blocks must start zeroed (on every rank, before the first call of any
 * rank), be 1024-byte aligned, and every rank must make the same sequence of calls.
 * The kernels themselves move the data and order it: a producer stores into every rank's block, fences at system
 * scope and writes the step number into a flag word of every block; a consumer polls the flag words of its own
 * block before it first touches the data of a rank.  No barrier kernel, no host synchronisation.
 *   ntx_peer_l2norm : ntx_l2norm whose rows land in rows [rank*2n, (rank+1)*2n) of the step's key buffer of every
 *                     block.  Replaces F.normalize (loss.py:33-34) followed by the cross-rank embedding gather of
 *                     BASELINE.json's north_star (the reference itself has no gather: distributed_utils.py:8-24
 *                     is only init_ddp / cleanup).  Counts this rank's steps (a word in its block; the count picks
 *                     the key buffer, so replaying a captured CUDA graph alternates buffers) and writes the
 *                     step number to *step (device memory), which the later calls of the same step take.
 *   ntx_peer_fwd    : ntx_fwd of this rank's rows against all keys of the step (loss.py:36-61 on the gathered
 *                     batch); its finalize also stores the row sums into every block.
 *   ntx_peer_post_weights : stores this rank's row weights (w: the upstream gradient, one value or one per row
 *                     as in ntx_bwd, unscaled) into every block: the first call of a backward.
 *   ntx_peer_bwd    : ntx_bwd that reads every key's row sum and weight from this rank's own block (w_stride as
 *                     posted, on every rank alike; w_scale as in ntx_bwd); call it after ntx_peer_post_weights
 *                     of the same step.
 * A rank may run ahead of another by less than one step: its forward waits for every rank's rows of the same
 * step, which a rank sends only after its previous backward.  At most two forwards may be outstanding without
 * their backward (two key buffers). */
#define NTX_MAX_PEERS 8
typedef struct NtxPeerBlocks {
    void* base[NTX_MAX_PEERS];
    void* multicast; /* optional: multicast (NVSwitch / NVLS) mapping of all the blocks, or NULL.  With it a producer
                      * issues ONE multimem.st per datum and the switch replicates it into every rank's block, instead
                      * of one store per rank over NVLink */
    int32_t world;
    int32_t rank;
} NtxPeerBlocks;
size_t ntx_peer_block_bytes(int q_rows, int ld, int rows_dtype, int world);
int ntx_peer_l2norm(const void* view0, const void* view1, int n, int d, int in_dtype, const NtxPeerBlocks* blocks,
                    int ld, int rows_dtype, float* inv_norm, float row_scale, uint32_t* step, void* stream);
int ntx_peer_fwd(const NtxPeerBlocks* blocks, int n_local, int d, int ld, int dtype, float temperature,
                 float row_scale, int loss_dtype, const uint32_t* step, float* rowsum, float* pos_logit, float* row_loss, float* loss_sum,
                 void* workspace, size_t workspace_bytes, void* stream);
int ntx_peer_post_weights(const NtxPeerBlocks* blocks, int n_local, int ld, int dtype, const uint32_t* step,
                          const void* w, int w_stride, int w_dtype, void* stream);
int ntx_peer_bwd(const NtxPeerBlocks* blocks, int n_local, int d, int ld, int dtype, float temperature,
                 float row_scale, const uint32_t* step, const float* rowsum_q, int w_stride, float w_scale, const void* view0,
                 const void* view1, int in_dtype, const float* inv_norm, void* dview0, void* dview1, void* workspace,
                 size_t workspace_bytes, void* stream);

/* Forward.  For each of this rank's q_rows = 2*n_local rows:
 *   rowsum[r]   = sum over keys c != self of exp(s_rc - 1/tau),  s = <row r, key c> / tau
 *   pos_logit[r]= s at the positive column
 *   row_loss[r] = log(rowsum[r]) + 1/tau - pos_logit[r]     (layout (2, n_local): loss.py:55-58)
 *   loss_sum[0] = sum of row_loss                             (loss.py:60-61)
 *   loss_sum[1] = loss_sum[0] / q_rows                        (loss.py:65, "mean")
 *                 (two floats, or two bf16 values when layout->loss_dtype == NTX_BF16)
 * Replaces loss.py:36-61: the three similarity GEMMs (:42,43,49), the boolean-mask
 * diagonal removal (:38,46-47), the concatenations (:51-52) and both cross-entropies
 * (:56-61); the 2N x 2N logits are never written to memory.
 * q_rows_ptr/k_rows_ptr: normalised rows ([q_rows x ld], [k_rows x ld]) of `dtype`.
 * Scratch: with workspace_bytes >= NtxPlan.fwd_ws_bytes and NtxPlan.fwd_symmetric the kernel exploits
 * S = S^T (keys are the rows themselves): each off-diagonal tile feeds the row sums of its rows and,
 * through a ones-vector MMA over the bf16 tile, the row sums of its columns.  With at least
 * fwd_ws_bytes_min every tile is computed.  Both are deterministic; they differ from each other in the
 * last bits (bf16 rounding of the column contributions, inside the 1e-2 contract). */
int ntx_fwd(const void* q_rows_ptr, const void* k_rows_ptr, int q_rows, int k_rows, int d, int ld,
            int dtype, const NtxLayout* layout, float temperature, float* rowsum,
            float* pos_logit, float* row_loss, float* loss_sum, void* workspace,
            size_t workspace_bytes, void* stream);

/* Backward: gradient of  sum_r w_r * row_loss_r  with respect to view0/view1 for this rank's
 * rows, including the terms in which its rows act as keys of other rows.
 *   w_r = w_q[r * w_stride] * w_scale for this rank's rows, w_c = w_k[c * w_stride] * w_scale for keys:
 *     reduction mean / sum: w_stride = 0, w_q = w_k = the upstream gradient (one device float),
 *                           w_scale = 1 / q_rows (mean, loss.py:65) or 1 (sum);
 *     reduction none      : w_stride = 1, w_q = upstream gradient laid out (2, n_local), w_scale = 1.
 *   w_dtype: NTX_F32 or NTX_BF16 -- the upstream gradient is read in the dtype the loss was returned in.
 *   rowsum_q / rowsum_k: the forward's rowsum of this rank's rows / of every key (with world > 1 the
 *   caller all-gathers rowsum and, for w_stride = 1, the weights).
 * Replaces the autograd graph of loss.py:33-65 (softmax - onehot, un-concatenate, index_put through
 * the mask, six GEMMs, normalisation Jacobian).
 * dview0/dview1 are (n_local x d) in `in_dtype`, like view0/view1. */
int ntx_bwd(const void* q_rows_ptr, const void* k_rows_ptr, int q_rows, int k_rows, int d, int ld,
            int dtype, const NtxLayout* layout, float temperature, const float* rowsum_q,
            const float* rowsum_k, const void* w_q, const void* w_k, int w_stride, int w_dtype, float w_scale,
            const void* view0, const void* view1, int in_dtype, const float* inv_norm, void* dview0,
            void* dview1, void* workspace, size_t workspace_bytes, void* stream);

/* ntx_bwd for callers that hold no un-normalised inputs (the rows came from ntx_project_l2norm): the gradient
 * with respect to the projections z = [h0 ; h1] . W^T, as one (q_rows x d) matrix dz in the rows' dtype.  The
 * normalisation Jacobian is taken from the normalised rows and inv_norm:
 *   dz_r = inv_norm_r * (g_r - zhat_r <zhat_r, g_r>),  g = d loss / d zhat.
 * The caller finishes the layer with two library GEMMs: dh = dz . W, dW = dz^T . [h0 ; h1]. */
int ntx_bwd_rows(const void* q_rows_ptr, const void* k_rows_ptr, int q_rows, int k_rows, int d, int ld, int dtype,
                 const NtxLayout* layout, float temperature, const float* rowsum_q, const float* rowsum_k,
                 const void* w_q, const void* w_k, int w_stride, int w_dtype, float w_scale, const float* inv_norm,
                 void* dz, void* workspace, size_t workspace_bytes, void* stream);

/* ntx_fwd without a caller-supplied scratch buffer.  This entry point has no counterpart in the
 * reference.  It is for hosts that hold nothing but their input and output buffers and cannot
 * obtain device scratch themselves: a plain C / ctypes caller that has no device allocator of
 * its own, or code running where cudaMalloc is not an option because it synchronises the whole
 * device and is illegal while a stream is being captured.  The partial row sums, the finalize
 * ticket and the block sums live in a fixed array inside the library instead
 * (NTX_STATIC_SCRATCH_BYTES of device memory on every GPU that loads the library, held until
 * the process exits).  Same arguments, outputs and arithmetic as ntx_fwd.
 *  - Problems whose scratch need (NtxPlan.fwd_ws_bytes_min) exceeds NTX_STATIC_SCRATCH_BYTES are
 *    refused with NTX_E_WORKSPACE and a message naming both numbers; use ntx_fwd for those.
 *  - The first call on each device looks the array's device address up once (no allocation);
 *    make that first call before any stream capture begins on that device.
 *  - All calls on a device share the one array: calls whose kernels can overlap on the GPU
 *    (different streams, no ordering between them) race on it.  Calls ordered on one stream are
 *    safe; nothing in the array needs clearing by the caller between calls.
 *  - Allocates nothing and frees nothing, like every other entry point here. */
#define NTX_STATIC_SCRATCH_BYTES (1u << 20)
int ntx_fwd_static(const void* q_rows_ptr, const void* k_rows_ptr, int q_rows, int k_rows, int d, int ld,
                   int dtype, const NtxLayout* layout, float temperature, float* rowsum,
                   float* pos_logit, float* row_loss, float* loss_sum, void* stream);

/* Backward of the layer ntx_project_l2norm fuses in front of the loss (Linear(h -> d, bias=False), model.py:69;
 * SURVEY.md section 8f, rank 1), given dz = d loss / d z from ntx_bwd_rows ((2n x d) bf16, rows [0, n) = first view):
 *   dh0 = dz[:n] . W,  dh1 = dz[n:] . W      (n x hdim, bf16)
 *   dweight = dz^T . [h0 ; h1]               (d x hdim, bf16, fp32 accumulation, deterministic)
 * One tcgen05 kernel (each dz tile feeds both products from shared memory) plus a small reduction of the per-CTA
 * dweight partials; replaces the two GEMMs autograd derives for that Linear.  d <= 128 and a multiple of 8, hdim a
 * multiple of 64, n * d a multiple of 8.  Workspace: ntx_project_bwd_workspace_bytes (sm_count <= 0: ask the device). */
size_t ntx_project_bwd_workspace_bytes(int n, int hdim, int sm_count);
int ntx_project_bwd(const void* dz, const void* h0, const void* h1, const void* weight, int n, int hdim, int d,
                    void* dh0, void* dh1, void* dweight, void* workspace, size_t workspace_bytes, void* stream);

/* Fused LARC + SGD-momentum optimizer step for fp32 parameters, all tensors in one launch (SURVEY.md section 8f,
 * rank 3).  Replaces optimizer.step() of main.py:116 for the optimizer built at main.py:85-94:
 * apex.parallel.LARC(torch.optim.SGD(momentum, nesterov=False), trust_coefficient=0.001, clip=False) --
 * per tensor: a = trust * ||p|| / (||g|| + wd * ||p|| + eps) when both norms are non-zero (clip != 0:
 * a = min(a / lr, 1)), g' = (g + wd * p) * a, else g' = g;  m = momentum * m + g';  p -= lr * m.
 * `table` (n_tensors entries) and `chunks` (n_chunks pairs {tensor index, chunk index within that tensor}, each
 * chunk = chunk_elems elements, listed tensor by tensor in order) live in DEVICE memory and are built once by the
 * caller; a tensor's momentum buffer must start at zero (that reproduces SGD's first step).  The workspace
 * (ntx_larc_workspace_bytes) must be zeroed once before the first call and is left ready for the next one.
 * Deterministic: chunk sums are added in a fixed order.  One cooperative launch; no host synchronisation (the
 * reference syncs twice per tensor through `norm != 0`). */
typedef struct NtxLarcParam {
    float* param;
    const float* grad;
    float* momentum;
    int64_t numel;
    float weight_decay;
    int32_t first_chunk; /* index of this tensor's first entry in `chunks` */
    int32_t n_chunks;    /* its number of entries: ceil(numel / chunk_elems) */
    int32_t reserved;
} NtxLarcParam;
size_t ntx_larc_workspace_bytes(int n_tensors, int n_chunks);
int ntx_larc_sgd_step(const NtxLarcParam* table, int n_tensors, const int32_t* chunks, int n_chunks, int chunk_elems,
                      float lr, float momentum, float trust_coefficient, float eps, int clip, void* workspace,
                      size_t workspace_bytes, void* stream);

/* SimCLR's two-view augmentation for a batch of uint8 images already on the device.
 * Replaces SimCLRTransforms.__call__ (dataset.py:42-50) applied per image by the data-loader workers:
 * RandomResizedCrop(out_size) -> RandomHorizontalFlip(0.5) -> RandomApply([ColorJitter(0.8 s, 0.8 s, 0.8 s, 0.2 s)], 0.8)
 * -> RandomGrayscale(0.2) -> ToTensor (dataset.py:19-39), drawn twice per image.
 *   images : (batch, height, width, 3) uint8, HWC as CIFAR stores them; height, width <= 64
 *   view0 / view1 : (batch, 3, out_size, out_size) float32 in [0, 1]
 *   seed, offset : the random parameters of (image i, view v) come from Philox4x32-10 keyed by seed with counter
 *                  (2 i + v, offset); pass a new offset on every call
 *   params_out : optional (batch, 2, 16) floats: top, left, h, w, flip, jitter applied, brightness, contrast,
 *                saturation, hue factors, grayscale, the order of the four adjustments; NULL to skip
 * The pixel arithmetic follows PIL's 8-bit paths (antialiased bilinear resize, Image.blend truncation, 8-bit HSV),
 * so replaying the written parameters through torchvision gives the same images up to the last bit of a few pixels. */
int ntx_simclr_augment(const void* images, int batch, int height, int width, float strength, unsigned long long seed,
                       unsigned long long offset, int out_size, float* view0, float* view1, float* params_out,
                       void* stream);

/* Test entry: the self and positive column of every row, computed on the device by
 * the same helper the kernels use (bit-exact contract on loss.py:37-38 indexing). */
int ntx_debug_indices(const NtxLayout* layout, int32_t* self_col, int32_t* pos_col, void* stream);

/* Measurement hook (bench.py): the next ntx_fwd (which = 0) or ntx_bwd (which = 1) call made
 * by this host thread records ev_start just before and ev_stop just after its MAIN kernel
 * (not the finalize kernels) on the call's stream, then disarms itself.  ev_* are
 * cudaEvent_t handles owned by the caller.  No counterpart in the reference, which has no
 * profiling hooks. */
int ntx_profile_next(int which, void* ev_start, void* ev_stop);

#ifdef __cplusplus
}
#endif
#endif /* NTXENT_B200_H */
