/* simclr_b200.h -- C ABI of libsimclr_b200.so: the sm_100a (B200) implementation of the two
 * data-parallel hot paths of Dyfine/SimCLR_pytorch_distributed.
 *
 *   path 1  SupConLoss.forward / autograd backward      /root/reference/losses.py:17-93
 *           + the gather / normalise glue that feeds it  /root/reference/main_supcon.py:266-286
 *   path 2  the nn.SyncBatchNorm layers of --syncBN      /root/reference/main_supcon.py:222-224
 *           (arithmetic: torch/nn/modules/_functions.py:10-209, ATen Normalization.cuh)
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in `_host`; the library never
 *     allocates, frees or keeps caller memory; scratch ("workspace") is passed in by the caller,
 *     sized by the matching *_workspace_bytes() query, and must be zero-filled ONCE when it is
 *     created (kernels leave it zeroed again when they finish);
 *   - every launch goes to the `stream` argument (a cudaStream_t passed as void*); nothing here
 *     synchronises the device or the host;
 *   - functions return SCB_OK (0) or a negative code; scb_last_error() gives the message of the
 *     last failure on the calling thread;
 *   - dtypes are the SCB_* codes below; all statistics / accumulators are float32.
 *
 * There is no CPU fallback: with no CUDA device the compute entry points fail with SCB_ERR_CUDA.
 */
#ifndef SIMCLR_B200_H_
#define SIMCLR_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SCB_OK 0
#define SCB_ERR_ARG (-1)
#define SCB_ERR_CUDA (-2)
#define SCB_ERR_UNSUPPORTED (-3)

/* element types */
#define SCB_F32 0
#define SCB_BF16 1
#define SCB_F16 2

/* positive-pair rule (losses.py:39-50) */
#define SCB_MASK_SIMCLR 0   /* labels=None, mask=None: eye(B)                 losses.py:42-43 */
#define SCB_MASK_LABELS 1   /* eq(labels, labels.T)                           losses.py:44-48 */
#define SCB_MASK_EXPLICIT 2 /* user [B,B] float mask, may be asymmetric       losses.py:49-50 */

/* contrast_mode (losses.py:54-61) */
#define SCB_MODE_ALL 0
#define SCB_MODE_ONE 1

/* arithmetic engines of the loss */
#define SCB_ENGINE_FP32 0   /* SIMT fp32 tiles: any D, any mask; the exact-mode path            */
#define SCB_ENGINE_BF16 1   /* tcgen05.mma kind::f16 (bf16 operands, fp32 TMEM accumulators)    */
#define SCB_ENGINE_BF16X3 2 /* same kernels, operands split hi+lo (3 MMAs): ~fp32-accurate      */

/* flags of the batch-norm apply kernels */
#define SCB_BN_RELU 1       /* y = max(y, 0) fused into the apply (resnet_big.py:26,58,59,111)   */
#define SCB_BN_ADD 2        /* y = bn(x) + z: the residual add of resnet_big.py:27-30,60-63;
                               with SCB_BN_RELU the ReLU comes after the add and the backward
                               gates dy by the SAVED output (y > 0) and also returns dz          */
#define SCB_BN_PDL 4        /* scb_bn_apply (training mode, no residual) and scb_bn_backward_apply only: the
                               caller vouches that the tensors the pass READS (x; x, dy, y) were complete
                               before the PREVIOUS kernel of the stream started -- true when that kernel is
                               scb_bn_stats / scb_bn_backward_reduce on the same tensors.  The pass is then
                               launched with programmatic stream serialisation and starts filling its
                               shared-memory ring while its predecessor is still in its reduction tail. */

const char* scb_last_error(void);
int scb_version(void);
/* compute capability of the current device as major*10+minor (100 on B200), negative on error */
int scb_device_cc(void);
/* id of the CUDA-graph capture `stream` is part of (0: not capturing, negative: error).  The host side of the
 * peer-memory exchanges uses it to restart its buffer rotation at every capture, so that a replayed graph never
 * re-uses a buffer its peers may still be reading (gather.py, sync_batchnorm.py) */
int64_t scb_stream_capture_id(void* stream);
/* number of kernels this library has launched in this process so far (bench bookkeeping) */
int64_t scb_launch_count(void);
/* diagnostics only, effective in builds with -DSCB_TRACE=1 (tools/build_trace_lib.sh): a device buffer of
 * 6*1024 int64 that the tensor-core kernels fill with (event id, %globaltimer ns) pairs per role of CTA 0 and
 * with the entry / exit times of every CTA (tools/trace_fwd.py, tools/trace_bwd.py read it); NULL = off */
int scb_debug_set_trace(void* device_buffer);
/* measurement aid (bench.py): two cudaEvent_t that the tensor-core backward records on its stream right
 * before and right after its main kernel, so that the kernel can be timed alone; NULL, NULL = off */
int scb_debug_set_events(void* start, void* stop);

/* ======================================================================== path 1: the loss
 *
 * Row order.  The contrast matrix K holds one embedding per row.  Rows are grouped in
 * `n_seg` segments (one per rank; 1 on a single GPU), each segment view-major:
 *      row = q*(n_views*seg) + v*seg + b        sample id = q*seg + b
 * With n_seg == 1 this is exactly the reference's `cat(unbind(features, 1))` (losses.py:53).
 * Anchors are all rows (`all`) or the v == 0 rows (`one`), losses.py:54-59.
 *
 * row_layout = SCB_ROWS_SAMPLE_MAJOR keeps the samples' own order inside a segment instead,
 *      row = q*(n_views*seg) + b*n_views + v    sample id = q*seg + b
 * which is the memory order of a contiguous `features[bsz, n_views, dim]`: a bf16 feature tensor can then
 * be handed over as K without the stacking copy (the loss is a sum over anchors, so the order of the rows
 * changes nothing but the order of floating-point additions).  `all` mode only.
 */
#define SCB_ROWS_VIEW_MAJOR 0
#define SCB_ROWS_SAMPLE_MAJOR 1

typedef struct scb_supcon_problem {
    const void* k;          /* [n_k, ldk] contrast matrix, dtype k_dtype (F32 for ENGINE_FP32, BF16 else) */
    const void* k_lo;       /* ENGINE_BF16X3 only: low-order halves, same shape (k ~= k + k_lo)   */
    const int64_t* labels;  /* [n_seg*seg] class id per sample id, or NULL                        */
    const float* mask;      /* [n_seg*seg, n_seg*seg] explicit mask, or NULL                      */
    int32_t k_dtype;
    int32_t n_k;            /* rows of k = n_seg * n_views * seg                                  */
    int32_t dim;            /* D, logical feature length                                          */
    int32_t ldk;            /* row stride of k in elements (>= dim; zero padded up to ldk)        */
    int32_t n_views;        /* V                                                                  */
    int32_t seg;            /* samples per segment                                                */
    int32_t n_seg;          /* W                                                                  */
    int32_t mask_kind;      /* SCB_MASK_*                                                         */
    int32_t mode;           /* SCB_MODE_*                                                         */
    int32_t engine;         /* SCB_ENGINE_*                                                       */
    int32_t scalar_dtype;   /* SCB_F32/BF16/F16: dtype of the loss scalar handed to autograd, i.e. of
                               out_loss[1] (forward) and of *grad_out (backward)                   */
    float inv_temperature;  /* 1/T                  losses.py:64-66                               */
    float loss_scale;       /* T / base_temperature losses.py:90                                  */
    int32_t row_layout;     /* SCB_ROWS_*                                                         */
    void* label_cache;      /* optional, SCB_MASK_LABELS on the tensor-core engines: a device buffer of
                               scb_supcon_label_cache_bytes(p) bytes that scb_supcon_forward fills with the
                               label bookkeeping (class representative / positive count per row, per-class
                               row sums) and that the backward of the SAME problem then re-uses, so that
                               the labels are scanned once per step; NULL: each pass recomputes what it
                               needs in the workspace                                                 */
} scb_supcon_problem;

/* Replaces main_supcon.py:283 (`F.normalize`, optional) + losses.py:52-53 (view-major stacking)
 * + the dtype cast: features [n_samples, n_views, dim] (element strides given) -> rows
 * [v*n_samples + b] of out_k (row stride ldk, columns dim..ldk-1 zero filled).
 * out_k_lo (BF16X3) receives bf16(x - bf16(x)); out_inv_norm [n_views*n_samples] receives
 * 1/max(||z||, eps) (1.0 when normalize == 0); both may be NULL. */
int scb_supcon_prep(const void* features, int feat_dtype, int n_samples, int n_views, int dim,
                    int64_t stride_sample, int64_t stride_view,
                    int normalize, float eps,
                    void* out_k, void* out_k_lo, int k_dtype, int ldk,
                    float* out_inv_norm, void* stream);

/* bytes of zero-initialised scratch the forward / backward of `p` need for `n_rows` local rows */
int64_t scb_supcon_workspace_bytes(const scb_supcon_problem* p, int n_rows);
/* size of p->label_cache (0 when the problem has no labels or runs on ENGINE_FP32); the label bookkeeping
 * replaces the [B, B] mask of losses.py:44-48 and the `mask.repeat` / `mask * logits_mask` of :72-80 */
int64_t scb_supcon_label_cache_bytes(const scb_supcon_problem* p);

/* Replaces losses.py:64-87 for the anchors among local rows [row_begin, row_begin+n_rows)
 * (a whole number of segments' worth of rows of ONE segment on a multi-rank run, everything on
 * one GPU).  Never materialises the [A, n_k] matrix.  Outputs, indexed by (row - row_begin):
 *   out_lse    log sum_{j != i} exp(s_ij)            (natural log, s = k_i . k_j / T)
 *   out_possum sum_j M_ij s_ij        out_npos  sum_j M_ij
 *   out_loss (TWO floats): [0] = (1/n_anchor_total) * sum over the LOCAL anchors of -(T/Tb)(possum/npos - lse)
 *                 and [1] = the same value rounded to p->scalar_dtype, stored in the low-order bytes
 *                 (losses.py:87-91; on one GPU this is the loss, across ranks the sum of the
 *                 ranks' values is)
 * Non-anchor local rows (mode `one`, v > 0) get lse = 0, npos = 1, possum = 0. */
int scb_supcon_forward(const scb_supcon_problem* p, int row_begin, int n_rows, int n_anchor_total,
                       float* out_lse, float* out_possum, float* out_npos, float* out_loss,
                       void* workspace, int64_t workspace_bytes, void* stream);

/* Replaces the autograd backward of losses.py:52-91 for the local rows: with
 *   G_ij = coef * (exp(s_ij - lse_i) [i != j] - M_ij / npos_i),  coef = *grad_out * (T/Tb) / n_anchor_total
 *   out_dk[row - row_begin, :] = (1/T) * sum_j ( [i anchor] G_ij + [j anchor] G_ji ) k_j
 * lse_all / npos_all are indexed by GLOBAL row (length n_k; gathered across ranks by the caller).
 * grad_out is a device scalar of dtype p->scalar_dtype (the upstream gradient of the loss; declared
 * float* for the common case), so no host sync and no conversion kernel is needed.
 * out_dk is float32 [n_rows, ld_dk]. */
int scb_supcon_backward(const scb_supcon_problem* p, int row_begin, int n_rows,
                        const float* lse_all, const float* npos_all, const float* grad_out,
                        int n_anchor_total, float* out_dk, int ld_dk,
                        void* workspace, int64_t workspace_bytes, void* stream);

/* Replaces the backward of main_supcon.py:283 (optional) and of losses.py:52-53: turns
 * dK rows [v*n_samples + b] into d features [n_samples, n_views, dim] of dtype feat_dtype:
 *   g = dk_row;  if normalize: g = (g - zh (zh . g)) * inv_norm,  zh = z * inv_norm. */
int scb_supcon_finish(const float* dk, int ld_dk, const void* features, int feat_dtype,
                      int n_samples, int n_views, int dim,
                      int64_t stride_sample, int64_t stride_view,
                      int normalize, const float* inv_norm,
                      void* out_dfeatures, int64_t out_stride_sample, int64_t out_stride_view,
                      void* stream);

/* Feature-norm side outputs of the training loop, /root/reference/main_supcon.py:298-317: from the
 * 1 / ||z_i|| vector that scb_supcon_prep leaves behind (n = n_views * n_samples rows of the UN-normalised
 * embeddings) computes out_terms[0..4] = norm_mean, norm_var (:300-301), loss_sec = mean (||z|| - record)^2 with
 * the moving average `record[0]` updated first (:304-309; record[1] != 0 once it holds a value), loss_l2reg =
 * mean ||z||^2 (:314), and sec_weight * loss_sec + l2reg_weight * loss_l2reg (what :311-317 add to the loss).
 * Everything stays on the device: no pass over the embeddings and none of the five .item() syncs of :327-333. */
int scb_feature_norm_terms(const float* inv_norm, int n, float* record, float momentum, float sec_weight,
                           float l2reg_weight, float* out_terms, void* stream);
/* The autograd backward of those two terms, added into d features [n_samples, n_views, dim]:
 *   d features += *grad_out * (2/n) * ( sec_weight (||z|| - record[0]) / ||z|| + l2reg_weight ) * z */
int scb_feature_norm_backward(const void* features, int feat_dtype, int n_samples, int n_views, int dim,
                              int64_t stride_sample, int64_t stride_view, const float* inv_norm,
                              const float* record, float sec_weight, float l2reg_weight,
                              const void* grad_out, int grad_dtype, void* dfeatures,
                              int64_t out_stride_sample, int64_t out_stride_view, void* stream);

/* scb_supcon_backward followed by scb_supcon_finish in one call for the tensor-core engines (two
 * launches instead of three, no [n_rows, ldk] fp32 round trip): the autograd backward of
 * losses.py:52-91 plus that of main_supcon.py:283, written straight into d features.
 * The fp32 engine returns SCB_ERR_UNSUPPORTED: use the two separate calls there. */
int scb_supcon_backward_features(const scb_supcon_problem* p, int row_begin, int n_rows,
                                 const float* lse_all, const float* npos_all, const float* grad_out,
                                 int n_anchor_total, const void* features, int feat_dtype, int n_samples,
                                 int64_t stride_sample, int64_t stride_view, int normalize,
                                 const float* inv_norm, void* out_dfeatures, int64_t out_stride_sample,
                                 int64_t out_stride_view, void* workspace, int64_t workspace_bytes, void* stream);

/* ---- fused exchanges of the row-sharded loss over peer memory (NVLink / NVSwitch) ------------------
 * Replace the two all-gathers around the loss (`gather_from_all`, main_supcon.py:268-279, and the
 * exchange of the per-row statistics the sharded backward needs) by stores issued from the PRODUCING
 * kernels themselves: the kernel that computes a rank's slice writes it into every rank's copy of the
 * gathered buffer (its own and its peers', mapped with symmetric / IPC memory), and its last CTA then
 * raises a flag in every rank's flag array.  The consumer side is scb_peer_wait, a one-warp kernel that
 * blocks the stream until the flags of all ranks show the current exchange.
 *
 * Flags hold exchange counters: `state` (local, 2 words, zeroed once) carries this rank's CTA ticket
 * and counter; every rank must issue the same sequence of exchanges on a channel.  Nothing here is
 * baked into launch parameters, so the calls can be captured in CUDA graphs.  The caller rotates >= 2
 * generations of destination buffers (a rank can run at most one exchange ahead of its slowest peer). */
typedef struct scb_peer_channel {
    void* dst[16];        /* destination buffers, index = rank (dst[my_rank] is the local copy)          */
    uint32_t* flag[16];   /* flag[w] = address, in rank w's flag array, of the word that belongs to THIS
                             rank (rank w's scb_peer_wait reads its own array)                          */
    uint32_t* state;      /* local scratch: [0] CTA ticket (self re-arming), [1] exchange counter        */
    int32_t n_dst;        /* world size W <= 16                                                          */
} scb_peer_channel;

/* scb_supcon_prep whose output rows [v*n_samples + b] go to row (row_offset + that) of EVERY
 * ch->dst[w] ([n_k, ldk] matrices of dtype k_dtype), followed by the flag signal. */
int scb_supcon_prep_peers(const void* features, int feat_dtype, int n_samples, int n_views, int dim,
                          int64_t stride_sample, int64_t stride_view, int normalize, float eps,
                          const scb_peer_channel* ch, int64_t row_offset, int k_dtype, int ldk,
                          float* out_inv_norm, void* stream);

/* scb_supcon_forward whose per-row results go to every rank: ch->dst[w] is a float array
 * [lse (n_k) | npos (n_k) | loss share of rank r at 2*n_k + 4*r], filled at this rank's rows / slot.
 * out_possum stays local. */
int scb_supcon_forward_peers(const scb_supcon_problem* p, int row_begin, int n_rows, int n_anchor_total,
                             float* out_possum, const scb_peer_channel* ch, void* workspace,
                             int64_t workspace_bytes, void* stream);

/* Blocks the stream until flags_local[w] >= state[1] for w < n (all ranks have delivered the current
 * exchange of this channel).  If loss_shares != NULL it then adds the n values loss_shares[4*w] into
 * out_loss[0] (float) and out_loss[1] (same value in scalar_dtype, low-order bytes). */
int scb_peer_wait(const uint32_t* flags_local, const uint32_t* state, int n, const float* loss_shares,
                  float* out_loss, int scalar_dtype, void* stream);

/* Projection-head tail: out = normalise(h W^T + bias) in one tcgen05 kernel.
 * Replaces `F.normalize(self.head[2](feat), dim=1)`, /root/reference/networks/resnet_big.py:168-172,177-181 (the
 * second F.normalize of main_supcon.py:283 is then the identity).  h [m, k] and w [n, k] are bf16, row strides
 * ld_h / ld_w in elements (multiples of 8); n <= 128, a multiple of 4; bias [n] fp32 or NULL.  out [m, n] in
 * out_dtype with row stride ld_out; bf16 rows with n in {64, 128} are what the loss takes in place
 * (SCB_ROWS_SAMPLE_MAJOR).  inv_norm [m] (may be NULL) receives 1 / max(||row||, eps) for the backward.
 * normalize == 0: plain Linear.  Products are bf16 x bf16 accumulated in fp32 (what autocast runs).
 * n_views > 1: the m input rows are view-major ([n_views][m / n_views], the encoder ran on cat([view0, view1]),
 * main_supcon.py:259-262) and output row b * n_views + v receives input row v * (m / n_views) + b -- the stacking of
 * main_supcon.py:285-286 done by the store, so `out` is the loss's [bsz, n_views, n] tensor.  inv_norm stays in input
 * row order. */
int scb_head_forward(const void* h, int64_t ld_h, const void* w, int64_t ld_w, const float* bias, int64_t m, int k,
                     int n, int n_views, void* out, int out_dtype, int64_t ld_out, float* inv_norm, int normalize,
                     float eps, void* stream);
/* Backward of the row normalisation: out = (g - y (g . y)) * inv_norm per row; g is the gradient with respect to the
 * normalised rows y (what scb_supcon_backward_features returns); with n_views > 1 g and y are in the sample-major
 * row order scb_head_forward wrote and `out` is back in input row order.  The two weight / input GEMMs of the Linear's
 * backward are plain library GEMMs on `out` (cuBLAS through torch.matmul in head.py). */
int scb_head_normalize_backward(const void* g, int g_dtype, int64_t ld_g, const void* y, int y_dtype, int64_t ld_y,
                                const float* inv_norm, int64_t m, int n, void* out, int out_dtype, int64_t ld_out,
                                int n_views, void* stream);

/* Peer-memory exchange of the row-sharded loss (replaces the `gather_from_all` / all_gather calls of
 * main_supcon.py:268-279 on one NVLink/NVSwitch box).  Copies up to 4 byte ranges of `src` into EVERY
 * destination buffer (the caller's own and its peers', mapped into this process with symmetric / IPC
 * memory): segment s copies seg[3s+2] bytes from src + seg[3s] to dst_ptrs[w] + seg[3s+1] for all w.
 * Pointers, offsets and sizes are multiples of 4 bytes (16-byte copies when all are multiples of 16).
 * `seg` and `dst_ptrs` are HOST arrays read during the call.  The stores travel over NVLink as plain 16-byte writes; the caller orders them against the peers'
 * reads with a cross-rank stream barrier after this kernel (gather.py uses the symmetric-memory
 * handle's barrier).  n_dst <= 16. */
int scb_peer_bcast(const void* src, const int64_t* seg, int n_seg, const void* const* dst_ptrs, int n_dst,
                   void* stream);

/* ================================================================== path 2: SyncBatchNorm
 *
 * Layout: x is [outer, C, inner] with element strides (stride_outer, 1*inner... ) described by
 * `channels_last`: 1 -> memory is [M, C] with C contiguous (NHWC; M = N*H*W),
 *                  0 -> memory is [N, C, HW] contiguous (NCHW; inner = HW).
 */

/* param_dtype (SCB_F32 / SCB_BF16 / SCB_F16) in the entry points below is the element type of weight, bias,
 * running_mean, running_var, grad_weight and grad_bias -- they are declared float* for the common case and read /
 * written in the module's own dtype, so that a model converted with `.to(torch.bfloat16)` needs no conversion kernels
 * around the passes (torch's own batch norm takes mixed dtypes the same way).  Statistics, saved mean / invstd and the
 * exchanged messages are always fp32. */

/* bytes of zero-initialised scratch for the two reduction kernels below (every kernel leaves it zeroed).  It has
 * two parts, [2C floats + tickets][4 x 2C floats + ticket]: channels-last tensors take the bulk-copy ring kernels
 * (csrc/bn_ring.cu), which use the second part; a caller that passes only the first (8 * C rounded up to 16, plus 16
 * bytes) gets the tiled kernels, same results.  The base must be 16-byte aligned. */
int64_t scb_bn_workspace_bytes(int channels);

/* Replaces torch.batch_norm_stats (_functions.py:39): per-channel statistics of this rank's
 * slice.  out_stats is the (2C+1)-float message the ranks exchange (_functions.py:41-79):
 *   out_stats[0:C] = mean, out_stats[C:2C] = M2 = sum (x-mean)^2, out_stats[2C] = count. */
int scb_bn_stats(const void* x, int dtype, int channels_last, int64_t outer, int channels,
                 int64_t inner, float* out_stats, void* workspace, int64_t workspace_bytes,
                 void* stream);

/* No counterpart in the reference.  Same result as scb_bn_stats, for callers that cannot allocate
 * device memory where they need the statistics -- e.g. a host framework whose memory pool already
 * owns the whole GPU, so that a cudaMalloc for scratch would fail or compete with it.  The scratch
 * is a fixed array built into the library (about 33 KiB of device memory on every GPU that loads
 * the library, kept until the process exits), so the entry point calls no cudaMalloc / cudaFree /
 * allocating routine; the caller still passes x and out_stats.
 *   - at most SCB_BN_NOALLOC_MAX_CHANNELS channels; more is rejected with SCB_ERR_ARG;
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress with
 *     that device current); make that first call before any stream capture;
 *   - every call uses the same array (the part it needs is cleared first, on `stream`): calls that
 *     overlap on the GPU -- different streams, or two host threads -- race with each other. */
#define SCB_BN_NOALLOC_MAX_CHANNELS 4096
int scb_bn_stats_noalloc(const void* x, int dtype, int channels_last, int64_t outer, int channels,
                         int64_t inner, float* out_stats, void* stream);

/* Replaces torch.batch_norm_gather_stats_with_counts + torch.batch_norm_elemt
 * (_functions.py:106-122; Normalization.cuh:458-502): merges `n_ranks` gathered messages
 * stats_all [n_ranks, 2C+1] (Chan, rank order, empty ranks skipped), updates running_mean /
 * running_var (momentum; unbiased global variance) when non-NULL, writes save_mean/save_invstd
 * [C], save_count[0] = global element count per channel (kept on the device so the backward
 * needs no host sync, cf. the count tensor of _functions.py:117) and
 *   y = (x - mean) * invstd * weight + bias  (+ z with SCB_BN_ADD)  (then ReLU with SCB_BN_RELU).
 * z is the residual operand of resnet_big.py:27-30,60-63 (same shape and layout as x; NULL without SCB_BN_ADD).
 * With stats_all == NULL (eval mode, batchnorm.py:791-828) running_mean / running_var are used
 * as the statistics and are not modified.  weight/bias may be NULL (affine=False). 
 * num_batches_tracked (may be NULL): the module's int64 step counter (batchnorm.py:763-764), incremented by one in
 * training mode (stats_all != NULL) by the same kernel, so that the module launches nothing else per call. */
int scb_bn_apply(const void* x, const void* z, void* y, int dtype, int channels_last, int64_t outer, int channels,
                 int64_t inner, const float* stats_all, int n_ranks,
                 const float* weight, const float* bias,
                 float* running_mean, float* running_var, float momentum, float eps,
                 float* save_mean, float* save_invstd, float* save_count, long long* num_batches_tracked,
                 int flags, int param_dtype, void* stream);

/* Replaces torch.batch_norm_backward_reduce (_functions.py:139-153):
 *   out_sums[0:C] = sum dy, out_sums[C:2C] = sum dy (x - mean)   (the 2C floats all-reduced),
 *   grad_weight = out_sums[C:2C] * invstd, grad_bias = out_sums[0:C]  (either may be NULL).
 * With SCB_BN_RELU, dy is first gated by (y > 0): y recomputed from x (weight/bias needed), or, with
 * SCB_BN_ADD as well, read from the saved output `y` (NULL otherwise). */
int scb_bn_backward_reduce(const void* dy, const void* x, const void* y, int dtype, int channels_last,
                           int64_t outer, int channels, int64_t inner,
                           const float* mean, const float* invstd,
                           const float* weight, const float* bias,
                           float* out_sums, float* grad_weight, float* grad_bias, int flags, int param_dtype,
                           void* workspace, int64_t workspace_bytes, void* stream);

/* Replaces torch.batch_norm_backward_elemt (_functions.py:170-179):
 *   dx = (dy - sum_dy/n - (x - mean) invstd^2 sum_dy_xmu/n) * weight * invstd,
 * sums = the all-reduced [2C] vector, n_total[0] = global element count per channel (device).
 * dy is gated as in scb_bn_backward_reduce; with SCB_BN_RELU|SCB_BN_ADD the gated dy is also written to dz
 * (gradient of the residual operand; may be NULL). */
int scb_bn_backward_apply(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype,
                          int channels_last, int64_t outer, int channels, int64_t inner,
                          const float* mean, const float* invstd,
                          const float* weight, const float* bias,
                          const float* sums, const float* n_total, int flags, int param_dtype, void* stream);

/* Single-launch variants (csrc/bn_cluster.cu): the same arithmetic as scb_bn_stats + scb_bn_apply, resp.
 * scb_bn_backward_reduce + scb_bn_backward_apply (torch/nn/modules/_functions.py:39-122 and :139-179 at world
 * size 1, i.e. the F.batch_norm fallback of torch/nn/modules/batchnorm.py:818-828) with x (and dy) read ONCE and
 * kept in registers.  A thread-block cluster owns a slab of channels: per-channel partial sums meet over
 * distributed shared memory and a cluster barrier -- no scratch, no global atomics, no grid-wide barrier, no
 * assumption that the whole grid is resident.  Channels-last [rows, channels] only, and only when the tensor
 * fits the register file of one resident wave: scb_bn_fused_supported(which = 0 forward / 1 backward, flags)
 * says so beforehand, otherwise they return SCB_ERR_UNSUPPORTED and the caller runs the two-kernel path.
 * With a workspace (the scb_bn_workspace_bytes scratch; may be NULL) tensors whose share per SM fits shared memory
 * take the resident passes of csrc/bn_ring.cu instead: one CTA per SM bulk-copies its contiguous rows into shared
 * memory, the per-channel sums meet in the workspace behind a grid barrier, and the rows are normalised out of shared
 * memory.  That kernel is launched cooperatively (all CTAs resident at once), which is what makes its barrier legal. */
int scb_bn_fused_supported(int dtype, int channels_last, int64_t outer, int channels, int64_t inner, int which,
                           int flags);
/* diagnostics: how many clusters of `cluster_size` CTAs of these kernels the current device holds at once */
int scb_bn_cluster_occupancy(int cluster_size);
int scb_bn_forward_fused(const void* x, const void* z, void* y, int dtype, int64_t rows, int channels,
                         const float* weight, const float* bias, float* running_mean, float* running_var,
                         float momentum, float eps, float* save_mean, float* save_invstd, float* save_count,
                         long long* num_batches_tracked, int flags, int param_dtype, void* workspace,
                         int64_t workspace_bytes, void* stream);
int scb_bn_backward_fused(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype, int64_t rows,
                          int channels, const float* mean, const float* invstd, const float* weight,
                          const float* bias, const float* n_total, float* out_sums, float* grad_weight,
                          float* grad_bias, int flags, int param_dtype, void* workspace, int64_t workspace_bytes,
                          void* stream);
/* The reductions alone through the same cluster scheme, for several ranks (the exchange of
 * _functions.py:41-79 / :155-165 sits between them and scb_bn_apply / scb_bn_backward_apply):
 * scb_bn_stats_fused writes the (2C+1)-float message of scb_bn_stats, scb_bn_backward_reduce_fused the
 * outputs of scb_bn_backward_reduce; neither needs scratch. */
int scb_bn_stats_fused(const void* x, int dtype, int64_t rows, int channels, float* out_stats, void* stream);
int scb_bn_backward_reduce_fused(const void* dy, const void* x, const void* y, int dtype, int64_t rows, int channels,
                                 const float* mean, const float* invstd, const float* weight, const float* bias,
                                 const float* n_total, float* out_sums, float* grad_weight, float* grad_bias,
                                 int flags, int param_dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SIMCLR_B200_H_ */


