// Tensor-core (tcgen05 / TMEM / TMA) engine of the loss: host-side entry points.
#pragma once
#include "common.cuh"

#include <cuda.h>

namespace scb {

int64_t tc_workspace_bytes(const scb_supcon_problem* p, int n_rows);

int64_t tc_label_cache_bytes(const scb_supcon_problem* p);

// forward for the anchors among the local rows in ONE launch: the per-CTA row partials are merged inside the kernel
// (last CTA of a row block), which also writes the scalar loss; ch != nullptr sends the per-row results and the
// loss share to every rank instead (scb_supcon_forward_peers)
int tc_forward(const scb_supcon_problem* p, int row_begin, int n_rows, int n_anchor_total, float* out_lse,
               float* out_possum, float* out_npos, float* out_loss, void* workspace, const scb_peer_channel* ch,
               cudaStream_t stream);

int tc_backward(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                const float* npos_all, const float* grad_out, int n_anchor_total, float* out_dk,
                int ld_dk, void* workspace, cudaStream_t stream);

// backward + un-stack / normalise-backward / cast: one launch when a row block belongs to one or two CTAs (cluster
// pair, rows finished inside the kernel), else main kernel + fused reduce-finish
int tc_backward_features(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                         const float* grad_out, int n_anchor_total, const void* features, int feat_dtype,
                         int n_samples, int64_t stride_s, int64_t stride_v, int normalize, const float* inv_norm,
                         void* out, int64_t ostride_s, int64_t ostride_v, void* workspace, cudaStream_t stream);

// TMA descriptor of a [rows, cols] bf16 matrix (row stride ld elements), [128 x 64] boxes, 128-byte swizzle
int tc_make_tmap_2d(const void* ptr, int64_t rows, int64_t cols, int64_t ld, CUtensorMap* out);

// diagnostics (-DSCB_TRACE=1 builds): device buffer of 6 x 1024 int64 for (event id, %globaltimer) pairs of
// CTA 0's roles and the entry / exit time of every CTA; nullptr switches tracing off
void tc_set_trace(long long* p);
// measurement aid: CUDA events recorded around the backward main kernel (nullptr = off)
void tc_set_events(void* start, void* stop);

}  // namespace scb
