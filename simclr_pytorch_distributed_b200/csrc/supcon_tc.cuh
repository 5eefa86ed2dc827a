// Tensor-core (tcgen05 / TMEM / TMA) engine of the loss: host-side entry points.
#pragma once
#include "common.cuh"

namespace scb {

int64_t tc_workspace_bytes(const scb_supcon_problem* p, int n_rows);

// writes per-row partials [n_split][4][n_rows] = (max, sum exp, possum, npos) like the fp32 engine
// (slots of a row block = the CTAs whose tile span overlaps it: q tiles per CTA, ct column tiles)
int tc_forward(const scb_supcon_problem* p, int row_begin, int n_rows, float* part, int* n_split, int* q,
               int* ct, cudaStream_t stream);

int tc_backward(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                const float* npos_all, const float* grad_out, int n_anchor_total, float* out_dk,
                int ld_dk, void* workspace, cudaStream_t stream);

// backward + un-stack / normalise-backward / cast in two launches (main kernel + fused reduce-finish)
int tc_backward_features(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                         const float* grad_out, int n_anchor_total, const void* features, int feat_dtype,
                         int n_samples, int64_t stride_s, int64_t stride_v, int normalize, const float* inv_norm,
                         void* out, int64_t ostride_s, int64_t ostride_v, void* workspace, cudaStream_t stream);

// diagnostics (-DSCB_TRACE=1 builds): device buffer of 6 x 1024 int64 for (event id, %globaltimer) pairs of
// CTA 0's roles and the entry / exit time of every CTA; nullptr switches tracing off
void tc_set_trace(long long* p);
// measurement aid: CUDA events recorded around the backward main kernel (nullptr = off)
void tc_set_events(void* start, void* stop);

}  // namespace scb
