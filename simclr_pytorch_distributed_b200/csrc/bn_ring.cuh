// SyncBatchNorm, channels-last streaming passes fed by bulk copies (bn_ring.cu).  Host entry points used by the
// C-ABI functions of bn.cu; each returns SCB_OK, an error, or kRingUnsupported (the caller then takes its own kernels).
#pragma once
#include "common.cuh"

namespace scb {

constexpr int kRingUnsupported = 1;
constexpr int kRingReplicas = 4;          // copies of the per-channel accumulators (spreads the L2 atomics)

// scratch of the tiled / NCHW kernels of bn.cu: [2C] float accumulators (16-byte rounded) + tickets; the ring kernels'
// scratch follows it in the caller's workspace (scb_bn_workspace_bytes)
__host__ __device__ inline int64_t legacy_bn_ws_bytes(int C) { return ((int64_t)2 * C * 4 + 15) / 16 * 16 + 16; }

// bytes of the ring kernels' scratch (accumulator replicas + ticket); zero on entry, zero again on exit
__host__ __device__ inline int64_t ring_ws_bytes(int C) { return (int64_t)kRingReplicas * 2 * C * 4 + 16; }

int ring_stats(const void* x, int dtype, int64_t M, int C, float* out_stats, void* ws, int64_t ws_bytes,
               cudaStream_t stream);
int ring_apply(const void* x, const void* z, void* y, int dtype, int64_t M, int C, const float* stats_all, int n_ranks,
               const float* weight, const float* bias, float* running_mean, float* running_var, float momentum,
               float eps, float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags,
               int param_dtype, cudaStream_t stream);
int ring_backward_reduce(const void* dy, const void* x, const void* y, int dtype, int64_t M, int C, const float* mean,
                         const float* invstd, const float* weight, const float* bias, float* out_sums,
                         float* grad_weight, float* grad_bias, int gate, int param_dtype, void* ws, int64_t ws_bytes,
                         cudaStream_t stream);
int ring_backward_apply(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype, int64_t M, int C,
                        const float* mean, const float* invstd, const float* weight, const float* bias,
                        const float* sums, const float* n_total, int gate, int flags, int param_dtype,
                        cudaStream_t stream);

// Single-launch passes with the tensor resident in shared memory and a grid barrier (cooperative launch); the scratch
// is the ring kernels' (ring_ws_bytes).  `which`: 0 forward, 1 backward.
bool ring_resident_supported(int dtype, int64_t M, int C, int which, int flags);
int ring_forward_resident(const void* x, const void* z, void* y, int dtype, int64_t M, int C, const float* weight,
                          const float* bias, float* running_mean, float* running_var, float momentum, float eps,
                          float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags,
                          int param_dtype, void* ws, int64_t ws_bytes, cudaStream_t stream);
int ring_backward_resident(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype, int64_t M, int C,
                           const float* mean, const float* invstd, const float* weight, const float* bias,
                           const float* n_total, float* out_sums, float* grad_weight, float* grad_bias, int gate,
                           int param_dtype, void* ws, int64_t ws_bytes, cudaStream_t stream);

}  // namespace scb
