// SyncBatchNorm, channels-last: the four streaming passes fed by a ring of bulk copies.
//
// Replaces, for [M, C] channels-last activations with C * sizeof(T) a multiple of 16 bytes and C <= 2048:
//   torch.batch_norm_stats                      (torch/nn/modules/_functions.py:39)   -> kOpStats
//   batch_norm_gather_stats_with_counts + batch_norm_elemt   (:96-122)               -> kOpApply
//   torch.batch_norm_backward_reduce            (:144)                                -> kOpBwdReduce
//   torch.batch_norm_backward_elemt             (:170)                                -> kOpBwdApply
//
// Why a ring.  A pass over a tensor of tens of megabytes is a pure HBM stream.  Written as "every thread loads a
// batch of rows, uses them, loads the next batch" the whole GPU moves in lock step and the memory system idles for
// one DRAM latency per batch (measured on B200: 3.1 TB/s of 6.5 for the statistics pass at 67 MB, ncu long-scoreboard
// stalls on the first use of every batch, profiles/r2_r2o_bn_big.md).  Here one CTA per SM owns a contiguous range of
// rows and a producer warp keeps up to eight 16 KiB `cp.async.bulk` copies per tensor in flight into a shared-memory
// ring (128-192 KiB in flight per SM at all times, no registers tied up); 16 consumer warps take the stages off the
// ring through mbarriers.  Rows are contiguous in memory, so a stage is ONE copy per tensor whatever C is, and a
// thread reads 16 bytes of a fixed channel group from every (512 / vectors-per-row)-th row: conflict-free
// shared-memory reads and a fixed set of channels per thread.
//   * statistics / backward-reduce: per-thread partial sums -> shared memory -> one float atomic per channel and CTA
//     into one of kRingReplicas accumulator copies (148 CTAs instead of the ~440 of the tiled kernels: a quarter of
//     the same-line atomic traffic that used to be a ~5 us tail) -> the last CTA (ticket) adds the copies.
//   * apply / backward-apply walk their row range back to front (what the reduce pass before them read last is what
//     the L2 still holds) and store straight from registers, 16 bytes per thread, coalesced.
#include "bn_ring.cuh"
#include "tc_ptx.cuh"

#include <stdlib.h>

namespace scb {
using namespace ptx;

constexpr int kRingConsumers = 512;
constexpr int kRingThreads = kRingConsumers + 32;       // + the producer warp
constexpr int kRingMaxStages = 8;
constexpr int kRingChunkBytes = 16384;
constexpr int kRingMaxC = 2048;
constexpr int kRingDataBudget = 192 * 1024;
// Widest column tile in 16-byte vectors.  512 = never cut a row: measured on B200, 1 KiB tiles (one bulk copy per row
// piece instead of one per stage) took [512,1024,8,8] bf16 from 110 us to 179 us per forward+backward -- small bulk
// copies cost far more than the atomics they were meant to save.  The tiling code stays for rows wider than 8 KiB.
constexpr int kRingTileVecs = 512;

enum { kOpStats = 0, kOpApply = 1, kOpBwdReduce = 2, kOpBwdApply = 3 };

// Timeline instrumentation of the resident passes (tools/trace_bn.py): only with -DSCB_TRACE=1 (tools/build_trace_lib.sh)
#ifndef SCB_TRACE
#define SCB_TRACE 0
#endif
#if SCB_TRACE
__device__ long long* g_ring_trace = nullptr;          // [CTA][8] globaltimer stamps
__device__ __forceinline__ void ring_stamp(int slot, float dep = 0.f) {
    if (g_ring_trace != nullptr && threadIdx.x == 0) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "f"(dep) : "memory");
        g_ring_trace[blockIdx.x * 8 + slot] = t;
    }
}
extern "C" int scb_debug_set_ring_trace(long long* p) {
    return cudaMemcpyToSymbol(g_ring_trace, &p, sizeof(p)) == cudaSuccess ? 0 : -1;
}
#else
#define ring_stamp(...) ((void)0)
#endif

struct RingArgs {
    const void* in[3];        // stats / apply: x (, z);   backward: x, dy (, saved y)
    void* out[2];             // apply: y;   backward apply: dx (, dz)
    int64_t M;
    int64_t rows_per_cta;
    int C, vpr, rpp, chunk_rows, n_stage;
    int ncol;                 // column tiles: a CTA owns vpr 16-byte vectors of every row (vpr * ncol = whole row)
    int pdt;                  // dtype code of weight / bias / running statistics / parameter gradients (ldp / stp)
    int pdl;                  // apply passes: launched with programmatic stream serialisation, the ring is filled
                              // BEFORE griddepcontrol.wait (the caller vouches that the inputs were complete before
                              // the previous kernel of the stream started: SCB_BN_PDL)
    int n_ranks, flags;
    float momentum, eps;
    const float *stats_all, *weight, *bias, *mean, *invstd, *sums, *n_total;
    float *running_mean, *running_var, *save_mean, *save_invstd, *save_count;
    long long* nbt;           // num_batches_tracked, incremented by the apply / resident forward pass (may be NULL)
    float* acc;               // [kRingReplicas][2C]
    unsigned* counter;
    unsigned* gbar;           // resident kernels: [arrived, left] of the grid barrier
    float *out0, *grad_weight, *grad_bias;      // out0: out_stats (statistics) / out_sums (backward reduce)
};

__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_smem), "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(bar) : "memory");
}

__device__ __forceinline__ void red_add_v4(float* p, const float4& v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(__cvta_generic_to_global(p)), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

// Ticket of the "last CTA finalises" protocol.  The CTA's reductions were issued by many threads: each fences its own,
// the barrier orders them before thread 0, and thread 0's ticket is an acq_rel atomic -- a release that is cumulative
// over everything ordered before it through the barrier (a relaxed atomicAdd behind per-thread fences is not a
// release pattern for the OTHER threads' writes), and the acquire the finalising CTA needs.
__device__ __forceinline__ unsigned ticket_acq_rel(unsigned* counter) {
    unsigned t;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], %2;" : "=r"(t) : "l"(counter), "r"(1u) : "memory");
    return t;
}
__device__ __forceinline__ bool ring_last_cta(unsigned* counter, unsigned total, unsigned* s_flag) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) *s_flag = (ticket_acq_rel(counter) == total - 1u) ? 1u : 0u;
    __syncthreads();
    return *s_flag != 0u;
}

// Merge of the ranks' (mean, M2, n) messages for channel c, or the running statistics in eval mode
// (Normalization.cuh:458-502 restated on M2; same arithmetic as merged_stats of bn.cu).
__device__ __forceinline__ void ring_merged_stats(const float* stats_all, int n_ranks, int C, int c,
                                                  const float* running_mean, const float* running_var, float eps,
                                                  float& mean, float& invstd, float& m2, float& n, int pdt) {
    if (stats_all == nullptr) {
        mean = ldp(running_mean, pdt, c, 0.f);
        invstd = rsqrtf(ldp(running_var, pdt, c, 1.f) + eps);
        m2 = 0.f; n = 1.f;
        return;
    }
    const int stride = 2 * C + 1;
    n = 0.f; mean = 0.f; m2 = 0.f;
    for (int r = 0; r < n_ranks; ++r) {
        const float* s = stats_all + (int64_t)r * stride;
        chan_merge(n, mean, m2, s[2 * C], s[c], s[C + c]);
    }
    invstd = rsqrtf(m2 / n + eps);
}

template <typename T, int kOp, int kGate, bool kAdd>
__global__ void __launch_bounds__(kRingThreads, 1)
bn_ring_kernel(const __grid_constant__ RingArgs a) {
    constexpr int VEC = Vec16<T>::N;
    constexpr int NL = (kOp == kOpStats) ? 1 : (kOp == kOpApply ? (kAdd ? 2 : 1) : (kGate == 2 ? 3 : 2));
    constexpr bool kReverse = (kOp == kOpApply || kOp == kOpBwdApply);
    constexpr bool kReduces = (kOp == kOpStats || kOp == kOpBwdReduce);
    extern __shared__ __align__(128) uint8_t smem[];
    const int C = a.C, vpr = a.vpr, rpp = a.rpp, n_stage = a.n_stage;
    const int row_bytes = vpr * 16;                       // of this CTA's column tile
    const int64_t full_row_bytes = (int64_t)row_bytes * a.ncol;
    const int tile_x = blockIdx.x % a.ncol, group = blockIdx.x / a.ncol;
    const int CW = vpr * VEC;                             // channels of the tile
    const int c_base = tile_x * CW;
    const int chunk_bytes = a.chunk_rows * row_bytes;
    uint8_t* ring = smem;                                                     // [n_stage][NL][chunk_bytes]
    uint64_t* bars = reinterpret_cast<uint64_t*>(ring + (size_t)n_stage * NL * chunk_bytes);
    unsigned* s_flag = reinterpret_cast<unsigned*>(bars + 2 * kRingMaxStages);
    float* table = reinterpret_cast<float*>(bars + 2 * kRingMaxStages + 2);   // apply: [2][CW] scale, offset
    const uint32_t bar0 = smem_u32(bars);
    auto full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto empty = [&](int s) { return bar0 + 8u * (uint32_t)(kRingMaxStages + s); };
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < n_stage; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), kRingConsumers / 32); }
        fence_barrier_init();
    }
    __syncthreads();                                   // barriers initialised: the producer may start

    const int64_t r_beg = (int64_t)group * a.rows_per_cta;
    const int64_t r_end = min(a.M, r_beg + a.rows_per_cta);
    const int n_chunks = r_end > r_beg ? (int)((r_end - r_beg + a.chunk_rows - 1) / a.chunk_rows) : 0;

    float s1[VEC], s2[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
    const int t = threadIdx.x;
    const int col = t % vpr, rip = t / vpr;
    const bool act = t < rpp * vpr;                       // consumer threads that own a column (all of them when
                                                          // vectors-per-row is a power of two)
    if (warp == kRingConsumers / 32) {
        // ------------------------------------------------------------ producer
        // whole rows (ncol == 1): a stage is one contiguous copy per tensor, issued by lane 0; column tiles: one copy
        // per row piece, the lanes share them.  With a.pdl the copies start while the previous kernel of the stream
        // (the reduce pass over the same tensors) is still in its tail: nothing the producer reads is written by it.
        if (!a.pdl) pdl_wait();
        for (int k = 0; k < n_chunks; ++k) {
            const int s = k % n_stage;
            const int kk = kReverse ? n_chunks - 1 - k : k;
            const int64_t row0 = r_beg + (int64_t)kk * a.chunk_rows;
            const int rows = (int)min((int64_t)a.chunk_rows, r_end - row0);
            if (lane == 0) {
                mbar_wait(empty(s), ((k / n_stage) & 1) ^ 1);
                mbar_arrive_expect_tx(full(s), (uint32_t)(rows * row_bytes) * NL);
            }
            __syncwarp();
            if (a.ncol == 1) {
                if (lane == 0) {
#pragma unroll
                    for (int q = 0; q < NL; ++q)
                        bulk_g2s(smem_u32(ring + ((size_t)s * NL + q) * chunk_bytes),
                                 reinterpret_cast<const uint8_t*>(a.in[q]) + row0 * row_bytes, (uint32_t)(rows * row_bytes),
                                 full(s));
                }
            } else {
                for (int i = lane; i < rows * NL; i += 32) {
                    const int q = i / rows, r = i - q * rows;
                    bulk_g2s(smem_u32(ring + ((size_t)s * NL + q) * chunk_bytes + (size_t)r * row_bytes),
                             reinterpret_cast<const uint8_t*>(a.in[q]) + (row0 + r) * full_row_bytes +
                                 (int64_t)tile_x * row_bytes,
                             (uint32_t)row_bytes, full(s));
                }
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------ consumers
        pdl_wait();                                       // the previous kernel's results (statistics / sums) are final
        if (kOp == kOpApply) {
            for (int j = t; j < CW; j += kRingConsumers) {
                const int c = c_base + j;
                float mean, invstd, m2, n;
                ring_merged_stats(a.stats_all, a.n_ranks, C, c, a.running_mean, a.running_var, a.eps, mean, invstd, m2, n, a.pdt);
                const float sc = ldp(a.weight, a.pdt, c, 1.f) * invstd;
                table[j] = sc;
                table[CW + j] = ldp(a.bias, a.pdt, c, 0.f) - mean * sc;
                if (group == 0) {
                    if (a.save_mean) a.save_mean[c] = mean;
                    if (a.save_invstd) a.save_invstd[c] = invstd;
                    if (a.stats_all != nullptr) {
                        if (a.running_mean) stp(a.running_mean, a.pdt, c, (1.f - a.momentum) * ldp(a.running_mean, a.pdt, c, 0.f) + a.momentum * mean);
                        if (a.running_var) stp(a.running_var, a.pdt, c, (1.f - a.momentum) * ldp(a.running_var, a.pdt, c, 0.f) + a.momentum * (m2 / (n - 1.f)));
                    }
                    if (c == 0 && a.save_count) a.save_count[0] = n;
                    if (c == 0 && a.nbt && a.stats_all != nullptr) a.nbt[0] += 1;
                }
            }
            named_bar_sync(1, kRingConsumers);
        }
        const int c0 = c_base + col * VEC;               // first of this thread's channels
        // per-thread channel constants
        float k0[VEC], k1[VEC], k2[VEC], k3[VEC], bb[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) { k0[i] = 0.f; k1[i] = 0.f; k2[i] = 0.f; k3[i] = 0.f; bb[i] = 0.f; }
        if (act) {
            if (kOp == kOpStats) {
                // common shift K_c = x[0, c]: S1 = sum (x - K), S2 = sum (x - K)^2 add across threads and CTAs and
                // carry no E[x^2] - E[x]^2 cancellation
                load_row_vec_cached<T, VEC>(reinterpret_cast<const T*>(a.in[0]) + c0, k0);
            } else if (kOp == kOpApply) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) { k1[i] = table[col * VEC + i]; k2[i] = table[CW + col * VEC + i]; }
            } else {
                const float inv_n = (kOp == kOpBwdApply) ? 1.f / a.n_total[0] : 0.f;
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const int c = c0 + i;
                    const float is = a.invstd[c];
                    k0[i] = a.mean[c];
                    if (kGate == 1 || kOp == kOpBwdApply) k1[i] = ldp(a.weight, a.pdt, c, 1.f) * is;
                    if (kGate == 1) bb[i] = ldp(a.bias, a.pdt, c, 0.f);
                    if (kOp == kOpBwdApply) {
                        k2[i] = a.sums[c] * inv_n;
                        k3[i] = is * is * a.sums[C + c] * inv_n;
                    }
                }
            }
        }
        const bool relu = (a.flags & SCB_BN_RELU) != 0;
        for (int k = 0; k < n_chunks; ++k) {
            const int s = k % n_stage;
            const int kk = kReverse ? n_chunks - 1 - k : k;
            const int64_t row0 = r_beg + (int64_t)kk * a.chunk_rows;
            const int rows = (int)min((int64_t)a.chunk_rows, r_end - row0);
            mbar_wait(full(s), (k / n_stage) & 1);
            const uint8_t* st0 = ring + (size_t)s * NL * chunk_bytes;
            if (act) {
#pragma unroll 2
                for (int p = rip; p < rows; p += rpp) {
                    const int off = (p * vpr + col) * 16;
                    const uint4 r0 = *reinterpret_cast<const uint4*>(st0 + off);
                    const T* e0 = reinterpret_cast<const T*>(&r0);
                    if (kOp == kOpStats) {
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            const float d = to_f32(e0[i]) - k0[i];
                            s1[i] += d;
                            s2[i] = fmaf(d, d, s2[i]);
                        }
                    } else if (kOp == kOpApply) {
                        uint4 rz = make_uint4(0u, 0u, 0u, 0u);
                        if (kAdd) rz = *reinterpret_cast<const uint4*>(st0 + chunk_bytes + off);
                        const T* ez = reinterpret_cast<const T*>(&rz);
                        uint4 o;
                        T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            float v = fmaf(to_f32(e0[i]), k1[i], k2[i]);
                            if (kAdd) v += to_f32(ez[i]);
                            oe[i] = from_f32<T>(relu ? fmaxf(v, 0.f) : v);
                        }
                        stg_stream16(reinterpret_cast<uint8_t*>(a.out[0]) + (row0 + p) * full_row_bytes +
                                         (int64_t)tile_x * row_bytes + col * 16, o);
                    } else {
                        const uint4 r1 = *reinterpret_cast<const uint4*>(st0 + chunk_bytes + off);
                        const T* e1 = reinterpret_cast<const T*>(&r1);
                        uint4 r2 = make_uint4(0u, 0u, 0u, 0u);
                        if (kGate == 2) r2 = *reinterpret_cast<const uint4*>(st0 + 2 * chunk_bytes + off);
                        const T* e2 = reinterpret_cast<const T*>(&r2);
                        uint4 o, oz;
                        T* oe = reinterpret_cast<T*>(&o);
                        T* oze = reinterpret_cast<T*>(&oz);
#pragma unroll
                        for (int i = 0; i < VEC; ++i) {
                            const float xm = to_f32(e0[i]) - k0[i];
                            float g = to_f32(e1[i]);
                            bool off_gate = false;
                            if (kGate == 1) off_gate = !(fmaf(xm, k1[i], bb[i]) > 0.f);
                            if (kGate == 2) off_gate = !(to_f32(e2[i]) > 0.f);
                            if (off_gate) g = 0.f;
                            if (kOp == kOpBwdReduce) {
                                s1[i] += g;
                                s2[i] = fmaf(g, xm, s2[i]);
                            } else {
                                oe[i] = from_f32<T>((g - k2[i] - xm * k3[i]) * k1[i]);
                                if (kGate == 2) oze[i] = off_gate ? from_f32<T>(0.f) : e1[i];
                            }
                        }
                        if (kOp == kOpBwdApply) {
                            const int64_t goff = (row0 + p) * full_row_bytes + (int64_t)tile_x * row_bytes + col * 16;
                            stg_stream16(reinterpret_cast<uint8_t*>(a.out[0]) + goff, o);
                            if (kGate == 2 && a.out[1] != nullptr)
                                stg_stream16(reinterpret_cast<uint8_t*>(a.out[1]) + goff, oz);
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty(s));
        }
        if (kReduces) {
            pdl_launch_dependents();                      // rows done: the apply pass behind may take this SM as we leave
            // CTA sums: the rpp threads that share a column meet in shared memory (the ring is idle now)
            named_bar_sync(1, kRingConsumers);
            float* red = reinterpret_cast<float*>(ring);          // [rpp][2][CW]
            if (act) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    red[(rip * 2 + 0) * CW + col * VEC + i] = s1[i];
                    red[(rip * 2 + 1) * CW + col * VEC + i] = s2[i];
                }
            }
            named_bar_sync(1, kRingConsumers);
            float* my_acc = a.acc + (size_t)(group % kRingReplicas) * 2 * C;
            // four channels per thread and ONE 16-byte reduction: a quarter of the L2 atomic operations
            for (int j = 4 * t; j < 2 * CW; j += 4 * kRingConsumers) {
                const int which = j >= CW ? 1 : 0, c = j - which * CW;
                float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int r = 0; r < rpp; ++r) {
                    const float4 v = *reinterpret_cast<const float4*>(red + (r * 2 + which) * CW + c);
                    tot.x += v.x; tot.y += v.y; tot.z += v.z; tot.w += v.w;
                }
                red_add_v4(my_acc + which * C + c_base + c, tot);
            }
        }
    }
    if (!kReduces) return;
    if (ring_last_cta(a.counter, gridDim.x, s_flag)) {
        const float n = (float)a.M;
        for (int c = threadIdx.x; c < C; c += kRingThreads) {
            float u = 0.f, v = 0.f;
#pragma unroll
            for (int r = 0; r < kRingReplicas; ++r) {
                u += __ldcg(a.acc + (size_t)r * 2 * C + c);
                v += __ldcg(a.acc + (size_t)r * 2 * C + C + c);
                a.acc[(size_t)r * 2 * C + c] = 0.f;
                a.acc[(size_t)r * 2 * C + C + c] = 0.f;
            }
            if (kOp == kOpStats) {
                const float k = to_f32(reinterpret_cast<const T*>(a.in[0])[c]);
                a.out0[c] = k + u / n;
                a.out0[C + c] = fmaxf(v - u * u / n, 0.f);
            } else {
                a.out0[c] = u;
                a.out0[C + c] = v;
                if (a.grad_weight) stp(a.grad_weight, a.pdt, c, v * a.invstd[c]);
                if (a.grad_bias) stp(a.grad_bias, a.pdt, c, u);
            }
        }
        if (threadIdx.x == 0) {
            if (kOp == kOpStats) a.out0[2 * C] = n;
            *a.counter = 0u;
        }
    }
}

// =================================================================== resident single-launch passes
// For tensors whose share per SM fits shared memory (x alone: up to ~24 MB; x and dy: ~12 MB -- BASELINE cfg 2 is
// 8 MiB), statistics and normalisation are ONE launch that reads x (and dy) exactly once:
//   * one CTA per SM owns a contiguous range of rows; the producer issues the bulk copies of the WHOLE range at once
//     (57 KiB per SM forward, 114 KiB backward at cfg 2: every byte of the pass is in flight from the first
//     microsecond, fully coalesced, no registers);
//   * the consumers accumulate per-channel sums as the chunks land, reduce them inside the CTA and add them to the
//     accumulator copies with 16-byte reductions;
//   * the grid meets at a barrier in global memory; then every CTA reads the 2C totals, builds its coefficient table
//     and normalises the rows still sitting in its shared memory, storing coalesced.
// The barrier is legal because the kernel is launched COOPERATIVELY (cudaLaunchAttributeCooperative): the driver
// starts the grid only when all of its CTAs are resident at once, so no CTA can wait for one that has not been
// scheduled (round 1 spun in a plain launch; that hazard is what this attribute removes).  The wait is bounded all
// the same: a protocol bug must surface as a launch failure, never as a hung GPU.
// Against the cluster kernels of bn_cluster.cu (which need no barrier): those are forced by the cluster-size limits of
// the chip to 16-byte-wide channel slabs at C = 256, i.e. every lane of a load in a different 128-byte line.
constexpr int kResMaxChunks = 16;
constexpr int kResSmemMax = 227 * 1024;

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <typename T, bool kBwd, int kGate, bool kAdd>
__global__ void __launch_bounds__(kRingThreads, 1)
bn_resident_kernel(const __grid_constant__ RingArgs a) {
    constexpr int VEC = Vec16<T>::N;
    constexpr int NL = kBwd ? (kGate == 2 ? 3 : 2) : (kAdd ? 2 : 1);
    extern __shared__ __align__(128) uint8_t smem[];
    const int C = a.C, vpr = a.vpr, rpp = a.rpp;
    const int row_bytes = vpr * 16;
    const int chunk_bytes = a.chunk_rows * row_bytes;
    uint8_t* data = smem;                                                          // [n_stage][NL][chunk_bytes]
    float* red = reinterpret_cast<float*>(data + (size_t)a.n_stage * NL * chunk_bytes);   // [rpp][2][C]
    float* table = red + (size_t)rpp * 2 * C;                                      // [2][C]
    uint64_t* bars = reinterpret_cast<uint64_t*>(table + 2 * C);                   // full[kResMaxChunks]
    unsigned* s_flag = reinterpret_cast<unsigned*>(bars + kResMaxChunks);
    const uint32_t bar0 = smem_u32(bars);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t r_beg = (int64_t)blockIdx.x * a.rows_per_cta;
    const int64_t r_end = min(a.M, r_beg + a.rows_per_cta);
    const int n_chunks = r_end > r_beg ? (int)((r_end - r_beg + a.chunk_rows - 1) / a.chunk_rows) : 0;

    ring_stamp(0);
    if (threadIdx.x == 0) {
        for (int s = 0; s < n_chunks; ++s) mbar_init(bar0 + 8u * (uint32_t)s, 1);
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == kRingConsumers / 32) {
        if (lane == 0) {
            for (int k = 0; k < n_chunks; ++k) {
                const int64_t row0 = r_beg + (int64_t)k * a.chunk_rows;
                const uint32_t bytes = (uint32_t)(min((int64_t)a.chunk_rows, r_end - row0) * row_bytes);
                mbar_arrive_expect_tx(bar0 + 8u * (uint32_t)k, bytes * NL);
#pragma unroll
                for (int q = 0; q < NL; ++q)
                    bulk_g2s(smem_u32(data + ((size_t)k * NL + q) * chunk_bytes),
                             reinterpret_cast<const uint8_t*>(a.in[q]) + row0 * row_bytes, bytes, bar0 + 8u * (uint32_t)k);
            }
        }
        return;                                      // the data stay in shared memory: the CTA lives on in its consumers
    }

    const int t = threadIdx.x;
    const int col = t % vpr, rip = t / vpr;
    const bool act = t < rpp * vpr;
    const int c0 = col * VEC;
    float s1[VEC], s2[VEC], k0[VEC], k1[VEC], bb[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; k0[i] = 0.f; k1[i] = 0.f; bb[i] = 0.f; }
    if (act) {
        if (!kBwd) {
            load_row_vec_cached<T, VEC>(reinterpret_cast<const T*>(a.in[0]) + c0, k0);      // common shift K_c = x[0, c]
        } else {
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const int c = c0 + i;
                k0[i] = a.mean[c];
                k1[i] = ldp(a.weight, a.pdt, c, 1.f) * a.invstd[c];
                if (kGate == 1) bb[i] = ldp(a.bias, a.pdt, c, 0.f);
            }
        }
    }
    // ---- phase 1: sums, as the chunks land
    for (int k = 0; k < n_chunks; ++k) {
        const int64_t row0 = r_beg + (int64_t)k * a.chunk_rows;
        const int rows = (int)min((int64_t)a.chunk_rows, r_end - row0);
        mbar_wait(bar0 + 8u * (uint32_t)k, 0);
        const uint8_t* st0 = data + (size_t)k * NL * chunk_bytes;
        if (act) {
#pragma unroll 2
            for (int p = rip; p < rows; p += rpp) {
                const int off = (p * vpr + col) * 16;
                const uint4 r0 = *reinterpret_cast<const uint4*>(st0 + off);
                const T* e0 = reinterpret_cast<const T*>(&r0);
                if (!kBwd) {
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float d = to_f32(e0[i]) - k0[i];
                        s1[i] += d;
                        s2[i] = fmaf(d, d, s2[i]);
                    }
                } else {
                    const uint4 r1 = *reinterpret_cast<const uint4*>(st0 + chunk_bytes + off);
                    const T* e1 = reinterpret_cast<const T*>(&r1);
                    uint4 r2 = make_uint4(0u, 0u, 0u, 0u);
                    if (kGate == 2) r2 = *reinterpret_cast<const uint4*>(st0 + 2 * chunk_bytes + off);
                    const T* e2 = reinterpret_cast<const T*>(&r2);
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float xm = to_f32(e0[i]) - k0[i];
                        float g = to_f32(e1[i]);
                        if (kGate == 1 && !(fmaf(xm, k1[i], bb[i]) > 0.f)) g = 0.f;
                        if (kGate == 2 && !(to_f32(e2[i]) > 0.f)) g = 0.f;
                        s1[i] += g;
                        s2[i] = fmaf(g, xm, s2[i]);
                    }
                }
            }
        }
    }
    ring_stamp(1, s1[0] + s2[0]);                    // all chunks landed, local sums done
    // ---- CTA sums -> accumulator copies
    if (act) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            red[(rip * 2 + 0) * C + c0 + i] = s1[i];
            red[(rip * 2 + 1) * C + c0 + i] = s2[i];
        }
    }
    named_bar_sync(1, kRingConsumers);
    {
        float* my_acc = a.acc + (size_t)(blockIdx.x % kRingReplicas) * 2 * C;
        for (int j = 4 * t; j < 2 * C; j += 4 * kRingConsumers) {
            const int which = j >= C ? 1 : 0, c = j - which * C;
            float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int r = 0; r < rpp; ++r) {
                const float4 v = *reinterpret_cast<const float4*>(red + (r * 2 + which) * C + c);
                tot.x += v.x; tot.y += v.y; tot.z += v.z; tot.w += v.w;
            }
            red_add_v4(my_acc + j, tot);
        }
    }
    ring_stamp(2);                                   // reductions issued
    // ---- grid barrier (cooperative launch: every CTA is resident)
    __threadfence();
    named_bar_sync(1, kRingConsumers);
    if (t == 0) {
        ticket_acq_rel(a.gbar);                      // release, cumulative over the CTA's reductions (see ring_last_cta)
        unsigned long long t0 = 0;
        for (unsigned spins = 1; ld_acquire_gpu(a.gbar) < gridDim.x; ++spins) {
            if ((spins & 1023u) == 0u) {
                const unsigned long long now = wall_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > 2000000000ull) __trap();      // 2 s
            }
        }
    }
    named_bar_sync(1, kRingConsumers);
    ring_stamp(3);                                   // through the grid barrier
    // ---- totals -> coefficient table; CTA 0 also writes the per-channel outputs
    {
        const float n = kBwd ? a.n_total[0] : (float)a.M;
        for (int c = t; c < C; c += kRingConsumers) {
            float u = 0.f, v = 0.f;
#pragma unroll
            for (int r = 0; r < kRingReplicas; ++r) {
                u += __ldcg(a.acc + (size_t)r * 2 * C + c);
                v += __ldcg(a.acc + (size_t)r * 2 * C + C + c);
            }
            if (!kBwd) {
                const float mean = to_f32(reinterpret_cast<const T*>(a.in[0])[c]) + u / n;
                const float m2 = fmaxf(v - u * u / n, 0.f);
                const float invstd = rsqrtf(m2 / n + a.eps);
                const float sc = ldp(a.weight, a.pdt, c, 1.f) * invstd;
                table[c] = sc;
                table[C + c] = ldp(a.bias, a.pdt, c, 0.f) - mean * sc;
                if (blockIdx.x == 0) {
                    if (a.save_mean) a.save_mean[c] = mean;
                    if (a.save_invstd) a.save_invstd[c] = invstd;
                    if (a.running_mean) stp(a.running_mean, a.pdt, c, (1.f - a.momentum) * ldp(a.running_mean, a.pdt, c, 0.f) + a.momentum * mean);
                    if (a.running_var) stp(a.running_var, a.pdt, c, (1.f - a.momentum) * ldp(a.running_var, a.pdt, c, 0.f) + a.momentum * (m2 / (n - 1.f)));
                    if (c == 0 && a.save_count) a.save_count[0] = n;
                    if (c == 0 && a.nbt) a.nbt[0] += 1;
                }
            } else {
                const float is = a.invstd[c];
                table[c] = u / n;
                table[C + c] = is * is * v / n;
                if (blockIdx.x == 0) {
                    if (a.out0) { a.out0[c] = u; a.out0[C + c] = v; }
                    if (a.grad_weight) stp(a.grad_weight, a.pdt, c, v * is);
                    if (a.grad_bias) stp(a.grad_bias, a.pdt, c, u);
                }
            }
        }
    }
    named_bar_sync(1, kRingConsumers);
    ring_stamp(4);                                   // table built
    // ---- phase 2: normalise the resident rows
    float q2[VEC], q3[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { q2[i] = act ? table[c0 + i] : 0.f; q3[i] = act ? table[C + c0 + i] : 0.f; }
    const bool relu = (a.flags & SCB_BN_RELU) != 0;
    for (int k = 0; k < n_chunks; ++k) {
        const int64_t row0 = r_beg + (int64_t)k * a.chunk_rows;
        const int rows = act ? (int)min((int64_t)a.chunk_rows, r_end - row0) : 0;
        const uint8_t* st0 = data + (size_t)k * NL * chunk_bytes;
#pragma unroll 2
        for (int p = rip; p < rows; p += rpp) {
            const int off = (p * vpr + col) * 16;
            const int64_t goff = (row0 + p) * row_bytes + col * 16;
            const uint4 r0 = *reinterpret_cast<const uint4*>(st0 + off);
            const T* e0 = reinterpret_cast<const T*>(&r0);
            uint4 o, oz;
            T* oe = reinterpret_cast<T*>(&o);
            T* oze = reinterpret_cast<T*>(&oz);
            if (!kBwd) {
                uint4 rz = make_uint4(0u, 0u, 0u, 0u);
                if (kAdd) rz = *reinterpret_cast<const uint4*>(st0 + chunk_bytes + off);
                const T* ez = reinterpret_cast<const T*>(&rz);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float v = fmaf(to_f32(e0[i]), q2[i], q3[i]);
                    if (kAdd) v += to_f32(ez[i]);
                    oe[i] = from_f32<T>(relu ? fmaxf(v, 0.f) : v);
                }
                stg_stream16(reinterpret_cast<uint8_t*>(a.out[0]) + goff, o);
            } else {
                const uint4 r1 = *reinterpret_cast<const uint4*>(st0 + chunk_bytes + off);
                const T* e1 = reinterpret_cast<const T*>(&r1);
                uint4 r2 = make_uint4(0u, 0u, 0u, 0u);
                if (kGate == 2) r2 = *reinterpret_cast<const uint4*>(st0 + 2 * chunk_bytes + off);
                const T* e2 = reinterpret_cast<const T*>(&r2);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float xm = to_f32(e0[i]) - k0[i];
                    float g = to_f32(e1[i]);
                    bool off_gate = false;
                    if (kGate == 1) off_gate = !(fmaf(xm, k1[i], bb[i]) > 0.f);
                    if (kGate == 2) off_gate = !(to_f32(e2[i]) > 0.f);
                    if (off_gate) g = 0.f;
                    oe[i] = from_f32<T>((g - q2[i] - xm * q3[i]) * k1[i]);
                    if (kGate == 2) oze[i] = off_gate ? from_f32<T>(0.f) : e1[i];
                }
                stg_stream16(reinterpret_cast<uint8_t*>(a.out[0]) + goff, o);
                if (kGate == 2 && a.out[1] != nullptr) stg_stream16(reinterpret_cast<uint8_t*>(a.out[1]) + goff, oz);
            }
        }
    }
    ring_stamp(5);
    // ---- leave (off the critical path): the last CTA through re-arms the scratch; everybody read the totals long ago
    if (t == 0) *s_flag = (atomicAdd(a.gbar + 1, 1u) == gridDim.x - 1u) ? 1u : 0u;
    named_bar_sync(1, kRingConsumers);
    if (*s_flag != 0u) {
        for (int j = t; j < kRingReplicas * 2 * C; j += kRingConsumers) a.acc[j] = 0.f;
        if (t == 0) { a.gbar[0] = 0u; a.gbar[1] = 0u; }
    }
}

// ------------------------------------------------------------------ host side
static bool ring_enabled() {
    static const bool on = [] { const char* e = getenv("SCB_BN_RING"); return e == nullptr || e[0] != '0'; }();
    return on;
}

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// false: the tensor does not fit the scheme (row not a whole number of 16-byte vectors, too wide, or too few rows to be
// worth a persistent grid)
static bool ring_geom(int64_t M, int C, int es, int n_loads, size_t table_bytes, RingArgs* a, int* grid, size_t* smem) {
    if (!ring_enabled()) return false;
    const int64_t row_bytes = (int64_t)C * es;
    if (C > kRingMaxC || row_bytes % 16 != 0 || row_bytes / 16 > kRingConsumers || M < 1) return false;
    // rows wider than kRingTileVecs vectors are cut into column tiles (see kRingTileVecs)
    int vpr = (int)(row_bytes / 16), ncol = 1;
    while (vpr > kRingTileVecs && vpr % 2 == 0) { vpr /= 2; ncol *= 2; }
    const int64_t tile_bytes = (int64_t)vpr * 16;
    const int rpp = kRingConsumers / vpr;
    int chunk_rows = (int)(kRingChunkBytes / tile_bytes) / rpp * rpp;
    if (chunk_rows < rpp) chunk_rows = rpp;
    const size_t chunk_bytes = (size_t)chunk_rows * tile_bytes;
    int n_stage = (int)(kRingDataBudget / (chunk_bytes * n_loads));
    if (n_stage > kRingMaxStages) n_stage = kRingMaxStages;
    if (n_stage < 2) return false;
    // the reduction scratch [rpp][2][C] floats reuses the ring
    if ((size_t)rpp * 2 * (C / ncol) * sizeof(float) > (size_t)n_stage * n_loads * chunk_bytes) return false;
    const int groups_max = sm_count() / ncol > 0 ? sm_count() / ncol : 1;
    int64_t g = ceil_div(M, (int64_t)chunk_rows);
    if (g > groups_max) g = groups_max;
    int64_t rpc = ceil_div(ceil_div(M, g), (int64_t)rpp) * rpp;
    g = ceil_div(M, rpc);
    a->M = M; a->C = C; a->vpr = vpr; a->rpp = rpp; a->chunk_rows = chunk_rows; a->n_stage = n_stage;
    a->ncol = ncol;
    a->rows_per_cta = rpc;
    *grid = (int)g * ncol;
    *smem = (size_t)n_stage * n_loads * chunk_bytes + (2 * kRingMaxStages + 2) * sizeof(uint64_t) + table_bytes;
    return true;
}

template <typename Kern>
static int ring_launch(Kern kern, RingArgs a, int grid, size_t smem, cudaStream_t stream) {
    SCB_CUDA_OK(set_max_dynamic_smem(kern, (int)smem));
    a.pdl = (a.pdl && pdl_enabled()) ? 1 : 0;
    if (a.pdl) {
        SCB_CUDA_OK(launch_pdl(kern, dim3((unsigned)grid), dim3(kRingThreads), smem, stream, a));
    } else {
        kern<<<grid, kRingThreads, smem, stream>>>(a);
        SCB_CUDA_OK(cudaGetLastError());
    }
    count_launch();
    return SCB_OK;
}

int ring_stats(const void* x, int dtype, int64_t M, int C, float* out_stats, void* ws, int64_t ws_bytes,
               cudaStream_t stream) {
    if (ws_bytes < ring_ws_bytes(C) || !al16(x)) return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    const int es = dtype == SCB_F32 ? 4 : 2;
    if (!ring_geom(M, C, es, 1, 0, &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x;
    a.acc = (float*)ws;
    a.counter = (unsigned*)((char*)ws + (size_t)kRingReplicas * 2 * C * 4);
    a.out0 = out_stats;
    SCB_DISPATCH_FLOAT(dtype, T, { return ring_launch(bn_ring_kernel<T, kOpStats, 0, false>, a, grid, smem, stream); });
    return SCB_OK;
}

int ring_apply(const void* x, const void* z, void* y, int dtype, int64_t M, int C, const float* stats_all, int n_ranks,
               const float* weight, const float* bias, float* running_mean, float* running_var, float momentum,
               float eps, float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags,
               int param_dtype, cudaStream_t stream) {
    const bool add = z != nullptr;
    if (!al16(x) || !al16(y) || (add && !al16(z))) return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    const int es = dtype == SCB_F32 ? 4 : 2;
    if (!ring_geom(M, C, es, add ? 2 : 1, (size_t)2 * C * sizeof(float), &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x; a.in[1] = z; a.out[0] = y;
    a.stats_all = stats_all; a.n_ranks = n_ranks; a.weight = weight; a.bias = bias;
    a.running_mean = running_mean; a.running_var = running_var; a.momentum = momentum; a.eps = eps;
    a.save_mean = save_mean; a.save_invstd = save_invstd; a.save_count = save_count; a.nbt = nbt; a.flags = flags;
    a.pdt = param_dtype;
    // the ring may fill before the statistics pass has fully retired only if that pass came before this launch in the
    // stream (training mode) and the caller vouches for the inputs; the residual operand is not covered by that
    a.pdl = ((flags & SCB_BN_PDL) && stats_all != nullptr && !add) ? 1 : 0;
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (add) return ring_launch(bn_ring_kernel<T, kOpApply, 0, true>, a, grid, smem, stream);
        return ring_launch(bn_ring_kernel<T, kOpApply, 0, false>, a, grid, smem, stream);
    });
    return SCB_OK;
}

int ring_backward_reduce(const void* dy, const void* x, const void* y, int dtype, int64_t M, int C, const float* mean,
                         const float* invstd, const float* weight, const float* bias, float* out_sums,
                         float* grad_weight, float* grad_bias, int gate, int param_dtype, void* ws, int64_t ws_bytes,
                         cudaStream_t stream) {
    if (ws_bytes < ring_ws_bytes(C) || !al16(x) || !al16(dy) || (gate == 2 && !al16(y))) return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    const int es = dtype == SCB_F32 ? 4 : 2;
    if (!ring_geom(M, C, es, gate == 2 ? 3 : 2, 0, &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x; a.in[1] = dy; a.in[2] = y;
    a.mean = mean; a.invstd = invstd; a.weight = weight; a.bias = bias;
    a.acc = (float*)ws;
    a.counter = (unsigned*)((char*)ws + (size_t)kRingReplicas * 2 * C * 4);
    a.out0 = out_sums; a.grad_weight = grad_weight; a.grad_bias = grad_bias; a.pdt = param_dtype;
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (gate == 2) return ring_launch(bn_ring_kernel<T, kOpBwdReduce, 2, false>, a, grid, smem, stream);
        if (gate == 1) return ring_launch(bn_ring_kernel<T, kOpBwdReduce, 1, false>, a, grid, smem, stream);
        return ring_launch(bn_ring_kernel<T, kOpBwdReduce, 0, false>, a, grid, smem, stream);
    });
    return SCB_OK;
}

int ring_backward_apply(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype, int64_t M, int C,
                        const float* mean, const float* invstd, const float* weight, const float* bias,
                        const float* sums, const float* n_total, int gate, int flags, int param_dtype,
                        cudaStream_t stream) {
    if (!al16(x) || !al16(dy) || !al16(dx) || (gate == 2 && !al16(y)) || (dz != nullptr && !al16(dz)))
        return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    const int es = dtype == SCB_F32 ? 4 : 2;
    if (!ring_geom(M, C, es, gate == 2 ? 3 : 2, 0, &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x; a.in[1] = dy; a.in[2] = y; a.out[0] = dx; a.out[1] = dz;
    a.mean = mean; a.invstd = invstd; a.weight = weight; a.bias = bias; a.sums = sums; a.n_total = n_total;
    a.pdl = (flags & SCB_BN_PDL) ? 1 : 0;
    a.pdt = param_dtype;
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (gate == 2) return ring_launch(bn_ring_kernel<T, kOpBwdApply, 2, false>, a, grid, smem, stream);
        if (gate == 1) return ring_launch(bn_ring_kernel<T, kOpBwdApply, 1, false>, a, grid, smem, stream);
        return ring_launch(bn_ring_kernel<T, kOpBwdApply, 0, false>, a, grid, smem, stream);
    });
    return SCB_OK;
}

// ------------------------------------------------------------------ resident passes: host side
static bool resident_geom(int64_t M, int C, int es, int n_loads, RingArgs* a, int* grid, size_t* smem) {
    if (!ring_enabled()) return false;
    const int64_t row_bytes = (int64_t)C * es;
    if (C > kRingMaxC || row_bytes % 16 != 0 || row_bytes / 16 > kRingConsumers || M < 2) return false;
    const int vpr = (int)(row_bytes / 16);
    const int rpp = kRingConsumers / vpr;
    const int sms = sm_count();
    int64_t g = ceil_div(M, (int64_t)rpp);
    if (g > sms) g = sms;
    const int64_t rpc = ceil_div(ceil_div(M, g), (int64_t)rpp) * rpp;
    g = ceil_div(M, rpc);
    int chunk_rows = (int)(kRingChunkBytes / row_bytes) / rpp * rpp;
    if (chunk_rows < rpp) chunk_rows = rpp;
    while (ceil_div(rpc, (int64_t)chunk_rows) > kResMaxChunks) chunk_rows *= 2;
    const int n_chunks = (int)ceil_div(rpc, (int64_t)chunk_rows);
    const size_t bytes = (size_t)n_chunks * n_loads * chunk_rows * row_bytes + (size_t)rpp * 2 * C * sizeof(float) +
                         (size_t)2 * C * sizeof(float) + (kResMaxChunks + 2) * sizeof(uint64_t);
    if (bytes > (size_t)kResSmemMax) return false;
    int coop = 0, dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev) != cudaSuccess ||
        !coop)
        return false;
    a->M = M; a->C = C; a->vpr = vpr; a->rpp = rpp; a->chunk_rows = chunk_rows; a->n_stage = n_chunks; a->ncol = 1;
    a->rows_per_cta = rpc;
    *grid = (int)g;
    *smem = bytes;
    return true;
}

bool ring_resident_supported(int dtype, int64_t M, int C, int which, int flags) {
    RingArgs a = {};
    int grid; size_t smem;
    const bool add = (flags & SCB_BN_ADD) != 0, relu = (flags & SCB_BN_RELU) != 0;
    const int n_loads = which == 0 ? (add ? 2 : 1) : ((relu && add) ? 3 : 2);
    return resident_geom(M, C, dtype == SCB_F32 ? 4 : 2, n_loads, &a, &grid, &smem);
}

template <typename Kern>
static int resident_launch(Kern kern, const RingArgs& a, int grid, size_t smem, cudaStream_t stream) {
    SCB_CUDA_OK(set_max_dynamic_smem(kern, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kRingThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;        // all CTAs resident at once: the in-kernel grid barrier is legal
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t err = cudaLaunchKernelEx(&cfg, kern, a);
    if (err == cudaErrorCooperativeLaunchTooLarge || err == cudaErrorLaunchOutOfResources) {
        // fewer CTAs can be resident at once than the grid needs (MPS / MIG limits, a smaller part): the barrier
        // would not be legal, the caller takes the cluster or ring passes instead
        cudaGetLastError();
        return kRingUnsupported;
    }
    SCB_CUDA_OK(err);
    count_launch();
    return SCB_OK;
}

static void resident_ws(RingArgs* a, void* ws, int C) {
    a->acc = (float*)ws;
    a->counter = (unsigned*)((char*)ws + (size_t)kRingReplicas * 2 * C * 4);
    a->gbar = a->counter + 1;
}

int ring_forward_resident(const void* x, const void* z, void* y, int dtype, int64_t M, int C, const float* weight,
                          const float* bias, float* running_mean, float* running_var, float momentum, float eps,
                          float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags,
                          int param_dtype, void* ws, int64_t ws_bytes, cudaStream_t stream) {
    const bool add = (flags & SCB_BN_ADD) != 0;
    if (ws == nullptr || ws_bytes < ring_ws_bytes(C) || !al16(x) || !al16(y) || (add && !al16(z))) return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    if (!resident_geom(M, C, dtype == SCB_F32 ? 4 : 2, add ? 2 : 1, &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x; a.in[1] = z; a.out[0] = y;
    a.weight = weight; a.bias = bias; a.running_mean = running_mean; a.running_var = running_var;
    a.momentum = momentum; a.eps = eps; a.save_mean = save_mean; a.save_invstd = save_invstd; a.save_count = save_count;
    a.nbt = nbt;
    a.pdt = param_dtype;
    a.flags = flags;
    resident_ws(&a, ws, C);
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (add) return resident_launch(bn_resident_kernel<T, false, 0, true>, a, grid, smem, stream);
        return resident_launch(bn_resident_kernel<T, false, 0, false>, a, grid, smem, stream);
    });
    return SCB_OK;
}

int ring_backward_resident(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype, int64_t M, int C,
                           const float* mean, const float* invstd, const float* weight, const float* bias,
                           const float* n_total, float* out_sums, float* grad_weight, float* grad_bias, int gate,
                           int param_dtype, void* ws, int64_t ws_bytes, cudaStream_t stream) {
    if (ws == nullptr || ws_bytes < ring_ws_bytes(C) || !al16(x) || !al16(dy) || !al16(dx) || (gate == 2 && !al16(y)) ||
        (dz != nullptr && !al16(dz)))
        return kRingUnsupported;
    RingArgs a = {};
    int grid; size_t smem;
    if (!resident_geom(M, C, dtype == SCB_F32 ? 4 : 2, gate == 2 ? 3 : 2, &a, &grid, &smem)) return kRingUnsupported;
    a.in[0] = x; a.in[1] = dy; a.in[2] = y; a.out[0] = dx; a.out[1] = dz;
    a.mean = mean; a.invstd = invstd; a.weight = weight; a.bias = bias; a.n_total = n_total;
    a.out0 = out_sums; a.grad_weight = grad_weight; a.grad_bias = grad_bias; a.pdt = param_dtype;
    resident_ws(&a, ws, C);
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (gate == 2) return resident_launch(bn_resident_kernel<T, true, 2, false>, a, grid, smem, stream);
        if (gate == 1) return resident_launch(bn_resident_kernel<T, true, 1, false>, a, grid, smem, stream);
        return resident_launch(bn_resident_kernel<T, true, 0, false>, a, grid, smem, stream);
    });
    return SCB_OK;
}

}  // namespace scb
