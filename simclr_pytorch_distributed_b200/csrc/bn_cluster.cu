// SyncBatchNorm, single-launch passes for activations that fit the register file of one resident wave
// (BASELINE cfg 2: [256,256,8,8] bf16 = 8 MiB): statistics and normalisation in ONE kernel, x (and dy) read once.
//
// Replaces torch.batch_norm_stats + batch_norm_gather_stats_with_counts + batch_norm_elemt, resp.
// batch_norm_backward_reduce + batch_norm_backward_elemt (torch/nn/modules/_functions.py:39-122, :139-179).
//
// Structure: a thread-block CLUSTER owns a slab of channels end to end.
//   * the [M, C] channels-last matrix is cut into C / CW channel slabs (CW channels = TX 16-byte vectors);
//     the CS CTAs of a cluster split the M rows of their slab; a thread keeps <= 8 rows x 16 bytes of x
//     (and of dy) in registers -- every load of the pass is in flight at once;
//   * per-channel partial sums: registers -> warp shuffles -> shared memory -> each CTA stores its 2*CW
//     partials into EVERY CTA of the cluster over distributed shared memory -> barrier.cluster -> every CTA
//     adds the CS partials in rank order (deterministic, identical everywhere);
//   * the rows still sitting in registers are normalised and stored.
// No global atomics, no grid-wide barrier and no co-residency assumption between clusters: a cluster that
// cannot be scheduled yet (MPS / MIG limits, a concurrent NCCL kernel holding SMs) simply runs later.
#include "common.cuh"
#include "bn_ring.cuh"

#include <stdlib.h>
#include <utility>

namespace scb {

constexpr int kClThreads = 512;    // one CTA per SM; 128 registers per thread: up to 16 rows x 16 bytes stay in registers
constexpr int kClRows = 8;         // rows a thread keeps in registers (per tensor); 16 in the large forward variant
constexpr int kClMaxCS = 16;

// Timeline instrumentation (tools/trace_bn.py): compiled in only with -DSCB_TRACE=1 (tools/build_trace_lib.sh).
#ifndef SCB_TRACE
#define SCB_TRACE 0
#endif
#if SCB_TRACE
__device__ long long* g_bn_trace = nullptr;            // [CTA][8] globaltimer stamps
__device__ __forceinline__ void bn_stamp(int slot, float dep = 0.f) {
    if (g_bn_trace != nullptr && threadIdx.x == 0) {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "f"(dep) : "memory");     // `dep`: not before this value exists
        g_bn_trace[blockIdx.x * 8 + slot] = t;
    }
}
extern "C" int scb_debug_set_bn_trace(long long* p) {
    return cudaMemcpyToSymbol(g_bn_trace, &p, sizeof(p)) == cudaSuccess ? 0 : -1;
}
#else
#define bn_stamp(...) ((void)0)
#endif

__device__ __forceinline__ uint32_t cl_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cl_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cl_store_remote(uint32_t local_addr, uint32_t rank, float v) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
}
__device__ __forceinline__ void cl_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// Makes a register quadruple opaque to the optimiser: conversions of the raw rows done before the reduction are
// not kept alive across it (the converted copy would double the register footprint), they are redone afterwards.
__device__ __forceinline__ void cl_opaque(uint4& v) {
    asm volatile("" : "+r"(v.x), "+r"(v.y), "+r"(v.z), "+r"(v.w));
}

struct ClGeom {
    int TX;              // threads (16-byte vectors) across a slab row; power of two <= 32
    int CS;              // CTAs per cluster (1 = no cluster)
    int n_slabs;
    int64_t rows_per_cta;
};

// CTA-level sum of a[VEC], b[VEC] over the threads with the same tx, then all-gather + sum over the cluster.
// Returns, in threads j < 2*CW (CW = TX*VEC), the cluster total of element j (j < CW: a, else b); other threads 0.
// sm: [warps][2*CW] floats, gath: [CS][2*CW] floats.
template <int VEC>
__device__ __forceinline__ float cluster_channel_sums(float (&a)[VEC], float (&b)[VEC], float* sm, float* gath,
                                                      int TX, int CS, uint32_t rank) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int CW = TX * VEC;
    // lanes with equal (lane % TX) hold the same channels
    for (int o = 16; o >= TX; o >>= 1) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            a[i] += __shfl_xor_sync(0xffffffffu, a[i], o);
            b[i] += __shfl_xor_sync(0xffffffffu, b[i], o);
        }
    }
    if (lane < TX) {
        float* dst = sm + warp * 2 * CW + lane * VEC;
#pragma unroll
        for (int i = 0; i < VEC; ++i) { dst[i] = a[i]; dst[CW + i] = b[i]; }
    }
    __syncthreads();
    float tot = 0.f;
    const int j = threadIdx.x;
    if (j < 2 * CW) {
#pragma unroll 8
        for (int w = 0; w < kClThreads / 32; ++w) tot += sm[w * 2 * CW + j];
        if (CS > 1) {
            const uint32_t local = cl_smem_u32(gath + rank * 2 * CW + j);
            for (int r = 0; r < CS; ++r) cl_store_remote(local, (uint32_t)r, tot);
        }
    }
    if (CS > 1) {
        cl_sync();
        if (j < 2 * CW) {
            tot = 0.f;
            for (int r = 0; r < CS; ++r) tot += gath[r * 2 * CW + j];      // rank order: same bits in every CTA
        }
    }
    return tot;
}

// ------------------------------------------------------------------ forward
// kMode 0: statistics + apply (one rank);  kMode 1: statistics only -> out_stats message (several ranks)
template <typename T, int kMode, bool kAdd, int ROWS>
__global__ void __launch_bounds__(kClThreads, 1)
bn_fwd_cluster_kernel(const T* __restrict__ x, const T* __restrict__ z, T* __restrict__ y, int64_t M, int C,
                      const ClGeom g, const float* __restrict__ weight, const float* __restrict__ bias,
                      float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                      float* save_invstd, float* save_count, long long* nbt, float* out_stats, int flags, int pdt) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float smf[];
    const int TX = g.TX, CS = g.CS, CW = TX * VEC, TY = kClThreads / TX;
    float* sm = smf;                                   // [warps][2*CW]
    float* gath = sm + (kClThreads / 32) * 2 * CW;     // [CS][2*CW]
    float* ab = gath + kClMaxCS * 2 * CW;              // [2][CW]
    const uint32_t rank = (CS > 1) ? cl_ctarank() : 0u;
    const int slab = blockIdx.x / CS;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = slab * CW + tx * VEC;
    const int64_t r_begin = (int64_t)rank * g.rows_per_cta;
    const int64_t r_end = min(M, r_begin + g.rows_per_cta);
    const int64_t r0 = r_begin + ty;
    const int64_t step = (int64_t)TY * C;
    bn_stamp(0);
    uint4 raw[ROWS], rz[ROWS];
    float s1[VEC], s2[VEC], k0[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
    int nrow = 0;
    {
        const T* p = x + r0 * C + c0;
#pragma unroll
        for (int u = 0; u < ROWS; ++u)
            if (r0 + (int64_t)u * TY < r_end) { raw[u] = ldg_stream16(p + u * step); nrow = u + 1; }
        if (kAdd && kMode == 0) {
            const T* pz = z + r0 * C + c0;
#pragma unroll
            for (int u = 0; u < ROWS; ++u)
                if (u < nrow) rz[u] = ldg_stream16(pz + u * step);
        }
        load_row_vec_cached<T, VEC>(x + c0, k0);       // common shift K_c = x[0, c]: no E[x^2] - E[x]^2 cancellation
        bn_stamp(1);
#pragma unroll
        for (int u = 0; u < ROWS; ++u) {
            if (u < nrow) {
                const T* e = reinterpret_cast<const T*>(&raw[u]);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float d = to_f32(e[i]) - k0[i];
                    s1[i] += d;
                    s2[i] = fmaf(d, d, s2[i]);
                }
            }
        }
    }
    bn_stamp(2, s1[0] + s2[0]);
    const float tot = cluster_channel_sums<VEC>(s1, s2, sm, gath, TX, CS, rank);
    bn_stamp(3, tot);
    // threads j < CW own channel slab*CW + j: tot = S1 there, S2 sits in thread j + CW -> through shared memory
    if ((int)threadIdx.x < 2 * CW) ab[threadIdx.x] = tot;
    __syncthreads();
    const float n = (float)M;
    float av = 0.f, bv = 0.f;
    if ((int)threadIdx.x < CW) {
        const int c = slab * CW + threadIdx.x;
        const float S1 = ab[threadIdx.x], S2 = ab[CW + threadIdx.x];
        const float mean = to_f32(x[c]) + S1 / n;
        const float m2 = fmaxf(S2 - S1 * S1 / n, 0.f);
        if (kMode == 1) {
            if (rank == 0) {
                out_stats[c] = mean;
                out_stats[C + c] = m2;
                if (c == 0) out_stats[2 * C] = n;
            }
        } else {
            const float invstd = rsqrtf(m2 / n + eps);
            av = ldp(weight, pdt, c, 1.f) * invstd;
            bv = ldp(bias, pdt, c, 0.f) - mean * av;
            if (rank == 0) {
                if (save_mean) save_mean[c] = mean;
                if (save_invstd) save_invstd[c] = invstd;
                if (running_mean) stp(running_mean, pdt, c, (1.f - momentum) * ldp(running_mean, pdt, c, 0.f) + momentum * mean);
                if (running_var) stp(running_var, pdt, c, (1.f - momentum) * ldp(running_var, pdt, c, 0.f) + momentum * (m2 / (n - 1.f)));
                if (c == 0 && save_count) save_count[0] = n;
                if (c == 0 && nbt) nbt[0] += 1;              // num_batches_tracked (batchnorm.py:763-764)
            }
        }
    }
    if (kMode == 1) return;
    __syncthreads();                                   // everybody has read S1 / S2
    if ((int)threadIdx.x < CW) { ab[threadIdx.x] = av; ab[CW + threadIdx.x] = bv; }
    __syncthreads();
    float a[VEC], b[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { a[i] = ab[tx * VEC + i]; b[i] = ab[CW + tx * VEC + i]; }
#pragma unroll
    for (int u = 0; u < ROWS; ++u) { cl_opaque(raw[u]); if (kAdd) cl_opaque(rz[u]); }
    const bool relu = (flags & SCB_BN_RELU) != 0;
    T* q = y + r0 * C + c0;
#pragma unroll
    for (int u = 0; u < ROWS; ++u) {
        if (u < nrow) {
            const T* e = reinterpret_cast<const T*>(&raw[u]);
            const T* ez = reinterpret_cast<const T*>(&rz[u]);
            uint4 o;
            T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                float v = fmaf(to_f32(e[i]), a[i], b[i]);
                if (kAdd) v += to_f32(ez[i]);
                oe[i] = from_f32<T>(relu ? fmaxf(v, 0.f) : v);
            }
            stg_stream16(q + u * step, o);
        }
    }
    bn_stamp(4);
}

// ------------------------------------------------------------------ backward
// kGate 0: plain; 1: ReLU, gate recomputed from x (y = (x - mean) a + b > 0); 2: gate read from the saved output y
// (y > 0: ReLU after a residual add), the gated dy is also written to dz (the gradient of the residual input).
// kMode 0: reduce + apply (one rank); 1: reduce only -> out_sums, dgamma, dbeta (several ranks)
template <typename T, int kMode, int kGate>
__global__ void __launch_bounds__(kClThreads, 1)
bn_bwd_cluster_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout,
                      T* __restrict__ dx, T* __restrict__ dz, int64_t M, int C, const ClGeom g,
                      const float* __restrict__ mean, const float* __restrict__ invstd,
                      const float* __restrict__ weight, const float* __restrict__ bias,
                      const float* __restrict__ n_total, float* out_sums, float* grad_weight, float* grad_bias, int pdt) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float smf[];
    const int TX = g.TX, CS = g.CS, CW = TX * VEC, TY = kClThreads / TX;
    float* sm = smf;
    float* gath = sm + (kClThreads / 32) * 2 * CW;
    float* ab = gath + kClMaxCS * 2 * CW;
    const uint32_t rank = (CS > 1) ? cl_ctarank() : 0u;
    const int slab = blockIdx.x / CS;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = slab * CW + tx * VEC;
    const int64_t r_begin = (int64_t)rank * g.rows_per_cta;
    const int64_t r_end = min(M, r_begin + g.rows_per_cta);
    const int64_t r0 = r_begin + ty;
    const int64_t step = (int64_t)TY * C;
    bn_stamp(0);
    uint4 rx[kClRows], rd[kClRows];
    unsigned gated[kClRows];               // bit i set: element i of the row is gated off (dy dropped)
    float s1[VEC], s2[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
    int nrow = 0;
    {
        const T* px = x + r0 * C + c0;
        const T* pd = dy + r0 * C + c0;
#pragma unroll
        for (int u = 0; u < kClRows; ++u)
            if (r0 + (int64_t)u * TY < r_end) {
                rx[u] = ldg_stream16(px + u * step);
                rd[u] = ldg_stream16(pd + u * step);
                nrow = u + 1;
            }
        bn_stamp(1);
#pragma unroll
        for (int u = 0; u < kClRows; ++u) gated[u] = 0u;
        if (kGate == 2) {
            const T* py = yout + r0 * C + c0;
#pragma unroll
            for (int u = 0; u < kClRows; ++u) {
                if (u < nrow) {
                    const uint4 ry = ldg_stream16(py + u * step);
                    const T* ey = reinterpret_cast<const T*>(&ry);
#pragma unroll
                    for (int i = 0; i < VEC; ++i)
                        if (!(to_f32(ey[i]) > 0.f)) gated[u] |= 1u << i;
                }
            }
        }
        float mu[VEC], k1[VEC], bb[VEC];
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            mu[i] = mean[c0 + i];
            // the scale is only needed before the reduction when the gate is recomputed from x
            k1[i] = (kGate == 1) ? ldp(weight, pdt, c0 + i, 1.f) * invstd[c0 + i] : 0.f;
            bb[i] = (kGate == 1) ? ldp(bias, pdt, c0 + i, 0.f) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < kClRows; ++u) {
            if (u < nrow) {
                const T* ex = reinterpret_cast<const T*>(&rx[u]);
                const T* ed = reinterpret_cast<const T*>(&rd[u]);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float xm = to_f32(ex[i]) - mu[i];
                    float gv = to_f32(ed[i]);
                    if (kGate == 1 && !(fmaf(xm, k1[i], bb[i]) > 0.f)) gated[u] |= 1u << i;
                    if (kGate != 0 && ((gated[u] >> i) & 1u)) gv = 0.f;
                    s1[i] += gv;
                    s2[i] = fmaf(gv, xm, s2[i]);
                }
            }
        }
    }
    bn_stamp(2, s1[0] + s2[0]);
    const float tot = cluster_channel_sums<VEC>(s1, s2, sm, gath, TX, CS, rank);
    bn_stamp(3, tot);
    if ((int)threadIdx.x < 2 * CW) ab[threadIdx.x] = tot;
    __syncthreads();
    float v2 = 0.f, v3 = 0.f;
    if ((int)threadIdx.x < CW) {
        const int c = slab * CW + threadIdx.x;
        const float sd = ab[threadIdx.x], sx = ab[CW + threadIdx.x];
        const float is = invstd[c];
        const float inv_n = 1.f / n_total[0];
        v2 = sd * inv_n;
        v3 = is * is * sx * inv_n;
        if (rank == 0) {
            if (out_sums) { out_sums[c] = sd; out_sums[C + c] = sx; }
            if (grad_weight) stp(grad_weight, pdt, c, sx * is);
            if (grad_bias) stp(grad_bias, pdt, c, sd);
        }
    }
    if (kMode == 1) return;
    __syncthreads();
    if ((int)threadIdx.x < CW) { ab[threadIdx.x] = v2; ab[CW + threadIdx.x] = v3; }
    __syncthreads();
#pragma unroll
    for (int u = 0; u < kClRows; ++u) { cl_opaque(rx[u]); cl_opaque(rd[u]); }
    // dx = (g - k2 - (x - mu) k3) k1  =  g A + x B + D   with  A = k1, B = -k3 k1, D = (mu k3 - k2) k1:
    // three coefficient vectors instead of four, and no subtraction of the mean per element
    float cA[VEC], cB[VEC], cD[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const float k2 = ab[tx * VEC + i], k3 = ab[CW + tx * VEC + i];
        const float kk = ldp(weight, pdt, c0 + i, 1.f) * invstd[c0 + i];
        cA[i] = kk;
        cB[i] = -k3 * kk;
        cD[i] = (mean[c0 + i] * k3 - k2) * kk;
    }
    T* po = dx + r0 * C + c0;
    T* pz = (kGate == 2 && dz != nullptr) ? dz + r0 * C + c0 : nullptr;
#pragma unroll
    for (int u = 0; u < kClRows; ++u) {
        if (u < nrow) {
            const T* ex = reinterpret_cast<const T*>(&rx[u]);
            const T* ed = reinterpret_cast<const T*>(&rd[u]);
            uint4 o, oz;
            T* oe = reinterpret_cast<T*>(&o);
            T* oze = reinterpret_cast<T*>(&oz);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const bool off = (kGate != 0) && ((gated[u] >> i) & 1u);
                const float gv = off ? 0.f : to_f32(ed[i]);
                oe[i] = from_f32<T>(fmaf(to_f32(ex[i]), cB[i], fmaf(gv, cA[i], cD[i])));
                if (kGate == 2) oze[i] = off ? from_f32<T>(0.f) : ed[i];
            }
            stg_stream16(po + u * step, o);
            if (kGate == 2 && pz != nullptr) stg_stream16(pz + u * step, oz);
        }
    }
    bn_stamp(4);
}

// ------------------------------------------------------------------ backward, rows parked in shared memory
// Same ownership and reduction as bn_bwd_cluster_kernel, but a thread's rows of x and dy sit in shared memory (16-byte
// asynchronous copies, all in flight at once, no registers tied up), so a CTA holds 12 rows per thread instead of 8:
// cfg 2 then fits 16 slabs of 32 bytes x clusters of 6 in one wave instead of 32 slabs of 16 bytes x 4, i.e. every
// load and store moves whole 32-byte sectors.  kGate 0 / 1 (the gate is recomputed from x in both passes).
constexpr int kClSmemRows = 12;

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(cl_smem_u32(smem_dst)), "l"(gsrc) : "memory");
}

template <typename T, int kGate>
__global__ void __launch_bounds__(kClThreads, 1)
bn_bwd_cluster_smem_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t M, int C,
                           const ClGeom g, const float* __restrict__ mean, const float* __restrict__ invstd,
                           const float* __restrict__ weight, const float* __restrict__ bias,
                           const float* __restrict__ n_total, float* out_sums, float* grad_weight, float* grad_bias,
                           int pdt) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float smf[];
    const int TX = g.TX, CS = g.CS, CW = TX * VEC, TY = kClThreads / TX;
    float* sm = smf;
    float* gath = sm + (kClThreads / 32) * 2 * CW;
    float* ab = gath + kClMaxCS * 2 * CW;
    uint4* park = reinterpret_cast<uint4*>(smf + ((kClThreads / 32 + kClMaxCS + 1) * 2 * CW + 3) / 4 * 4);
    uint4* sx = park + threadIdx.x;                              // row u of this thread: sx[u * kClThreads]
    uint4* sd = park + kClSmemRows * kClThreads + threadIdx.x;
    const uint32_t rank = (CS > 1) ? cl_ctarank() : 0u;
    const int slab = blockIdx.x / CS;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = slab * CW + tx * VEC;
    const int64_t r_begin = (int64_t)rank * g.rows_per_cta;
    const int64_t r_end = min(M, r_begin + g.rows_per_cta);
    const int64_t r0 = r_begin + ty;
    const int64_t step = (int64_t)TY * C;
    bn_stamp(0);
    int nrow = 0;
    {
        const T* px = x + r0 * C + c0;
        const T* pd = dy + r0 * C + c0;
#pragma unroll
        for (int u = 0; u < kClSmemRows; ++u)
            if (r0 + (int64_t)u * TY < r_end) {
                cp_async16(sx + u * kClThreads, px + u * step);
                cp_async16(sd + u * kClThreads, pd + u * step);
                nrow = u + 1;
            }
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    bn_stamp(1);
    float s1[VEC], s2[VEC], mu[VEC], k1[VEC], bb[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        s1[i] = 0.f; s2[i] = 0.f;
        mu[i] = mean[c0 + i];
        k1[i] = ldp(weight, pdt, c0 + i, 1.f) * invstd[c0 + i];
        bb[i] = (kGate == 1) ? ldp(bias, pdt, c0 + i, 0.f) : 0.f;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");         // a thread reads only what it copied itself
#pragma unroll 4
    for (int u = 0; u < nrow; ++u) {
        const uint4 vx = sx[u * kClThreads], vd = sd[u * kClThreads];
        const T* ex = reinterpret_cast<const T*>(&vx);
        const T* ed = reinterpret_cast<const T*>(&vd);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float xm = to_f32(ex[i]) - mu[i];
            float gv = to_f32(ed[i]);
            if (kGate == 1 && !(fmaf(xm, k1[i], bb[i]) > 0.f)) gv = 0.f;
            s1[i] += gv;
            s2[i] = fmaf(gv, xm, s2[i]);
        }
    }
    bn_stamp(2, s1[0] + s2[0]);
    const float tot = cluster_channel_sums<VEC>(s1, s2, sm, gath, TX, CS, rank);
    bn_stamp(3, tot);
    if ((int)threadIdx.x < 2 * CW) ab[threadIdx.x] = tot;
    __syncthreads();
    float v2 = 0.f, v3 = 0.f;
    if ((int)threadIdx.x < CW) {
        const int c = slab * CW + threadIdx.x;
        const float sdy = ab[threadIdx.x], sxy = ab[CW + threadIdx.x];
        const float is = invstd[c];
        const float inv_n = 1.f / n_total[0];
        v2 = sdy * inv_n;
        v3 = is * is * sxy * inv_n;
        if (rank == 0) {
            if (out_sums) { out_sums[c] = sdy; out_sums[C + c] = sxy; }
            if (grad_weight) stp(grad_weight, pdt, c, sxy * is);
            if (grad_bias) stp(grad_bias, pdt, c, sdy);
        }
    }
    __syncthreads();
    if ((int)threadIdx.x < CW) { ab[threadIdx.x] = v2; ab[CW + threadIdx.x] = v3; }
    __syncthreads();
    // dx = g A + x B + D with A = k1, B = -k3 k1, D = (mu k3 - k2) k1 (see bn_bwd_cluster_kernel); the gate test
    // y > 0 is (x - mu) k1 + beta > 0, recomputed exactly as in the first pass
    float cB[VEC], cD[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const float k2 = ab[tx * VEC + i], k3 = ab[CW + tx * VEC + i];
        cB[i] = -k3 * k1[i];
        cD[i] = (mu[i] * k3 - k2) * k1[i];
    }
    T* po = dx + r0 * C + c0;
#pragma unroll 4
    for (int u = 0; u < nrow; ++u) {
        const uint4 vx = sx[u * kClThreads], vd = sd[u * kClThreads];
        const T* ex = reinterpret_cast<const T*>(&vx);
        const T* ed = reinterpret_cast<const T*>(&vd);
        uint4 o;
        T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float xv = to_f32(ex[i]);
            float gv = to_f32(ed[i]);
            if (kGate == 1 && !(fmaf(xv - mu[i], k1[i], bb[i]) > 0.f)) gv = 0.f;
            oe[i] = from_f32<T>(fmaf(xv, cB[i], fmaf(gv, k1[i], cD[i])));
        }
        stg_stream16(po + u * step, o);
    }
    bn_stamp(4);
}

static size_t cl_smem_park_bytes(int TX, int vec) {
    return ((size_t)(kClThreads / 32 + kClMaxCS + 1) * 2 * TX * vec + 3) / 4 * 4 * sizeof(float) +
           (size_t)2 * kClSmemRows * kClThreads * sizeof(uint4);
}

// =================================================================== streaming cluster kernels
// Same ownership (a cluster owns a channel slab end to end, sums meet over distributed shared memory), but the rows
// are STREAMED twice instead of being held in registers: pass 1 accumulates the sums, pass 2 walks the same rows
// back to front (what pass 1 touched last is what the L2 still holds) and writes the result.  No capacity limit, so
// the geometry is free: the slabs may be uneven (C / 8 columns split over 7 clusters of 16 CTAs on a B200, the most
// it holds at once), which keeps the pieces of a row wide (64..80 bytes at C = 256) and one wave resident.
// Used for tensors that fit the L2 (the second pass then costs no DRAM traffic).
struct ClStream {
    int S;               // slabs = clusters
    int CS;              // CTAs per cluster
    int TX;              // threads across a slab row (power of two, >= widest slab in 16-byte vectors, <= 32)
    int cv;              // 16-byte vectors per row (C / VEC)
    int64_t rows_per_cta;
};
constexpr int kStreamU = 4;      // rows in flight per thread and tensor

template <typename T, bool kAdd>
__global__ void __launch_bounds__(kClThreads, 1)
bn_fwd_stream_kernel(const T* __restrict__ x, const T* __restrict__ z, T* __restrict__ y, int64_t M, int C,
                     const ClStream g, const float* __restrict__ weight, const float* __restrict__ bias,
                     float* running_mean, float* running_var, float momentum, float eps, float* save_mean,
                     float* save_invstd, float* save_count, long long* nbt, int flags, int pdt) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float smf[];
    const int TX = g.TX, CS = g.CS, CW = TX * VEC, TY = kClThreads / TX;
    float* sm = smf;
    float* gath = sm + (kClThreads / 32) * 2 * CW;
    float* ab = gath + kClMaxCS * 2 * CW;
    const uint32_t rank = (CS > 1) ? cl_ctarank() : 0u;
    const int slab = blockIdx.x / CS;
    const int v_lo = (int)(((int64_t)slab * g.cv) / g.S), v_hi = (int)(((int64_t)(slab + 1) * g.cv) / g.S);
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const bool active = tx < v_hi - v_lo;
    const int c0 = (v_lo + (active ? tx : 0)) * VEC;
    const int64_t r_begin = (int64_t)rank * g.rows_per_cta;
    const int64_t r_end = min(M, r_begin + g.rows_per_cta);
    float s1[VEC], s2[VEC], k0[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; }
    load_row_vec_cached<T, VEC>(x + c0, k0);                    // common shift K_c = x[0, c]
    if (active) {
        for (int64_t rk = r_begin + ty; rk < r_end; rk += (int64_t)kStreamU * TY) {
            uint4 raw[kStreamU];
#pragma unroll
            for (int u = 0; u < kStreamU; ++u)
                if (rk + (int64_t)u * TY < r_end) raw[u] = ldg_stream16(x + (rk + (int64_t)u * TY) * C + c0);
#pragma unroll
            for (int u = 0; u < kStreamU; ++u) {
                if (rk + (int64_t)u * TY < r_end) {
                    const T* e = reinterpret_cast<const T*>(&raw[u]);
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float d = to_f32(e[i]) - k0[i];
                        s1[i] += d;
                        s2[i] = fmaf(d, d, s2[i]);
                    }
                }
            }
        }
    }
    const float tot = cluster_channel_sums<VEC>(s1, s2, sm, gath, TX, CS, rank);
    if ((int)threadIdx.x < 2 * CW) ab[threadIdx.x] = tot;
    __syncthreads();
    const float n = (float)M;
    float av = 0.f, bv = 0.f;
    if ((int)threadIdx.x < CW && (int)threadIdx.x < (v_hi - v_lo) * VEC) {
        const int c = v_lo * VEC + threadIdx.x;
        const float S1 = ab[threadIdx.x], S2 = ab[CW + threadIdx.x];
        const float mean = to_f32(x[c]) + S1 / n;
        const float m2 = fmaxf(S2 - S1 * S1 / n, 0.f);
        const float invstd = rsqrtf(m2 / n + eps);
        av = ldp(weight, pdt, c, 1.f) * invstd;
        bv = ldp(bias, pdt, c, 0.f) - mean * av;
        if (rank == 0) {
            if (save_mean) save_mean[c] = mean;
            if (save_invstd) save_invstd[c] = invstd;
            if (running_mean) stp(running_mean, pdt, c, (1.f - momentum) * ldp(running_mean, pdt, c, 0.f) + momentum * mean);
            if (running_var) stp(running_var, pdt, c, (1.f - momentum) * ldp(running_var, pdt, c, 0.f) + momentum * (m2 / (n - 1.f)));
            if (c == 0 && save_count) save_count[0] = n;
            if (c == 0 && nbt) nbt[0] += 1;
        }
    }
    __syncthreads();
    if ((int)threadIdx.x < CW) { ab[threadIdx.x] = av; ab[CW + threadIdx.x] = bv; }
    __syncthreads();
    if (!active) return;
    float a[VEC], b[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { a[i] = ab[tx * VEC + i]; b[i] = ab[CW + tx * VEC + i]; }
    const bool relu = (flags & SCB_BN_RELU) != 0;
    const int64_t chunk = (int64_t)kStreamU * TY;
    for (int64_t k = (r_end - r_begin + chunk - 1) / chunk - 1; k >= 0; --k) {
        const int64_t rk = r_begin + k * chunk + ty;
        uint4 raw[kStreamU], rz[kStreamU];
#pragma unroll
        for (int u = 0; u < kStreamU; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                raw[u] = ldg_stream16(x + r * C + c0);
                if (kAdd) rz[u] = ldg_stream16(z + r * C + c0);
            }
        }
#pragma unroll
        for (int u = 0; u < kStreamU; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                const T* e = reinterpret_cast<const T*>(&raw[u]);
                const T* ez = reinterpret_cast<const T*>(&rz[u]);
                uint4 o;
                T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float v = fmaf(to_f32(e[i]), a[i], b[i]);
                    if (kAdd) v += to_f32(ez[i]);
                    oe[i] = from_f32<T>(relu ? fmaxf(v, 0.f) : v);
                }
                stg_stream16(y + r * C + c0, o);
            }
        }
    }
}

template <typename T, int kGate>
__global__ void __launch_bounds__(kClThreads, 1)
bn_bwd_stream_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout,
                     T* __restrict__ dx, T* __restrict__ dz, int64_t M, int C, const ClStream g,
                     const float* __restrict__ mean, const float* __restrict__ invstd,
                     const float* __restrict__ weight, const float* __restrict__ bias,
                     const float* __restrict__ n_total, float* out_sums, float* grad_weight, float* grad_bias, int pdt) {
    constexpr int VEC = Vec16<T>::N;
    extern __shared__ float smf[];
    const int TX = g.TX, CS = g.CS, CW = TX * VEC, TY = kClThreads / TX;
    float* sm = smf;
    float* gath = sm + (kClThreads / 32) * 2 * CW;
    float* ab = gath + kClMaxCS * 2 * CW;
    const uint32_t rank = (CS > 1) ? cl_ctarank() : 0u;
    const int slab = blockIdx.x / CS;
    const int v_lo = (int)(((int64_t)slab * g.cv) / g.S), v_hi = (int)(((int64_t)(slab + 1) * g.cv) / g.S);
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const bool active = tx < v_hi - v_lo;
    const int c0 = (v_lo + (active ? tx : 0)) * VEC;
    const int64_t r_begin = (int64_t)rank * g.rows_per_cta;
    const int64_t r_end = min(M, r_begin + g.rows_per_cta);
    float s1[VEC], s2[VEC], mu[VEC], k1[VEC], bb[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        s1[i] = 0.f; s2[i] = 0.f;
        mu[i] = mean[c0 + i];
        k1[i] = ldp(weight, pdt, c0 + i, 1.f) * invstd[c0 + i];
        bb[i] = (kGate == 1) ? ldp(bias, pdt, c0 + i, 0.f) : 0.f;
    }
    if (active) {
        for (int64_t rk = r_begin + ty; rk < r_end; rk += (int64_t)kStreamU * TY) {
            uint4 rx[kStreamU], rd[kStreamU], ry[kStreamU];
#pragma unroll
            for (int u = 0; u < kStreamU; ++u) {
                const int64_t r = rk + (int64_t)u * TY;
                if (r < r_end) {
                    rx[u] = ldg_stream16(x + r * C + c0);
                    rd[u] = ldg_stream16(dy + r * C + c0);
                    if (kGate == 2) ry[u] = ldg_stream16(yout + r * C + c0);
                }
            }
#pragma unroll
            for (int u = 0; u < kStreamU; ++u) {
                if (rk + (int64_t)u * TY < r_end) {
                    const T* ex = reinterpret_cast<const T*>(&rx[u]);
                    const T* ed = reinterpret_cast<const T*>(&rd[u]);
                    const T* ey = reinterpret_cast<const T*>(&ry[u]);
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float xm = to_f32(ex[i]) - mu[i];
                        float gv = to_f32(ed[i]);
                        if (kGate == 1 && !(fmaf(xm, k1[i], bb[i]) > 0.f)) gv = 0.f;
                        if (kGate == 2 && !(to_f32(ey[i]) > 0.f)) gv = 0.f;
                        s1[i] += gv;
                        s2[i] = fmaf(gv, xm, s2[i]);
                    }
                }
            }
        }
    }
    const float tot = cluster_channel_sums<VEC>(s1, s2, sm, gath, TX, CS, rank);
    if ((int)threadIdx.x < 2 * CW) ab[threadIdx.x] = tot;
    __syncthreads();
    float v2 = 0.f, v3 = 0.f;
    if ((int)threadIdx.x < CW && (int)threadIdx.x < (v_hi - v_lo) * VEC) {
        const int c = v_lo * VEC + threadIdx.x;
        const float sd = ab[threadIdx.x], sx = ab[CW + threadIdx.x];
        const float is = invstd[c];
        const float inv_n = 1.f / n_total[0];
        v2 = sd * inv_n;
        v3 = is * is * sx * inv_n;
        if (rank == 0) {
            if (out_sums) { out_sums[c] = sd; out_sums[C + c] = sx; }
            if (grad_weight) stp(grad_weight, pdt, c, sx * is);
            if (grad_bias) stp(grad_bias, pdt, c, sd);
        }
    }
    __syncthreads();
    if ((int)threadIdx.x < CW) { ab[threadIdx.x] = v2; ab[CW + threadIdx.x] = v3; }
    __syncthreads();
    if (!active) return;
    float cB[VEC], cD[VEC];                                    // dx = g k1 + x B + D  (see bn_bwd_cluster_kernel)
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const float k2 = ab[tx * VEC + i], k3 = ab[CW + tx * VEC + i];
        cB[i] = -k3 * k1[i];
        cD[i] = (mu[i] * k3 - k2) * k1[i];
    }
    const int64_t chunk = (int64_t)kStreamU * TY;
    for (int64_t k = (r_end - r_begin + chunk - 1) / chunk - 1; k >= 0; --k) {
        const int64_t rk = r_begin + k * chunk + ty;
        uint4 rx[kStreamU], rd[kStreamU], ry[kStreamU];
#pragma unroll
        for (int u = 0; u < kStreamU; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                rx[u] = ldg_stream16(x + r * C + c0);
                rd[u] = ldg_stream16(dy + r * C + c0);
                if (kGate == 2) ry[u] = ldg_stream16(yout + r * C + c0);
            }
        }
#pragma unroll
        for (int u = 0; u < kStreamU; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                const T* ex = reinterpret_cast<const T*>(&rx[u]);
                const T* ed = reinterpret_cast<const T*>(&rd[u]);
                const T* ey = reinterpret_cast<const T*>(&ry[u]);
                uint4 o, oz;
                T* oe = reinterpret_cast<T*>(&o);
                T* oze = reinterpret_cast<T*>(&oz);
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float xv = to_f32(ex[i]);
                    bool off = false;
                    if (kGate == 1) off = !(fmaf(xv - mu[i], k1[i], bb[i]) > 0.f);
                    if (kGate == 2) off = !(to_f32(ey[i]) > 0.f);
                    const float gv = off ? 0.f : to_f32(ed[i]);
                    oe[i] = from_f32<T>(fmaf(xv, cB[i], fmaf(gv, k1[i], cD[i])));
                    if (kGate == 2) oze[i] = off ? from_f32<T>(0.f) : ed[i];
                }
                stg_stream16(dx + r * C + c0, o);
                if (kGate == 2 && dz != nullptr) stg_stream16(dz + r * C + c0, oz);
            }
        }
    }
}

// ------------------------------------------------------------------ host side: geometry
static bool cl_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// Largest cluster size the device accepts for kernels of this shape (16 needs the non-portable opt-in and a GPC
// with 16 free SMs); probed once per device with the occupancy API on a representative kernel.
template <typename Kern>
static int cl_probe_max_cs(Kern kern, size_t smem) {
    for (int cs = kClMaxCS; cs > 1; cs >>= 1) {
        if (cs > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(cs);
        cfg.blockDim = dim3(kClThreads);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n = 0;
        if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) == cudaSuccess && n >= 1) return cs;
        cudaGetLastError();
    }
    return 1;
}

static size_t cl_smem_bytes(int TX, int vec) {
    return (size_t)(kClThreads / 32 + kClMaxCS + 1) * 2 * TX * vec * sizeof(float);
}

static int cl_max_cs() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 1;
    if (cached[dev] == 0) {
        const char* e = getenv("SCB_BN_MAX_CLUSTER");
        const int cap = e ? atoi(e) : kClMaxCS;
        auto kern = bn_fwd_cluster_kernel<__nv_bfloat16, 0, false, 8>;
        int m = cl_probe_max_cs(kern, cl_smem_bytes(4, 8));
        if (cap >= 1 && m > cap) m = cap;
        cached[dev] = m < 1 ? 1 : m;
    }
    return cached[dev];
}

// Clusters of `cs` CTAs of these kernels that the device can hold at once (occupancy API; cached per device).  On
// a B200 not every GPC has 16 free SMs: measured 7 clusters of 16, i.e. an 8-cluster grid runs a second wave.
static int cl_max_active(int cs) {
    static int cached[64][kClMaxCS + 1] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || cs < 1 || cs > kClMaxCS) return 1;
    if (cached[dev][cs] == 0) {
        auto kern = bn_bwd_cluster_kernel<__nv_bfloat16, 0, 0>;
        const size_t smem = cl_smem_bytes(4, 8);
        int n = 0;
        if (cs == 1) {
            n = sm_count();
        } else {
            if (cs > 8) cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(cs * 64);
            cfg.blockDim = dim3(kClThreads);
            cfg.dynamicSmemBytes = smem;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeClusterDimension;
            attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
        }
        cached[dev][cs] = n > 0 ? n : -1;
    }
    return cached[dev][cs] > 0 ? cached[dev][cs] : 0;
}

// CTAs of one cluster size that can be resident at once: the grid should not exceed it (a second wave doubles the
// time of a latency-bound pass)
static int cl_wave_cap(int cs) { return cl_max_active(cs) * cs; }

// picks (TX, CS) so that the tensor fits the registers of ONE resident wave; false if it does not
static bool cl_geom(int64_t M, int C, int es, int rows, ClGeom* out, int64_t* out_cost = nullptr) {
    const int vec = 16 / es;
    if (C % vec != 0 || M < 2) return false;
    const int cv = C / vec;
    const int max_cs = cl_max_cs();
    bool found = false;
    int64_t best_cost = 0;
    // cluster sizes need not be powers of two: 7 CTAs x 16 slabs fit one wave where 8 x 16 do not (SCB_BN_ANY_CLUSTER=0:
    // powers of two only)
    static const bool any_cs = [] { const char* e = getenv("SCB_BN_ANY_CLUSTER"); return e == nullptr || e[0] != '0'; }();
    for (int cs = max_cs; cs >= 1; cs = (any_cs && cs <= 8) ? cs - 1 : cs >> 1) {
        for (int tx = 32; tx >= 1; tx >>= 1) {
            if (cv % tx != 0) continue;
            const int ty = kClThreads / tx;
            const int64_t rpc = ceil_div(M, cs);
            if (rpc > (int64_t)rows * ty) continue;
            const int64_t grid = (int64_t)(cv / tx) * cs;
            if (grid > cl_wave_cap(cs)) continue;
            if (cl_smem_bytes(tx, vec) > 200 * 1024) continue;
            // cost ~ bytes per CTA; narrow pieces (< 64 bytes per row) waste DRAM sectors
            const int64_t cost = rpc * tx * (tx * 16 >= 64 ? 4 : (tx * 16 >= 32 ? 5 : 8));
            if (!found || cost < best_cost) {
                found = true; best_cost = cost;
                out->TX = tx; out->CS = cs; out->n_slabs = cv / tx; out->rows_per_cta = rpc;
            }
        }
    }
    if (found && out_cost != nullptr) *out_cost = best_cost;
    return found;
}

// Streaming kernels: as many clusters as the device holds at once, each owning an (uneven) slab of 16-byte columns.
// Returns false when the row is too wide for 32 threads per slab row with the clusters available.
static bool cl_stream_geom(int64_t M, int C, int es, ClStream* out) {
    const int vec = 16 / es;
    if (C % vec != 0 || M < 2) return false;
    const int cv = C / vec;
    bool found = false;
    double best = 0.0;
    for (int cs = cl_max_cs(); cs >= 1; cs >>= 1) {
        int S = cl_max_active(cs);
        if (S > cv) S = cv;
        if (S < 1) continue;
        const int wmax = (int)ceil_div(cv, S);
        if (wmax > 32) continue;
        int tx = 1;
        while (tx < wmax) tx <<= 1;
        const int64_t rpc = ceil_div(M, cs);
        const int piece = (cv / S) * 16;                                    // bytes of a row one slab reads
        const double cost = (double)rpc * tx * (piece >= 64 ? 1.0 : (piece >= 32 ? 1.25 : 2.0));
        if (!found || cost < best) {
            found = true; best = cost;
            out->S = S; out->CS = cs; out->TX = tx; out->cv = cv; out->rows_per_cta = rpc;
        }
    }
    return found;
}

// which single-launch implementation a tensor takes: 1 = rows held in registers, 2 = rows streamed twice (second
// pass out of the L2), 0 = neither (too large: the multi-kernel path).  SCB_BN_FUSED_MODE=reg|stream forces one.
static int cl_pick(int64_t M, int C, int es, int which, int flags) {
    static const int forced = [] {
        const char* e = getenv("SCB_BN_FUSED_MODE");
        if (e == nullptr) return 0;
        return e[0] == 'r' ? 1 : (e[0] == 's' ? 2 : (e[0] == 'c' ? 3 : 0));
    }();
    ClGeom g;
    ClStream gs;
    const bool add = (flags & SCB_BN_ADD) != 0;
    const int rows = (which == 0 && !add) ? 2 * kClRows : kClRows;
    const bool reg_ok = cl_geom(M, C, es, rows, &g);
    const double bytes = (double)M * C * es;
    // the second pass must come out of the L2: x (+ z) and y forward, x, dy (+ y) and dx backward.  Beyond ~56 MB of
    // resident tensors the two-launch ring kernels (bn_ring.cu) are faster: measured on B200, fwd+bwd in a CUDA graph,
    // [512,512,8,8] bf16 (32 MiB) streamed 79.5 us / ring 61.0 us; [512,256,8,8] (16 MiB) 41.8 / 41.7;
    // [512,1024,4,4] (16 MiB, wide slabs) 31.3 / 42.9
    const double l2_budget = 56e6;
    const double resident = which == 0 ? bytes * (add ? 3 : 2) : bytes * (((flags & SCB_BN_RELU) && add) ? 4 : 3);
    const bool stream_ok = resident <= l2_budget && cl_stream_geom(M, C, es, &gs);
    if (forced == 1) return reg_ok ? 1 : 0;
    if (forced == 2) return stream_ok ? 2 : 0;
    if (forced == 3) return 0;
    // measured on B200 (tools/time_bn.py, fwd+bwd in a CUDA graph): [256,256,8,8] bf16 registers 23.6 us / streamed
    // 26.5 us, [512,512,4,4] 17.3 / 26.5 (the streamed passes issue their rows in four dependent batches), while
    // [512,256,8,8], which only the streamed kernels take, gains over the multi-kernel path (43.4 us against 47.3)
    if (reg_ok) return 1;
    return stream_ok ? 2 : 0;
}

template <typename... KArgs, typename... Args>
static cudaError_t cl_launch(void (*kernel)(KArgs...), const ClGeom& g, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaError_t e = set_max_dynamic_smem(kernel, (int)smem);
    if (e != cudaSuccess) return e;
    // (no shared-memory carve-out preference here: with the maximum carve-out the register-resident passes, whose loads
    // are tracked by the L1, slow down by 30-60 %: cfg 2 forward 10.1 -> 13.5 us, backward 14.0 -> 22.1 us)
    if (g.CS > 8 && !attr_done((const void*)kernel, (int)cudaFuncAttributeNonPortableClusterSizeAllowed, 1)) {
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(g.n_slabs * g.CS));
    cfg.blockDim = dim3(kClThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)g.CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g.CS > 1 ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

template <typename... KArgs, typename... Args>
static cudaError_t cl_launch_stream(void (*kernel)(KArgs...), const ClStream& g, size_t smem, cudaStream_t stream,
                                    Args&&... args) {
    ClGeom gg;
    gg.TX = g.TX; gg.CS = g.CS; gg.n_slabs = g.S; gg.rows_per_cta = g.rows_per_cta;
    return cl_launch(kernel, gg, smem, stream, std::forward<Args>(args)...);
}

static int cl_check(const void* x, int dtype, int64_t rows, int channels) {
    SCB_REQUIRE(x != nullptr, "bn fused: null tensor pointer");
    SCB_REQUIRE(dtype == SCB_F32 || dtype == SCB_BF16 || dtype == SCB_F16, "bn fused: unsupported dtype %d", dtype);
    SCB_REQUIRE(channels > 0, "SyncBatchNorm number of input channels should be non-zero");
    SCB_REQUIRE(rows >= 0 && rows < (int64_t)1 << 40, "bn fused: bad row count");
    return SCB_OK;
}

template <typename T, int kMode, int kGate>
static cudaError_t cl_launch_bwd(const ClGeom& g, cudaStream_t stream, const void* dy, const void* x, const void* y,
                                 void* dx, void* dz, int64_t M, int C, const float* mean, const float* invstd,
                                 const float* weight, const float* bias, const float* n_total, float* out_sums,
                                 float* grad_weight, float* grad_bias, int pdt) {
    return cl_launch(bn_bwd_cluster_kernel<T, kMode, kGate>, g, cl_smem_bytes(g.TX, Vec16<T>::N), stream, (const T*)dy,
                     (const T*)x, (const T*)y, (T*)dx, (T*)dz, M, C, g, mean, invstd, weight, bias, n_total, out_sums,
                     grad_weight, grad_bias, pdt);
}

}  // namespace scb

using namespace scb;

// diagnostics: clusters of `cluster_size` CTAs of the fused kernels the current device can hold at once
extern "C" int scb_bn_cluster_occupancy(int cluster_size) { return cl_max_active(cluster_size); }

// which = 0: forward (with or without a residual operand, see `flags`), 1: backward
// The shared-memory-resident pass with a grid barrier (bn_ring.cu, cooperative launch) against the cluster passes of
// this file: SCB_BN_FUSED_MODE=coop|reg|stream forces one family; by default the resident pass is taken whenever the
// tensor fits (measurements in DESIGN.md section 4).
static bool cl_use_resident(int dtype, int64_t M, int C, int which, int flags) {
    static const int forced = [] {
        const char* e = getenv("SCB_BN_FUSED_MODE");
        if (e == nullptr) return 0;
        return e[0] == 'r' ? 1 : (e[0] == 's' ? 2 : (e[0] == 'c' ? 3 : 0));
    }();
    if (forced == 1 || forced == 2) return false;
    // register-resident cluster passes first where the tensor fits them: measured on B200 (fwd+bwd, CUDA graph)
    // [256,256,8,8] bf16 23.9 us against 24.9 us resident, [512,512,4,4] 17.6 against 25.1; the resident pass wins
    // where only the streamed cluster pass or the ring passes are left ([512,256,8,8] forward 15.9 against 18.3 / 19.5)
    // (the 16-rows-per-thread forward variant of the register pass is only kept where its slabs are at least 32 bytes
    // wide: [512,1024,4,4] 11.3 us against 17.4 us resident, but [512,256,8,8] with 16-byte slabs 19.5 against 15.9)
    if (forced == 0) {
        const int es = dtype == SCB_F32 ? 4 : 2;
        ClGeom g;
        if (cl_geom(M, C, es, kClRows, &g)) return false;
        if (which == 0 && !(flags & SCB_BN_ADD) && cl_geom(M, C, es, 2 * kClRows, &g) && g.TX >= 2) return false;
    }
    return ring_resident_supported(dtype, M, C, which, flags);
}

extern "C" int scb_bn_fused_supported(int dtype, int channels_last, int64_t outer, int channels, int64_t inner,
                                      int which, int flags) {
    if (!channels_last || channels <= 0) return 0;
    if (dtype != SCB_F32 && dtype != SCB_BF16 && dtype != SCB_F16) return 0;
    const int es = (dtype == SCB_F32) ? 4 : 2;
    if (cl_use_resident(dtype, outer * inner, channels, which, flags)) return 1;
    return cl_pick(outer * inner, channels, es, which, flags) != 0 ? 1 : 0;
}

#define SCB_CL_UNSUPPORTED(what)                                                             \
    do { set_error(what ": tensor does not fit the registers of one resident wave (or is not 16-byte aligned)"); \
         return SCB_ERR_UNSUPPORTED; } while (0)

extern "C" int scb_bn_forward_fused(const void* x, const void* z, void* y, int dtype, int64_t rows, int channels,
                                    const float* weight, const float* bias, float* running_mean,
                                    float* running_var, float momentum, float eps, float* save_mean,
                                    float* save_invstd, float* save_count, long long* num_batches_tracked, int flags,
                                    int param_dtype, void* workspace,
                                    int64_t workspace_bytes, void* stream_) {
    if (int e = cl_check(x, dtype, rows, channels)) return e;
    SCB_REQUIRE(y != nullptr, "bn_forward_fused: null output");
    const bool add = (flags & SCB_BN_ADD) != 0;
    SCB_REQUIRE(!add || z != nullptr, "bn_forward_fused: SCB_BN_ADD needs the residual operand z");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t M = rows;
    const int C = channels;
    if (!cl_aligned16(x) || !cl_aligned16(y) || (add && !cl_aligned16(z))) SCB_CL_UNSUPPORTED("bn_forward_fused");
    if (cl_use_resident(dtype, M, C, 0, flags) && workspace != nullptr) {
        const int r = ring_forward_resident(x, add ? z : nullptr, y, dtype, M, C, weight, bias, running_mean, running_var,
                                            momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, flags, param_dtype,
                                            (char*)workspace + legacy_bn_ws_bytes(C),
                                            workspace_bytes - legacy_bn_ws_bytes(C), stream);
        if (r != kRingUnsupported) return r;
    }
    const int pick = cl_pick(M, C, dtype == SCB_F32 ? 4 : 2, 0, flags);
    if (pick == 0) SCB_CL_UNSUPPORTED("bn_forward_fused");
    if (pick == 2) {
        SCB_DISPATCH_FLOAT(dtype, T, {
            constexpr int VEC = Vec16<T>::N;
            ClStream g;
            if (!cl_stream_geom(M, C, sizeof(T), &g)) SCB_CL_UNSUPPORTED("bn_forward_fused");
            const size_t smem = cl_smem_bytes(g.TX, VEC);
            const T* xp = (const T*)x; const T* zp = (const T*)z; T* yp = (T*)y;
            if (add)
                SCB_CUDA_OK(cl_launch_stream(bn_fwd_stream_kernel<T, true>, g, smem, stream, xp, zp, yp, M, C, g, weight, bias,
                                             running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, flags, param_dtype));
            else
                SCB_CUDA_OK(cl_launch_stream(bn_fwd_stream_kernel<T, false>, g, smem, stream, xp, zp, yp, M, C, g, weight, bias,
                                             running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, flags, param_dtype));
        });
        count_launch();
        return SCB_OK;
    }
    SCB_DISPATCH_FLOAT(dtype, T, {
        constexpr int VEC = Vec16<T>::N;
        ClGeom g;
        int64_t cost = 0, cost2 = 0;
        bool ok = cl_geom(M, C, sizeof(T), kClRows, &g, &cost);
        int nrows = kClRows;
        if (!add) {          // twice the rows per thread: wider slabs may fit one wave (cfg 2: 16 slabs of 32 bytes x 7 CTAs)
            ClGeom g2;
            if (cl_geom(M, C, sizeof(T), 2 * kClRows, &g2, &cost2) && (!ok || cost2 < cost)) {
                g = g2; ok = true; nrows = 2 * kClRows;
            }
        }
        if (!ok) SCB_CL_UNSUPPORTED("bn_forward_fused");
        const size_t smem = cl_smem_bytes(g.TX, VEC);
        const T* xp = (const T*)x; const T* zp = (const T*)z; T* yp = (T*)y;
        float* none = nullptr;
        if (add)
            SCB_CUDA_OK(cl_launch(bn_fwd_cluster_kernel<T, 0, true, kClRows>, g, smem, stream, xp, zp, yp, M, C, g, weight, bias,
                                  running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, none, flags, param_dtype));
        else if (nrows == kClRows)
            SCB_CUDA_OK(cl_launch(bn_fwd_cluster_kernel<T, 0, false, kClRows>, g, smem, stream, xp, zp, yp, M, C, g, weight, bias,
                                  running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, none, flags, param_dtype));
        else
            SCB_CUDA_OK(cl_launch(bn_fwd_cluster_kernel<T, 0, false, 2 * kClRows>, g, smem, stream, xp, zp, yp, M, C, g, weight, bias,
                                  running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, none, flags, param_dtype));
    });
    count_launch();
    return SCB_OK;
}

// statistics only, through the same cluster reduction: the (2C+1)-float message of scb_bn_stats without scratch
extern "C" int scb_bn_stats_fused(const void* x, int dtype, int64_t rows, int channels, float* out_stats,
                                  void* stream_) {
    if (int e = cl_check(x, dtype, rows, channels)) return e;
    SCB_REQUIRE(out_stats != nullptr, "bn_stats_fused: null output");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t M = rows;
    const int C = channels;
    if (!cl_aligned16(x)) SCB_CL_UNSUPPORTED("bn_stats_fused");
    SCB_DISPATCH_FLOAT(dtype, T, {
        constexpr int VEC = Vec16<T>::N;
        ClGeom g;
        if (!cl_geom(M, C, sizeof(T), 2 * kClRows, &g)) SCB_CL_UNSUPPORTED("bn_stats_fused");
        const size_t smem = cl_smem_bytes(g.TX, VEC);
        const T* xp = (const T*)x; const T* zp = nullptr; T* yp = nullptr;
        float* none = nullptr; const float* cnone = nullptr;
        SCB_CUDA_OK(cl_launch(bn_fwd_cluster_kernel<T, 1, false, 2 * kClRows>, g, smem, stream, xp, zp, yp, M, C, g, cnone, cnone,
                              none, none, 0.f, 0.f, none, none, none, (long long*)nullptr, out_stats, 0, SCB_F32));
    });
    count_launch();
    return SCB_OK;
}

static int bn_backward_cluster(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype,
                               int64_t rows, int channels, const float* mean, const float* invstd,
                               const float* weight, const float* bias, const float* n_total, float* out_sums,
                               float* grad_weight, float* grad_bias, int flags, int param_dtype, int reduce_only,
                               void* workspace, int64_t workspace_bytes, void* stream_) {
    if (int e = cl_check(x, dtype, rows, channels)) return e;
    SCB_REQUIRE(dy && mean && invstd && n_total, "bn_backward_fused: null pointer");
    SCB_REQUIRE(reduce_only || dx != nullptr, "bn_backward_fused: null dx");
    const int gate = (flags & SCB_BN_RELU) ? ((flags & SCB_BN_ADD) ? 2 : 1) : 0;
    SCB_REQUIRE(gate != 2 || y != nullptr, "bn_backward_fused: SCB_BN_RELU|SCB_BN_ADD needs the saved output y");
    if (gate != 2) dz = nullptr;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t M = rows;
    const int C = channels;
    if (!cl_aligned16(x) || !cl_aligned16(dy) || (dx && !cl_aligned16(dx)) || (gate == 2 && !cl_aligned16(y)) ||
        (dz && !cl_aligned16(dz)))
        SCB_CL_UNSUPPORTED("bn_backward_fused");
    if (!reduce_only && workspace != nullptr && cl_use_resident(dtype, M, C, 1, flags)) {
        const int r = ring_backward_resident(dy, x, y, dx, dz, dtype, M, C, mean, invstd, weight, bias, n_total, out_sums,
                                             grad_weight, grad_bias, gate, param_dtype, (char*)workspace + legacy_bn_ws_bytes(C),
                                             workspace_bytes - legacy_bn_ws_bytes(C), stream);
        if (r != kRingUnsupported) return r;
    }
    if (!reduce_only) {
        const int pick = cl_pick(M, C, dtype == SCB_F32 ? 4 : 2, 1, flags);
        if (pick == 0) SCB_CL_UNSUPPORTED("bn_backward_fused");
        if (pick == 2) {
            SCB_DISPATCH_FLOAT(dtype, T, {
                constexpr int VEC = Vec16<T>::N;
                ClStream g;
                if (!cl_stream_geom(M, C, sizeof(T), &g)) SCB_CL_UNSUPPORTED("bn_backward_fused");
                const size_t smem = cl_smem_bytes(g.TX, VEC);
                const T* dyp = (const T*)dy; const T* xp = (const T*)x; const T* yp = (const T*)y;
                T* dxp = (T*)dx; T* dzp = (T*)dz;
                cudaError_t err;
                if (gate == 2) err = cl_launch_stream(bn_bwd_stream_kernel<T, 2>, g, smem, stream, dyp, xp, yp, dxp, dzp, M, C, g, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
                else if (gate == 1) err = cl_launch_stream(bn_bwd_stream_kernel<T, 1>, g, smem, stream, dyp, xp, yp, dxp, dzp, M, C, g, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
                else err = cl_launch_stream(bn_bwd_stream_kernel<T, 0>, g, smem, stream, dyp, xp, yp, dxp, dzp, M, C, g, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
                SCB_CUDA_OK(err);
            });
            count_launch();
            return SCB_OK;
        }
    }
    SCB_DISPATCH_FLOAT(dtype, T, {
        ClGeom g;
        if (!cl_geom(M, C, sizeof(T), kClRows, &g)) SCB_CL_UNSUPPORTED("bn_backward_fused");
        cudaError_t err;
        // 16-byte slabs in registers, but 32 bytes or more with 12 rows per thread parked in shared memory: take those
        ClGeom gp;
        static const bool park_on = [] { const char* e = getenv("SCB_BN_PARK"); return e == nullptr || e[0] != '0'; }();
        if (park_on && !reduce_only && gate != 2 && g.TX == 1 && cl_geom(M, C, sizeof(T), kClSmemRows, &gp) && gp.TX >= 2 &&
            cl_smem_park_bytes(gp.TX, Vec16<T>::N) <= 220 * 1024) {
            const size_t smem = cl_smem_park_bytes(gp.TX, Vec16<T>::N);
            if (gate == 1)
                err = cl_launch(bn_bwd_cluster_smem_kernel<T, 1>, gp, smem, stream, (const T*)dy, (const T*)x, (T*)dx, M, C, gp,
                                mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
            else
                err = cl_launch(bn_bwd_cluster_smem_kernel<T, 0>, gp, smem, stream, (const T*)dy, (const T*)x, (T*)dx, M, C, gp,
                                mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
        } else
        if (reduce_only) {
            if (gate == 2) err = cl_launch_bwd<T, 1, 2>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
            else if (gate == 1) err = cl_launch_bwd<T, 1, 1>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
            else err = cl_launch_bwd<T, 1, 0>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
        } else {
            if (gate == 2) err = cl_launch_bwd<T, 0, 2>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
            else if (gate == 1) err = cl_launch_bwd<T, 0, 1>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
            else err = cl_launch_bwd<T, 0, 0>(g, stream, dy, x, y, dx, dz, M, C, mean, invstd, weight, bias, n_total, out_sums, grad_weight, grad_bias, param_dtype);
        }
        SCB_CUDA_OK(err);
    });
    count_launch();
    return SCB_OK;
}

extern "C" int scb_bn_backward_fused(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype,
                                     int64_t rows, int channels, const float* mean, const float* invstd,
                                     const float* weight, const float* bias, const float* n_total, float* out_sums,
                                     float* grad_weight, float* grad_bias, int flags, int param_dtype, void* workspace,
                                     int64_t workspace_bytes, void* stream_) {
    return bn_backward_cluster(dy, x, y, dx, dz, dtype, rows, channels, mean, invstd, weight, bias, n_total, out_sums,
                               grad_weight, grad_bias, flags, param_dtype, 0, workspace, workspace_bytes, stream_);
}

extern "C" int scb_bn_backward_reduce_fused(const void* dy, const void* x, const void* y, int dtype, int64_t rows,
                                            int channels, const float* mean, const float* invstd,
                                            const float* weight, const float* bias, const float* n_total,
                                            float* out_sums, float* grad_weight, float* grad_bias, int flags,
                                            int param_dtype, void* stream_) {
    SCB_REQUIRE(out_sums != nullptr, "bn_backward_reduce_fused: null out_sums");
    return bn_backward_cluster(dy, x, y, nullptr, nullptr, dtype, rows, channels, mean, invstd, weight, bias, n_total,
                               out_sums, grad_weight, grad_bias, flags, param_dtype, 1, nullptr, 0, stream_);
}
