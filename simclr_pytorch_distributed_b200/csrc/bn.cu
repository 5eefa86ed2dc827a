// SyncBatchNorm hot path for sm_100a: HBM-bound streaming kernels.
//
// Replaces the five ATen ops nn.SyncBatchNorm runs per layer (torch/nn/modules/_functions.py:39,
// 106,122,144,170; kernels in ATen/native/cuda/Normalization.cuh:970-1433) with four launches:
//   bn_stats          x -> (mean, M2, count)                      1 read of x
//   bn_apply          merge ranks' stats + running stats + y      1 read of x, 1 write of y
//   bn_backward_reduce  dy, x -> (sum dy, sum dy*(x-mean)), dgamma, dbeta   2 reads
//   bn_backward_apply   dy, x -> dx                                2 reads, 1 write
// = 8 tensor passes per forward+backward, the algorithmic figure of SURVEY.md section 8(d).
//
// Layout in HBM: channels-last activations are a row-major [M, C] matrix (M = N*H*W).  A thread
// owns VEC = 16/sizeof(T) consecutive channels and walks down the rows, so every global access is
// a 16-byte vector and a warp touches 512 contiguous bytes (C >= 256) or several whole rows
// (C < 256).  Partial sums go warp -> shared memory -> one float atomic per channel per CTA into
// a zeroed accumulator; the last CTA to finish (ticket counter) turns the accumulators into the
// result and zeroes them again, so the scratch is reusable without a memset.
#include "common.cuh"

#include <map>
#include <mutex>
#include <utility>

#include "bn_ring.cuh"

namespace scb {

constexpr int kBnThreads = 256;
constexpr int kUnroll = 4;
constexpr int kStatsUnroll = 8;     // rows in flight per thread in the read-only statistics pass

// ------------------------------------------------------------------ NHWC launch geometry
struct NhwcGeom {
    int vec, tx, ty, grid_x;
    int64_t rows_per_cta, grid_y;
};

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

static int nhwc_vec(int C, int elem_size, bool aligned16) {
    const int vmax = 16 / elem_size;
    return (aligned16 && C % vmax == 0) ? vmax : 1;
}

static NhwcGeom nhwc_geom(int64_t M, int C, int elem_size, bool aligned16, int target_ctas,
                          int threads = kBnThreads) {
    NhwcGeom g;
    int vmax = 16 / elem_size;
    g.vec = (aligned16 && C % vmax == 0) ? vmax : 1;
    int cv = C / g.vec;
    g.tx = next_pow2(cv) < 32 ? next_pow2(cv) : 32;
    g.ty = threads / g.tx;
    g.grid_x = (int)ceil_div(cv, g.tx);
    int64_t want_y = target_ctas / g.grid_x; if (want_y < 1) want_y = 1;
    int64_t chunk = (int64_t)g.ty * kUnroll;
    int64_t rpc = ceil_div(ceil_div(M, want_y), chunk) * chunk;
    if (rpc < chunk) rpc = chunk;
    g.rows_per_cta = rpc;
    g.grid_y = ceil_div(M, rpc); if (g.grid_y < 1) g.grid_y = 1;
    return g;
}

// Cross-CTA sums go through float atomics in L2, and hundreds of CTAs adding into the same few cache
// lines serialise there (~27 cycles per 128-byte warp-op per line).  The accumulators are therefore
// replicated R times (CTA y uses replica y % R) and the R copies are added by whoever finalises.
__host__ __device__ inline int bn_replicas(int C) {
    (void)C;
    return 1;       // measured on B200: replication does not pay (the latency chain per CTA dominates)
}
// scratch layout: [R][2C] float accumulators, then two unsigned tickets (at a 16-byte aligned
// offset).  Zero on entry, zero again on exit.
__host__ __device__ inline int64_t bn_ws_counter_offset(int C) {
    return ((int64_t)bn_replicas(C) * 2 * C * 4 + 15) / 16 * 16;
}

// Block-level: sum `v[VEC]` over the ty rows of the CTA and atomically add into acc[c].
template <int VEC>
__device__ __forceinline__ void cta_reduce_to_global(float (&a)[VEC], float (&b)[VEC], float* sm,
                                                     int tx, int ty, int TX, int TY, int c0, int C,
                                                     float* acc_a, float* acc_b) {
    // sm: [2][TY][TX*VEC]
    const int W = TX * VEC;
    float* sa = sm;
    float* sb = sm + TY * W;
    if constexpr (VEC % 4 == 0) {      // 16-byte shared stores: scalar ones would be 8-way bank conflicted
#pragma unroll
        for (int i = 0; i < VEC; i += 4) {
            *reinterpret_cast<float4*>(sa + ty * W + tx * VEC + i) = make_float4(a[i], a[i + 1], a[i + 2], a[i + 3]);
            *reinterpret_cast<float4*>(sb + ty * W + tx * VEC + i) = make_float4(b[i], b[i + 1], b[i + 2], b[i + 3]);
        }
    } else {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            sa[ty * W + tx * VEC + i] = a[i];
            sb[ty * W + tx * VEC + i] = b[i];
        }
    }
    __syncthreads();
    const int tile_c = c0 - tx * VEC;         // first channel of this CTA's tile
    for (int t = threadIdx.x; t < 2 * W; t += blockDim.x) {
        const int which = t / W, col = t - which * W;
        const float* s = which ? sb : sa;
        float tot = 0.f;
        for (int r = 0; r < TY; ++r) tot += s[r * W + col];
        const int c = tile_c + col;
        if (c < C) atomicAdd((which ? acc_b : acc_a) + c, tot);
    }
}

// returns true in every thread of the last CTA of the grid to get here
__device__ __forceinline__ bool last_cta_done(unsigned* counter, unsigned total) {
    __shared__ unsigned s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        // acq_rel: a release cumulative over the whole CTA's atomics (ordered before it by the barrier) and the
        // acquire the finalising CTA needs -- a relaxed atomicAdd behind per-thread fences is neither
        unsigned t;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], %2;" : "=r"(t) : "l"(counter), "r"(1u) : "memory");
        s_last = (t == total - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0u;
}

// ------------------------------------------------------------------ B1: statistics (NHWC)
// Shifted sums: every CTA uses the same per-channel shift K_c = x[0, c], accumulates
// S1 = sum (x - K), S2 = sum (x - K)^2, which add across CTAs; mean = K + S1/n,
// M2 = S2 - S1^2/n.  The shift removes the E[x^2]-E[x]^2 cancellation.
template <typename T, int VEC>
__global__ void __launch_bounds__(kBnThreads)
bn_stats_nhwc_kernel(const T* __restrict__ x, int64_t M, int C, int TX, int64_t rows_per_cta,
                     float* __restrict__ acc, unsigned* counter, float* __restrict__ out_stats) {
    extern __shared__ float sm[];
    const int TY = kBnThreads / TX;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = (blockIdx.x * TX + tx) * VEC;
    const bool active = c0 < C;
    float s1[VEC], s2[VEC], shift[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; shift[i] = 0.f; }
    if (active) {
        // The thread accumulates around its own first row (no dependent load in front of the
        // stream); the common reference K_c = x[0, c] is fetched alongside and only used at the
        // end to re-centre the two sums so that they add across threads and CTAs.
        float k0[VEC];
        load_row_vec_cached<T, VEC>(x + c0, k0);
        const int64_t r_end = min(M, (int64_t)(blockIdx.y + 1) * rows_per_cta);
        int64_t r = (int64_t)blockIdx.y * rows_per_cta + ty;
        const int64_t r_first = r;
        const T* p = x + r * C + c0;
        const int64_t step = (int64_t)TY * C;
        if (r < r_end) load_row_vec<T, VEC>(p, shift);
        for (; r + (kStatsUnroll - 1) * TY < r_end; r += kStatsUnroll * TY, p += kStatsUnroll * step) {
            float v[kStatsUnroll][VEC];
#pragma unroll
            for (int u = 0; u < kStatsUnroll; ++u) load_row_vec<T, VEC>(p + u * step, v[u]);
#pragma unroll
            for (int u = 0; u < kStatsUnroll; ++u)
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float d = v[u][i] - shift[i];
                    s1[i] += d;
                    s2[i] = fmaf(d, d, s2[i]);
                }
        }
        for (; r < r_end; r += TY, p += step) {
            float v[VEC];
            load_row_vec<T, VEC>(p, v);
#pragma unroll
            for (int i = 0; i < VEC; ++i) {
                const float d = v[i] - shift[i];
                s1[i] += d;
                s2[i] = fmaf(d, d, s2[i]);
            }
        }
        // re-centre on K_c:  sum (x-K) = s1 + n d,  sum (x-K)^2 = s2 + 2 d s1 + n d^2,  d = shift - K
        const float nt = r_first < r_end ? (float)((r_end - r_first + TY - 1) / TY) : 0.f;
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float d = shift[i] - k0[i];
            s2[i] = s2[i] + d * (2.f * s1[i] + nt * d);
            s1[i] = fmaf(nt, d, s1[i]);
        }
    }
    const int R = bn_replicas(C);
    float* my_acc = acc + (size_t)(blockIdx.y % R) * 2 * C;
    cta_reduce_to_global<VEC>(s1, s2, sm, tx, ty, TX, TY, c0, C, my_acc, my_acc + C);
    if (last_cta_done(counter, gridDim.x * gridDim.y)) {
        const float n = (float)M;
        for (int c = threadIdx.x; c < C; c += kBnThreads) {
            float a = 0.f, b = 0.f;
            for (int rr = 0; rr < R; ++rr) {
                a += __ldcg(acc + (size_t)rr * 2 * C + c);
                b += __ldcg(acc + (size_t)rr * 2 * C + C + c);
                acc[(size_t)rr * 2 * C + c] = 0.f;
                acc[(size_t)rr * 2 * C + C + c] = 0.f;
            }
            const float k = to_f32(x[c]);
            out_stats[c] = k + a / n;
            out_stats[C + c] = fmaxf(b - a * a / n, 0.f);
        }
        if (threadIdx.x == 0) { out_stats[2 * C] = n; *counter = 0u; }
    }
}

// ------------------------------------------------------------------ per-channel coefficients
// Merge the ranks' (mean, M2, n) messages for channel c (Normalization.cuh:458-502 restated
// on M2 instead of 1/invstd^2 - eps), or take the running statistics in eval mode.
__device__ __forceinline__ void merged_stats(const float* __restrict__ stats_all, int n_ranks, int C,
                                             int c, const float* running_mean,
                                             const float* running_var, float eps,
                                             float& mean, float& invstd, float& m2, float& n, int pdt) {
    if (stats_all == nullptr) {
        mean = ldp(running_mean, pdt, c, 0.f);
        invstd = rsqrtf(ldp(running_var, pdt, c, 1.f) + eps);
        m2 = 0.f; n = 1.f;      // eval: statistics are constants, the backward's 1/n terms vanish
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

// ------------------------------------------------------------------ B2+B3: apply (NHWC)
// y = x * a_c + b_c (+ z) (ReLU).  The CTA walks its rows from the LAST chunk to the first: the statistics pass
// before it read the same rows front to back, so the tail of every CTA's range is what the L2 still holds.
template <typename T, int VEC, bool kAdd>
__global__ void __launch_bounds__(kBnThreads)
bn_apply_nhwc_kernel(const T* __restrict__ x, const T* __restrict__ z, T* __restrict__ y, int64_t M, int C, int TX,
                     int64_t rows_per_cta, const float* __restrict__ stats_all, int n_ranks,
                     const float* __restrict__ weight, const float* __restrict__ bias,
                     float* running_mean, float* running_var, float momentum, float eps,
                     float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags, int pdt) {
    extern __shared__ float sm[];           // [2][TX*VEC]: scale a, offset b
    const int TY = kBnThreads / TX;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int W = TX * VEC;
    const int tile_c0 = blockIdx.x * W;
    const int c0 = tile_c0 + tx * VEC;
    for (int t = threadIdx.x; t < W; t += kBnThreads) {
        const int c = tile_c0 + t;
        float a = 0.f, b = 0.f;
        if (c < C) {
            float mean, invstd, m2, n;
            merged_stats(stats_all, n_ranks, C, c, running_mean, running_var, eps, mean, invstd, m2, n, pdt);
            const float g = ldp(weight, pdt, c, 1.f);
            a = g * invstd;
            b = ldp(bias, pdt, c, 0.f) - mean * a;
            if (blockIdx.y == 0) {
                if (save_mean) save_mean[c] = mean;
                if (save_invstd) save_invstd[c] = invstd;
                if (stats_all != nullptr) {
                    if (running_mean) stp(running_mean, pdt, c, (1.f - momentum) * ldp(running_mean, pdt, c, 0.f) + momentum * mean);
                    if (running_var) stp(running_var, pdt, c, (1.f - momentum) * ldp(running_var, pdt, c, 0.f) + momentum * (m2 / (n - 1.f)));
                }
                if (c == 0 && save_count) save_count[0] = n;
                if (c == 0 && nbt && stats_all != nullptr) nbt[0] += 1;
            }
        }
        sm[t] = a;
        sm[W + t] = b;
    }
    __syncthreads();
    if (c0 >= C) return;
    float a[VEC], b[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { a[i] = sm[tx * VEC + i]; b[i] = sm[W + tx * VEC + i]; }
    const bool relu = (flags & SCB_BN_RELU) != 0;
    const int64_t r_beg = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r_end = min(M, r_beg + rows_per_cta);
    const int64_t chunk = (int64_t)kUnroll * TY;
    for (int64_t k = (r_end - r_beg + chunk - 1) / chunk - 1; k >= 0; --k) {
        const int64_t rk = r_beg + k * chunk + ty;
        float v[kUnroll][VEC], vz[kUnroll][VEC];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                load_row_vec<T, VEC>(x + r * C + c0, v[u]);
                if (kAdd) load_row_vec<T, VEC>(z + r * C + c0, vz[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    float o = fmaf(v[u][i], a[i], b[i]);
                    if (kAdd) o += vz[u][i];
                    v[u][i] = relu ? fmaxf(o, 0.f) : o;
                }
                store_row_vec<T, VEC>(y + r * C + c0, v[u]);
            }
        }
    }
}

// gate of the fused ReLU in the backward passes:
//   kGate 0: none;  1: ReLU directly after the normalisation, y > 0 recomputed from x;
//   2: ReLU after a residual add (SCB_BN_RELU | SCB_BN_ADD): y > 0 read from the saved output
template <int kGate>
__device__ __forceinline__ float gate_dy(float g, float xm, float a, float b, float yv) {
    if (kGate == 1) return (fmaf(xm, a, b) > 0.f) ? g : 0.f;
    if (kGate == 2) return (yv > 0.f) ? g : 0.f;
    return g;
}

// ------------------------------------------------------------------ B4: backward reduce (NHWC)
template <typename T, int VEC, int kGate>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_nhwc_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout, int64_t M,
                          int C, int TX, int64_t rows_per_cta, const float* __restrict__ mean,
                          const float* __restrict__ invstd, const float* __restrict__ weight,
                          const float* __restrict__ bias, float* __restrict__ acc, unsigned* counter,
                          float* __restrict__ out_sums, float* grad_weight, float* grad_bias, int pdt) {
    extern __shared__ float sm[];
    const int TY = kBnThreads / TX;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = (blockIdx.x * TX + tx) * VEC;
    const bool active = c0 < C;
    float s1[VEC], s2[VEC], mu[VEC], a[VEC], b[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) { s1[i] = 0.f; s2[i] = 0.f; mu[i] = 0.f; a[i] = 0.f; b[i] = 0.f; }
    if (active) {
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            mu[i] = mean[c0 + i];
            if (kGate == 1) {   // y = (x - mu) * a + b, gate dy by y > 0
                a[i] = ldp(weight, pdt, c0 + i, 1.f) * invstd[c0 + i];
                b[i] = ldp(bias, pdt, c0 + i, 0.f);
            }
        }
        const int64_t r_end = min(M, (int64_t)(blockIdx.y + 1) * rows_per_cta);
        constexpr int U = 4;   // 4 rows x 2 (3) tensors = 8 (12) vector loads in flight
        for (int64_t rk = (int64_t)blockIdx.y * rows_per_cta + ty; rk < r_end; rk += (int64_t)U * TY) {
            float vx[U][VEC], vd[U][VEC], vy[U][VEC];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = rk + (int64_t)u * TY;
                if (r < r_end) {
                    load_row_vec<T, VEC>(x + r * C + c0, vx[u]);
                    load_row_vec<T, VEC>(dy + r * C + c0, vd[u]);
                    if (kGate == 2) load_row_vec<T, VEC>(yout + r * C + c0, vy[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t r = rk + (int64_t)u * TY;
                if (r < r_end) {
#pragma unroll
                    for (int i = 0; i < VEC; ++i) {
                        const float xm = vx[u][i] - mu[i];
                        const float g = gate_dy<kGate>(vd[u][i], xm, a[i], b[i], kGate == 2 ? vy[u][i] : 0.f);
                        s1[i] += g;
                        s2[i] = fmaf(g, xm, s2[i]);
                    }
                }
            }
        }
    }
    const int R = bn_replicas(C);
    float* my_acc = acc + (size_t)(blockIdx.y % R) * 2 * C;
    cta_reduce_to_global<VEC>(s1, s2, sm, tx, ty, TX, TY, c0, C, my_acc, my_acc + C);
    if (last_cta_done(counter, gridDim.x * gridDim.y)) {
        for (int c = threadIdx.x; c < C; c += kBnThreads) {
            float sd = 0.f, sx = 0.f;
            for (int rr = 0; rr < R; ++rr) {
                sd += __ldcg(acc + (size_t)rr * 2 * C + c);
                sx += __ldcg(acc + (size_t)rr * 2 * C + C + c);
                acc[(size_t)rr * 2 * C + c] = 0.f;
                acc[(size_t)rr * 2 * C + C + c] = 0.f;
            }
            out_sums[c] = sd;
            out_sums[C + c] = sx;
            if (grad_weight) stp(grad_weight, pdt, c, sx * invstd[c]);
            if (grad_bias) stp(grad_bias, pdt, c, sd);
        }
        if (threadIdx.x == 0) *counter = 0u;
    }
}

// ------------------------------------------------------------------ B5: backward apply (NHWC)
// rows walked back to front (see bn_apply_nhwc_kernel); kGate 2 also writes dz = gated dy, the gradient of the
// residual operand
template <typename T, int VEC, int kGate>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_nhwc_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout,
                         T* __restrict__ dx, T* __restrict__ dz, int64_t M, int C, int TX, int64_t rows_per_cta,
                         const float* __restrict__ mean, const float* __restrict__ invstd,
                         const float* __restrict__ weight, const float* __restrict__ bias,
                         const float* __restrict__ sums, const float* __restrict__ n_total, int pdt) {
    const int TY = kBnThreads / TX;
    const int tx = threadIdx.x % TX, ty = threadIdx.x / TX;
    const int c0 = (blockIdx.x * TX + tx) * VEC;
    if (c0 >= C) return;
    const float inv_n = 1.f / n_total[0];
    // dx = (dy - k2 - (x - mu) * k3) * k1
    float mu[VEC], k1[VEC], k2[VEC], k3[VEC], bb[VEC];
#pragma unroll
    for (int i = 0; i < VEC; ++i) {
        const int c = c0 + i;
        const float is = invstd[c];
        mu[i] = mean[c];
        k1[i] = ldp(weight, pdt, c, 1.f) * is;
        k2[i] = sums[c] * inv_n;
        k3[i] = is * is * sums[C + c] * inv_n;
        bb[i] = ldp(bias, pdt, c, 0.f);
    }
    const int64_t r_beg = (int64_t)blockIdx.y * rows_per_cta;
    const int64_t r_end = min(M, r_beg + rows_per_cta);
    constexpr int U = 4;
    const int64_t chunk = (int64_t)U * TY;
    for (int64_t k = (r_end - r_beg + chunk - 1) / chunk - 1; k >= 0; --k) {
        const int64_t rk = r_beg + k * chunk + ty;
        float vx[U][VEC], vd[U][VEC], vy[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
                load_row_vec<T, VEC>(x + r * C + c0, vx[u]);
                load_row_vec<T, VEC>(dy + r * C + c0, vd[u]);
                if (kGate == 2) load_row_vec<T, VEC>(yout + r * C + c0, vy[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t r = rk + (int64_t)u * TY;
            if (r < r_end) {
#pragma unroll
                for (int i = 0; i < VEC; ++i) {
                    const float xm = vx[u][i] - mu[i];
                    const float g = gate_dy<kGate>(vd[u][i], xm, k1[i], bb[i], kGate == 2 ? vy[u][i] : 0.f);
                    vd[u][i] = g;
                    vx[u][i] = (g - k2[i] - xm * k3[i]) * k1[i];
                }
                store_row_vec<T, VEC>(dx + r * C + c0, vx[u]);
                if (kGate == 2 && dz != nullptr) store_row_vec<T, VEC>(dz + r * C + c0, vd[u]);
            }
        }
    }
}

// ================================================================== NCHW ([N, C, inner]) family
// One CTA per (channel, slice of N); threads walk the contiguous `inner` run of each sample.
// This is the layout the reference model actually runs (no channels_last anywhere in
// main_supcon.py), kept correct and vectorised but not the tuned path (BASELINE cfg 2 is NHWC).
template <typename T, int VEC, typename F>
__device__ __forceinline__ void nchw_foreach(int64_t n_begin, int64_t n_end, int C, int64_t inner,
                                             int c, F&& f) {
    // element index space of this CTA: (n, i) with i a multiple of VEC
    const int64_t per_n = inner / VEC;
    const int64_t total = (n_end - n_begin) * per_n;
    for (int64_t e = threadIdx.x; e < total; e += kBnThreads) {
        const int64_t n = n_begin + e / per_n;
        const int64_t i = (e % per_n) * VEC;
        f((n * C + c) * inner + i);
    }
}

__device__ __forceinline__ float block_sum(float v, float* sm) {
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    __syncthreads();
    if (l == 0) sm[w] = v;
    __syncthreads();
    float t = (threadIdx.x < kBnThreads / 32) ? sm[threadIdx.x] : 0.f;
    if (w == 0) t = warp_sum(t);
    return t;   // valid in warp 0
}

template <typename T, int VEC>
__global__ void __launch_bounds__(kBnThreads)
bn_stats_nchw_kernel(const T* __restrict__ x, int64_t N, int C, int64_t inner, int64_t n_per_cta,
                     float* __restrict__ acc, unsigned* counter, float* __restrict__ out_stats) {
    __shared__ float sm[32];
    const int c = blockIdx.x;
    const float shift = to_f32(x[(int64_t)c * inner]);
    const int64_t n0 = (int64_t)blockIdx.y * n_per_cta, n1 = min(N, n0 + n_per_cta);
    float s1 = 0.f, s2 = 0.f;
    nchw_foreach<T, VEC>(n0, n1, C, inner, c, [&](int64_t off) {
        float v[VEC];
        load_row_vec<T, VEC>(x + off, v);
#pragma unroll
        for (int i = 0; i < VEC; ++i) { const float d = v[i] - shift; s1 += d; s2 = fmaf(d, d, s2); }
    });
    s1 = block_sum(s1, sm);
    s2 = block_sum(s2, sm);
    if (threadIdx.x == 0) { atomicAdd(acc + c, s1); atomicAdd(acc + C + c, s2); }
    if (last_cta_done(counter, gridDim.x * gridDim.y)) {
        const float n = (float)(N * inner);
        for (int cc = threadIdx.x; cc < C; cc += kBnThreads) {
            const float a = __ldcg(acc + cc), b = __ldcg(acc + C + cc);
            const float k = to_f32(x[(int64_t)cc * inner]);
            out_stats[cc] = k + a / n;
            out_stats[C + cc] = fmaxf(b - a * a / n, 0.f);
            acc[cc] = 0.f; acc[C + cc] = 0.f;
        }
        if (threadIdx.x == 0) { out_stats[2 * C] = n; *counter = 0u; }
    }
}

template <typename T, int VEC>
__global__ void __launch_bounds__(kBnThreads)
bn_apply_nchw_kernel(const T* __restrict__ x, const T* __restrict__ z, T* __restrict__ y, int64_t N, int C, int64_t inner,
                     int64_t n_per_cta, const float* __restrict__ stats_all, int n_ranks,
                     const float* __restrict__ weight, const float* __restrict__ bias,
                     float* running_mean, float* running_var, float momentum, float eps,
                     float* save_mean, float* save_invstd, float* save_count, long long* nbt, int flags, int pdt) {
    const int c = blockIdx.x;
    float mean, invstd, m2, n;
    merged_stats(stats_all, n_ranks, C, c, running_mean, running_var, eps, mean, invstd, m2, n, pdt);
    const float a = ldp(weight, pdt, c, 1.f) * invstd;
    const float b = ldp(bias, pdt, c, 0.f) - mean * a;
    __syncthreads();   // every thread has read the running stats before thread 0 rewrites them
    if (blockIdx.y == 0 && threadIdx.x == 0) {
        if (save_mean) save_mean[c] = mean;
        if (save_invstd) save_invstd[c] = invstd;
        if (stats_all != nullptr) {
            if (running_mean) stp(running_mean, pdt, c, (1.f - momentum) * ldp(running_mean, pdt, c, 0.f) + momentum * mean);
            if (running_var) stp(running_var, pdt, c, (1.f - momentum) * ldp(running_var, pdt, c, 0.f) + momentum * (m2 / (n - 1.f)));
        }
        if (c == 0 && save_count) save_count[0] = n;
        if (c == 0 && nbt && stats_all != nullptr) nbt[0] += 1;
    }
    const bool relu = (flags & SCB_BN_RELU) != 0;
    const int64_t n0 = (int64_t)blockIdx.y * n_per_cta, n1 = min(N, n0 + n_per_cta);
    nchw_foreach<T, VEC>(n0, n1, C, inner, c, [&](int64_t off) {
        float v[VEC], vz[VEC];
        load_row_vec<T, VEC>(x + off, v);
        if (z != nullptr) load_row_vec<T, VEC>(z + off, vz);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            float o = fmaf(v[i], a, b);
            if (z != nullptr) o += vz[i];
            v[i] = relu ? fmaxf(o, 0.f) : o;
        }
        store_row_vec<T, VEC>(y + off, v);
    });
}

template <typename T, int VEC>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_nchw_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout, int64_t N, int C,
                          int64_t inner, int64_t n_per_cta, const float* __restrict__ mean,
                          const float* __restrict__ invstd, const float* __restrict__ weight,
                          const float* __restrict__ bias, float* __restrict__ acc, unsigned* counter,
                          float* __restrict__ out_sums, float* grad_weight, float* grad_bias, int flags, int pdt) {
    __shared__ float sm[32];
    const int c = blockIdx.x;
    const float mu = mean[c];
    const bool relu = (flags & SCB_BN_RELU) != 0;
    const bool by_y = relu && (flags & SCB_BN_ADD) != 0;          // gate read from the saved output
    const float a = relu ? ldp(weight, pdt, c, 1.f) * invstd[c] : 0.f;
    const float b = (relu) ? ldp(bias, pdt, c, 0.f) : 0.f;
    const int64_t n0 = (int64_t)blockIdx.y * n_per_cta, n1 = min(N, n0 + n_per_cta);
    float s1 = 0.f, s2 = 0.f;
    nchw_foreach<T, VEC>(n0, n1, C, inner, c, [&](int64_t off) {
        float vx[VEC], vd[VEC], vy[VEC];
        load_row_vec<T, VEC>(x + off, vx);
        load_row_vec<T, VEC>(dy + off, vd);
        if (by_y) load_row_vec<T, VEC>(yout + off, vy);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float xm = vx[i] - mu;
            float g = vd[i];
            if (by_y) g = (vy[i] > 0.f) ? g : 0.f;
            else if (relu) g = (fmaf(xm, a, b) > 0.f) ? g : 0.f;
            s1 += g; s2 = fmaf(g, xm, s2);
        }
    });
    s1 = block_sum(s1, sm);
    s2 = block_sum(s2, sm);
    if (threadIdx.x == 0) { atomicAdd(acc + c, s1); atomicAdd(acc + C + c, s2); }
    if (last_cta_done(counter, gridDim.x * gridDim.y)) {
        for (int cc = threadIdx.x; cc < C; cc += kBnThreads) {
            const float sd = __ldcg(acc + cc), sx = __ldcg(acc + C + cc);
            out_sums[cc] = sd; out_sums[C + cc] = sx;
            if (grad_weight) stp(grad_weight, pdt, cc, sx * invstd[cc]);
            if (grad_bias) stp(grad_bias, pdt, cc, sd);
            acc[cc] = 0.f; acc[C + cc] = 0.f;
        }
        if (threadIdx.x == 0) *counter = 0u;
    }
}

template <typename T, int VEC>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_nchw_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ yout,
                         T* __restrict__ dx, T* __restrict__ dz, int64_t N, int C, int64_t inner, int64_t n_per_cta,
                         const float* __restrict__ mean, const float* __restrict__ invstd,
                         const float* __restrict__ weight, const float* __restrict__ bias,
                         const float* __restrict__ sums, const float* __restrict__ n_total, int flags, int pdt) {
    const int c = blockIdx.x;
    const bool relu = (flags & SCB_BN_RELU) != 0;
    const float inv_n = 1.f / n_total[0];
    const float is = invstd[c], mu = mean[c];
    const float k1 = ldp(weight, pdt, c, 1.f) * is;
    const float k2 = sums[c] * inv_n;
    const float k3 = is * is * sums[C + c] * inv_n;
    const float bb = ldp(bias, pdt, c, 0.f);
    const int64_t n0 = (int64_t)blockIdx.y * n_per_cta, n1 = min(N, n0 + n_per_cta);
    const bool by_y = relu && (flags & SCB_BN_ADD) != 0;
    nchw_foreach<T, VEC>(n0, n1, C, inner, c, [&](int64_t off) {
        float vx[VEC], vd[VEC], vy[VEC];
        load_row_vec<T, VEC>(x + off, vx);
        load_row_vec<T, VEC>(dy + off, vd);
        if (by_y) load_row_vec<T, VEC>(yout + off, vy);
#pragma unroll
        for (int i = 0; i < VEC; ++i) {
            const float xm = vx[i] - mu;
            float g = vd[i];
            if (by_y) g = (vy[i] > 0.f) ? g : 0.f;
            else if (relu) g = (fmaf(xm, k1, bb) > 0.f) ? g : 0.f;
            vd[i] = g;
            vx[i] = (g - k2 - xm * k3) * k1;
        }
        store_row_vec<T, VEC>(dx + off, vx);
        if (by_y && dz != nullptr) store_row_vec<T, VEC>(dz + off, vd);
    });
}

// ------------------------------------------------------------------ host-side dispatch
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// CTAs of `kern` that are resident at once on the current device: the streaming kernels are launched as exactly one
// resident wave (a grid sized for 4 CTAs per SM when only 3 fit runs a second, nearly empty wave and takes twice as
// long).  Cached per (kernel, device).
template <typename Kern>
static int resident_ctas(Kern kern, size_t smem) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, int> cache;
    int dev = 0;
    cudaGetDevice(&dev);
    const std::pair<const void*, int> key((const void*)kern, dev);
    {
        std::lock_guard<std::mutex> lock(mu);
        auto it = cache.find(key);
        if (it != cache.end()) return it->second;
    }
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kBnThreads, smem) != cudaSuccess || occ < 1) {
        cudaGetLastError();
        occ = 2;
    }
    if (occ > 4) occ = 4;
    const int n = occ * sm_count();
    std::lock_guard<std::mutex> lock(mu);
    cache[key] = n;
    return n;
}

struct NchwGeom { int vec; int64_t n_per_cta, grid_y; };
static NchwGeom nchw_geom(int64_t N, int C, int64_t inner, int elem_size, bool al16, int target_ctas) {
    NchwGeom g;
    int vmax = 16 / elem_size;
    g.vec = (al16 && inner % vmax == 0) ? vmax : 1;
    int64_t want_y = target_ctas / C; if (want_y < 1) want_y = 1;
    if (want_y > N) want_y = N;
    g.n_per_cta = ceil_div(N, want_y);
    g.grid_y = ceil_div(N, g.n_per_cta);
    return g;
}

static int check_bn_common(const void* x, int dtype, int64_t outer, int channels, int64_t inner) {
    SCB_REQUIRE(x != nullptr, "bn: null tensor pointer");
    SCB_REQUIRE(dtype == SCB_F32 || dtype == SCB_BF16 || dtype == SCB_F16, "bn: unsupported dtype %d", dtype);
    SCB_REQUIRE(channels > 0, "SyncBatchNorm number of input channels should be non-zero");
    SCB_REQUIRE(outer >= 0 && inner >= 1, "bn: bad shape outer=%lld inner=%lld", (long long)outer, (long long)inner);
    SCB_REQUIRE(outer * inner < (int64_t)1 << 40, "bn: tensor too large");
    return SCB_OK;
}

}  // namespace scb

using namespace scb;

// scratch of the tiled / NCHW kernels: [2C] float accumulators + tickets
static int64_t legacy_ws_bytes(int channels) { return legacy_bn_ws_bytes(channels); }   // bn_replicas() == 1

// [scratch of the tiled kernels][scratch of the ring kernels (bn_ring.cu)]; a caller that passes only the first part
// (scb_bn_stats_noalloc does) gets the tiled kernels
extern "C" int64_t scb_bn_workspace_bytes(int channels) {
    if (channels <= 0) return 0;
    return legacy_ws_bytes(channels) + ring_ws_bytes(channels);
}

#define SCB_VEC_SWITCH(vecv, VEC_, ...)                                   \
    if ((vecv) == 1) { constexpr int VEC_ = 1; __VA_ARGS__; }            \
    else { constexpr int VEC_ = Vec16<T>::N; __VA_ARGS__; }

extern "C" int scb_bn_stats(const void* x, int dtype, int channels_last, int64_t outer, int channels,
                            int64_t inner, float* out_stats, void* workspace, int64_t workspace_bytes,
                            void* stream_) {
    if (int e = check_bn_common(x, dtype, outer, channels, inner)) return e;
    SCB_REQUIRE(out_stats && workspace, "bn_stats: null output/workspace");
    SCB_REQUIRE(workspace_bytes >= legacy_ws_bytes(channels), "bn_stats: workspace too small");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int C = channels;
    if (outer * inner == 0) {   // empty slice still sends a zero message (_functions.py:50-57)
        SCB_CUDA_OK(cudaMemsetAsync(out_stats, 0, (2 * C + 1) * sizeof(float), stream));
        return SCB_OK;
    }
    if (channels_last) {
        const int r = ring_stats(x, dtype, outer * inner, C, out_stats, (char*)workspace + legacy_ws_bytes(C),
                                 workspace_bytes - legacy_ws_bytes(C), stream);
        if (r != kRingUnsupported) return r;
    }
    float* acc = (float*)workspace;
    unsigned* counter = (unsigned*)((char*)workspace + bn_ws_counter_offset(C));
    const int target = 4 * sm_count();   // one resident wave: every load of the pass is in flight at once
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (channels_last) {
            const int64_t M = outer * inner;
            SCB_VEC_SWITCH(nhwc_vec(C, sizeof(T), aligned16(x)), VEC, {
                auto kern = bn_stats_nhwc_kernel<T, VEC>;
                const size_t smem = 2 * kBnThreads * VEC * sizeof(float);
                NhwcGeom g = nhwc_geom(M, C, sizeof(T), aligned16(x), resident_ctas(kern, smem));
                dim3 grid(g.grid_x, (unsigned)g.grid_y);
                kern<<<grid, kBnThreads, smem, stream>>>((const T*)x, M, C, g.tx, g.rows_per_cta, acc, counter, out_stats);
            });
        } else {
            NchwGeom g = nchw_geom(outer, C, inner, sizeof(T), aligned16(x), target);
            dim3 grid(C, (unsigned)g.grid_y);
            SCB_VEC_SWITCH(g.vec, VEC, (bn_stats_nchw_kernel<T, VEC><<<grid, kBnThreads, 0, stream>>>(
                (const T*)x, outer, C, inner, g.n_per_cta, acc, counter, out_stats)));
        }
    });
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

extern "C" int scb_bn_apply(const void* x, const void* z, void* y, int dtype, int channels_last, int64_t outer,
                            int channels, int64_t inner, const float* stats_all, int n_ranks,
                            const float* weight, const float* bias, float* running_mean,
                            float* running_var, float momentum, float eps, float* save_mean,
                            float* save_invstd, float* save_count, long long* num_batches_tracked, int flags,
                            int param_dtype, void* stream_) {
    if (int e = check_bn_common(x, dtype, outer, channels, inner)) return e;
    SCB_REQUIRE(y != nullptr, "bn_apply: null output");
    SCB_REQUIRE(stats_all != nullptr || (running_mean && running_var),
                "bn_apply: eval mode needs running_mean and running_var");
    SCB_REQUIRE(stats_all == nullptr || n_ranks >= 1, "bn_apply: n_ranks must be >= 1");
    const bool add = (flags & SCB_BN_ADD) != 0;
    SCB_REQUIRE(!add || z != nullptr, "bn_apply: SCB_BN_ADD needs the residual operand z");
    if (!add) z = nullptr;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int C = channels;
    if (outer * inner == 0) return SCB_OK;
    if (channels_last) {
        const int r = ring_apply(x, z, y, dtype, outer * inner, C, stats_all, n_ranks, weight, bias, running_mean,
                                 running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, flags, param_dtype,
                                 stream);
        if (r != kRingUnsupported) return r;
    }
    const int target = 4 * sm_count();
    const bool al = aligned16(x) && aligned16(y) && (!add || aligned16(z));
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (channels_last) {
            const int64_t M = outer * inner;
if (add) {
                SCB_VEC_SWITCH(nhwc_vec(C, sizeof(T), al), VEC, {
                    auto kern = bn_apply_nhwc_kernel<T, VEC, true>;
                    const size_t smem = 2 * 32 * VEC * sizeof(float);
                    NhwcGeom g = nhwc_geom(M, C, sizeof(T), al, resident_ctas(kern, smem));
                    dim3 grid(g.grid_x, (unsigned)g.grid_y);
                    kern<<<grid, kBnThreads, smem, stream>>>((const T*)x, (const T*)z, (T*)y, M, C, g.tx, g.rows_per_cta,
                        stats_all, n_ranks, weight, bias, running_mean, running_var, momentum, eps, save_mean,
                        save_invstd, save_count, num_batches_tracked, flags, param_dtype);
                });
            } else {
                SCB_VEC_SWITCH(nhwc_vec(C, sizeof(T), al), VEC, {
                    auto kern = bn_apply_nhwc_kernel<T, VEC, false>;
                    const size_t smem = 2 * 32 * VEC * sizeof(float);
                    NhwcGeom g = nhwc_geom(M, C, sizeof(T), al, resident_ctas(kern, smem));
                    dim3 grid(g.grid_x, (unsigned)g.grid_y);
                    kern<<<grid, kBnThreads, smem, stream>>>((const T*)x, (const T*)z, (T*)y, M, C, g.tx, g.rows_per_cta,
                        stats_all, n_ranks, weight, bias, running_mean, running_var, momentum, eps, save_mean,
                        save_invstd, save_count, num_batches_tracked, flags, param_dtype);
                });
            }
        } else {
            NchwGeom g = nchw_geom(outer, C, inner, sizeof(T), al, target);
            dim3 grid(C, (unsigned)g.grid_y);
            SCB_VEC_SWITCH(g.vec, VEC, (bn_apply_nchw_kernel<T, VEC><<<grid, kBnThreads, 0, stream>>>(
                (const T*)x, (const T*)z, (T*)y, outer, C, inner, g.n_per_cta, stats_all, n_ranks, weight, bias,
                running_mean, running_var, momentum, eps, save_mean, save_invstd, save_count, num_batches_tracked, flags,
                param_dtype)));
        }
    });
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

// gate mode of the backward kernels from the flags of the forward
static int gate_mode(int flags) { return (flags & SCB_BN_RELU) ? ((flags & SCB_BN_ADD) ? 2 : 1) : 0; }

#define SCB_GATE_SWITCH(gatev, GATE_, ...)                                \
    if ((gatev) == 2) { constexpr int GATE_ = 2; __VA_ARGS__; }          \
    else if ((gatev) == 1) { constexpr int GATE_ = 1; __VA_ARGS__; }     \
    else { constexpr int GATE_ = 0; __VA_ARGS__; }

extern "C" int scb_bn_backward_reduce(const void* dy, const void* x, const void* y, int dtype, int channels_last,
                                      int64_t outer, int channels, int64_t inner, const float* mean,
                                      const float* invstd, const float* weight, const float* bias,
                                      float* out_sums, float* grad_weight, float* grad_bias, int flags,
                                      int param_dtype, void* workspace, int64_t workspace_bytes, void* stream_) {
    if (int e = check_bn_common(x, dtype, outer, channels, inner)) return e;
    SCB_REQUIRE(dy && mean && invstd && out_sums && workspace, "bn_backward_reduce: null pointer");
    SCB_REQUIRE(workspace_bytes >= legacy_ws_bytes(channels), "bn_backward_reduce: workspace too small");
    const int gate = gate_mode(flags);
    SCB_REQUIRE(gate != 2 || y != nullptr, "bn_backward_reduce: SCB_BN_RELU|SCB_BN_ADD needs the saved output y");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int C = channels;
    if (outer * inner == 0) {
        SCB_CUDA_OK(cudaMemsetAsync(out_sums, 0, 2 * C * sizeof(float), stream));
        if (grad_weight) SCB_CUDA_OK(cudaMemsetAsync(grad_weight, 0, C * sizeof(float), stream));
        if (grad_bias) SCB_CUDA_OK(cudaMemsetAsync(grad_bias, 0, C * sizeof(float), stream));
        return SCB_OK;
    }
    if (channels_last) {
        const int r = ring_backward_reduce(dy, x, y, dtype, outer * inner, C, mean, invstd, weight, bias, out_sums,
                                           grad_weight, grad_bias, gate, param_dtype, (char*)workspace + legacy_ws_bytes(C),
                                           workspace_bytes - legacy_ws_bytes(C), stream);
        if (r != kRingUnsupported) return r;
    }
    float* acc = (float*)workspace;
    unsigned* counter = (unsigned*)((char*)workspace + bn_ws_counter_offset(C));
    const int target = 4 * sm_count();   // one resident wave: every load of the pass is in flight at once
    const bool al = aligned16(x) && aligned16(dy) && (gate != 2 || aligned16(y));
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (channels_last) {
            const int64_t M = outer * inner;
            SCB_GATE_SWITCH(gate, GATE, {
                SCB_VEC_SWITCH(nhwc_vec(C, sizeof(T), al), VEC, {
                    auto kern = bn_bwd_reduce_nhwc_kernel<T, VEC, GATE>;
                    const size_t smem = 2 * kBnThreads * VEC * sizeof(float);
                    NhwcGeom g = nhwc_geom(M, C, sizeof(T), al, resident_ctas(kern, smem));
                    dim3 grid(g.grid_x, (unsigned)g.grid_y);
                    kern<<<grid, kBnThreads, smem, stream>>>((const T*)dy, (const T*)x, (const T*)y, M, C, g.tx,
                        g.rows_per_cta, mean, invstd, weight, bias, acc, counter, out_sums, grad_weight, grad_bias, param_dtype);
                });
            });
        } else {
            NchwGeom g = nchw_geom(outer, C, inner, sizeof(T), al, target);
            dim3 grid(C, (unsigned)g.grid_y);
            SCB_VEC_SWITCH(g.vec, VEC, (bn_bwd_reduce_nchw_kernel<T, VEC><<<grid, kBnThreads, 0, stream>>>(
                (const T*)dy, (const T*)x, (const T*)y, outer, C, inner, g.n_per_cta, mean, invstd, weight, bias, acc,
                counter, out_sums, grad_weight, grad_bias, flags, param_dtype)));
        }
    });
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

extern "C" int scb_bn_backward_apply(const void* dy, const void* x, const void* y, void* dx, void* dz, int dtype,
                                     int channels_last, int64_t outer, int channels, int64_t inner,
                                     const float* mean, const float* invstd, const float* weight,
                                     const float* bias, const float* sums, const float* n_total,
                                     int flags, int param_dtype, void* stream_) {
    if (int e = check_bn_common(x, dtype, outer, channels, inner)) return e;
    SCB_REQUIRE(dy && dx && mean && invstd && sums && n_total, "bn_backward_apply: null pointer");
    const int gate = gate_mode(flags);
    SCB_REQUIRE(gate != 2 || y != nullptr, "bn_backward_apply: SCB_BN_RELU|SCB_BN_ADD needs the saved output y");
    if (gate != 2) dz = nullptr;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int C = channels;
    if (outer * inner == 0) return SCB_OK;
    if (channels_last) {
        const int r = ring_backward_apply(dy, x, y, dx, dz, dtype, outer * inner, C, mean, invstd, weight, bias, sums,
                                          n_total, gate, flags, param_dtype, stream);
        if (r != kRingUnsupported) return r;
    }
    const int target = 4 * sm_count();
    const bool al = aligned16(x) && aligned16(dy) && aligned16(dx) && (gate != 2 || aligned16(y)) &&
                    (dz == nullptr || aligned16(dz));
    SCB_DISPATCH_FLOAT(dtype, T, {
        if (channels_last) {
            const int64_t M = outer * inner;
            SCB_GATE_SWITCH(gate, GATE, {
                SCB_VEC_SWITCH(nhwc_vec(C, sizeof(T), al), VEC, {
                    auto kern = bn_bwd_apply_nhwc_kernel<T, VEC, GATE>;
                    NhwcGeom g = nhwc_geom(M, C, sizeof(T), al, resident_ctas(kern, 0));
                    dim3 grid(g.grid_x, (unsigned)g.grid_y);
                    kern<<<grid, kBnThreads, 0, stream>>>((const T*)dy, (const T*)x, (const T*)y, (T*)dx, (T*)dz, M, C,
                        g.tx, g.rows_per_cta, mean, invstd, weight, bias, sums, n_total, param_dtype);
                });
            });
        } else {
            NchwGeom g = nchw_geom(outer, C, inner, sizeof(T), al, target);
            dim3 grid(C, (unsigned)g.grid_y);
            SCB_VEC_SWITCH(g.vec, VEC, (bn_bwd_apply_nchw_kernel<T, VEC><<<grid, kBnThreads, 0, stream>>>(
                (const T*)dy, (const T*)x, (const T*)y, (T*)dx, (T*)dz, outer, C, inner, g.n_per_cta, mean, invstd, weight,
                bias, sums, n_total, flags, param_dtype)));
        }
    });
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

// ================================================================== allocation-free statistics
// Some hosts of this library are not allowed to ask CUDA for memory at the point where they call
// batch-norm statistics: an embedding framework whose own pool has already claimed the whole
// device (a cudaMalloc for our scratch would fail or fight that pool), or a call site where
// allocating is simply against the rules of the surrounding system.  For them the scratch of the
// statistics kernel lives inside the library image: a fixed array in device memory that every GPU
// loading the library carries until the process exits.  Nothing here calls cudaMalloc, cudaFree or
// any routine that allocates.
//
// kNoAllocMaxChannels bounds the array (2 floats per channel plus the ticket word).  Its name is a
// host-side stand-in only: the device address is looked up once per device with
// cudaGetSymbolAddress while that device is current, and that address is what the kernel receives.
constexpr int kNoAllocMaxChannels = SCB_BN_NOALLOC_MAX_CHANNELS;
constexpr size_t kNoAllocScratchBytes = (size_t)2 * kNoAllocMaxChannels * sizeof(float) + 256;   // >= legacy_ws_bytes(c) for c <= max
static __device__ __align__(256) unsigned char g_bn_noalloc_scratch[kNoAllocScratchBytes];

#include <mutex>
namespace {
std::mutex g_noalloc_mutex;
void* g_noalloc_ptr[64] = {nullptr};   // device address of the array, per device ordinal
}  // namespace

extern "C" int scb_bn_stats_noalloc(const void* x, int dtype, int channels_last, int64_t outer, int channels,
                                    int64_t inner, float* out_stats, void* stream_) {
    SCB_REQUIRE(channels >= 1 && channels <= kNoAllocMaxChannels,
                "bn_stats_noalloc: %d channels exceed the built-in scratch (at most %d); use scb_bn_stats with "
                "caller-provided scratch instead", channels, kNoAllocMaxChannels);
    int dev = 0;
    SCB_CUDA_OK(cudaGetDevice(&dev));
    SCB_REQUIRE(dev >= 0 && dev < 64, "bn_stats_noalloc: device ordinal %d out of range", dev);
    void* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (g_noalloc_ptr[dev] == nullptr) {
            // first call on this device: resolve the symbol to this device's copy of the array
            void* ptr = nullptr;
            SCB_CUDA_OK(cudaGetSymbolAddress((void**)&ptr, g_bn_noalloc_scratch));
            g_noalloc_ptr[dev] = ptr;
        }
        scratch = g_noalloc_ptr[dev];
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    // The array is shared by every call and keeps whatever the last one left behind; the kernel needs
    // its accumulators and ticket to start at zero, so clear the part this call uses.
    const int64_t need = legacy_ws_bytes(channels);      // the built-in array holds the tiled kernels' scratch only
    SCB_CUDA_OK(cudaMemsetAsync(scratch, 0, (size_t)need, stream));
    return scb_bn_stats(x, dtype, channels_last, outer, channels, inner, out_stats, scratch, need, stream_);
}

