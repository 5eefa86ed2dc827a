// SupCon / NT-Xent loss: layout kernels shared by every engine (prep, finalize, finish) and the
// fp32 SIMT engine (SCB_ENGINE_FP32): the exact-arithmetic path that accepts any feature length
// and any mask (user [B,B] float masks included).  The tensor-core engine lives in supcon_tc.cu.
//
// Reference being replaced: /root/reference/losses.py:52-91 and its autograd backward; closed
// form in SURVEY.md Appendix A.  Nothing of size [A, n_k] is ever written to memory.
#include "common.cuh"
#include "supcon_common.cuh"

namespace scb {

// ------------------------------------------------------------------ prep (normalise + stack + cast)
template <typename TI, typename TO>
__device__ __forceinline__ void prep_row_impl(const TI* feat, int n_samples, int dim, int64_t stride_s,
                                              int64_t stride_v, int normalize, float eps, TO* __restrict__ out_k,
                                              __nv_bfloat16* __restrict__ out_lo, int ldk,
                                              float* __restrict__ out_inv_norm, int warp, int lane);
template <typename TI, typename TO>
__device__ __forceinline__ void prep_row(const TI* feat, int n_samples, int dim, int64_t stride_s,
                                         int64_t stride_v, int normalize, float eps, TO* __restrict__ out_k,
                                         __nv_bfloat16* __restrict__ out_lo, int ldk, float* __restrict__ out_inv_norm,
                                         int warp, int lane) {
    prep_row_impl<TI, TO>(feat, n_samples, dim, stride_s, stride_v, normalize, eps, out_k, out_lo, ldk, out_inv_norm,
                          warp, lane);
}

template <typename TI, typename TO>
__global__ void __launch_bounds__(256)
supcon_prep_kernel(const TI* feat, int n_samples, int n_views, int dim,
                   int64_t stride_s, int64_t stride_v, int normalize, float eps,
                   TO* __restrict__ out_k, __nv_bfloat16* __restrict__ out_lo, int ldk,
                   float* __restrict__ out_inv_norm, const PeerOut po, int64_t row_offset) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int n_rows = n_samples * n_views;
    pdl_launch_dependents();
    pdl_wait();
    if (po.n_dst > 0) {
        // peer variant: the row goes to every rank's copy of the gathered matrix, then the flags are raised
        // (grid-stride over the rows: the CTA ticket of peer_signal must stay a few hundred atomics)
        const int n_warps = (gridDim.x * blockDim.x) >> 5;
        for (int row = warp; row < n_rows; row += n_warps)
            for (int w = 0; w < po.n_dst; ++w)
                prep_row<TI, TO>(feat, n_samples, dim, stride_s, stride_v, normalize, eps,
                                 reinterpret_cast<TO*>(po.dst[w]) + row_offset * ldk, nullptr, ldk,
                                 w == 0 ? out_inv_norm : nullptr, row, lane);
        peer_signal(po);
        return;
    }
    if (warp >= n_rows) return;
    prep_row<TI, TO>(feat, n_samples, dim, stride_s, stride_v, normalize, eps, out_k, out_lo, ldk, out_inv_norm, warp, lane);
}

template <typename TI, typename TO>
__device__ __forceinline__ void prep_row_impl(const TI* feat, int n_samples, int dim, int64_t stride_s,
                                              int64_t stride_v, int normalize, float eps, TO* __restrict__ out_k,
                                              __nv_bfloat16* __restrict__ out_lo, int ldk,
                                              float* __restrict__ out_inv_norm, int warp, int lane) {
    const int v = warp / n_samples, b = warp - v * n_samples;     // out row = v*n_samples + b
    const TI* src = feat + (int64_t)b * stride_s + (int64_t)v * stride_v;
    TO* dst = out_k + (int64_t)warp * ldk;
    // fast path (the common D = 128): a lane owns 4 consecutive elements, one vector access each way
    const bool vec4 = (ldk == 128) && (dim == 128) && !out_lo &&
                      ((reinterpret_cast<uintptr_t>(src) & (4 * sizeof(TI) - 1)) == 0) &&
                      ((reinterpret_cast<uintptr_t>(dst) & (4 * sizeof(TO) - 1)) == 0);
    if (vec4) {
        float x[4];
        if constexpr (sizeof(TI) == 4) {
            const float4 t = *reinterpret_cast<const float4*>(src + lane * 4);
            x[0] = t.x; x[1] = t.y; x[2] = t.z; x[3] = t.w;
        } else {
            const uint2 t = *reinterpret_cast<const uint2*>(src + lane * 4);
            const TI* e = reinterpret_cast<const TI*>(&t);
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = to_f32(e[i]);
        }
        float inv = 1.f;
        if (normalize) {
            const float ss = warp_sum(x[0] * x[0] + x[1] * x[1] + x[2] * x[2] + x[3] * x[3]);
            inv = 1.f / fmaxf(sqrtf(ss), eps);
        }
        if (lane == 0 && out_inv_norm) out_inv_norm[warp] = inv;
        if constexpr (sizeof(TO) == 4) {
            *reinterpret_cast<float4*>(dst + lane * 4) = make_float4(x[0] * inv, x[1] * inv, x[2] * inv, x[3] * inv);
        } else {
            uint2 o;
            TO* e = reinterpret_cast<TO*>(&o);
#pragma unroll
            for (int i = 0; i < 4; ++i) e[i] = from_f32<TO>(x[i] * inv);
            *reinterpret_cast<uint2*>(dst + lane * 4) = o;
        }
        return;
    }
    float inv = 1.f;
    if (normalize) {
        float ss = 0.f;
        for (int d = lane; d < dim; d += 32) { const float x = to_f32(src[d]); ss = fmaf(x, x, ss); }
        ss = warp_sum(ss);
        inv = 1.f / fmaxf(sqrtf(ss), eps);                         // F.normalize: x / max(||x||, eps)
    }
    if (lane == 0 && out_inv_norm) out_inv_norm[warp] = inv;
    for (int d = lane; d < ldk; d += 32) {
        const float x = d < dim ? to_f32(src[d]) * inv : 0.f;
        const TO hi = from_f32<TO>(x);
        dst[d] = hi;
        if (out_lo) out_lo[(int64_t)warp * ldk + d] = __float2bfloat16_rn(x - to_f32(hi));
    }
}

// ------------------------------------------------------------------ finish (un-stack + normalise bwd)
template <typename TF>
__global__ void __launch_bounds__(256)
supcon_finish_kernel(const float* dk, int ld_dk, const TF* feat,
                     int n_samples, int n_views, int dim, int64_t stride_s, int64_t stride_v,
                     int normalize, const float* inv_norm, TF* __restrict__ out,
                     int64_t ostride_s, int64_t ostride_v) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int n_rows = n_samples * n_views;
    pdl_launch_dependents();
    pdl_wait();
    if (warp >= n_rows) return;
    const int v = warp / n_samples, b = warp - v * n_samples;
    const float* g = dk + (int64_t)warp * ld_dk;
    TF* dst = out + (int64_t)b * ostride_s + (int64_t)v * ostride_v;
    // fast path (D = 128): 4 consecutive elements per lane, vector loads/stores
    if (dim == 128 && (ld_dk & 3) == 0 && (reinterpret_cast<uintptr_t>(dk) & 15) == 0 &&
        (reinterpret_cast<uintptr_t>(dst) & (4 * sizeof(TF) - 1)) == 0 &&
        (!normalize || (reinterpret_cast<uintptr_t>(feat + (int64_t)b * stride_s + (int64_t)v * stride_v) &
                        (4 * sizeof(TF) - 1)) == 0)) {
        const float4 gv = *reinterpret_cast<const float4*>(g + lane * 4);
        float o[4] = {gv.x, gv.y, gv.z, gv.w};
        if (normalize) {
            const TF* src = feat + (int64_t)b * stride_s + (int64_t)v * stride_v;
            const float inv = inv_norm[warp];
            float z[4];
            if constexpr (sizeof(TF) == 4) {
                const float4 t = *reinterpret_cast<const float4*>(src + lane * 4);
                z[0] = t.x; z[1] = t.y; z[2] = t.z; z[3] = t.w;
            } else {
                const uint2 t = *reinterpret_cast<const uint2*>(src + lane * 4);
                const TF* e = reinterpret_cast<const TF*>(&t);
#pragma unroll
                for (int i = 0; i < 4; ++i) z[i] = to_f32(e[i]);
            }
            float dot = 0.f;
#pragma unroll
            for (int i = 0; i < 4; ++i) { z[i] *= inv; dot = fmaf(o[i], z[i], dot); }
            dot = warp_sum(dot);
#pragma unroll
            for (int i = 0; i < 4; ++i) o[i] = (o[i] - z[i] * dot) * inv;
        }
        if constexpr (sizeof(TF) == 4) {
            *reinterpret_cast<float4*>(dst + lane * 4) = make_float4(o[0], o[1], o[2], o[3]);
        } else {
            uint2 pk;
            TF* e = reinterpret_cast<TF*>(&pk);
#pragma unroll
            for (int i = 0; i < 4; ++i) e[i] = from_f32<TF>(o[i]);
            *reinterpret_cast<uint2*>(dst + lane * 4) = pk;
        }
        return;
    }
    if (!normalize) {
        for (int d = lane; d < dim; d += 32) dst[d] = from_f32<TF>(g[d]);
        return;
    }
    const TF* src = feat + (int64_t)b * stride_s + (int64_t)v * stride_v;
    const float inv = inv_norm[warp];
    float dot = 0.f;
    for (int d = lane; d < dim; d += 32) dot = fmaf(g[d], to_f32(src[d]) * inv, dot);
    dot = warp_sum(dot);
    for (int d = lane; d < dim; d += 32) {
        const float zh = to_f32(src[d]) * inv;
        dst[d] = from_f32<TF>((g[d] - zh * dot) * inv);
    }
}

// ------------------------------------------------------------------ finalize (merge column splits)
// partials: [n_split][4][n_rows] = (running max m, sum exp(s - m), possum, npos) per local row.
// One row per thread; every CTA leaves its partial loss in `scratch`, the last CTA to finish
// (ticket counter, self-resetting) adds them in CTA order -> deterministic loss, no atomics on
// floating-point data.  scratch: [0] ticket counter, doubles from byte 16 on (zero on entry).
constexpr int kFinThreads = 256;
constexpr int kFinMaxCtas = 480;

__global__ void __launch_bounds__(kFinThreads)
supcon_finalize_kernel(const float* part, int n_split_in, int flat_q, int flat_ct, int n_rows,
                       int row_begin,
                       RowMap map, int mode, float loss_scale, double inv_anchors,
                       float* __restrict__ out_lse,
                       float* __restrict__ out_possum, float* __restrict__ out_npos,
                       float* __restrict__ out_loss_sum, void* scratch, int scalar_dtype, const PeerOut po,
                       int n_k, int my_slot) {
    __shared__ double red[32];
    __shared__ unsigned s_last;
    unsigned* ticket = reinterpret_cast<unsigned*>(scratch);
    double* cta_sum = reinterpret_cast<double*>(reinterpret_cast<char*>(scratch) + 16);
    double local = 0.0;
    pdl_launch_dependents();
    pdl_wait();
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += gridDim.x * blockDim.x) {
        float lse = 0.f, ps = 0.f, np = 1.f;
        if (map.is_anchor(row_begin + r, mode)) {
            int n_split = n_split_in;
            if (flat_q > 0) {      // slots = CTAs whose span [c*q, (c+1)*q) meets this row block's tiles
                const int rb = r >> 8;
                n_split = 2 * (((rb + 1) * flat_ct - 1) / flat_q - (rb * flat_ct) / flat_q + 1);
            }
            // slots are merged eight at a time so that all their loads are in flight together
            float m = -INFINITY, sum = 0.f;
            ps = 0.f; np = 0.f;
            for (int s0 = 0; s0 < n_split; s0 += 8) {
                float vm[8], vs[8], vp[8], vn[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const bool ok = s0 + u < n_split;
                    const float* q = part + ((int64_t)(s0 + u) * 4) * n_rows + r;
                    vm[u] = ok ? q[0] : -INFINITY;
                    vs[u] = ok ? q[n_rows] : 0.f;
                    vp[u] = ok ? q[2 * (int64_t)n_rows] : 0.f;
                    vn[u] = ok ? q[3 * (int64_t)n_rows] : 0.f;
                }
                float cm = vm[0];
#pragma unroll
                for (int u = 1; u < 8; ++u) cm = fmaxf(cm, vm[u]);
                const float nm = fmaxf(m, cm);
                if (nm > -INFINITY) {
                    float acc = (m > -INFINITY) ? sum * expf(m - nm) : 0.f;
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (vm[u] > -INFINITY) acc += vs[u] * expf(vm[u] - nm);
                    sum = acc;
                    m = nm;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) { ps += vp[u]; np += vn[u]; }
            }
            lse = m + logf(sum);
            local += (double)(-loss_scale * (ps / np - lse));        // losses.py:87,90 (0/0 -> nan)
        }
        out_possum[r] = ps;
        if (po.n_dst > 0) {
            // every rank's [lse (n_k) | npos (n_k) | loss shares] array gets this rank's rows
            for (int w = 0; w < po.n_dst; ++w) {
                float* d = reinterpret_cast<float*>(po.dst[w]);
                d[row_begin + r] = lse;
                d[n_k + row_begin + r] = np;
            }
        } else {
            out_lse[r] = lse; out_npos[r] = np;
        }
    }
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x < 32) {
        double t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) {
            cta_sum[blockIdx.x] = t;
            __threadfence();
            s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
        }
    }
    __syncthreads();
    if (s_last && threadIdx.x < 32) {
        // last CTA: lanes fetch the partials in parallel, fixed-shape shuffle tree -> deterministic
        __threadfence();
        double t = 0.0;
        for (unsigned b = threadIdx.x; b < gridDim.x; b += 32) {
            t += __ldcg(cta_sum + b);
            cta_sum[b] = 0.0;
        }
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) {
            const float loss = (float)(t * inv_anchors);
            if (po.n_dst > 0) {
                // this rank's share of the global mean goes to every rank; scb_peer_wait adds them up
                for (int w = 0; w < po.n_dst; ++w) reinterpret_cast<float*>(po.dst[w])[2 * n_k + 4 * my_slot] = loss;
            } else {
                out_loss_sum[0] = loss;
                out_loss_sum[1] = 0.f;             // [1]: the loss in the caller's scalar dtype (low-order bytes)
                if (scalar_dtype == SCB_BF16) *reinterpret_cast<__nv_bfloat16*>(out_loss_sum + 1) = __float2bfloat16_rn(loss);
                else if (scalar_dtype == SCB_F16) *reinterpret_cast<__half*>(out_loss_sum + 1) = __float2half_rn(loss);
                else out_loss_sum[1] = loss;
            }
            *ticket = 0u;
        }
    }
    if (po.n_dst > 0) {
        // The loss share is written by the last CTA of the ticket above, which is not necessarily the last
        // CTA to reach peer_signal: order it with the same fence + ticket pattern (peer_signal fences every
        // CTA's stores, and the signalling CTA is the last to take ITS ticket).
        peer_signal(po);
    }
}

// ------------------------------------------------------------------ fp32 engine: forward
// CTA = 32 local rows x a strided set of 32-column tiles; warp ty owns row ty, lane tx a column.
__global__ void __launch_bounds__(1024)
supcon_fwd_fp32_kernel(const float* K, int ldk, int dim, int n_k, RowMap map,
                       MaskSpec msk, int mode, float inv_T, int row_begin, int n_rows,
                       float* __restrict__ part) {
    __shared__ float xs[32][33], ys[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int i = row_begin + blockIdx.x * 32 + ty;
    const bool row_ok = (blockIdx.x * 32 + ty) < n_rows && map.is_anchor(i, mode);
    const int sid_i = map.sid(i < n_k ? i : 0);
    float m = -INFINITY, sum = 0.f, possum = 0.f, npos = 0.f;
    const int n_tiles = (n_k + 31) / 32;
    for (int ct = blockIdx.y; ct < n_tiles; ct += gridDim.y) {
        const int j0 = ct * 32;
        float acc = 0.f;
        for (int d0 = 0; d0 < dim; d0 += 32) {
            const int d = d0 + tx;
            const int ri = row_begin + blockIdx.x * 32 + ty, rj = j0 + ty;
            xs[ty][tx] = (ri < n_k && d < dim) ? K[(int64_t)ri * ldk + d] : 0.f;
            ys[ty][tx] = (rj < n_k && d < dim) ? K[(int64_t)rj * ldk + d] : 0.f;
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 32; ++k) acc = fmaf(xs[ty][k], ys[tx][k], acc);
            __syncthreads();
        }
        const int j = j0 + tx;
        const bool ok = row_ok && j < n_k && j != i;
        const float s = acc * inv_T;
        const float mv = ok ? msk.value(sid_i, map.sid(j < n_k ? j : 0)) : 0.f;
        const float tm = warp_max(ok ? s : -INFINITY);
        if (tm > -INFINITY) {
            const float nm = fmaxf(m, tm);
            const float e = ok ? expf(s - nm) : 0.f;
            sum = sum * (m > -INFINITY ? expf(m - nm) : 0.f) + warp_sum(e);
            m = nm;
        }
        possum += warp_sum(mv * (ok ? s : 0.f));
        npos += warp_sum(mv);
    }
    const int r = blockIdx.x * 32 + ty;
    if (tx == 0 && r < n_rows) {
        const int64_t base = (int64_t)blockIdx.y * 4 * n_rows + r;
        part[base] = m;
        part[base + n_rows] = sum;
        part[base + 2 * (int64_t)n_rows] = possum;
        part[base + 3 * (int64_t)n_rows] = npos;
    }
}

// ------------------------------------------------------------------ fp32 engine: backward
// H_ij = [i anchor] G_ij + [j anchor] G_ji, dK_i = (1/T) sum_j H_ij k_j; each CTA owns 32 local
// rows and a 128-wide slice of the feature axis (blockIdx.z), so there are no atomics.
__global__ void __launch_bounds__(1024)
supcon_bwd_fp32_kernel(const float* K, int ldk, int dim, int n_k, RowMap map,
                       MaskSpec msk, int mode, float inv_T, float coef_scale,
                       const float* grad_out, int grad_dtype, int row_begin, int n_rows,
                       const float* lse_all, const float* npos_all,
                       float* __restrict__ out_dk, int ld_dk) {
    __shared__ float xs[32][33], ys[32][33], hs[32][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int r = blockIdx.x * 32 + ty;
    const int i = row_begin + r;
    const bool row_ok = r < n_rows;
    const bool anchor_i = row_ok && map.is_anchor(i, mode);
    const int sid_i = map.sid(i < n_k ? i : 0);
    const float lse_i = anchor_i ? lse_all[i] : 0.f;
    const float inv_np_i = anchor_i ? 1.f / npos_all[i] : 0.f;
    const float coef = load_scalar(grad_out, grad_dtype) * coef_scale;
    const int dbeg = blockIdx.z * 128;
    float accd[4] = {0.f, 0.f, 0.f, 0.f};
    const int n_tiles = (n_k + 31) / 32;
    for (int ct = 0; ct < n_tiles; ++ct) {
        const int j0 = ct * 32;
        float acc = 0.f;
        for (int d0 = 0; d0 < dim; d0 += 32) {
            const int d = d0 + tx;
            const int rj = j0 + ty;
            xs[ty][tx] = (i < n_k && row_ok && d < dim) ? K[(int64_t)i * ldk + d] : 0.f;
            ys[ty][tx] = (rj < n_k && d < dim) ? K[(int64_t)rj * ldk + d] : 0.f;
            __syncthreads();
#pragma unroll
            for (int k = 0; k < 32; ++k) acc = fmaf(xs[ty][k], ys[tx][k], acc);
            __syncthreads();
        }
        const int j = j0 + tx;
        float h = 0.f;
        if (row_ok && j < n_k && j != i) {
            const float s = acc * inv_T;
            const int sid_j = map.sid(j);
            if (anchor_i) h += expf(s - lse_i) - msk.value(sid_i, sid_j) * inv_np_i;
            if (map.is_anchor(j, mode)) h += expf(s - lse_all[j]) - msk.value(sid_j, sid_i) / npos_all[j];
            h *= coef * inv_T;
        }
        hs[ty][tx] = h;
        __syncthreads();
        const int jn = min(32, n_k - j0);
        for (int c = 0; c < jn; ++c) {
            const float hv = hs[ty][c];
            const float* kj = K + (int64_t)(j0 + c) * ldk + dbeg + tx;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (dbeg + tx + 32 * k < dim) accd[k] = fmaf(hv, kj[32 * k], accd[k]);
        }
        __syncthreads();
    }
    if (row_ok) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int d = dbeg + tx + 32 * k;
            if (d < dim) out_dk[(int64_t)r * ld_dk + d] = accd[k];
        }
    }
}

// ------------------------------------------------------------------ consumer side of a peer exchange
__global__ void __launch_bounds__(32)
peer_wait_kernel(const unsigned* flags, const unsigned* state, int n,
                 const float* loss_shares, float* __restrict__ out_loss, int scalar_dtype) {
    pdl_launch_dependents();
    pdl_wait();
    const unsigned want = state[1];               // bumped by this rank's own producer, earlier in the stream
    if ((int)threadIdx.x < n) {
        unsigned long long t0 = 0;
        for (unsigned spins = 1;; ++spins) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + threadIdx.x) : "memory");
            if ((int)(v - want) >= 0) break;
            if ((spins & 1023u) == 0u) {              // a peer that died must not hang this GPU for ever
                unsigned long long now;
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > 30000000000ull) __trap();   // 30 s
            }
        }
    }
    __syncwarp();
    if (loss_shares != nullptr && threadIdx.x == 0) {
        float loss = 0.f;
        for (int w = 0; w < n; ++w) loss += __ldcg(loss_shares + 4 * w);      // fixed order: identical on all ranks
        out_loss[0] = loss;
        out_loss[1] = 0.f;
        if (scalar_dtype == SCB_BF16) *reinterpret_cast<__nv_bfloat16*>(out_loss + 1) = __float2bfloat16_rn(loss);
        else if (scalar_dtype == SCB_F16) *reinterpret_cast<__half*>(out_loss + 1) = __float2half_rn(loss);
        else out_loss[1] = loss;
    }
}

// ------------------------------------------------------------------ peer-memory broadcast
struct PeerBcastArgs {
    char* dst[16];
    long long src_off[4], dst_off[4], bytes[4];
    int n_dst, n_seg;
};

template <typename W>      // uint4 when every offset and size is a multiple of 16 bytes, else 32-bit words
__global__ void __launch_bounds__(256)
peer_bcast_kernel(const char* src, const PeerBcastArgs a) {
    pdl_launch_dependents();
    pdl_wait();
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    constexpr long long B = sizeof(W);
    for (int s = 0; s < a.n_seg; ++s) {
        const char* from = src + a.src_off[s];
        for (long long i = tid * B; i < a.bytes[s]; i += stride * B) {
            const W v = *reinterpret_cast<const W*>(from + i);
            for (int w = 0; w < a.n_dst; ++w) *reinterpret_cast<W*>(a.dst[w] + a.dst_off[s] + i) = v;
        }
    }
    // no fence here: the kernel boundary completes the stores before the caller's cross-rank barrier starts
}

// ------------------------------------------------------------------ host side
int fp32_forward(const scb_supcon_problem* p, int row_begin, int n_rows, float* part, int n_split,
                 cudaStream_t stream) {
    dim3 grid((unsigned)ceil_div(n_rows, 32), (unsigned)n_split), block(32, 32);
    supcon_fwd_fp32_kernel<<<grid, block, 0, stream>>>(
        (const float*)p->k, p->ldk, p->dim, p->n_k, RowMap::of(p), MaskSpec::of(p), p->mode,
        p->inv_temperature, row_begin, n_rows, part);
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

int fp32_backward(const scb_supcon_problem* p, int row_begin, int n_rows, const float* lse_all,
                  const float* npos_all, const float* grad_out, int n_anchor_total, float* out_dk,
                  int ld_dk, cudaStream_t stream) {
    dim3 grid((unsigned)ceil_div(n_rows, 32), 1, (unsigned)ceil_div(p->dim, 128)), block(32, 32);
    supcon_bwd_fp32_kernel<<<grid, block, 0, stream>>>(
        (const float*)p->k, p->ldk, p->dim, p->n_k, RowMap::of(p), MaskSpec::of(p), p->mode,
        p->inv_temperature, p->loss_scale / (float)n_anchor_total, grad_out, p->scalar_dtype, row_begin, n_rows,
        lse_all, npos_all, out_dk, ld_dk);
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

int launch_finalize(const scb_supcon_problem* p, const float* part, int n_split, int flat_q, int flat_ct,
                    int row_begin, int n_rows, double inv_anchors, float* out_lse, float* out_possum, float* out_npos,
                    float* out_loss_sum, void* scratch, cudaStream_t stream, const scb_peer_channel* ch) {
    int ctas = (int)ceil_div(n_rows, kFinThreads);
    if (ctas > kFinMaxCtas) ctas = kFinMaxCtas;
    prefer_max_smem_carveout(supcon_finalize_kernel);
    const PeerOut po = ch ? PeerOut::of(ch) : PeerOut::none();
    const int my_slot = row_begin / n_rows;          // rank: every rank holds n_rows consecutive rows
    SCB_CUDA_OK(launch_pdl(supcon_finalize_kernel, dim3(ctas), dim3(kFinThreads), 0, stream, part, n_split, flat_q,
                           flat_ct, n_rows, row_begin, RowMap::of(p), p->mode, p->loss_scale, inv_anchors, out_lse,
                           out_possum, out_npos, out_loss_sum, scratch, p->scalar_dtype, po, p->n_k, my_slot));
    count_launch();
    return SCB_OK;
}

}  // namespace scb

using namespace scb;

extern "C" int scb_supcon_prep(const void* features, int feat_dtype, int n_samples, int n_views,
                               int dim, int64_t stride_sample, int64_t stride_view, int normalize,
                               float eps, void* out_k, void* out_k_lo, int k_dtype, int ldk,
                               float* out_inv_norm, void* stream_) {
    SCB_REQUIRE(features && out_k, "supcon_prep: null pointer");
    SCB_REQUIRE(n_samples > 0 && n_views > 0 && dim > 0 && ldk >= dim, "supcon_prep: bad shape");
    SCB_REQUIRE(k_dtype == SCB_F32 || k_dtype == SCB_BF16, "supcon_prep: k_dtype must be F32 or BF16");
    SCB_REQUIRE(out_k_lo == nullptr || k_dtype == SCB_BF16, "supcon_prep: k_lo needs a BF16 k");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t rows = (int64_t)n_samples * n_views;
    const unsigned blocks = (unsigned)ceil_div(rows * 32, 256);
    SCB_DISPATCH_FLOAT(feat_dtype, TI, {
        prefer_max_smem_carveout(supcon_prep_kernel<TI, float>);
        prefer_max_smem_carveout(supcon_prep_kernel<TI, __nv_bfloat16>);
        if (k_dtype == SCB_F32)
            SCB_CUDA_OK(launch_pdl(supcon_prep_kernel<TI, float>, dim3(blocks), dim3(256), 0, stream,
                                   (const TI*)features, n_samples, n_views, dim, stride_sample, stride_view, normalize,
                                   eps, (float*)out_k, (__nv_bfloat16*)nullptr, ldk, out_inv_norm, PeerOut::none(),
                                   (int64_t)0));
        else
            SCB_CUDA_OK(launch_pdl(supcon_prep_kernel<TI, __nv_bfloat16>, dim3(blocks), dim3(256), 0, stream,
                                   (const TI*)features, n_samples, n_views, dim, stride_sample, stride_view, normalize,
                                   eps, (__nv_bfloat16*)out_k, (__nv_bfloat16*)out_k_lo, ldk, out_inv_norm,
                                   PeerOut::none(), (int64_t)0));
    });
    count_launch();
    return SCB_OK;
}

extern "C" int scb_supcon_finish(const float* dk, int ld_dk, const void* features, int feat_dtype,
                                 int n_samples, int n_views, int dim, int64_t stride_sample,
                                 int64_t stride_view, int normalize, const float* inv_norm,
                                 void* out_dfeatures, int64_t out_stride_sample,
                                 int64_t out_stride_view, void* stream_) {
    SCB_REQUIRE(dk && out_dfeatures, "supcon_finish: null pointer");
    SCB_REQUIRE(!normalize || (features && inv_norm), "supcon_finish: normalize needs features and inv_norm");
    SCB_REQUIRE(n_samples > 0 && n_views > 0 && dim > 0 && ld_dk >= dim, "supcon_finish: bad shape");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t rows = (int64_t)n_samples * n_views;
    const unsigned blocks = (unsigned)ceil_div(rows * 32, 256);
    SCB_DISPATCH_FLOAT(feat_dtype, TF, {
        prefer_max_smem_carveout(supcon_finish_kernel<TF>);
        SCB_CUDA_OK(launch_pdl(supcon_finish_kernel<TF>, dim3(blocks), dim3(256), 0, stream, dk, ld_dk,
                               (const TF*)features, n_samples, n_views, dim, stride_sample, stride_view, normalize,
                               inv_norm, (TF*)out_dfeatures, out_stride_sample, out_stride_view));
    });
    count_launch();
    return SCB_OK;
}

extern "C" int scb_peer_bcast(const void* src, const int64_t* seg, int n_seg, const void* const* dst_ptrs, int n_dst,
                              void* stream_) {
    SCB_REQUIRE(src && seg && dst_ptrs, "peer_bcast: null pointer");
    SCB_REQUIRE(n_seg >= 1 && n_seg <= 4 && n_dst >= 1 && n_dst <= 16, "peer_bcast: 1..4 segments, 1..16 destinations");
    PeerBcastArgs a;
    a.n_dst = n_dst;
    a.n_seg = n_seg;
    long long total = 0, bits = reinterpret_cast<uintptr_t>(src);
    for (int s = 0; s < n_seg; ++s) {
        a.src_off[s] = seg[3 * s]; a.dst_off[s] = seg[3 * s + 1]; a.bytes[s] = seg[3 * s + 2];
        SCB_REQUIRE(a.src_off[s] >= 0 && a.dst_off[s] >= 0 && a.bytes[s] >= 0,
                    "peer_bcast: offsets and sizes must be non-negative");
        bits |= a.src_off[s] | a.dst_off[s] | a.bytes[s];
        total += a.bytes[s];
    }
    for (int w = 0; w < n_dst; ++w) {
        SCB_REQUIRE(dst_ptrs[w] != nullptr, "peer_bcast: destination %d is null", w);
        bits |= reinterpret_cast<uintptr_t>(dst_ptrs[w]);
        a.dst[w] = (char*)const_cast<void*>(dst_ptrs[w]);
    }
    SCB_REQUIRE((bits & 3) == 0, "peer_bcast: pointers, offsets and sizes must be multiples of 4 bytes");
    if (total == 0) return SCB_OK;
    const bool wide = (bits & 15) == 0;
    long long blocks = ceil_div(total / (wide ? 16 : 4), 256);
    if (blocks > 4 * sm_count()) blocks = 4 * sm_count();
    if (blocks < 1) blocks = 1;
    if (wide)
        SCB_CUDA_OK(launch_pdl(peer_bcast_kernel<uint4>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream_,
                               (const char*)src, a));
    else
        SCB_CUDA_OK(launch_pdl(peer_bcast_kernel<uint32_t>, dim3((unsigned)blocks), dim3(256), 0, (cudaStream_t)stream_,
                               (const char*)src, a));
    count_launch();
    return SCB_OK;
}

static int check_channel(const scb_peer_channel* ch) {
    SCB_REQUIRE(ch && ch->state && ch->n_dst >= 1 && ch->n_dst <= 16, "peer channel: needs state and 1..16 destinations");
    for (int w = 0; w < ch->n_dst; ++w)
        SCB_REQUIRE(ch->dst[w] && ch->flag[w] && (reinterpret_cast<uintptr_t>(ch->dst[w]) & 15u) == 0,
                    "peer channel: destination %d is null or not 16-byte aligned", w);
    return SCB_OK;
}

extern "C" int scb_supcon_prep_peers(const void* features, int feat_dtype, int n_samples, int n_views, int dim,
                                     int64_t stride_sample, int64_t stride_view, int normalize, float eps,
                                     const scb_peer_channel* ch, int64_t row_offset, int k_dtype, int ldk,
                                     float* out_inv_norm, void* stream_) {
    SCB_REQUIRE(features, "supcon_prep_peers: null pointer");
    if (int e = check_channel(ch)) return e;
    SCB_REQUIRE(n_samples > 0 && n_views > 0 && dim > 0 && ldk >= dim && row_offset >= 0, "supcon_prep_peers: bad shape");
    SCB_REQUIRE(k_dtype == SCB_F32 || k_dtype == SCB_BF16, "supcon_prep_peers: k_dtype must be F32 or BF16");
    cudaStream_t stream = (cudaStream_t)stream_;
    const int64_t rows = (int64_t)n_samples * n_views;
    unsigned blocks = (unsigned)ceil_div(rows * 32, 256);
    if (blocks > 4u * (unsigned)sm_count()) blocks = 4u * (unsigned)sm_count();
    const PeerOut po = PeerOut::of(ch);
    SCB_DISPATCH_FLOAT(feat_dtype, TI, {
        if (k_dtype == SCB_F32)
            SCB_CUDA_OK(launch_pdl(supcon_prep_kernel<TI, float>, dim3(blocks), dim3(256), 0, stream,
                                   (const TI*)features, n_samples, n_views, dim, stride_sample, stride_view, normalize,
                                   eps, (float*)nullptr, (__nv_bfloat16*)nullptr, ldk, out_inv_norm, po, row_offset));
        else
            SCB_CUDA_OK(launch_pdl(supcon_prep_kernel<TI, __nv_bfloat16>, dim3(blocks), dim3(256), 0, stream,
                                   (const TI*)features, n_samples, n_views, dim, stride_sample, stride_view, normalize,
                                   eps, (__nv_bfloat16*)nullptr, (__nv_bfloat16*)nullptr, ldk, out_inv_norm, po,
                                   row_offset));
    });
    count_launch();
    return SCB_OK;
}

extern "C" int scb_peer_wait(const uint32_t* flags_local, const uint32_t* state, int n, const float* loss_shares,
                             float* out_loss, int scalar_dtype, void* stream_) {
    SCB_REQUIRE(flags_local && state && n >= 1 && n <= 16, "peer_wait: needs flags, state and 1..16 ranks");
    SCB_REQUIRE(loss_shares == nullptr || out_loss != nullptr, "peer_wait: loss_shares needs out_loss");
    SCB_CUDA_OK(launch_pdl(peer_wait_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream_, (const unsigned*)flags_local,
                           (const unsigned*)state, n, loss_shares, out_loss, scalar_dtype));
    count_launch();
    return SCB_OK;
}

// ================================================================== feature-norm side outputs (main_supcon.py:298-317)
// The training loop of the reference looks at the norms of the UN-normalised embeddings every step: their mean and
// variance (logged), an exponential moving average of the mean (`record_norm_mean`), and two optional regularisers,
//   loss_sec   = mean (||z_i|| - record_norm_mean)^2        (:304-312, the moving average updated first, detached)
//   loss_l2reg = mean ||z_i||^2                             (:314-317).
// The fused normalisation already has 1 / ||z_i|| (scb_supcon_prep's out_inv_norm), so one small CTA turns that
// vector into all of it -- no pass over the embeddings, no host synchronisation.
namespace scb {

__global__ void __launch_bounds__(1024)
feature_norm_terms_kernel(const float* inv_norm, int n, float* record, float momentum, float w_sec,
                          float w_l2, float* __restrict__ out) {
    __shared__ double red[3][32];
    pdl_launch_dependents();
    pdl_wait();
    double s1 = 0.0, s2 = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double v = 1.0 / (double)inv_norm[i];
        s1 += v; s2 += v * v;
    }
    auto block_sum = [&](double v, int slot) {
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[slot][threadIdx.x >> 5] = v;
        __syncthreads();
        double t = (threadIdx.x < 32) ? red[slot][threadIdx.x] : 0.0;
        if (threadIdx.x < 32)
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        return t;            // valid in warp 0
    };
    const double t1 = block_sum(s1, 0), t2 = block_sum(s2, 1);
    if (threadIdx.x == 0) {
        const double mean = t1 / n, msq = t2 / n;
        // record[0] = moving average, record[1] = 1 once it holds a value (`record_norm_mean is not None`, :304-307)
        const double rec = (record[1] != 0.f) ? (1.0 - momentum) * (double)record[0] + momentum * mean : mean;
        record[0] = (float)rec;
        record[1] = 1.f;
        const double var = msq - mean * mean;                       // mean (v - mean)^2
        const double sec = msq - 2.0 * rec * mean + rec * rec;      // mean (v - rec)^2
        out[0] = (float)mean;
        out[1] = (float)(var > 0.0 ? var : 0.0);
        out[2] = (float)sec;
        out[3] = (float)msq;
        out[4] = (float)(w_sec * sec + w_l2 * msq);                 // what the caller adds to the loss
    }
}

// d features += upstream * [ w_sec * (2/n) (||z|| - record) / ||z|| + w_l2 * (2/n) ] * z      (one warp per row)
template <typename TF>
__global__ void __launch_bounds__(256)
feature_norm_backward_kernel(const TF* feat, int n_samples, int n_views, int dim, int64_t stride_s,
                             int64_t stride_v, const float* inv_norm, const float* record,
                             float w_sec, float w_l2, const void* grad_out, int grad_dtype,
                             TF* __restrict__ dfeat, int64_t ostride_s, int64_t ostride_v) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    pdl_launch_dependents();
    pdl_wait();
    const int n = n_samples * n_views;
    if (warp >= n) return;
    const int v = warp / n_samples, b = warp - v * n_samples;          // inv_norm is indexed by stacked row v*n_samples + b
    const float norm = 1.f / inv_norm[warp];
    const float g = load_scalar(grad_out, grad_dtype) * (2.f / (float)n);
    const float c = g * (w_sec * (norm - record[0]) / norm + w_l2);
    const TF* src = feat + (int64_t)b * stride_s + (int64_t)v * stride_v;
    TF* dst = dfeat + (int64_t)b * ostride_s + (int64_t)v * ostride_v;
    for (int d = lane; d < dim; d += 32) dst[d] = from_f32<TF>(to_f32(dst[d]) + c * to_f32(src[d]));
}

}  // namespace scb

extern "C" int scb_feature_norm_terms(const float* inv_norm, int n, float* record, float momentum, float sec_weight,
                                      float l2reg_weight, float* out_terms, void* stream_) {
    SCB_REQUIRE(inv_norm && record && out_terms && n >= 1, "feature_norm_terms: null pointer or empty input");
    SCB_CUDA_OK(launch_pdl(feature_norm_terms_kernel, dim3(1), dim3(1024), 0, (cudaStream_t)stream_, inv_norm, n, record,
                           momentum, sec_weight, l2reg_weight, out_terms));
    count_launch();
    return SCB_OK;
}

extern "C" int scb_feature_norm_backward(const void* features, int feat_dtype, int n_samples, int n_views, int dim,
                                         int64_t stride_sample, int64_t stride_view, const float* inv_norm,
                                         const float* record, float sec_weight, float l2reg_weight,
                                         const void* grad_out, int grad_dtype, void* dfeatures,
                                         int64_t out_stride_sample, int64_t out_stride_view, void* stream_) {
    SCB_REQUIRE(features && inv_norm && record && grad_out && dfeatures, "feature_norm_backward: null pointer");
    SCB_REQUIRE(n_samples > 0 && n_views > 0 && dim > 0, "feature_norm_backward: bad shape");
    const int64_t rows = (int64_t)n_samples * n_views;
    const unsigned blocks = (unsigned)ceil_div(rows * 32, 256);
    SCB_DISPATCH_FLOAT(feat_dtype, TF, {
        SCB_CUDA_OK(launch_pdl(feature_norm_backward_kernel<TF>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream_,
                               (const TF*)features, n_samples, n_views, dim, stride_sample, stride_view, inv_norm, record,
                               sec_weight, l2reg_weight, grad_out, grad_dtype, (TF*)dfeatures, out_stride_sample,
                               out_stride_view));
    });
    count_launch();
    return SCB_OK;
}
