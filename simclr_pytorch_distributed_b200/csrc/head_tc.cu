// Projection-head tail: the last Linear of the MLP head, its bias, the row L2
// normalisation and the cast to the loss's operand type in ONE kernel.
//
// Replaces  feat = F.normalize(self.head[2](h), dim=1)      /root/reference/networks/resnet_big.py:168-172,177-181
// (and the second F.normalize of main_supcon.py:283, which is the identity on unit rows): out = normalise(h W^T + b)
// lands as rows of the contrast matrix -- bf16 [B * V, 128] is exactly what the loss kernels take in place
// (SCB_ROWS_SAMPLE_MAJOR), so there is no prep kernel and no fp32 round trip between the head and the loss.
//
// Structure: out tile [128 rows x N <= 128] = h[128, K] W[N, K]^T on tcgen05.mma (both operands K-major, TMA-staged
// 128-byte-swizzled [128 x 64] sub-tiles, fp32 accumulator in TMEM).  M is small (1024 rows per GPU in BASELINE cfg 3),
// so K is split over the CTAs of a thread-block cluster (up to 8): every CTA accumulates its K range, parks its
// partial tile in shared memory, and after barrier.cluster CTA q reads the q-th slice of rows out of EVERY CTA's
// shared memory (DSMEM), adds bias, takes the row norm (a warp per row) and stores the normalised row and 1/||row||.
#include "supcon_tc.cuh"
#include "tc_ptx.cuh"

namespace scb {
using namespace ptx;

constexpr int kHdThreads = 192;     // warp 0 TMA, warp 1 MMA + TMEM, warps 2..5 TMEM -> shared memory
constexpr int kHdStages = 4;
constexpr int kHdSub = 16384;       // one [128 x 64] bf16 sub-tile
constexpr int kHdPitch = 132;       // floats per staged row (128 + 4: 16-byte aligned, rows on different banks)

struct HeadParams {
    int M, N, chunks;               // rows, output features, 64-wide K chunks in total
    int splits;                     // CTAs per cluster = K splits
    const float* bias;
    void* out;
    int out_dtype;
    int64_t ld_out;
    float* inv_norm;
    int normalize;
    float eps;
    int n_views, rows_per_view;     // n_views > 1: input rows are view-major [V][B], output rows sample-major [B][V]
};

__device__ __forceinline__ uint32_t hd_mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ float4 hd_ld_cluster(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void hd_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t hd_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__global__ void __launch_bounds__(kHdThreads, 1)
head_tc_kernel(const __grid_constant__ CUtensorMap tmap_h, const __grid_constant__ CUtensorMap tmap_w,
               const __grid_constant__ HeadParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* stages = smem;                                                  // [kHdStages][A 16 KiB | B 16 KiB]
    uint64_t* bars = reinterpret_cast<uint64_t*>(stages + kHdStages * 2 * kHdSub);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kHdStages + 1);
    float* staged = reinterpret_cast<float*>(smem);                          // [128][kHdPitch], over the stages
    const uint32_t bar0 = smem_u32(bars);
    auto full = [&](int s) { return bar0 + 8u * (uint32_t)s; };
    auto empty = [&](int s) { return bar0 + 8u * (uint32_t)(kHdStages + s); };
    const uint32_t acc_full = bar0 + 8u * (uint32_t)(2 * kHdStages);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = p.splits > 1 ? hd_cluster_rank() : 0u;
    const int tile = blockIdx.x / p.splits;
    const int m0 = tile * 128;
    // this CTA's K chunks
    const int per = (p.chunks + p.splits - 1) / p.splits;
    const int kc0 = (int)rank * per, kc1 = min(p.chunks, kc0 + per);
    const int n_kc = max(kc1 - kc0, 0);

    if ((smem_u32(smem) & 1023u) != 0u) __trap();
    if (threadIdx.x == 0) {
        for (int s = 0; s < kHdStages; ++s) { mbar_init(full(s), 1); mbar_init(empty(s), 1); }
        mbar_init(acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(tmem_slot), 128);
    if (warp == 0 && lane == 0) { prefetch_tensormap(&tmap_h); prefetch_tensormap(&tmap_w); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        for (int i = 0; i < n_kc; ++i) {
            const int s = i % kHdStages;
            mbar_wait(empty(s), ((i / kHdStages) & 1) ^ 1);
            if (elect_one()) {
                mbar_arrive_expect_tx(full(s), 2 * kHdSub);
                tma_load_2d(smem_u32(stages + s * 2 * kHdSub), &tmap_h, (kc0 + i) * 64, m0, full(s));
                tma_load_2d(smem_u32(stages + s * 2 * kHdSub + kHdSub), &tmap_w, (kc0 + i) * 64, 0, full(s));
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        constexpr uint32_t idesc = idesc_bf16(128, 128, 0);
        for (int i = 0; i < n_kc; ++i) {
            const int s = i % kHdStages;
            mbar_wait(full(s), (i / kHdStages) & 1);
            tc_fence_after();
            const uint32_t a_lo = desc_lo(smem_u32(stages + s * 2 * kHdSub), 16);
            const uint32_t b_lo = desc_lo(smem_u32(stages + s * 2 * kHdSub + kHdSub), 16);
            if (elect_one()) {
#pragma unroll
                for (int k16 = 0; k16 < 4; ++k16) {
                    const uint32_t off = (uint32_t)(k16 * 32) >> 4;
                    if (i == 0 && k16 == 0) umma_bf16_lo<false>(tmem_base, a_lo + off, b_lo + off, idesc);
                    else umma_bf16_lo<true>(tmem_base, a_lo + off, b_lo + off, idesc);
                }
                umma_commit(empty(s));
                if (i == n_kc - 1) umma_commit(acc_full);
            }
            __syncwarp();
        }
    }
    // every thread: the accumulator is complete (and with it every TMA load has been consumed: the stages are free)
    if (n_kc > 0) mbar_wait(acc_full, 0);
    tc_fence_after();
    __syncthreads();
    if (warp >= 2) {
        const int q4 = warp & 3;                         // TMEM lane quarter this warp may touch
        const int r = q4 * 32 + lane;
        float* dst = staged + (size_t)r * kHdPitch;
#pragma unroll 1
        for (int c = 0; c < 128; c += 32) {
            uint32_t v[32];
            if (n_kc > 0) {
                tmem_ld32(tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)c, v);
                tmem_ld_wait_dep(v);
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = 0u;
            }
#pragma unroll
            for (int i = 0; i < 32; i += 4)
                *reinterpret_cast<uint4*>(dst + c + i) = make_uint4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        }
    }
    tc_fence_before();
    if (p.splits > 1) hd_cluster_sync(); else __syncthreads();

    // finish: rows [rank * rows_per, ...) of the tile, a warp per row, a lane per four columns
    const int rows_per = 128 / p.splits;
    const uint32_t staged_addr = smem_u32(staged);
    const int c = lane * 4;
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p.bias != nullptr && c < p.N) {
        b4.x = p.bias[c];
        if (c + 1 < p.N) b4.y = p.bias[c + 1];
        if (c + 2 < p.N) b4.z = p.bias[c + 2];
        if (c + 3 < p.N) b4.w = p.bias[c + 3];
    }
    for (int rr = warp; rr < rows_per; rr += kHdThreads / 32) {
        const int row = (int)rank * rows_per + rr;
        const int gm = m0 + row;
        if (gm >= p.M) break;                              // warp-uniform
        const uint32_t off = staged_addr + (uint32_t)((row * kHdPitch + c) * 4);
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q = 0; q < p.splits; ++q) {               // rank order: the same bits whoever finishes the row
            const float4 v = p.splits > 1 ? hd_ld_cluster(hd_mapa(off, (uint32_t)q))
                                          : *reinterpret_cast<const float4*>(staged + row * kHdPitch + c);
            a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        }
        a.x += b4.x; a.y += b4.y; a.z += b4.z; a.w += b4.w;
        if (c >= p.N) a = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c + 1 >= p.N) a.y = 0.f;
        if (c + 2 >= p.N) a.z = 0.f;
        if (c + 3 >= p.N) a.w = 0.f;
        float inv = 1.f;
        if (p.normalize) {
            const float ss = warp_sum(a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w);
            inv = 1.f / fmaxf(sqrtf(ss), p.eps);           // F.normalize: x / max(||x||, eps)
            a.x *= inv; a.y *= inv; a.z *= inv; a.w *= inv;
        }
        if (lane == 0 && p.inv_norm != nullptr) p.inv_norm[gm] = inv;
        // the loss wants the views of a sample next to each other (main_supcon.py:285-286 stacks them that way)
        const int64_t go = p.n_views > 1 ? (int64_t)(gm % p.rows_per_view) * p.n_views + gm / p.rows_per_view : gm;
        if (c < p.N) {                                      // N is a multiple of 4
            if (p.out_dtype == SCB_F32) {
                *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.out) + go * p.ld_out + c) = a;
            } else if (p.out_dtype == SCB_BF16) {
                __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
                uint2 o;
                o.x = *reinterpret_cast<uint32_t*>(&lo);
                o.y = *reinterpret_cast<uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.out) + go * p.ld_out + c) = o;
            } else {
                __half2 lo = __floats2half2_rn(a.x, a.y), hi = __floats2half2_rn(a.z, a.w);
                uint2 o;
                o.x = *reinterpret_cast<uint32_t*>(&lo);
                o.y = *reinterpret_cast<uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(reinterpret_cast<__half*>(p.out) + go * p.ld_out + c) = o;
            }
        }
    }
    // nobody leaves while a partner may still be reading its shared memory
    if (p.splits > 1) hd_cluster_sync(); else __syncthreads();
    if (warp == 1) tmem_dealloc(tmem_base, 128);
}

// backward of the row normalisation y = x / ||x||:  dx = (g - y (g . y)) / ||x||      (a warp per row, n <= 128)
__global__ void __launch_bounds__(256)
head_normalize_backward_kernel(const void* g, int g_dtype, int64_t ld_g, const void* y, int y_dtype, int64_t ld_y,
                               const float* inv_norm, int64_t m, int n, void* out, int out_dtype, int64_t ld_out,
                               int n_views) {
    const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);     // row of the Linear's input / of `out`
    const int lane = threadIdx.x & 31;
    if (row >= m) return;
    const int64_t rpv = m / n_views;
    const int64_t gr = n_views > 1 ? (row % rpv) * n_views + row / rpv : row;   // its row in g and y (sample-major)
    float gv[4], yv[4];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i;
        gv[i] = 0.f; yv[i] = 0.f;
        if (c < n) {
            gv[i] = g_dtype == SCB_F32 ? reinterpret_cast<const float*>(g)[gr * ld_g + c]
                  : g_dtype == SCB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g)[gr * ld_g + c])
                                        : __half2float(reinterpret_cast<const __half*>(g)[gr * ld_g + c]);
            yv[i] = y_dtype == SCB_F32 ? reinterpret_cast<const float*>(y)[gr * ld_y + c]
                  : y_dtype == SCB_BF16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(y)[gr * ld_y + c])
                                        : __half2float(reinterpret_cast<const __half*>(y)[gr * ld_y + c]);
            dot = fmaf(gv[i], yv[i], dot);
        }
    }
    dot = warp_sum(dot);
    const float inv = inv_norm[row];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = lane + 32 * i;
        if (c < n) {
            const float v = (gv[i] - yv[i] * dot) * inv;
            if (out_dtype == SCB_F32) reinterpret_cast<float*>(out)[row * ld_out + c] = v;
            else if (out_dtype == SCB_BF16) reinterpret_cast<__nv_bfloat16*>(out)[row * ld_out + c] = __float2bfloat16_rn(v);
            else reinterpret_cast<__half*>(out)[row * ld_out + c] = __float2half_rn(v);
        }
    }
}

}  // namespace scb

using namespace scb;

extern "C" int scb_head_normalize_backward(const void* g, int g_dtype, int64_t ld_g, const void* y, int y_dtype,
                                           int64_t ld_y, const float* inv_norm, int64_t m, int n, void* out,
                                           int out_dtype, int64_t ld_out, int n_views, void* stream_) {
    SCB_REQUIRE(g && y && inv_norm && out, "head_normalize_backward: null pointer");
    SCB_REQUIRE(m >= 0 && n >= 1 && n <= 128, "head_normalize_backward: n must be in 1..128");
    SCB_REQUIRE(n_views >= 1 && m % n_views == 0, "head_normalize_backward: m must be a multiple of n_views");
    if (m == 0) return SCB_OK;
    head_normalize_backward_kernel<<<(unsigned)((m + 7) / 8), 256, 0, (cudaStream_t)stream_>>>(
        g, g_dtype, ld_g, y, y_dtype, ld_y, inv_norm, m, n, out, out_dtype, ld_out, n_views);
    SCB_CUDA_OK(cudaGetLastError());
    count_launch();
    return SCB_OK;
}

extern "C" int scb_head_forward(const void* h, int64_t ld_h, const void* w, int64_t ld_w, const float* bias, int64_t m,
                                int k, int n, int n_views, void* out, int out_dtype, int64_t ld_out, float* inv_norm,
                                int normalize, float eps, void* stream_) {
    SCB_REQUIRE(h && w && out, "head_forward: null pointer");
    SCB_REQUIRE(m >= 0 && k >= 8 && n >= 4 && n <= 128 && n % 4 == 0, "head_forward: unsupported shape m=%lld k=%d n=%d "
                "(n must be a multiple of 4, at most 128)", (long long)m, k, n);
    SCB_REQUIRE(k % 8 == 0 && ld_h % 8 == 0 && ld_w % 8 == 0 && ld_h >= k && ld_w >= k,
                "head_forward: k and the row strides must be multiples of 8 elements (16-byte rows for TMA)");
    SCB_REQUIRE(out_dtype == SCB_F32 || out_dtype == SCB_BF16 || out_dtype == SCB_F16, "head_forward: bad out dtype");
    SCB_REQUIRE(n_views >= 1 && m % n_views == 0, "head_forward: m must be a multiple of n_views");
    SCB_REQUIRE(ld_out >= n && (ld_out * (out_dtype == SCB_F32 ? 4 : 2)) % (out_dtype == SCB_F32 ? 16 : 8) == 0,
                "head_forward: ld_out must keep rows vector-aligned");
    if (m == 0) return SCB_OK;
    cudaStream_t stream = (cudaStream_t)stream_;
    CUtensorMap th, tw;
    if (int e = tc_make_tmap_2d(h, m, k, ld_h, &th)) return e;
    if (int e = tc_make_tmap_2d(w, n, k, ld_w, &tw)) return e;
    HeadParams p = {};
    p.M = (int)m; p.N = n; p.chunks = (k + 63) / 64;
    const int tiles = (int)((m + 127) / 128);
    int splits = 1;
    while (splits < 8 && splits * 2 <= p.chunks && tiles * splits * 2 <= sm_count()) splits *= 2;
    p.splits = splits;
    p.bias = bias; p.out = out; p.out_dtype = out_dtype; p.ld_out = ld_out; p.inv_norm = inv_norm;
    p.normalize = normalize; p.eps = eps;
    p.n_views = n_views; p.rows_per_view = (int)(m / n_views);
    const size_t smem = (size_t)kHdStages * 2 * kHdSub + (2 * kHdStages + 2) * sizeof(uint64_t);
    static_assert((size_t)128 * kHdPitch * sizeof(float) <= (size_t)kHdStages * 2 * kHdSub, "staging fits over the stages");
    SCB_CUDA_OK(set_max_dynamic_smem(head_tc_kernel, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(tiles * splits));
    cfg.blockDim = dim3(kHdThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)splits; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = splits > 1 ? 1 : 0;
    SCB_CUDA_OK(cudaLaunchKernelEx(&cfg, head_tc_kernel, th, tw, p));
    count_launch();
    return SCB_OK;
}
