// Diagnostics: what a read-only pass and a read+write pass over a tensor of a given size can reach on this GPU,
// per launch, with inputs rotating through more memory than the L2 holds -- the yardstick for the batch-norm passes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/stream_bench tools/stream_bench.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

__device__ __forceinline__ uint4 ldg_nc(const void* p) {
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void stg_cs(void* p, uint4 v) {
    asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// (a) grid-stride read, U loads in flight per thread
template <int U>
__global__ void read_gs(const uint4* __restrict__ src, size_t n, unsigned* sink) {
    unsigned acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + (U - 1) * stride < n; i += U * stride) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldg_nc(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < U; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    for (; i < n; i += stride) { uint4 v = ldg_nc(src + i); acc += v.x ^ v.w; }
    if (acc == 0x12345678u) *sink = acc;
}
// (b) each CTA owns a contiguous range (like the batch-norm passes), U loads in flight per thread
template <int U>
__global__ void read_contig(const uint4* __restrict__ src, size_t n, unsigned* sink) {
    unsigned acc = 0;
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t beg = (size_t)blockIdx.x * per, end = beg + per < n ? beg + per : n;
    size_t i = beg + threadIdx.x;
    for (; i + (U - 1) * blockDim.x < end; i += U * blockDim.x) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldg_nc(src + i + u * blockDim.x);
#pragma unroll
        for (int u = 0; u < U; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    for (; i < end; i += blockDim.x) { uint4 v = ldg_nc(src + i); acc += v.x ^ v.w; }
    if (acc == 0x12345678u) *sink = acc;
}
// (c) copy, contiguous range per CTA
template <int U>
__global__ void copy_contig(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n) {
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t beg = (size_t)blockIdx.x * per, end = beg + per < n ? beg + per : n;
    size_t i = beg + threadIdx.x;
    for (; i + (U - 1) * blockDim.x < end; i += U * blockDim.x) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ldg_nc(src + i + u * blockDim.x);
#pragma unroll
        for (int u = 0; u < U; ++u) stg_cs(dst + i + u * blockDim.x, v[u]);
    }
    for (; i < end; i += blockDim.x) stg_cs(dst + i, ldg_nc(src + i));
}
// (d) bulk-copy ring read: 1 producer thread, 8 stages of 16 KiB, consumers xor the data
__global__ void __launch_bounds__(544, 1) read_ring(const uint8_t* __restrict__ src, size_t bytes, unsigned* sink, int n_stage, int chunk) {
    extern __shared__ __align__(128) uint8_t sm[];
    uint64_t* bars = (uint64_t*)(sm + (size_t)n_stage * chunk);
    const uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(bars);
    if (threadIdx.x == 0) {
        for (int s = 0; s < n_stage; ++s) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * s), "r"(1));
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar0 + 8 * (16 + s)), "r"(16));
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const size_t per = ((bytes + gridDim.x - 1) / gridDim.x + 15) / 16 * 16;
    const size_t beg = (size_t)blockIdx.x * per, end = beg + per < bytes ? beg + per : bytes;
    const int n_chunks = end > beg ? (int)((end - beg + chunk - 1) / chunk) : 0;
    auto wait = [&](uint32_t bar, uint32_t parity) {
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    };
    if (threadIdx.x >= 512) {
        if (threadIdx.x == 512) {
            for (int k = 0; k < n_chunks; ++k) {
                const int s = k % n_stage;
                wait(bar0 + 8 * (16 + s), ((k / n_stage) & 1) ^ 1);
                const size_t o = beg + (size_t)k * chunk;
                const uint32_t nb = (uint32_t)(end - o < (size_t)chunk ? end - o : chunk);
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8 * s), "r"(nb) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"((uint32_t)__cvta_generic_to_shared(sm + (size_t)s * chunk)), "l"(__cvta_generic_to_global(src + o)),
                             "r"(nb), "r"(bar0 + 8 * s) : "memory");
            }
        }
    } else {
        unsigned acc = 0;
        for (int k = 0; k < n_chunks; ++k) {
            const int s = k % n_stage;
            wait(bar0 + 8 * s, (k / n_stage) & 1);
            const size_t o = beg + (size_t)k * chunk;
            const int nb = (int)(end - o < (size_t)chunk ? end - o : chunk);
            for (int q = threadIdx.x * 16; q < nb; q += 512 * 16) {
                const uint4 v = *(const uint4*)(sm + (size_t)s * chunk + q);
                acc += v.x ^ v.y ^ v.z ^ v.w;
            }
            __syncwarp();
            if ((threadIdx.x & 31) == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar0 + 8 * (16 + s)) : "memory");
        }
        if (acc == 0x12345678u) *sink = acc;
    }
}

template <typename F>
static float time_us(F&& launch, int n_sets, int iters) {
    for (int i = 0; i < 3; ++i) launch(i % n_sets);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) launch(i % n_sets);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    if (cudaGetLastError() != cudaSuccess) return -1.f;
    return ms * 1e3f / iters;
}

int main() {
    const size_t sizes[] = {8u << 20, 16u << 20, 32u << 20, 64u << 20, 256u << 20};
    unsigned* sink;
    cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(read_ring, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (size_t bytes : sizes) {
        const int n_sets = (int)((768u << 20) / bytes) < 2 ? 2 : (int)((768u << 20) / bytes);
        std::vector<uint8_t*> src(n_sets), dst(n_sets);
        for (int i = 0; i < n_sets; ++i) { cudaMalloc(&src[i], bytes); cudaMalloc(&dst[i], bytes); cudaMemset(src[i], i + 1, bytes); }
        const size_t n = bytes / 16;
        const int iters = 40;
        auto rep = [&](const char* name, float us, double moved) { printf("%4zu MiB  %-34s %8.2f us  %7.0f GB/s\n", bytes >> 20, name, us, moved / us / 1e3); };
        rep("read grid-stride 1184x256 U8", time_us([&](int i) { read_gs<8><<<1184, 256>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read grid-stride 592x512 U8", time_us([&](int i) { read_gs<8><<<592, 512>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read contiguous 444x256 U8", time_us([&](int i) { read_contig<8><<<444, 256>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read contiguous 1184x256 U8", time_us([&](int i) { read_contig<8><<<1184, 256>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read contiguous 148x1024 U8", time_us([&](int i) { read_contig<8><<<148, 1024>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read contiguous 2368x256 U4", time_us([&](int i) { read_contig<4><<<2368, 256>>>((const uint4*)src[i], n, sink); }, n_sets, iters), (double)bytes);
        rep("read ring 148 CTAs 8x16K", time_us([&](int i) { read_ring<<<148, 544, 8 * 16384 + 512>>>(src[i], bytes, sink, 8, 16384); }, n_sets, iters), (double)bytes);
        rep("read ring 148 CTAs 6x32K", time_us([&](int i) { read_ring<<<148, 544, 6 * 32768 + 512>>>(src[i], bytes, sink, 6, 32768); }, n_sets, iters), (double)bytes);
        rep("read ring 148 CTAs 12x8K", time_us([&](int i) { read_ring<<<148, 544, 12 * 8192 + 512>>>(src[i], bytes, sink, 12, 8192); }, n_sets, iters), (double)bytes);
        rep("copy contiguous 444x256 U8 (r+w)", time_us([&](int i) { copy_contig<8><<<444, 256>>>((const uint4*)src[i], (uint4*)dst[i], n); }, n_sets, iters), 2.0 * bytes);
        rep("copy contiguous 1184x256 U4 (r+w)", time_us([&](int i) { copy_contig<4><<<1184, 256>>>((const uint4*)src[i], (uint4*)dst[i], n); }, n_sets, iters), 2.0 * bytes);
        rep("cudaMemcpyAsync D2D (r+w)", time_us([&](int i) { cudaMemcpyAsync(dst[i], src[i], bytes, cudaMemcpyDeviceToDevice); }, n_sets, iters), 2.0 * bytes);
        for (int i = 0; i < n_sets; ++i) { cudaFree(src[i]); cudaFree(dst[i]); }
    }
    return 0;
}
