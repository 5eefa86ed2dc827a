// Diagnostics: how load flavour and L2 eviction hints change a streaming pass on this GPU (single pass over a cold
// tensor, and a second pass over the same tensor right behind it).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/l2_bench tools/l2_bench.cu
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

// MODE 0: ld.global.nc.L1::no_allocate          1: ld.global.L1::no_allocate (coherent)
//      2: coherent + L2 evict_last hint          3: coherent + L2 evict_first hint
//      4: .nc + evict_first hint                 5: .nc + evict_last hint
template <int MODE>
__device__ __forceinline__ uint4 ld16(const void* p, unsigned long long pol) {
    uint4 v;
    if (MODE == 0) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 1) asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (MODE == 2 || MODE == 3) asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    return v;
}
template <int MODE>
__global__ void __launch_bounds__(1024) rd(const uint4* __restrict__ src, size_t n, unsigned* sink) {
    unsigned long long pol = 0;
    if (MODE == 2 || MODE == 5) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    if (MODE == 3 || MODE == 4) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    unsigned acc = 0;
    const size_t per = (n + gridDim.x - 1) / gridDim.x;
    const size_t beg = (size_t)blockIdx.x * per, end = beg + per < n ? beg + per : n;
    size_t i = beg + threadIdx.x;
    for (; i + 7 * blockDim.x < end; i += 8 * blockDim.x) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = ld16<MODE>(src + i + u * blockDim.x, pol);
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
    }
    for (; i < end; i += blockDim.x) { uint4 v = ld16<MODE>(src + i, pol); acc += v.x ^ v.w; }
    if (acc == 0x12345678u) *sink = acc;
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
    const size_t sizes[] = {8u << 20, 32u << 20, 64u << 20, 256u << 20};
    unsigned* sink;
    cudaMalloc(&sink, 4);
    for (size_t bytes : sizes) {
        const int n_sets = (int)((1024u << 20) / bytes) < 2 ? 2 : (int)((1024u << 20) / bytes);
        std::vector<uint8_t*> src(n_sets);
        for (int i = 0; i < n_sets; ++i) { cudaMalloc(&src[i], bytes); cudaMemset(src[i], i + 1, bytes); }
        const size_t n = bytes / 16;
        const int iters = 40;
        const char* names[6] = {"nc", "coherent", "coherent+evict_last", "coherent+evict_first", "nc+evict_first", "nc+evict_last"};
#define SINGLE(M) printf("%4zu MiB single pass %-22s %8.2f us\n", bytes >> 20, names[M], \
        time_us([&](int i) { rd<M><<<148, 1024>>>((const uint4*)src[i], n, sink); }, n_sets, iters));
        SINGLE(0) SINGLE(1) SINGLE(2) SINGLE(3) SINGLE(4) SINGLE(5)
#define PAIR(A, B) printf("%4zu MiB pair %-20s then %-20s %8.2f us\n", bytes >> 20, names[A], names[B], \
        time_us([&](int i) { rd<A><<<148, 1024>>>((const uint4*)src[i], n, sink); rd<B><<<148, 1024>>>((const uint4*)src[i], n, sink); }, n_sets, iters));
        PAIR(0, 0) PAIR(1, 1) PAIR(2, 3) PAIR(5, 4) PAIR(2, 1) PAIR(1, 3)
        for (int i = 0; i < n_sets; ++i) cudaFree(src[i]);
    }
    return 0;
}
