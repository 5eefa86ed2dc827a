// Micro-benchmark: the per-tile arithmetic of the backward epilogue (64 values per thread: fma, ex2, fma, mul,
// bf16 pack) with 1, 2 or 4 warps per SM sub-partition and nothing else on the SM.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/epi_bench tools/epi_bench.cu
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ long long gns() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

template <int MODE>
__global__ void k(const float* __restrict__ in, uint32_t* out, int iters, long long* ns) {
    extern __shared__ float cj[];
    for (int i = threadIdx.x; i < 64; i += blockDim.x) cj[i] = 0.001f * i;
    __syncthreads();
    float v[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = in[(threadIdx.x * 64 + i) & 1023];
    const float c2 = in[0] + 2.88f, lse = in[1] + 9.f, R = in[2] + 0.5f;
    uint32_t acc = 0;
    const long long t0 = gns();
    for (int it = 0; it < iters; ++it) {
        uint32_t w[32];
#pragma unroll
        for (int c4 = 0; c4 < 16; ++c4) {
            const float4 cc = *reinterpret_cast<const float4*>(cj + c4 * 4);
            const float cv[4] = {cc.x, cc.y, cc.z, cc.w};
            float pv[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float a = fmaf(v[c4 * 4 + u], c2, -lse);
                const float e = (MODE == 1) ? a * 1.0001f : ex2f(a);
                pv[u] = e * fmaf(R, cv[u], 1.f);
            }
            __nv_bfloat162 h0 = __floats2bfloat162_rn(pv[0], pv[1]), h1 = __floats2bfloat162_rn(pv[2], pv[3]);
            w[c4 * 2] = *reinterpret_cast<uint32_t*>(&h0);
            w[c4 * 2 + 1] = *reinterpret_cast<uint32_t*>(&h1);
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) acc ^= w[i];
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] += 1e-6f * (float)(acc & 1);   // keep the loop honest
    }
    const long long t1 = gns();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) ns[blockIdx.x] = t1 - t0;
}

template <int MODE>
void run(const char* name, int threads) {
    float* in; uint32_t* out; long long* ns;
    cudaMalloc(&in, 4096); cudaMemset(in, 0, 4096); cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&ns, 148 * 8);
    const int iters = 2000;
    cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    k<MODE><<<148, threads, 200 * 1024>>>(in, out, 10, ns);
    k<MODE><<<148, threads, 200 * 1024>>>(in, out, iters, ns);
    long long h[148]; cudaMemcpy(h, ns, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const double per_batch_ns = avg / iters;
    printf("%-24s %4d threads (%d warps/SMSP): %.0f ns per 64-value batch per warp-set = %.1f cycles per warp-MUFU slot (ideal 8)\n",
           name, threads, threads / 128, per_batch_ns, per_batch_ns * 1.965 / (64.0 * threads / 128));
    cudaFree(in); cudaFree(out); cudaFree(ns);
}

int main() {
    for (int th : {128, 256, 512}) run<0>("fma+ex2+fma+mul+pack", th);
    for (int th : {128, 256, 512}) run<1>("same without MUFU", th);
    return 0;
}
