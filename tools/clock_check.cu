// Diagnostics: SM clock (clock64 cycles per globaltimer ns) for an idle-ish spin and for a MUFU-heavy spin.
// nvcc -gencode arch=compute_100a,code=sm_100a -o tools/bin/clock_check tools/clock_check.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ long long gns() { long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

__global__ void spin(long long cycles, int heavy, long long* out, float* sink) {
    const long long c0 = clock64(), t0 = gns();
    float a = threadIdx.x * 1e-3f, acc = 0.f;
    while (clock64() - c0 < cycles) {
        if (heavy) {
#pragma unroll
            for (int i = 0; i < 64; ++i) { acc += exp2f(a); a = a * 0.999f + 1e-4f; }
        }
    }
    const long long c1 = clock64(), t1 = gns();
    if (threadIdx.x == 0) { out[2 * blockIdx.x] = c1 - c0; out[2 * blockIdx.x + 1] = t1 - t0; }
    if (acc == 123.456f) *sink = acc;
}

int main() {
    long long* d; float* s;
    cudaMalloc(&d, 148 * 2 * sizeof(long long)); cudaMalloc(&s, 4);
    long long h[296];
    const long long lens[4] = {40000, 200000, 2000000, 20000000};
    for (int heavy = 0; heavy < 2; ++heavy)
        for (int li = 0; li < 4; ++li)
            for (int rep = 0; rep < 3; ++rep) {
                spin<<<148, heavy ? 512 : 32>>>(lens[li], heavy, d, s);
                cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
                double mn = 1e9, mx = 0;
                for (int b = 0; b < 148; ++b) { double f = (double)h[2 * b] / (double)h[2 * b + 1]; mn = f < mn ? f : mn; mx = f > mx ? f : mx; }
                printf("heavy=%d cycles=%lld rep=%d: %.3f .. %.3f GHz (SM0: %lld cyc in %lld ns)\n", heavy, lens[li], rep, mn, mx, h[0], h[1]);
            }
    return 0;
}
