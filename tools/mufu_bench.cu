// Micro-benchmark: exp2 throughput per SM for the variants the loss epilogues could use.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/mufu_bench tools/mufu_bench.cu
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdint>

__device__ __forceinline__ float ex2f(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t ex2h2(uint32_t x) { uint32_t y; asm("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
// Cody-Waite + degree-3 polynomial on the FMA pipe (FA4-style), x <= 0
__device__ __forceinline__ float ex2poly(float x) {
    x = fmaxf(x, -126.f);
    const float fl = floorf(x);
    const float f = x - fl;                                 // [0, 1)
    float p = fmaf(f, 0.0555054f, 0.2402265f);
    p = fmaf(p, f, 0.6931472f);
    p = fmaf(p, f, 1.0f);
    return __int_as_float(__float_as_int(p) + ((int)fl << 23));
}

// the same without any conversion instruction (F2I / FRND run on the XU pipe next to MUFU): round-to-nearest
// through the 1.5*2^23 magic number, 2^f on [-0.5, 0.5] as a cubic, exponent added to the bits
__device__ __forceinline__ float ex2magic(float x) {
    x = fmaxf(x, -125.f);
    const float t = x + 12582912.f;
    const float f = x - (t - 12582912.f);                   // [-0.5, 0.5]
    float p = fmaf(f, 0.05550411f, 0.24022651f);
    p = fmaf(p, f, 0.69314718f);
    p = fmaf(p, f, 1.0f);
    return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
    float acc[8];
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { x[i] = -seed * (threadIdx.x % 7 + i) * 0.01f; acc[i] = 0.f; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) acc[i] += ex2f(x[i]);
            else if (MODE == 1) acc[i] += ex2poly(x[i]);
            else if (MODE == 2) {           // two values per MUFU op: pack, ex2.f16x2, unpack, add
                __half2 h = __floats2half2_rn(x[i], x[i] - 0.5f);
                uint32_t r = ex2h2(*reinterpret_cast<uint32_t*>(&h));
                float2 f = __half22float2(*reinterpret_cast<__half2*>(&r));
                acc[i] += f.x + f.y;
            } else if (MODE == 3) {         // mixed: 3 MUFU : 1 poly
                acc[i] += (i & 3) == 3 ? ex2poly(x[i]) : ex2f(x[i]);
            } else if (MODE == 4) {         // mixed 1:1
                acc[i] += (i & 1) ? ex2poly(x[i]) : ex2f(x[i]);
            } else if (MODE == 5) {
                acc[i] += ex2magic(x[i]);
            } else if (MODE == 6) {         // 3 MUFU : 1 magic poly
                acc[i] += (i & 3) == 3 ? ex2magic(x[i]) : ex2f(x[i]);
            } else if (MODE == 7) {         // 7 MUFU : 1 magic poly
                acc[i] += (i & 7) == 7 ? ex2magic(x[i]) : ex2f(x[i]);
            } else if (MODE == 8) {         // 1 : 1
                acc[i] += (i & 1) ? ex2magic(x[i]) : ex2f(x[i]);
            }
            x[i] -= 0.001f;
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += acc[i];
    if (s == 12345.678f) out[0] = s;
}

template <int MODE>
void run(const char* name, int per_elem) {
    float* out; cudaMalloc(&out, 4);
    int iters = 4096, sms = 148, ctas = sms * 8;
    k<MODE><<<ctas, 256>>>(out, 16, 1.f);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    k<MODE><<<ctas, 256>>>(out, iters, 1.f);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double elems = (double)ctas * 256 * iters * 8 * per_elem;
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("%-28s %.3f ms  %.2f Gexp/s  %.2f exp/clk/SM (at %d MHz nominal)\n", name, ms, elems / ms * 1e-6,
           elems / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000);
    cudaFree(out);
}

int main() {
    run<0>("MUFU ex2.f32", 1);
    run<1>("poly3 on FMA pipe", 1);
    run<2>("MUFU ex2.f16x2 (2/instr)", 2);
    run<3>("3 MUFU : 1 poly", 1);
    run<4>("1 MUFU : 1 poly", 1);
    run<5>("magic-number poly3", 1);
    run<6>("3 MUFU : 1 magic poly", 1);
    run<7>("7 MUFU : 1 magic poly", 1);
    run<8>("1 MUFU : 1 magic poly", 1);
    return 0;
}
