// Shared helpers for the sm_100a kernels behind the C-ABI in include/simclr_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/simclr_b200.h"

namespace scb {

// ---------------------------------------------------------------- error reporting
void set_error(const char* fmt, ...);

#define SCB_CUDA_OK(expr)                                                              \
    do {                                                                               \
        cudaError_t _e = (expr);                                                       \
        if (_e != cudaSuccess) {                                                       \
            scb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,               \
                           cudaGetErrorString(_e));                                    \
            return SCB_ERR_CUDA;                                                       \
        }                                                                              \
    } while (0)

#define SCB_REQUIRE(cond, ...)                                                         \
    do {                                                                               \
        if (!(cond)) {                                                                 \
            scb::set_error(__VA_ARGS__);                                               \
            return SCB_ERR_ARG;                                                        \
        }                                                                              \
    } while (0)

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

int sm_count();   // SMs of the current device (cached per device)
void count_launch(int n = 1);   // bookkeeping behind scb_launch_count()

// The loss pipeline alternates tiny helper kernels with the persistent tensor-core kernels, which need
// the maximum shared-memory carve-out; asking for the same carve-out everywhere avoids an L1/shared
// re-partition of every SM between consecutive launches.
// Function attributes stick (per device): set each (kernel, attribute, value) once instead of paying a driver
// call on every launch.  attr_done() returns true when this combination was already applied on the device.
bool attr_done(const void* kern, int attr, int value);   // api.cu

template <typename Kern>
inline void prefer_max_smem_carveout(Kern kern) {
    if (attr_done((const void*)kern, (int)cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared))
        return;
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
}
template <typename Kern>
inline cudaError_t set_max_dynamic_smem(Kern kern, int bytes) {
    if (attr_done((const void*)kern, (int)cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)) return cudaSuccess;
    return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

// ---------------------------------------------------------------- programmatic dependent launch
// Kernels of one call chain (prep -> forward -> finalize -> backward -> reduce) are launched with
// programmatic stream serialisation: the next grid is scheduled while the previous one still runs, does
// its prologue (barrier init, TMEM allocation, descriptor prefetch) and then blocks in pdl_wait() until the
// previous grid has completed and its memory is visible.  RULE: a kernel launched through launch_pdl()
// calls pdl_wait() in every thread before its first global-memory access, and reads what an earlier grid of
// the chain wrote through coherent loads only (no __ldg, no `const T* __restrict__` parameter: ld.global.nc
// promises data that does not change during the reading grid's lifetime, and that grid is already alive while
// its predecessor writes).  Without the launch attribute both instructions are no-ops.  SCB_PDL=0 in the environment leaves the attribute off (api.cu).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();   // api.cu

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- dtype plumbing
template <typename T> struct Vec16 { static constexpr int N = 16 / sizeof(T); };

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// Streaming 16-byte global accesses: data is touched once per pass, keep it out of L1.
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void stg_stream16(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
                 :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// load VEC consecutive elements (VEC == 16/sizeof(T) -> one 16-byte access, VEC == 1 -> scalar)
template <typename T, int VEC>
__device__ __forceinline__ void load_row_vec(const T* p, float (&v)[VEC]) {
    if constexpr (VEC == 1) {
        v[0] = to_f32(*p);
    } else {
        static_assert(VEC * sizeof(T) == 16, "vector path is 16 bytes");
        uint4 raw = ldg_stream16(p);
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int i = 0; i < VEC; ++i) v[i] = to_f32(e[i]);
    }
}
// same, through L1: for the few lines that EVERY thread of the grid reads (a common reference row, per-channel
// vectors) -- streamed with L1::no_allocate they would all queue up on the same L2 lines
template <typename T, int VEC>
__device__ __forceinline__ void load_row_vec_cached(const T* p, float (&v)[VEC]) {
    if constexpr (VEC == 1) {
        v[0] = to_f32(__ldg(p));
    } else {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(p));
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int i = 0; i < VEC; ++i) v[i] = to_f32(e[i]);
    }
}
template <typename T, int VEC>
__device__ __forceinline__ void store_row_vec(T* p, const float (&v)[VEC]) {
    if constexpr (VEC == 1) {
        *p = from_f32<T>(v[0]);
    } else {
        uint4 raw;
        T* e = reinterpret_cast<T*>(&raw);
#pragma unroll
        for (int i = 0; i < VEC; ++i) e[i] = from_f32<T>(v[i]);
        stg_stream16(p, raw);
    }
}

// Per-channel parameters (weight, bias, running statistics, their gradients) in the module's own dtype: the pointers
// are declared float* for the common fp32 case and read / written through these with the run-time dtype code, so a
// bf16 model needs no conversion kernels around the batch-norm passes.  A null pointer reads as `dflt`.
__device__ __forceinline__ float ldp(const float* p, int dt, int i, float dflt) {
    if (p == nullptr) return dflt;
    if (dt == SCB_F32) return p[i];
    if (dt == SCB_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    return __half2float(reinterpret_cast<const __half*>(p)[i]);
}
__device__ __forceinline__ void stp(float* p, int dt, int i, float v) {
    if (dt == SCB_F32) p[i] = v;
    else if (dt == SCB_BF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
    else reinterpret_cast<__half*>(p)[i] = __float2half_rn(v);
}

// a device scalar whose element type is only known at run time (SCB_F32 / SCB_BF16 / SCB_F16)
__device__ __forceinline__ float load_scalar(const void* p, int dtype) {
    if (dtype == SCB_BF16) return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(p));
    if (dtype == SCB_F16) return __half2float(*reinterpret_cast<const __half*>(p));
    return *reinterpret_cast<const float*>(p);
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Chan's pairwise merge of two (count, mean, M2) summaries; (n == 0) entries are neutral.
__device__ __forceinline__ void chan_merge(float& n, float& mean, float& m2,
                                           float n2, float mean2, float m22) {
    if (n2 <= 0.f) return;
    const float tot = n + n2;
    const float f = n2 / tot;
    const float delta = mean2 - mean;
    mean = fmaf(delta, f, mean);
    m2 = m2 + m22 + delta * delta * n * f;
    n = tot;
}

}  // namespace scb

// dtype dispatch: binds `T` inside the lambda-like body
#define SCB_DISPATCH_FLOAT(dtype, T, ...)                                              \
    switch (dtype) {                                                                   \
        case SCB_F32: { using T = float; __VA_ARGS__; break; }                         \
        case SCB_BF16: { using T = __nv_bfloat16; __VA_ARGS__; break; }                \
        case SCB_F16: { using T = __half; __VA_ARGS__; break; }                        \
        default: scb::set_error("unsupported dtype code %d", (int)(dtype)); return SCB_ERR_ARG; \
    }
