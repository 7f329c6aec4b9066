// Shared device/host helpers for the sm_100a kernels of libb200llm.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

typedef __nv_bfloat16 bf16;
typedef __nv_bfloat162 bf162;

extern "C" const char* b200_last_error();
void b200_set_error(const char* fmt, ...);

// Every entry point ends with this: no device sync, just the launch status.
#define B200_RETURN_LAUNCH_STATUS(what)                                     \
    do {                                                                    \
        cudaError_t e__ = cudaGetLastError();                               \
        if (e__ != cudaSuccess) {                                           \
            b200_set_error("%s: %s", what, cudaGetErrorString(e__));        \
            return (int)e__;                                                \
        }                                                                   \
        return 0;                                                           \
    } while (0)

#define B200_CHECK_CUDA(expr)                                               \
    do {                                                                    \
        cudaError_t e__ = (expr);                                           \
        if (e__ != cudaSuccess) {                                           \
            b200_set_error("%s failed: %s", #expr, cudaGetErrorString(e__)); \
            return (int)e__;                                                \
        }                                                                   \
    } while (0)

#define B200_REQUIRE(cond, ...)                                             \
    do {                                                                    \
        if (!(cond)) {                                                      \
            b200_set_error(__VA_ARGS__);                                    \
            return -1;                                                      \
        }                                                                   \
    } while (0)

static inline int b200_num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
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

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    bf162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
    bf162 v = *reinterpret_cast<bf162*>(&u);
    return __bfloat1622float2(v);
}

// streaming (read-once) 16-byte load / store that do not pollute L1
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint2 ldg_stream_u2(const uint2* p) {
    uint2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];"
                 : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
