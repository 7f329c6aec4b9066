// Shared device/host helpers for the sm_100a kernels of libb200llm.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

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

// Launch helper.  Programmatic dependent launch is OPT-IN (B200_PDL=1): measured on B200 it is neutral for the
// training step (818.7k vs 823.9k tokens/s, within noise) and slower for the decode step (11.6k vs 15.4k
// tokens/s at batch 32: early-resident dependents of the many tiny kernels get in the way), so it stays off.
// B200_PDL: 0 / unset = off, 1 = every kernel, 2 = only the kernels that ask for it (the CTA-pair GEMM, whose prologue -
// tensor-map prefetch, barrier init, TMEM allocation - is long enough to be worth overlapping with the previous kernel's tail)
static inline int b200_pdl_mode() {
    static int mode = -1;
    if (mode < 0) {
        const char* e = getenv("B200_PDL");
        mode = (e != nullptr && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 0;
    }
    return mode;
}

template <bool WANTS_PDL = false, typename... KArgs, typename... Args>
static inline cudaError_t b200_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      Args&&... args) {
    const int mode = b200_pdl_mode();
    const int pdl = (mode == 1 || (mode == 2 && WANTS_PDL)) ? 1 : 0;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

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

// ---- programmatic dependent launch (PDL) ------------------------------------------------------------
// With B200_PDL=1 every kernel is launched with programmaticStreamSerialization: its CTAs may become resident
// and run their prologue (barrier init, TMEM allocation, descriptor prefetch) while the previous kernel of the
// stream drains.  pdl_sync() is the point after which memory written (or still read) by earlier kernels may be
// touched: it lets OUR dependents start launching, then waits for all prerequisite grids to finish and flush.
__device__ __forceinline__ void pdl_sync() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
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
