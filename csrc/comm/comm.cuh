// Device-side helpers shared by the communication kernels (dp.cu, zero.cu, vp.cu): system-scope flags in peer
// memory and the NVSwitch multicast (multimem) loads / stores.
#pragma once
#include "common.cuh"

#define DP_MAX_WORLD 8

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// spin until *p >= target; gives up after ~20 s and raises the error flag instead of hanging
__device__ __forceinline__ bool wait_flag(const int* p, int target, int* err) {
    const long long t0 = clock64();
    while (ld_acquire_sys(p) < target) {
        __nanosleep(64);
        if (clock64() - t0 > 40000000000ll) {
            if (err) *err = 1;
            return false;
        }
    }
    return true;
}

// ---- multimem: `p` is an address inside a multicast mapping (cuMulticastBindMem of every rank's buffer) ----------
// one load that returns the SUM of the 16 bytes at this offset of every rank's copy, added inside the NVSwitch
__device__ __forceinline__ float4 multimem_ld_reduce_add_f4(const float* p) {
    float4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    return r;
}
// one store the NVSwitch replicates into every rank's copy
__device__ __forceinline__ void multimem_st_f4(float* p, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x4(bf16* p, uint2 v) {     // 4 bf16 = 8 bytes
    asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void multimem_st_bf16x8(bf16* p, uint4 v) {     // 8 bf16 = 16 bytes
    asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
