// Host-side TMA descriptor (CUtensorMap) construction with a small cache.
// The driver entry point is resolved through the runtime so the library links only libcudart.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

struct Key {
    const void* ptr;
    uint64_t d0, d1, ld_bytes;
    uint32_t b0, b1, elt, swz;
    bool operator==(const Key& o) const { return memcmp(this, &o, sizeof(Key)) == 0; }
};
struct KeyHash {
    size_t operator()(const Key& k) const {
        const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
        size_t h = 1469598103934665603ull;
        for (size_t i = 0; i < sizeof(Key) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
        return h;
    }
};

// 2-D row-major tensor [rows, cols] (cols contiguous), element size `elt` bytes (2 = bf16, 4 = fp32),
// row pitch ld_bytes, box = {box_cols, box_rows}, 128-byte swizzle.
inline int make_2d(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_bytes,
                   uint32_t box_cols, uint32_t box_rows, uint32_t elt, bool swizzle128 = true) {
    static std::mutex mu;
    static std::unordered_map<Key, CUtensorMap, KeyHash> cache;
    Key k;
    memset(&k, 0, sizeof(k));
    k.ptr = ptr; k.d0 = cols; k.d1 = rows; k.ld_bytes = ld_bytes; k.b0 = box_cols; k.b1 = box_rows; k.elt = elt;
    k.swz = swizzle128 ? 1 : 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(k);
        if (it != cache.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn fn = encode_fn();
    if (!fn) { b200_set_error("cuTensorMapEncodeTiled entry point not available"); return -1; }
    if ((uintptr_t)ptr % 16 != 0 || ld_bytes % 16 != 0) {
        b200_set_error("TMA: base pointer and row pitch must be 16-byte aligned (ptr=%p pitch=%llu)", ptr,
                       (unsigned long long)ld_bytes);
        return -1;
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld_bytes};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUtensorMapDataType dt = elt == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUresult r = fn(out, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        b200_set_error("cuTensorMapEncodeTiled failed: CUresult %d (rows=%llu cols=%llu pitch=%llu box=%ux%u)", (int)r,
                       (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_bytes, box_cols,
                       box_rows);
        return -1;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 65536) cache.clear();
    cache.emplace(k, *out);
    return 0;
}

// 3-D tensor [d2][d1][d0] (d0 contiguous), pitches in bytes for d1 and d2, box = {b0, b1, 1}, 128-byte swizzle.
// Used for per-batch attention tiles: out-of-range rows of one batch element are zero-filled on load and
// clipped on store instead of spilling into the next batch element.
inline int make_3d(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t pitch1,
                   uint64_t pitch2, uint32_t b0, uint32_t b1, uint32_t elt) {
    static std::mutex mu;
    static std::unordered_map<Key, CUtensorMap, KeyHash> cache;
    Key k;
    memset(&k, 0, sizeof(k));
    k.ptr = ptr; k.d0 = d0; k.d1 = d1 ^ (d2 << 40); k.ld_bytes = pitch1 ^ (pitch2 << 1); k.b0 = b0; k.b1 = b1; k.elt = elt;
    k.swz = 3;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find(k);
        if (it != cache.end()) { *out = it->second; return 0; }
    }
    EncodeTiledFn fn = encode_fn();
    if (!fn) { b200_set_error("cuTensorMapEncodeTiled entry point not available"); return -1; }
    if ((uintptr_t)ptr % 16 != 0 || pitch1 % 16 != 0 || pitch2 % 16 != 0) {
        b200_set_error("TMA: base pointer and pitches must be 16-byte aligned");
        return -1;
    }
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {pitch1, pitch2};
    cuuint32_t box[3] = {b0, b1, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUtensorMapDataType dt = elt == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUresult r = fn(out, dt, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        b200_set_error("cuTensorMapEncodeTiled(3d) failed: CUresult %d", (int)r);
        return -1;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (cache.size() > 65536) cache.clear();
    cache.emplace(k, *out);
    return 0;
}

}  // namespace tma
