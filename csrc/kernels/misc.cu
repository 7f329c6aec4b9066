// Error plumbing + small utility kernels (cast, column sums, L2 flush).
#include <cstdarg>
#include <mutex>

#include "b200llm.h"
#include "common.cuh"

static char g_err[1024] = "";
static std::mutex g_err_mu;

void b200_set_error(const char* fmt, ...) {
    std::lock_guard<std::mutex> lk(g_err_mu);
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* b200_last_error() { return g_err; }

extern "C" int b200_abi_version() { return B200_ABI_VERSION; }

// ------------------------------------------------------------------------------------------
// fp32 -> bf16 cast (n % 4 == 0 on the vector path, scalar tail)
// ------------------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, bf16* __restrict__ out, long long n) {
    pdl_sync();
    long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 4;
    long long stride = (long long)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(in + i));
        uint2 o = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
        *reinterpret_cast<uint2*>(out + i) = o;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (long long j = n & ~3ll; j < n; ++j) out[j] = __float2bfloat16_rn(in[j]);
}

extern "C" int b200_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t st) {
    if (n <= 0) return 0;
    B200_REQUIRE(((uintptr_t)in % 16 == 0) && ((uintptr_t)out % 8 == 0), "cast: unaligned pointers");
    int threads = 256;
    long long blocks = (n / 4 + threads - 1) / threads;
    long long cap = (long long)b200_num_sms() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    B200_CHECK_CUDA(b200_launch(cast_f32_bf16_kernel, dim3((int)blocks), dim3(threads), 0, st, in, (bf16*)out, n));
    B200_RETURN_LAUNCH_STATUS("cast_f32_bf16");
}

// ------------------------------------------------------------------------------------------
// Column sums of a row-major (M, N) bf16 matrix -> fp32 (N): bias gradients.
// Each CTA owns 64 columns x a slab of rows; partial sums are combined with fp32 RED.
// `out` must be zeroed by the caller unless accumulate != 0.
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) colsum_bf16_kernel(const bf16* __restrict__ a, float* __restrict__ out,
                                                          int M, int N, int ld, int rows_per_cta) {
    pdl_sync();
    // thread (ty, tx): tx in [0,8) covers 8 columns (one 16-byte load), ty in [0,32) strides rows
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int col0 = blockIdx.x * 64 + tx * 8;
    const int r0 = blockIdx.y * rows_per_cta;
    const int r1 = min(M, r0 + rows_per_cta);
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    if (col0 + 7 < N) {
        for (int r = r0 + ty; r < r1; r += 32) {
            uint4 u = ldg_stream_u4(reinterpret_cast<const uint4*>(a + (size_t)r * ld + col0));
            float2 f;
            f = unpack_bf16x2(u.x); acc[0] += f.x; acc[1] += f.y;
            f = unpack_bf16x2(u.y); acc[2] += f.x; acc[3] += f.y;
            f = unpack_bf16x2(u.z); acc[4] += f.x; acc[5] += f.y;
            f = unpack_bf16x2(u.w); acc[6] += f.x; acc[7] += f.y;
        }
    } else {
        for (int r = r0 + ty; r < r1; r += 32)
            for (int j = 0; j < 8; ++j)
                if (col0 + j < N) acc[j] += __bfloat162float(a[(size_t)r * ld + col0 + j]);
    }
    __shared__ float red[32][65];
#pragma unroll
    for (int j = 0; j < 8; ++j) red[ty][tx * 8 + j] = acc[j];
    __syncthreads();
    if (threadIdx.x < 64) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < 32; ++k) s += red[k][threadIdx.x];
        int c = blockIdx.x * 64 + threadIdx.x;
        if (c < N) atomicAdd(out + c, s);
    }
}

extern "C" int b200_colsum_bf16(const void* a, float* out, int M, int N, int ld, int accumulate,
                                cudaStream_t st) {
    B200_REQUIRE(ld % 8 == 0 && (uintptr_t)a % 16 == 0, "colsum: ld must be a multiple of 8 and a 16B aligned");
    if (!accumulate) B200_CHECK_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)N, st));
    int gx = (N + 63) / 64;
    int target = b200_num_sms() * 4;
    int gy = max(1, min((M + 255) / 256, (target + gx - 1) / gx));
    static int det = -1;
    if (det < 0) {
        const char* e = getenv("B200_DETERMINISTIC");      // one CTA per 64 columns walks every row: a single add per
        det = (e != nullptr && e[0] == '1') ? 1 : 0;       // output, in a fixed order (slower: fewer CTAs)
    }
    if (det) gy = 1;
    int rows = (M + gy - 1) / gy;
    rows = (rows + 31) / 32 * 32;
    gy = (M + rows - 1) / rows;
    B200_CHECK_CUDA(b200_launch(colsum_bf16_kernel, dim3(dim3(gx, gy)), dim3(256), 0, st, (const bf16*)a, out, M, N, ld, rows));
    B200_RETURN_LAUNCH_STATUS("colsum_bf16");
}

// ------------------------------------------------------------------------------------------
// L2 flush helper for benchmarks: overwrite a buffer larger than L2.
// ------------------------------------------------------------------------------------------
__global__ void fill_kernel(float4* p, long long n4, float v) {
    pdl_sync();
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n4; i += stride) p[i] = make_float4(v, v, v, v);
}

extern "C" int b200_l2_flush(void* buf, long long bytes, cudaStream_t st) {
    B200_CHECK_CUDA(b200_launch(fill_kernel, dim3(b200_num_sms() * 8), dim3(256), 0, st, (float4*)buf, bytes / 16, 0.f));
    B200_RETURN_LAUNCH_STATUS("l2_flush");
}


// Zero-fill as a memset node (DMA engine, capturable): what the training step uses for the scatter-add target of the
// embedding gradient and the loss accumulator, so that a captured step contains this library's kernels and memset
// nodes only.
extern "C" int b200_memset_zero(void* ptr, long long bytes, cudaStream_t st) {
    if (bytes <= 0) return 0;
    B200_CHECK_CUDA(cudaMemsetAsync(ptr, 0, (size_t)bytes, st));
    return 0;
}
