// Fused residual-add + LayerNorm, forward and backward.
//
// Forward replaces, per block of the reference, `x = x + branch` followed by `ln(x)`
// (reference src/models/transformer_block.py:44,46; transformer.py:70): one pass reads the
// fp32 residual stream and the bf16 branch output, writes the updated fp32 residual, the
// bf16 normalised activations for the GEMM that follows, and the row statistics.
// Backward fuses LN-backward with the residual-gradient accumulation, emits the bf16 copy of
// the residual gradient that the next GEMMs consume, and produces per-CTA partial sums for
// dgamma / dbeta (and, optionally, the column sums of the incoming residual gradient = the
// bias gradient of the Linear that fed this residual).
//
// Thread mapping: a row is handled by WPR warps, each lane holding VPL float4 (d = 128*VPL*WPR)
// in registers, so every element is read exactly once.  CTAs have 8 warps = 8/WPR row groups
// that synchronise among themselves with named barriers only.
#include "b200llm.h"
#include "common.cuh"

namespace {

template <int WPR>
__device__ __forceinline__ void group_bar(int group) {
    if (WPR > 1) asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "r"(WPR * 32) : "memory");
}

template <int WPR>
__device__ __forceinline__ float group_sum(float v, float* sred, int wir, int group, int lane) {
    v = warp_sum(v);
    if (WPR == 1) return v;
    if (lane == 0) sred[wir] = v;
    group_bar<WPR>(group);
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < WPR; ++i) t += sred[i];
    group_bar<WPR>(group);
    return t;
}

template <int VPL, int WPR>
__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* x, float* x_out, const bf16* __restrict__ branch,
                                                     const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, bf16* __restrict__ out,
                                                     float* __restrict__ out_f32, float* __restrict__ mean_out,
                                                     float* __restrict__ rstd_out, int M, float eps) {
    pdl_sync();
    constexpr int G = 8 / WPR;
    constexpr int D4 = 32 * WPR * VPL;  // float4 per row
    __shared__ float sred_all[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int group = warp / WPR, wir = warp % WPR;
    float* sred = sred_all + group * WPR;
    const int row = blockIdx.x * G + group;
    if (row >= M) return;  // whole group exits together

    float4 v[VPL];
    const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D4 * 4);
    float4* xo = reinterpret_cast<float4*>(x_out + (size_t)row * D4 * 4);  // may alias x
#pragma unroll
    for (int i = 0; i < VPL; ++i) v[i] = xr[wir * 32 + lane + i * 32 * WPR];
    if (branch != nullptr) {
        const uint2* br = reinterpret_cast<const uint2*>(branch + (size_t)row * D4 * 4);
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            uint2 u = ldg_stream_u2(br + wir * 32 + lane + i * 32 * WPR);
            float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y);
            v[i].x += a.x; v[i].y += a.y; v[i].z += b.x; v[i].w += b.y;
            xo[wir * 32 + lane + i * 32 * WPR] = v[i];
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    const float mean = group_sum<WPR>(s, sred, wir, group, lane) * (1.f / (D4 * 4));
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, e = v[i].w - mean;
        q += (a * a + b * b) + (c * c + e * e);
    }
    const float var = group_sum<WPR>(q, sred, wir, group, lane) * (1.f / (D4 * 4));
    const float rstd = rsqrtf(var + eps);
    if (wir == 0 && lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
    const float4* g4 = reinterpret_cast<const float4*>(gamma);
    const float4* b4 = reinterpret_cast<const float4*>(beta);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        const int c = wir * 32 + lane + i * 32 * WPR;
        float4 g = __ldg(g4 + c), b = __ldg(b4 + c);
        float4 y;
        y.x = (v[i].x - mean) * rstd * g.x + b.x;
        y.y = (v[i].y - mean) * rstd * g.y + b.y;
        y.z = (v[i].z - mean) * rstd * g.z + b.z;
        y.w = (v[i].w - mean) * rstd * g.w + b.w;
        if (out) reinterpret_cast<uint2*>(out + (size_t)row * D4 * 4)[c] =
                make_uint2(pack_bf16x2(y.x, y.y), pack_bf16x2(y.z, y.w));
        if (out_f32) reinterpret_cast<float4*>(out_f32 + (size_t)row * D4 * 4)[c] = y;
    }
}

template <int VPL, int WPR, bool COLSUM>
__global__ void __launch_bounds__(256, 2)
    ln_bwd_kernel(const bf16* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ mean,
                  const float* __restrict__ rstd, const float* __restrict__ gamma, const float* dx_in,
                  float* dx_out, bf16* __restrict__ dx_bf16, float* __restrict__ partial, int M) {
    pdl_sync();
    constexpr int G = 8 / WPR;
    constexpr int D4 = 32 * WPR * VPL;
    constexpr int NACC = COLSUM ? 3 : 2;
    __shared__ float sred_all[16];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int group = warp / WPR, wir = warp % WPR;
    float* sred = sred_all + group * WPR * 2;

    float4 acc_g[VPL], acc_b[VPL], acc_c[COLSUM ? VPL : 1];
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
        acc_g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        acc_b[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (COLSUM) acc_c[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const float4* g4 = reinterpret_cast<const float4*>(gamma);

    for (int row = blockIdx.x * G + group; row < M; row += gridDim.x * G) {
        const size_t base4 = (size_t)row * D4;
        const float mu = mean[row], rs = rstd[row];
        float4 xh[VPL], gd[VPL], pin[VPL];
        float s1 = 0.f, s2 = 0.f;
        // the incoming residual gradient is only needed after the row reduction, but loading it HERE keeps it in the
        // same memory round trip as x and dy (it was a second, serialised round trip per row: 71% -> of the HBM floor)
        if (dx_in != nullptr) {
#pragma unroll
            for (int i = 0; i < VPL; ++i)
                pin[i] = reinterpret_cast<const float4*>(dx_in)[base4 + wir * 32 + lane + i * 32 * WPR];  // may alias dx_out
        }
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = wir * 32 + lane + i * 32 * WPR;
            float4 xv = ldg_stream_f4(reinterpret_cast<const float4*>(x) + base4 + c);
            uint2 u = ldg_stream_u2(reinterpret_cast<const uint2*>(dy) + base4 + c);
            float2 d0 = unpack_bf16x2(u.x), d1 = unpack_bf16x2(u.y);
            float4 g = __ldg(g4 + c);
            xh[i].x = (xv.x - mu) * rs; xh[i].y = (xv.y - mu) * rs;
            xh[i].z = (xv.z - mu) * rs; xh[i].w = (xv.w - mu) * rs;
            acc_g[i].x += d0.x * xh[i].x; acc_g[i].y += d0.y * xh[i].y;
            acc_g[i].z += d1.x * xh[i].z; acc_g[i].w += d1.y * xh[i].w;
            acc_b[i].x += d0.x; acc_b[i].y += d0.y; acc_b[i].z += d1.x; acc_b[i].w += d1.y;
            gd[i].x = d0.x * g.x; gd[i].y = d0.y * g.y; gd[i].z = d1.x * g.z; gd[i].w = d1.y * g.w;
            s1 += (gd[i].x + gd[i].y) + (gd[i].z + gd[i].w);
            s2 += (gd[i].x * xh[i].x + gd[i].y * xh[i].y) + (gd[i].z * xh[i].z + gd[i].w * xh[i].w);
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (WPR > 1) {
            if (lane == 0) { sred[wir * 2] = s1; sred[wir * 2 + 1] = s2; }
            group_bar<WPR>(group);
            s1 = 0.f; s2 = 0.f;
#pragma unroll
            for (int i = 0; i < WPR; ++i) { s1 += sred[i * 2]; s2 += sred[i * 2 + 1]; }
            group_bar<WPR>(group);
        }
        const float c1 = s1 * (1.f / (D4 * 4)), c2 = s2 * (1.f / (D4 * 4));
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = wir * 32 + lane + i * 32 * WPR;
            float4 r;
            r.x = rs * (gd[i].x - c1 - xh[i].x * c2);
            r.y = rs * (gd[i].y - c1 - xh[i].y * c2);
            r.z = rs * (gd[i].z - c1 - xh[i].z * c2);
            r.w = rs * (gd[i].w - c1 - xh[i].w * c2);
            if (dx_in != nullptr) {
                const float4 p = pin[i];
                if (COLSUM) { acc_c[i].x += p.x; acc_c[i].y += p.y; acc_c[i].z += p.z; acc_c[i].w += p.w; }
                r.x += p.x; r.y += p.y; r.z += p.z; r.w += p.w;
            }
            reinterpret_cast<float4*>(dx_out)[base4 + c] = r;
            if (dx_bf16 != nullptr)
                reinterpret_cast<uint2*>(dx_bf16)[base4 + c] =
                    make_uint2(pack_bf16x2(r.x, r.y), pack_bf16x2(r.z, r.w));
        }
    }
    // CTA-level reduction of the G row-groups through shared memory, then ONE partial row per CTA:
    // partial[cta][NACC][d]
    extern __shared__ float s_part[];   // [G][NACC][D4*4]
    {
        float4* sp = reinterpret_cast<float4*>(s_part) + (size_t)group * NACC * D4;
#pragma unroll
        for (int i = 0; i < VPL; ++i) {
            const int c = wir * 32 + lane + i * 32 * WPR;
            sp[c] = acc_g[i];
            sp[D4 + c] = acc_b[i];
            if (COLSUM) sp[2 * D4 + c] = acc_c[i];
        }
    }
    __syncthreads();
    float4* pr = reinterpret_cast<float4*>(partial) + (size_t)blockIdx.x * NACC * D4;
    for (int e = threadIdx.x; e < NACC * D4; e += 256) {
        float4 t = reinterpret_cast<const float4*>(s_part)[e];
#pragma unroll
        for (int gi = 1; gi < G; ++gi) {
            const float4 u = reinterpret_cast<const float4*>(s_part)[(size_t)gi * NACC * D4 + e];
            t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
        }
        pr[e] = t;
    }
}

// out[k][c] (+)= sum_p partial[p][k][c];  CTA = 32 columns x 8 row lanes, grid = (d/32, nacc)
__global__ void __launch_bounds__(256) ln_bwd_finalize_kernel(const float* __restrict__ partial, int nparts, int nacc,
                                                              int d, float* __restrict__ o0, float* __restrict__ o1,
                                                              float* __restrict__ o2, int accumulate) {
    pdl_sync();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    const int k = blockIdx.y;
    float* o = k == 0 ? o0 : (k == 1 ? o1 : o2);
    __shared__ float red[8][33];
    float s = 0.f;
    if (c < d && o != nullptr) {
        const float* p = partial + (size_t)k * d + c;
        const size_t stride = (size_t)nacc * d;
        for (int i = ty; i < nparts; i += 8) s += p[(size_t)i * stride];
    }
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < d && o != nullptr) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += red[j][tx];
        o[c] = accumulate ? o[c] + t : t;
    }
}

struct LnCfg { int vpl, wpr; };
inline bool ln_cfg(int d, LnCfg* c) {
    switch (d) {
        case 128: *c = {1, 1}; return true;
        case 256: *c = {2, 1}; return true;
        case 384: *c = {3, 1}; return true;
        case 512: *c = {4, 1}; return true;
        case 768: *c = {3, 2}; return true;
        case 1024: *c = {4, 2}; return true;
        case 1536: *c = {3, 4}; return true;
        case 2048: *c = {4, 4}; return true;
        case 3072: *c = {3, 8}; return true;
        case 4096: *c = {4, 8}; return true;
        default: return false;
    }
}

}  // namespace

#define LN_DISPATCH(VPL_, WPR_, ...)                    \
    if (cfg.vpl == VPL_ && cfg.wpr == WPR_) {           \
        constexpr int VPL = VPL_, WPR = WPR_;           \
        __VA_ARGS__;                                    \
    }
#define LN_DISPATCH_ALL(...)                                                                         \
    LN_DISPATCH(1, 1, __VA_ARGS__) LN_DISPATCH(2, 1, __VA_ARGS__) LN_DISPATCH(3, 1, __VA_ARGS__)     \
    LN_DISPATCH(4, 1, __VA_ARGS__) LN_DISPATCH(3, 2, __VA_ARGS__) LN_DISPATCH(4, 2, __VA_ARGS__)     \
    LN_DISPATCH(3, 4, __VA_ARGS__) LN_DISPATCH(4, 4, __VA_ARGS__) LN_DISPATCH(3, 8, __VA_ARGS__)     \
    LN_DISPATCH(4, 8, __VA_ARGS__)

extern "C" int b200_ln_supported(int d) {
    LnCfg c;
    return ln_cfg(d, &c) ? 1 : 0;
}

extern "C" int b200_ln_fwd(const float* x, const void* branch, float* x_out, const float* gamma, const float* beta,
                           void* out_bf16, float* out_f32, float* mean, float* rstd, int M, int d, float eps,
                           cudaStream_t st) {
    B200_REQUIRE(branch == nullptr || x_out != nullptr, "ln_fwd: x_out is required when a branch is added");
    LnCfg cfg;
    B200_REQUIRE(ln_cfg(d, &cfg), "ln_fwd: unsupported width d=%d", d);
    if (M <= 0) return 0;
    const int G = 8 / cfg.wpr;
    const int grid = (M + G - 1) / G;
    LN_DISPATCH_ALL({
        B200_CHECK_CUDA(b200_launch(ln_fwd_kernel<VPL, WPR>, dim3(grid), dim3(256), 0, st, x, x_out, (const bf16*)branch,
                                    gamma, beta, (bf16*)out_bf16, out_f32, mean, rstd, M, eps));
    })
    B200_RETURN_LAUNCH_STATUS("ln_fwd");
}

// number of partial rows ln_bwd writes for (M, d): the caller sizes `partial` as
// nparts * (colsum ? 3 : 2) * d floats.
extern "C" int b200_ln_bwd_nparts(int M, int d) {
    LnCfg cfg;
    if (!ln_cfg(d, &cfg)) return -1;
    const int G = 8 / cfg.wpr;
    int grid = min((M + G - 1) / G, b200_num_sms() * 2);
    return max(grid, 1);
}

// main kernel of the backward: dx (+ residual gradient, + bf16 copy) and one partial row of (dgamma, dbeta[, colsum]) per CTA
static int ln_bwd_main(const void* dy, const float* x, const float* mean, const float* rstd, const float* gamma,
                       const float* dx_in, float* dx_out, void* dx_bf16, float* partial, bool cs, int M, int d,
                       int* grid_out, cudaStream_t st) {
    LnCfg cfg;
    B200_REQUIRE(ln_cfg(d, &cfg), "ln_bwd: unsupported width d=%d", d);
    const int G = 8 / cfg.wpr;
    const int grid = max(1, min((M + G - 1) / G, b200_num_sms() * 2));
    B200_REQUIRE(!cs || dx_in != nullptr, "ln_bwd: colsum needs dx_in");
    const int nacc = cs ? 3 : 2;
    const int smem = G * nacc * d * (int)sizeof(float);
    if (cs) {
        LN_DISPATCH_ALL({
            auto kern = ln_bwd_kernel<VPL, WPR, true>;
            static bool cfg_done = false;
            if (!cfg_done) { B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304)); cfg_done = true; }
            B200_CHECK_CUDA(b200_launch(kern, dim3(grid), dim3(256), smem, st, (const bf16*)dy, x, mean, rstd, gamma, dx_in, dx_out, (bf16*)dx_bf16, partial, M));
        })
    } else {
        LN_DISPATCH_ALL({
            auto kern = ln_bwd_kernel<VPL, WPR, false>;
            static bool cfg_done = false;
            if (!cfg_done) { B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304)); cfg_done = true; }
            B200_CHECK_CUDA(b200_launch(kern, dim3(grid), dim3(256), smem, st, (const bf16*)dy, x, mean, rstd, gamma, dx_in, dx_out, (bf16*)dx_bf16, partial, M));
        })
    }
    *grid_out = grid;
    return 0;
}

extern "C" int b200_ln_bwd(const void* dy, const float* x, const float* mean, const float* rstd,
                           const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16,
                           float* partial, float* dgamma, float* dbeta, float* colsum, int M, int d,
                           int accumulate, cudaStream_t st) {
    if (M <= 0) return 0;
    const bool cs = colsum != nullptr;
    int grid = 0;
    if (int rc = ln_bwd_main(dy, x, mean, rstd, gamma, dx_in, dx_out, dx_bf16, partial, cs, M, d, &grid, st)) return rc;
    const int nacc = cs ? 3 : 2;
    B200_CHECK_CUDA(b200_launch(ln_bwd_finalize_kernel, dim3(dim3((d + 31) / 32, nacc)), dim3(256), 0, st, partial, grid, nacc, d, dgamma, dbeta, colsum,
                                                                       accumulate));
    B200_RETURN_LAUNCH_STATUS("ln_bwd");
}

// Two-step form: the main kernel only (partial sums stay in `partial`), and ONE finalize launch for two LayerNorms of
// the same (M, d) - the two of a transformer block - instead of one tiny launch each (25 per step, 10 us apiece).
extern "C" int b200_ln_bwd_partial(const void* dy, const float* x, const float* mean, const float* rstd,
                                   const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16,
                                   float* partial, int with_colsum, int M, int d, cudaStream_t st) {
    if (M <= 0) return 0;
    int grid = 0;
    if (int rc = ln_bwd_main(dy, x, mean, rstd, gamma, dx_in, dx_out, dx_bf16, partial, with_colsum != 0, M, d, &grid, st)) return rc;
    B200_RETURN_LAUNCH_STATUS("ln_bwd_partial");
}

struct LnFin { const float* partial; float* o0; float* o1; float* o2; int nacc; };

__global__ void __launch_bounds__(256) ln_bwd_finalize2_kernel(LnFin a, LnFin b, int nparts, int d, int accumulate) {
    pdl_sync();
    const LnFin f = blockIdx.z == 0 ? a : b;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    const int k = blockIdx.y;
    if (k >= f.nacc) return;                                   // uniform per CTA
    float* o = k == 0 ? f.o0 : (k == 1 ? f.o1 : f.o2);
    __shared__ float red[8][33];
    float s = 0.f;
    if (c < d && o != nullptr) {
        const float* p = f.partial + (size_t)k * d + c;
        const size_t stride = (size_t)f.nacc * d;
        for (int i = ty; i < nparts; i += 8) s += p[(size_t)i * stride];
    }
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < d && o != nullptr) {
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) t += red[j][tx];
        o[c] = accumulate ? o[c] + t : t;
    }
}

extern "C" int b200_ln_bwd_finalize2(const float* partial_a, float* dgamma_a, float* dbeta_a, float* colsum_a,
                                     const float* partial_b, float* dgamma_b, float* dbeta_b, float* colsum_b, int M, int d,
                                     int accumulate, cudaStream_t st) {
    if (M <= 0) return 0;
    const int nparts = b200_ln_bwd_nparts(M, d);
    B200_REQUIRE(nparts > 0, "ln_bwd_finalize2: unsupported width d=%d", d);
    LnFin a{partial_a, dgamma_a, dbeta_a, colsum_a, colsum_a ? 3 : 2};
    LnFin b{partial_b, dgamma_b, dbeta_b, colsum_b, colsum_b ? 3 : 2};
    B200_CHECK_CUDA(b200_launch(ln_bwd_finalize2_kernel, dim3((d + 31) / 32, 3, 2), dim3(256), 0, st, a, b, nparts, d, accumulate));
    B200_RETURN_LAUNCH_STATUS("ln_bwd_finalize2");
}

// =============================================================================================
// LayerNorm backward for callers that must not allocate.
//
// b200_ln_bwd needs a scratch buffer for its per-CTA partial sums and leaves it to the caller to
// provide one.  Some callers cannot: a C/C++ host that handed the whole device to its own memory
// arena at start-up has no cudaMalloc left to give (it would fail, or fragment the arena), and a
// steady-state training loop may refuse cudaMalloc/cudaFree because they synchronise the whole
// device.  For them the scratch below is part of the library image: a fixed array in device global
// memory that every GPU loading this library carries until the process exits (6.94 MiB).
//
//   * capacity: LN_BWD_STATIC_FLOATS floats = 296 partial rows x 3 accumulators x width 2048.  The
//     kernel needs  b200_ln_bwd_nparts(M, d) * (colsum ? 3 : 2) * d  floats; bigger requests are
//     refused with an error, never truncated.
//   * the array's NAME is only a host-side handle.  Its device address is looked up once per device
//     with cudaGetSymbolAddress, with that device current, on the first call made on that device;
//     that first call should therefore happen before any stream capture starts.  Later calls reuse
//     the cached address and make no CUDA runtime call besides the kernel launches.
//   * nothing in the array has to be zero on entry: every call overwrites each partial row it later
//     reads, so the previous call's leftovers are never consumed.
//   * there is ONE array per GPU.  Two calls whose kernels can overlap on the same GPU (different
//     streams, or different host threads) race on it; serialise them on one stream.
// No cudaMalloc, cudaFree or any allocating library routine is called on this path.
// =============================================================================================
#define LN_BWD_STATIC_FLOATS (296 * 3 * 2048)
static __device__ __align__(256) float g_ln_bwd_static_scratch[LN_BWD_STATIC_FLOATS];

// Host threads: the per-device cache below is written without a lock.  Two threads that make the FIRST call on the same
// device at the same time both look the symbol up and store the same value (a benign same-value race on an aligned
// pointer); the array itself is what concurrent callers must not share (see above).
static float* ln_bwd_static_address() {
    static float* addr[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        b200_set_error("ln_bwd_static: cannot identify the current device");
        return nullptr;
    }
    if (addr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_ln_bwd_static_scratch);
        if (e != cudaSuccess || p == nullptr) {
            b200_set_error("ln_bwd_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return nullptr;
        }
        addr[dev] = static_cast<float*>(p);
    }
    return addr[dev];
}

extern "C" long long b200_ln_bwd_static_capacity() { return (long long)LN_BWD_STATIC_FLOATS; }

extern "C" int b200_ln_bwd_static(const void* dy, const float* x, const float* mean, const float* rstd,
                                  const float* gamma, const float* dx_in, float* dx_out, void* dx_bf16, float* dgamma,
                                  float* dbeta, float* colsum, int M, int d, int accumulate, cudaStream_t st) {
    if (M <= 0) return 0;                       // nothing to do (same early-out as b200_ln_bwd_partial)
    const int nparts = b200_ln_bwd_nparts(M, d);
    B200_REQUIRE(nparts > 0, "ln_bwd_static: unsupported width d=%d", d);
    const long long need = (long long)nparts * (colsum != nullptr ? 3 : 2) * d;
    B200_REQUIRE(need <= (long long)LN_BWD_STATIC_FLOATS,
                 "ln_bwd_static: M=%d d=%d needs %lld floats of scratch but the built-in array holds %lld "
                 "(use b200_ln_bwd with a caller-provided buffer)", M, d, need, (long long)LN_BWD_STATIC_FLOATS);
    float* scratch = ln_bwd_static_address();
    if (scratch == nullptr) return -1;
    return b200_ln_bwd(dy, x, mean, rstd, gamma, dx_in, dx_out, dx_bf16, scratch, dgamma, dbeta, colsum, M, d, accumulate,
                       st);
}
