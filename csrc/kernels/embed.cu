// Fused token + position embedding, forward and backward.
// Replaces two ATen gathers + an add (reference src/models/transformer.py:51-54) and
// embedding_dense_backward.  fp32 residual stream like the reference under autocast.
#include "b200llm.h"
#include "common.cuh"

// one warp per token row; d % 4 == 0
__global__ void __launch_bounds__(256) embed_fwd_kernel(const long long* __restrict__ idx,
                                                        const float* __restrict__ wte,
                                                        const float* __restrict__ wpe,
                                                        float* __restrict__ x, int M, int T, int d, int V,
                                                        int pos0, const int* __restrict__ pos_dev) {
    pdl_sync();
    if (pos_dev != nullptr) pos0 = *pos_dev;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int m = warp; m < M; m += nwarps) {
        long long tok = idx[m];
        tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
        const int t = pos0 + m % T;
        const float4* a = reinterpret_cast<const float4*>(wte + (size_t)tok * d);
        const float4* b = reinterpret_cast<const float4*>(wpe + (size_t)t * d);
        float4* o = reinterpret_cast<float4*>(x + (size_t)m * d);
        for (int c = lane; c < d / 4; c += 32) {
            float4 u = __ldg(a + c), v = __ldg(b + c);
            o[c] = make_float4(u.x + v.x, u.y + v.y, u.z + v.z, u.w + v.w);
        }
    }
}

extern "C" int b200_embed_fwd(const long long* idx, const float* wte, const float* wpe, float* x,
                              int B, int T, int d, int V, int pos0, const int* pos_dev, cudaStream_t st) {
    B200_REQUIRE(d % 4 == 0, "embed_fwd: d must be a multiple of 4");
    int M = B * T;
    int blocks = min((M + 7) / 8, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_fwd_kernel, dim3(max(blocks, 1)), dim3(256), 0, st, idx, wte, wpe, x, M, T, d, V, pos0, pos_dev));
    B200_RETURN_LAUNCH_STATUS("embed_fwd");
}

// Token-embedding gradient: scatter-add rows of dx into dwte with vector fp32 RED
// (red.global.add.v4.f32, sm_90+).  dwte must hold the running gradient (zeros on the first
// micro-step).  Position-embedding gradient: dwpe[t] (+)= sum_b dx[b, t] - a plain reduction.
__global__ void __launch_bounds__(256) embed_bwd_tok_kernel(const long long* __restrict__ idx,
                                                            const float* __restrict__ dx,
                                                            float* __restrict__ dwte, int M, int d, int V) {
    pdl_sync();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int m = warp; m < M; m += nwarps) {
        long long tok = idx[m];
        tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
        const float4* g = reinterpret_cast<const float4*>(dx + (size_t)m * d);
        float* o = dwte + (size_t)tok * d;
        for (int c = lane; c < d / 4; c += 32) {
            float4 v = ldg_stream_f4(g + c);
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + c * 4), "f"(v.x), "f"(v.y),
                         "f"(v.z), "f"(v.w)
                         : "memory");
        }
    }
}

__global__ void __launch_bounds__(256) embed_bwd_pos_kernel(const float* __restrict__ dx, float* __restrict__ dwpe,
                                                            int B, int T, int d, int accumulate) {
    pdl_sync();
    const int total4 = T * d / 4;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
        float4 acc = accumulate ? reinterpret_cast<float4*>(dwpe)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int b = 0; b < B; ++b) {
            float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(dx + (size_t)b * T * d) + i);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        reinterpret_cast<float4*>(dwpe)[i] = acc;
    }
}

extern "C" int b200_embed_bwd(const long long* idx, const float* dx, float* dwte, float* dwpe, int B, int T,
                              int d, int V, int accumulate, cudaStream_t st) {
    B200_REQUIRE(d % 4 == 0, "embed_bwd: d must be a multiple of 4");
    int M = B * T;
    int blocks = min((M + 7) / 8, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_bwd_tok_kernel, dim3(max(blocks, 1)), dim3(256), 0, st, idx, dx, dwte, M, d, V));
    int total4 = T * d / 4;
    int blocks2 = min((total4 + 255) / 256, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_bwd_pos_kernel, dim3(max(blocks2, 1)), dim3(256), 0, st, dx, dwpe, B, T, d, accumulate));
    B200_RETURN_LAUNCH_STATUS("embed_bwd");
}


// Deterministic token-embedding gradient (B200_DETERMINISTIC=1 in the engine): `order` is a STABLE argsort of the token
// ids (torch.sort), so equal tokens sit next to each other in position order.  One warp per position p that starts a
// run (sorted token differs from its predecessor) sums the run's dx rows in that fixed order and writes - not adds -
// the row, so the result does not depend on the order in which atomics land.  dwte must be zero (or hold the running
// sum when accumulate != 0) for the rows no token touches / before the add.
__global__ void __launch_bounds__(256) embed_bwd_tok_sorted_kernel(const long long* __restrict__ idx,
                                                                   const long long* __restrict__ order,
                                                                   const float* __restrict__ dx, float* __restrict__ dwte,
                                                                   int M, int d, int V, int accumulate) {
    pdl_sync();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int p = warp; p < M; p += nwarps) {
        const long long tok = idx[order[p]];
        if (p > 0 && idx[order[p - 1]] == tok) continue;          // not the start of a run
        const long long row = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
        float* o = dwte + (size_t)row * d;
        for (int c = lane; c < d / 4; c += 32) {
            float4 acc = accumulate ? reinterpret_cast<const float4*>(o)[c] : make_float4(0.f, 0.f, 0.f, 0.f);
            for (int q = p; q < M && idx[order[q]] == tok; ++q) {
                const float4 v = ldg_stream_f4(reinterpret_cast<const float4*>(dx + (size_t)order[q] * d) + c);
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            reinterpret_cast<float4*>(o)[c] = acc;
        }
    }
}

extern "C" int b200_embed_bwd_sorted(const long long* idx, const long long* order, const float* dx, float* dwte, float* dwpe,
                                     int B, int T, int d, int V, int accumulate, cudaStream_t st) {
    B200_REQUIRE(d % 4 == 0, "embed_bwd_sorted: d must be a multiple of 4");
    const int M = B * T;
    const int blocks = min((M + 7) / 8, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_bwd_tok_sorted_kernel, dim3(max(blocks, 1)), dim3(256), 0, st, idx, order, dx, dwte, M, d, V,
                                accumulate));
    const int total4 = T * d / 4;
    const int blocks2 = min((total4 + 255) / 256, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_bwd_pos_kernel, dim3(max(blocks2, 1)), dim3(256), 0, st, dx, dwpe, B, T, d, accumulate));
    B200_RETURN_LAUNCH_STATUS("embed_bwd_sorted");
}
