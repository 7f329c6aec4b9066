// Softmax cross-entropy over bf16 logits produced by the LM-head GEMM (whose epilogue already
// reduced each 128 x BN tile to per-row (max, sum exp) statistics).
//
// Replaces the reference's fp32 upcast of the whole (B*T, V) logits + log_softmax + nll_loss
// (reference src/models/transformer.py:74-77) and their autograd:
//   finalize : stats[n_tiles][M] -> lse[M], loss rows, loss sum        (reads M * n_tiles * 8 B)
//   grad     : logits <- (softmax - onehot) * scale in place (bf16), + column sums = d lm_head.bias
#include "b200llm.h"
#include "common.cuh"

// CTA = 32 rows x 8 tile-slices: thread (tx, ty) merges the statistics entries ty, ty + 8, ... of row tx (consecutive tx
// read consecutive float2 -> coalesced), the 8 partial (max, sum) pairs of a row are merged through shared memory.
// (One thread per row walking all ~400 entries serially was latency-bound: 66 us for 52 MB.)
__global__ void __launch_bounds__(256) ce_finalize_kernel(const float2* __restrict__ stats, int n_tiles,
                                                          const bf16* __restrict__ logits, long long ld,
                                                          const long long* __restrict__ targets,
                                                          float* __restrict__ lse, float* __restrict__ loss_rows,
                                                          float* __restrict__ loss_sum, int M, int V, float sum_scale) {
    pdl_sync();
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int m = blockIdx.x * 32 + tx;
    float mx = -INFINITY, sum = 0.f;
    if (m < M) {
        for (int t = ty; t < n_tiles; t += 8) {
            const float2 s = stats[(size_t)t * M + m];
            const float nm = fmaxf(mx, s.x);
            if (nm > -INFINITY) sum = sum * __expf(mx - nm) + s.y * __expf(s.x - nm);
            mx = nm;
        }
    }
    __shared__ float2 part[8][33];
    part[ty][tx] = make_float2(mx, sum);
    __syncthreads();
    float loss = 0.f;
    if (ty == 0 && m < M) {
#pragma unroll
        for (int k = 1; k < 8; ++k) {
            const float2 s = part[k][tx];
            const float nm = fmaxf(mx, s.x);
            if (nm > -INFINITY) sum = sum * __expf(mx - nm) + s.y * __expf(s.x - nm);
            mx = nm;
        }
        const float l = mx + __logf(sum);
        lse[m] = l;
        long long tgt = targets[m];
        tgt = tgt < 0 ? 0 : (tgt >= V ? V - 1 : tgt);
        loss = l - __bfloat162float(logits[(size_t)m * ld + tgt]);
        if (loss_rows) loss_rows[m] = loss;
    }
    if (ty == 0) {
        const float w = warp_sum(loss);
        if (tx == 0 && loss_sum) atomicAdd(loss_sum, w * sum_scale);
    }
}

extern "C" int b200_ce_finalize(const void* stats, int n_tiles, const void* logits, long long ld,
                                const long long* targets, float* lse, float* loss_rows, float* loss_sum, int M, int V,
                                float sum_scale, cudaStream_t st) {
    B200_CHECK_CUDA(b200_launch(ce_finalize_kernel, dim3((M + 31) / 32), dim3(256), 0, st, (const float2*)stats, n_tiles, (const bf16*)logits, ld, targets,
                                                        lse, loss_rows, loss_sum, M, V, sum_scale));
    B200_RETURN_LAUNCH_STATUS("ce_finalize");
}

// CTA = 256 threads x 8 columns = 2048 columns, ROWS rows.
constexpr int CE_ROWS = 128;

__global__ void __launch_bounds__(256) ce_grad_kernel(bf16* __restrict__ logits, long long ld,
                                                      const float* __restrict__ lse,
                                                      const long long* __restrict__ targets,
                                                      float* __restrict__ dbias, int M, int V, float scale) {
    pdl_sync();
    const int col0 = (blockIdx.x * 256 + threadIdx.x) * 8;
    const int r0 = blockIdx.y * CE_ROWS, r1 = min(M, r0 + CE_ROWS);
    __shared__ float s_lse[CE_ROWS];
    __shared__ int s_tgt[CE_ROWS];
    for (int i = threadIdx.x; i < r1 - r0; i += 256) {
        s_lse[i] = lse[r0 + i];
        s_tgt[i] = (int)targets[r0 + i];
    }
    __syncthreads();
    if (col0 >= V) return;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const bool full = col0 + 7 < V;
    for (int r = r0; r < r1; ++r) {
        bf16* p = logits + (size_t)r * ld + col0;
        const float l = s_lse[r - r0];
        const int tg = s_tgt[r - r0] - col0;
        float g[8];
        if (full) {
            const uint4 u = *reinterpret_cast<const uint4*>(p);
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                float2 f = unpack_bf16x2(w[t]);
                g[2 * t] = f.x;
                g[2 * t + 1] = f.y;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = col0 + j < V ? __bfloat162float(p[j]) : -INFINITY;
        }
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = (__expf(g[j] - l) - (j == tg ? 1.f : 0.f)) * scale;
            // accumulate the rounded value: that is what the wgrad GEMM will see
            v = __bfloat162float(__float2bfloat16_rn(v));
            g[j] = v;
            acc[j] += v;
        }
#pragma unroll
        for (int t = 0; t < 4; ++t) o[t] = pack_bf16x2(g[2 * t], g[2 * t + 1]);
        if (full) {
            *reinterpret_cast<uint4*>(p) = make_uint4(o[0], o[1], o[2], o[3]);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (col0 + j < V) p[j] = __float2bfloat16_rn(g[j]);
        }
    }
    if (dbias != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (col0 + j < V) atomicAdd(dbias + col0 + j, acc[j]);
    }
}

extern "C" int b200_ce_grad(void* logits, long long ld, const float* lse, const long long* targets, float* dbias,
                            int M, int V, float scale, int accumulate, cudaStream_t st) {
    B200_REQUIRE(ld % 8 == 0 && (uintptr_t)logits % 16 == 0, "ce_grad: ld must be a multiple of 8, logits 16B aligned");
    if (dbias != nullptr && !accumulate) B200_CHECK_CUDA(cudaMemsetAsync(dbias, 0, sizeof(float) * (size_t)V, st));
    dim3 grid((V + 2047) / 2048, (M + CE_ROWS - 1) / CE_ROWS);
    B200_CHECK_CUDA(b200_launch(ce_grad_kernel, dim3(grid), dim3(256), 0, st, (bf16*)logits, ld, lse, targets, dbias, M, V, scale));
    B200_RETURN_LAUNCH_STATUS("ce_grad");
}
