// Vocab-parallel LM head + cross-entropy across the ranks of one NVSwitch domain.
//
// Every rank owns a contiguous slice of the vocabulary (rows of lm_head.weight / entries of lm_head.bias).
// Per step, entirely over peer-mapped memory (no NCCL):
//   vp_gather_rows   pull every rank's final hidden states (and targets) -> hf_all [W*M, d]
//   (tcgen05 GEMM)   logits_slice = hf_all @ W_slice^T + b_slice, epilogue emits per-(row, tile) (max, sum exp)
//   vp_ce_finalize   each rank, for ITS OWN M rows: combines the statistics of ALL ranks' vocabulary slices
//                    (P2P loads), fetches the target logit from the slice that owns the target, writes the
//                    loss and pushes log-sum-exp of its rows to every rank (P2P stores)
//   (ce_grad)        dlogits_slice in place for all W*M rows, d bias slice
//   (tcgen05 GEMMs)  dW_slice = dlogits^T @ hf_all  (complete: no gradient all-reduce for the LM head)
//                    dhf_partial = dlogits @ W_slice
//   vp_reduce_rows   each rank sums the W partial dhf rows of its own tokens (P2P loads), scales, -> bf16
// Hand-shakes are system-scope flags in peer memory (same protocol as comm/dp.cu), keyed by the device-side
// step counter.  Replaces the reference's replicated lm_head + F.cross_entropy (src/models/transformer.py:71-77)
// and DDP's all-reduce of the largest gradient tensor of the model.
#include "b200llm.h"
#include "common.cuh"

#define VP_MAX_WORLD 8

namespace {

__device__ __forceinline__ void st_release_sys(int* p, int v) {
    asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_sys(const int* p) {
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

struct VpPtrs { void* p[VP_MAX_WORLD]; };

// every rank has reached this point of step `epoch`: sig[r] = rank r's pad, int[nslots][W]
__device__ __forceinline__ void vp_handshake(const VpPtrs& sig, int W, int rank, int slot, int epoch, int* err) {
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            for (int r = 0; r < W; ++r) st_release_sys((int*)sig.p[r] + slot * W + rank, epoch);
        }
        for (int r = 0; r < W; ++r) {
            const long long t0 = clock64();
            while (ld_acquire_sys((const int*)sig.p[rank] + slot * W + r) < epoch) {
                __nanosleep(64);
                if (clock64() - t0 > 40000000000ll) { if (err) *err = 2; break; }
            }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(256) vp_barrier_kernel(VpPtrs sig, int W, int rank, int slot, const int* epoch_dev, int* err) {
    pdl_sync();
    vp_handshake(sig, W, rank, slot, *epoch_dev, err);
}

// dst_all[src*M + i, :] = hf[src][i, :]   (16-byte vectors; own rows included), tgt_all likewise
__global__ void __launch_bounds__(256) vp_gather_rows_kernel(VpPtrs hf, VpPtrs tgt, bf16* __restrict__ dst_all,
                                                            long long* __restrict__ tgt_all, int M, int d, int W, int rank,
                                                            long long tgt_shift, VpPtrs sig, int slot, const int* epoch_dev, int* err) {
    pdl_sync();
    vp_handshake(sig, W, rank, slot, *epoch_dev, err);
    const long long vec_per_rank = (long long)M * d / 8;
    const long long total = vec_per_rank * W;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += stride) {
        const int src = (int)(i / vec_per_rank);
        const long long j = i - (long long)src * vec_per_rank;
        reinterpret_cast<uint4*>(dst_all)[i] = ldg_stream_u4(reinterpret_cast<const uint4*>(hf.p[src]) + j);
    }
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)M * W; i += stride) {
        const int src = (int)(i / M);
        tgt_all[i] = reinterpret_cast<const long long*>(tgt.p[src])[i - (long long)src * M] - tgt_shift;   // slice-local column
    }
}

// one thread per OWN row: global row g = rank*M + i
__global__ void __launch_bounds__(128) vp_ce_finalize_kernel(VpPtrs stats, VpPtrs logits, VpPtrs lse_all,
                                                            const long long* __restrict__ targets, float* __restrict__ loss_sum,
                                                            float* __restrict__ loss_rows, int M, int W, int rank, int V, int Vs,
                                                            int nt_local, long long ldl, float sum_scale, VpPtrs sig, int slot,
                                                            const int* epoch_dev, int* err) {
    pdl_sync();
    vp_handshake(sig, W, rank, slot, *epoch_dev, err);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const long long Mw = (long long)M * W;
    float loss = 0.f;
    if (i < M) {
        const long long g = (long long)rank * M + i;
        float mx = -INFINITY, sum = 0.f;
        for (int src = 0; src < W; ++src) {
            const float2* st = reinterpret_cast<const float2*>(stats.p[src]);
            for (int t = 0; t < nt_local; ++t) {
                const float2 s = st[(size_t)t * Mw + g];
                if (s.x == -INFINITY) continue;          // tile entirely outside this rank's slice
                const float nm = fmaxf(mx, s.x);
                sum = sum * __expf(mx - nm) + s.y * __expf(s.x - nm);
                mx = nm;
            }
        }
        const float l = mx + __logf(sum);
        long long tg = targets[i];
        tg = tg < 0 ? 0 : (tg >= V ? V - 1 : tg);
        const int owner = (int)(tg / Vs);
        const bf16 tl = reinterpret_cast<const bf16*>(logits.p[owner])[(size_t)g * ldl + (tg - (long long)owner * Vs)];
        loss = l - __bfloat162float(tl);
        if (loss_rows) loss_rows[i] = loss;
        for (int r = 0; r < W; ++r) reinterpret_cast<float*>(lse_all.p[r])[g] = l;
    }
    __shared__ float red[4];
    float w = warp_sum(loss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = w;
    __syncthreads();
    if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, (red[0] + red[1] + red[2] + red[3]) * sum_scale);
}

// out[i, :] = scale * sum_src partial[src][rank*M + i, :]
__global__ void __launch_bounds__(256) vp_reduce_rows_kernel(VpPtrs partial, bf16* __restrict__ out, int M, int d, int W, int rank,
                                                            float scale, VpPtrs sig, int slot, const int* epoch_dev, int* err) {
    pdl_sync();
    vp_handshake(sig, W, rank, slot, *epoch_dev, err);
    const long long nvec = (long long)M * d / 8;
    const long long base = (long long)rank * M * d / 8;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float acc[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] = 0.f;
        uint4 v[VP_MAX_WORLD];
#pragma unroll
        for (int s = 0; s < VP_MAX_WORLD; ++s)
            if (s < W) v[s] = ldg_stream_u4(reinterpret_cast<const uint4*>(partial.p[s]) + base + i);
#pragma unroll
        for (int s = 0; s < VP_MAX_WORLD; ++s) {
            if (s < W) {
                const uint32_t w4[4] = {v[s].x, v[s].y, v[s].z, v[s].w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 f = unpack_bf16x2(w4[k]);
                    acc[2 * k] += f.x;
                    acc[2 * k + 1] += f.y;
                }
            }
        }
        reinterpret_cast<uint4*>(out)[i] = make_uint4(pack_bf16x2(acc[0] * scale, acc[1] * scale),
                                                      pack_bf16x2(acc[2] * scale, acc[3] * scale),
                                                      pack_bf16x2(acc[4] * scale, acc[5] * scale),
                                                      pack_bf16x2(acc[6] * scale, acc[7] * scale));
    }
}

inline VpPtrs mk(void* const* h, int W) {
    VpPtrs p;
    memset(&p, 0, sizeof(p));
    for (int r = 0; r < W && r < VP_MAX_WORLD; ++r) p.p[r] = h[r];
    return p;
}

}  // namespace

extern "C" int b200_vp_barrier(void* const* h_sig, int world, int rank, int slot, const int* epoch_dev, int* err,
                               cudaStream_t st) {
    B200_REQUIRE(world <= VP_MAX_WORLD, "vp: world too large");
    B200_CHECK_CUDA(b200_launch(vp_barrier_kernel, dim3(1), dim3(32), 0, st, mk(h_sig, world), world, rank, slot, epoch_dev, err));
    B200_RETURN_LAUNCH_STATUS("vp_barrier");
}

extern "C" int b200_vp_gather_rows(void* const* h_hf, void* const* h_tgt, void* dst_all, long long* tgt_all, int M, int d,
                                   int world, int rank, long long tgt_shift, void* const* h_sig, int slot, const int* epoch_dev,
                                   int* err, cudaStream_t st) {
    B200_REQUIRE(world <= VP_MAX_WORLD && ((long long)M * d) % 8 == 0, "vp_gather_rows: bad shape");
    B200_CHECK_CUDA(b200_launch(vp_gather_rows_kernel, dim3(b200_num_sms() * 2), dim3(256), 0, st, mk(h_hf, world), mk(h_tgt, world), (bf16*)dst_all, tgt_all, M, d,
                                                              world, rank, tgt_shift, mk(h_sig, world), slot, epoch_dev, err));
    B200_RETURN_LAUNCH_STATUS("vp_gather_rows");
}

extern "C" int b200_vp_ce_finalize(void* const* h_stats, void* const* h_logits, void* const* h_lse_all, const long long* targets,
                                   float* loss_sum, float* loss_rows, int M, int world, int rank, int V, int Vs, int nt_local,
                                   long long ldl, float sum_scale, void* const* h_sig, int slot, const int* epoch_dev, int* err,
                                   cudaStream_t st) {
    B200_REQUIRE(world <= VP_MAX_WORLD, "vp: world too large");
    B200_CHECK_CUDA(b200_launch(vp_ce_finalize_kernel, dim3((M + 127) / 128), dim3(128), 0, st, mk(h_stats, world), mk(h_logits, world), mk(h_lse_all, world), targets,
                                                           loss_sum, loss_rows, M, world, rank, V, Vs, nt_local, ldl, sum_scale,
                                                           mk(h_sig, world), slot, epoch_dev, err));
    B200_RETURN_LAUNCH_STATUS("vp_ce_finalize");
}

extern "C" int b200_vp_reduce_rows(void* const* h_partial, void* out, int M, int d, int world, int rank, float scale,
                                   void* const* h_sig, int slot, const int* epoch_dev, int* err, cudaStream_t st) {
    B200_REQUIRE(world <= VP_MAX_WORLD && ((long long)M * d) % 8 == 0, "vp_reduce_rows: bad shape");
    B200_CHECK_CUDA(b200_launch(vp_reduce_rows_kernel, dim3(b200_num_sms() * 2), dim3(256), 0, st, mk(h_partial, world), (bf16*)out, M, d, world, rank, scale,
                                                              mk(h_sig, world), slot, epoch_dev, err));
    B200_RETURN_LAUNCH_STATUS("vp_reduce_rows");
}
