// Sharded data parallelism ("zero" mode): every rank keeps only its 1/W slice of the fp32 master weights, the AdamW
// moments and the bf16 weights; gradients exist only in a two-layer rolling window.
//
//  * zero_reduce_adamw: ONE kernel per gradient bucket.  Handshake (every rank has finished writing the bucket into
//    its gradient window) -> reduce-scatter: the owner pulls its slice, summed over the ranks, either with one
//    multimem.ld_reduce per 16 bytes (NVSwitch adds the W copies) or with W peer loads -> AdamW on the owned fp32
//    masters / moments -> the new bf16 values go into the owner's bf16 SHARD (what the peers' all-gather pulls inside
//    their GEMMs, gemm2.cu) and, for the few parameters consumed in fp32 (LayerNorm, biases), into every rank's small
//    replicated fp32 store (multimem.st or W peer stores) -> done flag.  Nothing full-size is written anywhere.
//  * zero_wait: gate on the done flags of a range of buckets (end of step; and, per bucket, before the gradient
//    window it lives in is overwritten by the layer two below).
//  * embed_fwd_sharded: token + position embedding whose fp32 rows are read from their owners' master shards over
//    NVLink (16 k rows x 3-8 KB per step) instead of every rank holding a pushed fp32 copy of the table.
//
// Replaces torch DDP's replicated parameters + Reducer + ncclAllReduce + separate AdamW (reference
// scripts/train_transformer.py:122-123,92,126).
#include "b200llm.h"
#include "common.cuh"
#include "comm.cuh"

#define ZR_MAX_SLOTS 16

struct ZrPeers {
    const float* grad[DP_MAX_WORLD];   // this bucket's gradient buffer on every rank (same layout everywhere)
    const float* mc_grad;              // multicast mapping of the same buffer (MC kernels)
    float* small[DP_MAX_WORLD];        // replicated fp32 store of the small fp32-consumed parameters, every rank
    float* mc_small;
    int* sig[DP_MAX_WORLD];            // signal pads: [2][W][NB] ints (ready | done)
};
struct ZrSlots {                       // parameter slots of the bucket, offsets RELATIVE to the bucket start
    long long end[ZR_MAX_SLOTS];       // ascending END offsets
    long long small_off[ZR_MAX_SLOTS]; // offset of the slot in the small fp32 store minus the slot's start, or LLONG_MIN
    int n;
};
#define ZR_NOT_SMALL (-(1ll << 62))

template <int W, bool MC, int UMC>
__global__ void __launch_bounds__(512) zero_reduce_adamw_kernel(ZrPeers P, int rank, int bucket, int nb,
                                                                const int* __restrict__ epoch_dev, long long lo, long long n,
                                                                float* __restrict__ master, float* __restrict__ m,
                                                                float* __restrict__ v, bf16* __restrict__ shard16,
                                                                const float* __restrict__ hyper, ZrSlots slots,
                                                                unsigned int* cta_counter, int* err,
                                                                float* __restrict__ gacc, int accumulate_only) {
    pdl_sync();
    const int epoch = *epoch_dev;
    __shared__ int s_peers_ok;
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            for (int r = 0; r < W; ++r) st_release_sys(P.sig[r] + rank * nb + bucket, epoch);
        }
        bool ok = true;
        for (int r = 0; r < W; ++r) ok = wait_flag(P.sig[rank] + r * nb + bucket, epoch, err) && ok;
        s_peers_ok = ok ? 1 : 0;
    }
    __syncthreads();
    if (!s_peers_ok) return;      // a peer never arrived: no update, no done flag (the host raises at check_errors)

    const float decay = hyper[0], b2 = hyper[1], w1 = hyper[2], w2 = hyper[3], step_size = hyper[4],
                bc2s = hyper[5], eps = hyper[6], gs = hyper[7];
    const long long n4 = n >> 2;
    // U independent pulls in flight per thread (see dp.cu: a multimem.ld_reduce is a switch round trip away)
    constexpr int U = MC ? UMC : (W <= 2 ? 4 : (W == 4 ? 2 : 1));
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long base = blockIdx.x * (long long)blockDim.x + threadIdx.x; base < n4; base += stride * U) {
        float4 sum[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = base + u * stride;
            if (i < n4) {
                const long long e = lo + i * 4;             // offset inside the bucket
                if (MC) {
                    sum[u] = multimem_ld_reduce_add_f4(P.mc_grad + e);
                } else {
                    float4 g[W];
#pragma unroll
                    for (int r = 0; r < W; ++r) g[r] = ldg_stream_f4(reinterpret_cast<const float4*>(P.grad[r] + e));
                    float4 s = g[0];
#pragma unroll
                    for (int r = 1; r < W; ++r) { s.x += g[r].x; s.y += g[r].y; s.z += g[r].z; s.w += g[r].w; }
                    sum[u] = s;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = base + u * stride;
            if (i >= n4) break;
            const long long e = lo + i * 4;
            float4 s = sum[u];
            if (gacc != nullptr) {
                // gradient accumulation over micro-steps: the reduced slice of every micro-step but the last is added to
                // a 1/W-sized fp32 accumulator (no update); the last one picks the accumulator up and clears it
                float4 ga = reinterpret_cast<float4*>(gacc)[i];
                s.x += ga.x; s.y += ga.y; s.z += ga.z; s.w += ga.w;
                reinterpret_cast<float4*>(gacc)[i] = accumulate_only ? s : make_float4(0.f, 0.f, 0.f, 0.f);
                if (accumulate_only) continue;
            }
            float4 pv = reinterpret_cast<float4*>(master)[i];
            float4 mv = reinterpret_cast<float4*>(m)[i];
            float4 vv = reinterpret_cast<float4*>(v)[i];
            float* pe = &pv.x; float* me = &mv.x; float* ve = &vv.x; const float* se = &s.x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {                   // same operation order as adamw.cu
                const float gk = se[k] * gs;
                pe[k] = pe[k] * decay;
                me[k] = fmaf(w1, gk - me[k], me[k]);
                ve[k] = fmaf(w2 * gk, gk, ve[k] * b2);
                const float denom = __fdiv_rn(__fsqrt_rn(ve[k]), bc2s) + eps;
                pe[k] = fmaf(-step_size, __fdiv_rn(me[k], denom), pe[k]);
            }
            reinterpret_cast<float4*>(master)[i] = pv;
            reinterpret_cast<float4*>(m)[i] = mv;
            reinterpret_cast<float4*>(v)[i] = vv;
            reinterpret_cast<uint2*>(shard16)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
            long long so = ZR_NOT_SMALL;
#pragma unroll 4
            for (int k = 0; k < slots.n; ++k) {
                if (e < slots.end[k]) { so = slots.small_off[k]; break; }
            }
            if (so != ZR_NOT_SMALL) {                       // LayerNorm / bias parameter: every rank reads it in fp32
                if (MC) {
                    multimem_st_f4(P.mc_small + (e + so), pv);
                } else {
#pragma unroll
                    for (int r = 0; r < W; ++r) *reinterpret_cast<float4*>(P.small[r] + (e + so)) = pv;
                }
            }
        }
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(cta_counter, 1u);
        if (prev == gridDim.x - 1) {
            *cta_counter = 0;
            __threadfence_system();
            for (int r = 0; r < W; ++r) st_release_sys(P.sig[r] + (W + rank) * nb + bucket, epoch);
        }
    }
}

// waits until every rank's done flag of buckets [b_lo, b_hi) has reached *epoch_dev - lag
__global__ void zero_wait_kernel(const int* sig, int W, int nb, int b_lo, int b_hi, const int* epoch_dev, int lag, int* err) {
    pdl_sync();
    const int target = *epoch_dev - lag;
    const int nbk = b_hi - b_lo;
    for (int i = threadIdx.x; i < W * nbk; i += blockDim.x) {
        const int r = i / nbk, b = b_lo + i % nbk;
        wait_flag(sig + (W + r) * nb + b, target, err);
    }
}

struct EmbPeers {
    const float* shard[DP_MAX_WORLD];   // rank r's fp32 master slice of the embedding bucket (element 0 = bucket offset r * chunk)
};

// one warp per token row; every float4 is read from the rank that owns it (chunk % 4 == 0)
__global__ void __launch_bounds__(256) embed_fwd_sharded_kernel(const long long* __restrict__ idx, EmbPeers P, long long chunk,
                                                                long long wte_off, long long wpe_off, float* __restrict__ x,
                                                                int M, int T, int d, int V) {
    pdl_sync();
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int m = warp; m < M; m += nwarps) {
        long long tok = idx[m];
        tok = tok < 0 ? 0 : (tok >= V ? V - 1 : tok);
        const long long ea = wte_off + tok * d, eb = wpe_off + (long long)(m % T) * d;
        float4* o = reinterpret_cast<float4*>(x + (size_t)m * d);
        for (int c = lane; c < d / 4; c += 32) {
            const long long a = ea + 4 * c, b = eb + 4 * c;
            const int ra = (int)(a / chunk), rb = (int)(b / chunk);
            const float4 u = ldg_stream_f4(reinterpret_cast<const float4*>(P.shard[ra] + (a - ra * chunk)));
            const float4 w = ldg_stream_f4(reinterpret_cast<const float4*>(P.shard[rb] + (b - rb * chunk)));
            o[c] = make_float4(u.x + w.x, u.y + w.y, u.z + w.z, u.w + w.w);
        }
    }
}

extern "C" int b200_zero_reduce_adamw(void* const* h_grad, void* mc_grad, void* const* h_small, void* mc_small,
                                      void* const* h_sig, int world, int rank, int bucket, int num_buckets,
                                      const int* epoch_dev, long long lo, long long n, float* master, float* m, float* v,
                                      void* shard16, const float* hyper, const long long* h_slot_end,
                                      const long long* h_slot_small_off, int n_slots, void* cta_counter, int* err, int ctas,
                                      float* gacc, int accumulate_only, cudaStream_t st) {
    B200_REQUIRE(world >= 1 && world <= DP_MAX_WORLD, "zero: world size %d not supported (max %d)", world, DP_MAX_WORLD);
    B200_REQUIRE(n_slots <= ZR_MAX_SLOTS, "zero: too many slots in one bucket");
    B200_REQUIRE(n % 4 == 0 && lo % 4 == 0, "zero: slice must be 16-byte aligned");
    if (n <= 0) ctas = 1;
    ZrPeers P;
    memset(&P, 0, sizeof(P));
    for (int r = 0; r < world; ++r) {
        P.grad[r] = (const float*)h_grad[r]; P.small[r] = (float*)h_small[r]; P.sig[r] = (int*)h_sig[r];
    }
    P.mc_grad = (const float*)mc_grad; P.mc_small = (float*)mc_small;
    const bool mc = mc_grad != nullptr && mc_small != nullptr && world > 1;
    ZrSlots S;
    memset(&S, 0, sizeof(S));
    S.n = n_slots;
    for (int k = 0; k < n_slots; ++k) { S.end[k] = h_slot_end[k]; S.small_off[k] = h_slot_small_off[k]; }
#define ZR_LAUNCH(W_)                                                                                               \
    if (mc)                                                                                                         \
        B200_CHECK_CUDA(b200_launch(zero_reduce_adamw_kernel<W_, true, 4>, dim3(ctas), dim3(512), 0, st, P, rank, bucket, num_buckets, \
                                    epoch_dev, lo, n, master, m, v, (bf16*)shard16, hyper, S, (unsigned int*)cta_counter, err, gacc, accumulate_only)); \
    else                                                                                                            \
        B200_CHECK_CUDA(b200_launch(zero_reduce_adamw_kernel<W_, false, 1>, dim3(ctas), dim3(512), 0, st, P, rank, bucket, num_buckets, \
                                    epoch_dev, lo, n, master, m, v, (bf16*)shard16, hyper, S, (unsigned int*)cta_counter, err, gacc, accumulate_only))
    switch (world) {
        case 1: ZR_LAUNCH(1); break;
        case 2: ZR_LAUNCH(2); break;
        case 4: ZR_LAUNCH(4); break;
        case 8: ZR_LAUNCH(8); break;
        default: b200_set_error("zero: world size must be 1, 2, 4 or 8"); return -1;
    }
#undef ZR_LAUNCH
    B200_RETURN_LAUNCH_STATUS("zero_reduce_adamw");
}

extern "C" int b200_zero_wait(const int* sig, int world, int num_buckets, int bucket_lo, int bucket_hi, const int* epoch_dev,
                              int lag, int* err, cudaStream_t st) {
    B200_REQUIRE(bucket_lo >= 0 && bucket_hi <= num_buckets && bucket_lo < bucket_hi, "zero_wait: bad bucket range");
    B200_CHECK_CUDA(b200_launch(zero_wait_kernel, dim3(1), dim3(256), 0, st, sig, world, num_buckets, bucket_lo, bucket_hi,
                                epoch_dev, lag, err));
    B200_RETURN_LAUNCH_STATUS("zero_wait");
}

extern "C" int b200_embed_fwd_sharded(const long long* idx, void* const* h_shard, int world, long long chunk, long long wte_off,
                                      long long wpe_off, float* x, int B, int T, int d, int V, cudaStream_t st) {
    B200_REQUIRE(d % 4 == 0 && chunk % 4 == 0 && wte_off % 4 == 0 && wpe_off % 4 == 0, "embed_fwd_sharded: alignment");
    B200_REQUIRE(world >= 1 && world <= DP_MAX_WORLD, "embed_fwd_sharded: world size");
    EmbPeers P;
    memset(&P, 0, sizeof(P));
    for (int r = 0; r < world; ++r) P.shard[r] = (const float*)h_shard[r];
    const int M = B * T;
    const int blocks = min((M + 7) / 8, b200_num_sms() * 8);
    B200_CHECK_CUDA(b200_launch(embed_fwd_sharded_kernel, dim3(max(blocks, 1)), dim3(256), 0, st, idx, P, chunk, wte_off, wpe_off,
                                x, M, T, d, V));
    B200_RETURN_LAUNCH_STATUS("embed_fwd_sharded");
}
