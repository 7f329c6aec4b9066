// Data-parallel communication over NVLink 5 / NVSwitch without NCCL on the gradient path.
//
//  * symmetric memory bootstrap: device buffers allocated with cudaMalloc, exported as CUDA IPC
//    handles, imported (peer-mapped) by every other rank of the node - 64-byte handles travel
//    over the torch.distributed store, nothing else does;
//  * dp_reduce_adamw: ONE kernel per gradient bucket that (1) handshakes with the peers through
//    system-scope flags in peer memory, (2) pulls this rank's 1/W slice of the bucket's fp32
//    gradients straight out of every peer's gradient buffer (P2P loads over NVLink), sums and
//    scales them, (3) runs the AdamW update on the slice's fp32 master weights and (sharded)
//    moments, (4) pushes the new bf16 shadow weights (and, for parameters that are consumed in
//    fp32 - embeddings, LayerNorm, biases - the fp32 values) into every peer's buffers (P2P
//    stores), (5) publishes a completion flag.  It runs on a side stream as soon as backward has
//    produced the bucket, overlapping the rest of backward.
//  * dp_wait_done: the next forward's gate - waits until every owner has delivered every bucket.
//
// Replaces torch DDP's Reducer + ncclAllReduce per 25 MiB bucket + a separate foreach AdamW
// (reference scripts/train_transformer.py:122-123,92).
#include "b200llm.h"
#include "common.cuh"
#include "comm.cuh"

#define DP_MAX_SLOTS 16

// ---------------------------------------------------------------------------------------------
// symmetric memory bootstrap (CUDA IPC)
// ---------------------------------------------------------------------------------------------
extern "C" int b200_ipc_alloc(void** ptr, long long bytes) {
    B200_CHECK_CUDA(cudaMalloc(ptr, (size_t)bytes));
    B200_CHECK_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
    return 0;
}
extern "C" int b200_ipc_free(void* ptr) {
    B200_CHECK_CUDA(cudaFree(ptr));
    return 0;
}
// handle_out: 64 bytes (host)
extern "C" int b200_ipc_export(void* ptr, void* h_handle_out) {
    cudaIpcMemHandle_t h;
    B200_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(h_handle_out, &h, sizeof(h));
    return 0;
}
extern "C" int b200_ipc_import(const void* h_handle, void** ptr_out) {
    cudaIpcMemHandle_t h;
    memcpy(&h, h_handle, sizeof(h));
    B200_CHECK_CUDA(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int b200_ipc_close(void* ptr) {
    B200_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}

struct DpPeers {
    float* grad[DP_MAX_WORLD];   // full flat fp32 gradient buffer of every rank
    bf16* wb[DP_MAX_WORLD];      // full flat bf16 shadow weights of every rank
    float* flat[DP_MAX_WORLD];   // full flat fp32 master weights of every rank
    int* sig[DP_MAX_WORLD];      // signal pad of every rank: [2][W][NB] ints (ready | done)
    float* mc_grad;              // multicast mappings of the three buffers (MC kernels only; null otherwise)
    bf16* mc_wb;
    float* mc_flat;
};
struct DpSlots {                 // slots of the bucket: element END offsets (ascending) + "peers need fp32"
    long long end[DP_MAX_SLOTS];
    int fp32[DP_MAX_SLOTS];
    int n;
};

// MC = true: the buffers are bound into NVSwitch multicast objects.  The pull is ONE multimem.ld_reduce per 16 bytes
// (the switch adds the W copies) instead of W peer loads, and each push is ONE multimem.st the switch replicates
// instead of W peer stores: NVLink carries 1/W-th of the reduce traffic into this GPU and one copy out of it.
template <int W, bool MC, int UMC>
__global__ void __launch_bounds__(512) dp_reduce_adamw_kernel(DpPeers P, int rank, int bucket, int nb, const int* __restrict__ epoch_dev,
                                                              long long start, long long n, float* __restrict__ m,
                                                              float* __restrict__ v, const float* __restrict__ hyper,
                                                              DpSlots slots, unsigned int* cta_counter, int* err) {
    pdl_sync();
    const int epoch = *epoch_dev;   // device-resident step counter: the same CUDA graph serves every step
    // ---- 1. handshake: every rank has finished writing this bucket's gradients
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
    // A peer never showed up (dead or diverged rank): its gradients are not there.  Leave the weights, the moments and
    // the done flag untouched - the error flag is up, dp_wait_done times out on the peers, and the host raises at its
    // next check_errors() - instead of applying an update built from incomplete sums.
    if (!s_peers_ok) return;

    const float decay = hyper[0], b2 = hyper[1], w1 = hyper[2], w2 = hyper[3], step_size = hyper[4],
                bc2s = hyper[5], eps = hyper[6], gs = hyper[7];
    float* p_master = P.flat[rank];
    const long long n4 = n >> 2;  // slices are multiples of 4 elements
    // U independent 16-byte pulls in flight per thread before the first use: a multimem.ld_reduce comes back after a
    // switch round trip (~4-5 us), so one load per thread leaves NVLink latency-bound (148 x 512 x 16 B = 1.2 MB in flight
    // against ~2-3 MB of bandwidth-delay product).  The per-peer path already has W loads per thread in flight.
    constexpr int U = MC ? UMC : (W <= 2 ? 4 : (W == 4 ? 2 : 1));
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long base = blockIdx.x * (long long)blockDim.x + threadIdx.x; base < n4; base += stride * U) {
        float4 sum[U];
        // ---- 2. pull + reduce
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long i = base + u * stride;
            if (i < n4) {
                const long long e = start + i * 4;
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
            const long long e = start + i * 4;
            const float4 s = sum[u];
            // ---- 3. AdamW on the owned slice (same operation order as adamw.cu)
            float4 pv = *reinterpret_cast<float4*>(p_master + e);
            float4 mv = reinterpret_cast<float4*>(m)[i];
            float4 vv = reinterpret_cast<float4*>(v)[i];
            float* pe = &pv.x; float* me = &mv.x; float* ve = &vv.x; const float* se = &s.x;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float gk = se[k] * gs;
                pe[k] = pe[k] * decay;
                me[k] = fmaf(w1, gk - me[k], me[k]);
                ve[k] = fmaf(w2 * gk, gk, ve[k] * b2);
                const float denom = __fdiv_rn(__fsqrt_rn(ve[k]), bc2s) + eps;
                pe[k] = fmaf(-step_size, __fdiv_rn(me[k], denom), pe[k]);
            }
            reinterpret_cast<float4*>(m)[i] = mv;
            reinterpret_cast<float4*>(v)[i] = vv;
            // ---- 4. push to every rank (own copy included)
            int need32 = 0;
#pragma unroll 4
            for (int k = 0; k < slots.n; ++k) {
                if (e < slots.end[k]) { need32 = slots.fp32[k]; break; }
            }
            const uint2 pb = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
            *reinterpret_cast<float4*>(p_master + e) = pv;
            if (need32 == 2) {
                // embedding tables: nobody consumes a bf16 copy, and embed_fwd reads the fp32 rows from their OWNER over
                // NVLink (b200_embed_fwd_sharded) - nothing to push.  This is the bucket backward finishes last, so its
                // reduce is the exposed one: no 206 MB of outgoing stores behind the last backward kernel.
                *reinterpret_cast<uint2*>(P.wb[rank] + e) = pb;
            } else if (MC) {
                multimem_st_bf16x4(P.mc_wb + e, pb);
                if (need32) multimem_st_f4(P.mc_flat + e, pv);
            } else {
#pragma unroll
                for (int r = 0; r < W; ++r) {
                    *reinterpret_cast<uint2*>(P.wb[r] + e) = pb;
                    if (need32 && r != rank) *reinterpret_cast<float4*>(P.flat[r] + e) = pv;
                }
            }
        }
    }
    // ---- 5. publish: all CTAs' stores are visible before the done flag goes up
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

__global__ void dp_wait_done_kernel(const int* sig, int W, int nb, const int* epoch_dev, int* err) {
    pdl_sync();
    const int epoch = *epoch_dev;
    const int i = threadIdx.x;
    if (i < W * nb) wait_flag(sig + W * nb + i, epoch, err);
}

extern "C" int b200_dp_reduce_adamw(void* const* h_grad, void* const* h_wb, void* const* h_flat, void* const* h_sig,
                                    int world, int rank, int bucket, int num_buckets, const int* epoch_dev, long long start,
                                    long long n, float* m, float* v, const float* hyper, const long long* h_slot_end,
                                    const int* h_slot_fp32, int n_slots, void* cta_counter, int* err, int ctas,
                                    void* mc_grad, void* mc_wb, void* mc_flat, cudaStream_t st) {
    B200_REQUIRE(world >= 1 && world <= DP_MAX_WORLD, "dp: world size %d not supported (max %d)", world, DP_MAX_WORLD);
    B200_REQUIRE(n_slots <= DP_MAX_SLOTS, "dp: too many slots in one bucket");
    B200_REQUIRE(n % 4 == 0 && start % 4 == 0, "dp: slice must be 16-byte aligned");
    if (n <= 0) ctas = 1;
    DpPeers P;
    memset(&P, 0, sizeof(P));
    for (int r = 0; r < world; ++r) {
        P.grad[r] = (float*)h_grad[r]; P.wb[r] = (bf16*)h_wb[r]; P.flat[r] = (float*)h_flat[r]; P.sig[r] = (int*)h_sig[r];
    }
    const bool mc = mc_grad != nullptr && mc_wb != nullptr && mc_flat != nullptr && world > 1;
    P.mc_grad = (float*)mc_grad; P.mc_wb = (bf16*)mc_wb; P.mc_flat = (float*)mc_flat;
    DpSlots S;
    memset(&S, 0, sizeof(S));
    S.n = n_slots;
    for (int k = 0; k < n_slots; ++k) { S.end[k] = h_slot_end[k]; S.fp32[k] = h_slot_fp32[k]; }
    // CTA size: 512 threads measured best (B200_DP_THREADS sweeps it).  Smaller CTAs (128 threads = 7.4K registers) can
    // share an SM with the backward's GEMM / attention CTAs, which 512 x 58 registers cannot, but the step got SLOWER
    // (2 GPUs: 18.7 ms at 128 threads vs 17.0 ms at 512): what is exposed is the last bucket (embeddings), whose
    // duration is set by the bytes the kernel keeps in flight, not by co-residency with compute.
    static int threads = 0;
    if (threads == 0) {
        const char* e = getenv("B200_DP_THREADS");
        threads = e ? atoi(e) : 512;
        if (threads < 32 || threads > 512 || threads % 32) threads = 512;
    }
    static int unroll = 0;     // multimem pulls in flight per thread: 4 (default) or 1 (B200_DP_UNROLL=1, for A/B runs)
    if (unroll == 0) {
        const char* e = getenv("B200_DP_UNROLL");
        unroll = (e != nullptr && atoi(e) == 1) ? 1 : 4;
    }
#define DP_LAUNCH(W_)                                                                                            \
    if (mc && unroll > 1)                                                                                        \
        B200_CHECK_CUDA(b200_launch(dp_reduce_adamw_kernel<W_, true, 4>, dim3(ctas), dim3(threads), 0, st, P, rank, bucket, num_buckets, epoch_dev, start, n, m, v, hyper, S, \
                                    (unsigned int*)cta_counter, err));                               \
    else if (mc)                                                                                                 \
        B200_CHECK_CUDA(b200_launch(dp_reduce_adamw_kernel<W_, true, 1>, dim3(ctas), dim3(threads), 0, st, P, rank, bucket, num_buckets, epoch_dev, start, n, m, v, hyper, S, \
                                    (unsigned int*)cta_counter, err));                               \
    else                                                                                                         \
        B200_CHECK_CUDA(b200_launch(dp_reduce_adamw_kernel<W_, false, 1>, dim3(ctas), dim3(threads), 0, st, P, rank, bucket, num_buckets, epoch_dev, start, n, m, v, hyper, S, \
                                    (unsigned int*)cta_counter, err))
    switch (world) {
        case 1: DP_LAUNCH(1); break;
        case 2: DP_LAUNCH(2); break;
        case 4: DP_LAUNCH(4); break;
        case 8: DP_LAUNCH(8); break;
        default: b200_set_error("dp: world size must be 1, 2, 4 or 8"); return -1;
    }
#undef DP_LAUNCH
    B200_RETURN_LAUNCH_STATUS("dp_reduce_adamw");
}

extern "C" int b200_dp_wait_done(const int* sig, int world, int num_buckets, const int* epoch_dev, int* err,
                                 cudaStream_t st) {
    B200_REQUIRE(world * num_buckets <= 1024, "dp_wait_done: too many (rank, bucket) pairs");
    B200_CHECK_CUDA(b200_launch(dp_wait_done_kernel, dim3(1), dim3(1024), 0, st, sig, world, num_buckets, epoch_dev, err));
    B200_RETURN_LAUNCH_STATUS("dp_wait_done");
}
