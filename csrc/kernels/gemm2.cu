// 2-CTA (cta_group::2) variant of the tcgen05 GEMM: a CTA PAIR (two SMs of one TPC, a cluster of 2)
// computes a 256 x BN tile.  Each CTA stages only ITS half of the operands - its 128 rows of A and
// its BN/2 rows of B - and the leader CTA issues tcgen05.mma.cta_group::2 (M = 256), which reads both
// halves out of both CTAs' shared memory.  Per MMA that is 1.5x fewer shared-memory bytes per CTA
// than the 1-CTA kernel (32 KiB instead of 48 KiB per 64-deep k-block at BN = 256), so the same
// 192 KiB of smem holds 6 pipeline stages instead of 4: the k-loop survives the L2 -> smem latency
// that the ncu profile of the 1-CTA kernel showed it was bound by (tensor pipe 52-70 % active).
//
//   TMA loads of both CTAs are credited to the LEADER's full barrier; the leader's MMA thread commits
//   with a multicast arrive to the empty / accumulator-full barriers of BOTH CTAs; each CTA's epilogue
//   warps drain their own 128 TMEM lanes and arrive (remotely for the follower) on the leader's
//   accumulator-empty barrier.  Epilogues, operand majors and split-K are those of gemm.cu.
#include <cuda.h>

#include "b200llm.h"
#include "common.cuh"
#include "ptx.cuh"
#include "tma.cuh"
#include "gemm_epilogue.cuh"

namespace {

using gemm::Params;
using gemm::AgParams;
using gemm::AgJob;
using gemm::EPI_BUF_BYTES;
constexpr int BK = 64;
constexpr int A_BYTES = 128 * BK * 2;   // this CTA's 128 rows of A
constexpr int NUM_THREADS = 192;       // TMA warp, MMA warp, 4 epilogue warps; + 4 more epilogue warps (or the 4 gather warps of the AG variant)

template <int BN> struct Cfg2 {
    static constexpr int B_BYTES = (BN / 2) * BK * 2;          // this CTA's half of the B tile
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = BN == 256 ? 6 : 8;
    static constexpr int TMEM_COLS = 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * EPI_BUF_BYTES + 256 + 1024 + 1024;
};

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// AG = true: 4 extra warps (6..9) pull the remote chunks of the layer's weights over NVLink into the local
// buffer while the tiles whose weights are already local run; the TMA producer waits per chunk.
template <int BN, bool A_MN, bool B_MN, bool OUT_F32, bool AG, bool COLSUM = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS + 128, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
             const __grid_constant__ CUtensorMap tmC, const Params p, const AgParams ag) {
    using C = Cfg2<BN>;
    constexpr int NG = AG ? 1 : 2;          // epilogue warp groups (the AG variant's warps 6..9 gather weights instead)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = smem_base + C::STAGES * C::STAGE_BYTES;
    const uint32_t bar_base = epi_base + 2 * EPI_BUF_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - ptx::smem_u32(smem_raw));
    float* s_bias = reinterpret_cast<float*>(smem_gen + (bar_base + 256 - smem_base));

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const uint32_t rank = ptx::cluster_ctarank();      // 0 = leader
    const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        if (!OUT_F32) ptx::prefetch_tmap(&tmC);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            ptx::mbar_init(full_bar(s), 1);      // leader's producer arrives (+ transaction bytes of both CTAs)
            ptx::mbar_init(empty_bar(s), 1);     // multicast commit from the leader's MMA thread
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(tfull_bar(s), 1);     // multicast commit
            ptx::mbar_init(tempty_bar(s), 8 * NG);   // epilogue warps x 2 CTAs (used in the leader only)
        }
        ptx::fence_mbar_init();
    }
    ptx::cluster_sync_all();                     // barriers of both CTAs exist before anyone signals them
    if (warp == 2) {
        ptx::tmem_alloc_2sm(tmem_slot, C::TMEM_COLS);
        ptx::tmem_relinquish_2sm();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    pdl_sync();   // prologue done: from here on we touch memory owned by earlier kernels
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base)), 0);

    const int ntiles = p.num_m * p.num_n;        // num_m counts 256-row tiles here
    const int nunits = ntiles * p.splits;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        // whole warp converged; one elected lane executes each wrapper (uniform operands, see ptx.cuh)
        {
            int stage = 0;
            uint32_t phase = 0;
            unsigned ready_mask = 0;                    // chunks already seen valid (AG)
            const int ag_gen = (AG && ag.wait_flags != nullptr) ? __shfl_sync(0xffffffffu, *ag.gen, 0) : 0;
            for (int u = pair; u < nunits; u += npairs) {
                const int tile = u % ntiles, split = u / ntiles;
                const int m0 = (tile % p.num_m) * 256 + (int)rank * 128;
                const int n0 = (((tile / p.num_m) + p.n_shift) % p.num_n) * BN + (int)rank * (BN / 2);
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
                if (AG && ag.wait_flags != nullptr && !B_MN) {
                    // rows [n0, n0 + BN/2) of the [N, K] weight -> element range of the window -> owning chunks
                    const long long e0 = ag.w_offset + (long long)min(n0, p.N) * p.K;
                    const long long e1 = ag.w_offset + (long long)min(n0 + BN / 2, p.N) * p.K;
                    if (e1 > e0) {
                        const int c0 = (int)(e0 / ag.chunk), c1 = (int)((e1 - 1) / ag.chunk);
                        for (int c = c0; c <= c1; ++c) {
                            if ((ready_mask >> c) & 1u) continue;
                            const long long t0 = clock64();
                            while (__shfl_sync(0xffffffffu, ld_acquire_gpu(ag.wait_flags + c), 0) < ag_gen) {
                                __nanosleep(32);
                                if (clock64() - t0 > 20000000000ll) {      // ~10 s: the gather warps are stuck (dead peer)
                                    if (lane == 0 && ag.err != nullptr) *ag.err = 1;
                                    break;
                                }
                            }
                            ready_mask |= 1u << c;
                        }
                        asm volatile("fence.proxy.async;" ::: "memory");   // gathered by generic stores, read by TMA
                    }
                }
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(empty_bar(stage), phase ^ 1);
                    if (rank == 0) ptx::mbar_arrive_expect_tx_w(full_bar(stage), 2 * C::STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sb = sa + A_BYTES;
                    if (!A_MN) {
                        ptx::tma_load_2d_2sm_w(sa, &tmA, full_bar(stage), kb * BK, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < 2; ++j)
                            ptx::tma_load_2d_2sm_w(sa + j * 8192, &tmA, full_bar(stage), m0 + j * 64, kb * BK);
                    }
                    if (!B_MN) {
                        ptx::tma_load_2d_2sm_w(sb, &tmB, full_bar(stage), kb * BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 128; ++j)
                            ptx::tma_load_2d_2sm_w(sb + j * 8192, &tmB, full_bar(stage), n0 + j * 64, kb * BK);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA only)
        // whole warp converged, one elected lane issues inside the wrappers (uniform operands, see ptx.cuh)
        if (rank == 0) {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(256, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            constexpr uint32_t A_LBO = A_MN ? 8192 : 16, B_LBO = B_MN ? 8192 : 16;
            constexpr uint32_t A_KSTEP = A_MN ? 2048 : 32, B_KSTEP = B_MN ? 2048 : 32;
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int u = pair; u < nunits; u += npairs) {
                const int split = u / ntiles;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
                ptx::mbar_wait(tempty_bar(acc), acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(full_bar(stage), phase);
                    ptx::tc_fence_after();
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sb = sa + A_BYTES;
                    const uint64_t adesc = ptx::umma_desc(sa, A_LBO, 1024);
                    const uint64_t bdesc = ptx::umma_desc(sb, B_LBO, 1024);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        ptx::mma_bf16_ss_2sm_w(tmem_d, adesc + (uint64_t)((k * A_KSTEP) >> 4),
                                             bdesc + (uint64_t)((k * B_KSTEP) >> 4), idesc,
                                             (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    ptx::mma_commit_2sm_w(empty_bar(stage), 3);    // slot free in both CTAs
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit_2sm_w(tfull_bar(acc), 3);          // accumulators complete in both CTAs
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp < 2 + 4 * NG) {
        // ------------------------------------------------------------------ epilogue (warps 2..5 [+ 6..9], both CTAs)
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int grp = (warp - 2) >> 2;        // epilogue group: takes the 64-column chunks grp, grp + NG, ...
        const bool store_thread = (threadIdx.x == 64 + 128 * grp);
        int acc = 0;
        uint32_t acc_phase = 0;
        int buf = NG == 2 ? grp : 0;
        const int et = (int)threadIdx.x - 64;
        auto n0_of = [&](int u) { return ((((u % ntiles) / p.num_m) + p.n_shift) % p.num_n) * BN; };
        float bias_pf[(BN + 127) / 128];
#pragma unroll
        for (int i = 0; i < (BN + 127) / 128; ++i) {
            const int j = et + 128 * NG * i;
            bias_pf[i] = (!OUT_F32 && p.bias != nullptr && pair < nunits && j < BN && n0_of(pair) + j < p.N)
                             ? __ldg(p.bias + n0_of(pair) + j) : 0.f;
        }
        for (int u = pair; u < nunits; u += npairs) {
            const int tile = u % ntiles;
            const int n_blk = ((tile / p.num_m) + p.n_shift) % p.num_n;
            const int m0 = (tile % p.num_m) * 256 + (int)rank * 128, n0 = n_blk * BN;
            const int n0_next = (u + npairs < nunits) ? n0_of(u + npairs) : -1;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
            const uint32_t te = ptx::mapa(tempty_bar(acc), 0);   // the leader's accumulator-empty barrier
            gemm::epilogue_tile<BN, OUT_F32, NG, COLSUM>(p, &tmC, tfull_bar(acc), acc_phase, taddr, m0, n0, n_blk, row, lane, epi_base,
                                                 buf, store_thread, s_bias, et, bias_pf, n0_next, grp,
                                                 [te]() { ptx::mbar_arrive_cluster(te); });
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (!OUT_F32 && store_thread) ptx::tma_store_wait_all<0>();
    } else if (AG) {
        // ------------------------------------------------------------------ gather warps (6..9, every CTA)
        const int gt = (int)threadIdx.x - NUM_THREADS;              // 0..127
        const int gen = ag.njobs > 0 ? *ag.gen : 0;
        for (int j = 0; j < ag.njobs; ++j) {
            const AgJob& job = ag.job[j];
            for (int s = 0; s < ag.world; ++s) {
                const int src_rank = (ag.rank + s) % ag.world;          // own chunk first (local copy), then the peers
                const long long c_lo = (long long)src_rank * job.chunk;
                const long long c_hi = c_lo + job.chunk < job.total ? c_lo + job.chunk : job.total;
                // 16-byte vectors, 4 loads in flight per thread, grid-strided over all CTAs' gather threads
                const long long nvec = c_hi > c_lo ? (c_hi - c_lo) / 8 : 0;
                const uint4* sv = reinterpret_cast<const uint4*>(job.src[src_rank]);
                uint4* dv = reinterpret_cast<uint4*>(job.dst + c_lo);
                const long long stride = (long long)gridDim.x * 128;
                long long i = (long long)blockIdx.x * 128 + gt;
                for (; i + 3 * stride < nvec; i += 4 * stride) {
                    const uint4 a = ldg_stream_u4(sv + i), b = ldg_stream_u4(sv + i + stride),
                                c = ldg_stream_u4(sv + i + 2 * stride), d = ldg_stream_u4(sv + i + 3 * stride);
                    dv[i] = a; dv[i + stride] = b; dv[i + 2 * stride] = c; dv[i + 3 * stride] = d;
                }
                for (; i < nvec; i += stride) dv[i] = ldg_stream_u4(sv + i);
                __threadfence();
                // named barriers of this kernel: 1 (+2 when there are two epilogue groups) = epilogue groups, 3 = bias staging,
                // 2 = these gather warps (the AG variant has ONE epilogue group, so id 2 is free here)
                asm volatile("bar.sync 2, 128;" ::: "memory");
                if (gt == 0) {
                    const unsigned prev = atomicAdd(job.counters + src_rank, 1u);
                    if (prev == gridDim.x - 1) {
                        job.counters[src_rank] = 0;
                        __threadfence();
                        st_release_gpu(job.flags + src_rank, gen);
                    }
                }
            }
        }
    }

    ptx::tc_fence_before();
    ptx::cluster_sync_all();                     // nobody frees TMEM / exits while the peer may still touch it
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_2sm(tmem_base, C::TMEM_COLS);
    }
}

template <int BN, bool A_MN, bool B_MN, bool OUT_F32, bool AG = false, bool COLSUM = false>
int launch2(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const Params& p, int grid, cudaStream_t st,
            const AgParams* ag = nullptr) {
    auto kern = gemm2_kernel<BN, A_MN, B_MN, OUT_F32, AG, COLSUM>;
    static bool configured = false;
    if (!configured) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2<BN>::SMEM_BYTES));
        configured = true;
    }
    AgParams none;
    memset(&none, 0, sizeof(none));
    B200_CHECK_CUDA(b200_launch<true>(kern, dim3(grid), dim3(NUM_THREADS + 128), Cfg2<BN>::SMEM_BYTES, st, tA, tB, tC, p, ag ? *ag : none));
    return 0;
}

template <int BN, bool OUT_F32>
int launch2_major(int a_mn, int b_mn, const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const Params& p,
                  int grid, cudaStream_t st) {
    if (!a_mn && !b_mn) return launch2<BN, false, false, OUT_F32>(tA, tB, tC, p, grid, st);
    if (!a_mn && b_mn) return launch2<BN, false, true, OUT_F32>(tA, tB, tC, p, grid, st);
    if (a_mn && b_mn) return launch2<BN, true, true, OUT_F32>(tA, tB, tC, p, grid, st);
    return launch2<BN, true, false, OUT_F32>(tA, tB, tC, p, grid, st);
}

}  // namespace

static int gemm_2cta_impl(const void* A, const void* B, void* Cout, int M, int N, int K, long long lda, long long ldb,
                          long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha, const float* bias,
                          int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out,
                          const void* relu_bits, int bn, int split_k, float* colsum, cudaStream_t st);

// Same contract as b200_gemm, tiles of 256 x bn (bn = 256 or 128) on CTA pairs.
extern "C" int b200_gemm_2cta(const void* A, const void* B, void* Cout, int M, int N, int K, long long lda, long long ldb,
                              long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha, const float* bias,
                              int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out,
                         const void* relu_bits, int bn, int split_k, cudaStream_t st) {
    return gemm_2cta_impl(A, B, Cout, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_f32, f32_mode, alpha, bias, relu, mask, ld_mask,
                          stats, relu_bits_out, relu_bits, bn, split_k, nullptr, st);
}

// Same, and colsum[N] (fp32, must hold the running sum - zero it for a fresh one) += the column sums of the bf16 output:
// the bias gradient of the Linear whose output gradient this dgrad GEMM produces, without a second pass over it.
extern "C" int b200_gemm_2cta_colsum(const void* A, const void* B, void* Cout, int M, int N, int K, long long lda,
                                     long long ldb, long long ldc, int a_mn, int b_mn, float alpha, const float* bias, int relu,
                                     void* relu_bits_out, const void* relu_bits, int bn, float* colsum, cudaStream_t st) {
    B200_REQUIRE(colsum != nullptr, "gemm_2cta_colsum: colsum is null");
    return gemm_2cta_impl(A, B, Cout, M, N, K, lda, ldb, ldc, a_mn, b_mn, 0, 0, alpha, bias, relu, nullptr, 0, nullptr,
                          relu_bits_out, relu_bits, bn, 1, colsum, st);
}

static int gemm_2cta_impl(const void* A, const void* B, void* Cout, int M, int N, int K, long long lda, long long ldb,
                          long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha, const float* bias,
                          int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out,
                          const void* relu_bits, int bn, int split_k, float* colsum, cudaStream_t st) {
    B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_2cta: empty problem");
    B200_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_2cta: lda/ldb must be multiples of 8 elements");
    B200_REQUIRE(out_f32 ? (ldc % 4 == 0 && N % 4 == 0) : (ldc % 8 == 0), "gemm_2cta: bad ldc / N alignment");
    B200_REQUIRE(mask == nullptr || (N % 8 == 0 && ld_mask % 8 == 0), "gemm_2cta: mask needs N, ld_mask multiples of 8");
    const int BN = bn > 0 ? bn : 256;
    B200_REQUIRE(BN == 256 || BN == 128, "gemm_2cta: bn must be 128 or 256");
    Params p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N; p.K = K;
    p.num_m = (M + 255) / 256;
    p.num_n = (N + BN - 1) / BN;
    p.num_kb = (K + BK - 1) / BK;
    const int sms = b200_num_sms();
    int splits = split_k > 0 ? split_k : 1;
    if (split_k == 0 && out_f32 && f32_mode == 2) {
        const int tiles = p.num_m * p.num_n;
        splits = max(1, min(p.num_kb / 4, (3 * (sms / 2)) / max(tiles, 1)));
    }
    splits = max(1, min(splits, p.num_kb));
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;
    p.splits = splits;
    B200_REQUIRE(splits == 1 || (out_f32 && f32_mode == 2), "gemm_2cta: split-K needs the atomic fp32 epilogue");
    p.alpha = alpha;
    p.bias = bias; p.relu = relu; p.mask = (const bf16*)mask; p.ld_mask = ld_mask; p.stats = (float2*)stats;
    p.relu_bits_out = (unsigned long long*)relu_bits_out; p.relu_bits = (const unsigned long long*)relu_bits;
    p.out_f32 = (float*)Cout; p.ldc = ldc; p.f32_mode = f32_mode;
    p.colsum = colsum;

    CUtensorMap tA, tB, tC;
    memset(&tC, 0, sizeof(tC));
    if (!a_mn) { if (tma::make_2d(&tA, A, M, K, lda * 2, 64, 128, 2)) return -1; }
    else       { if (tma::make_2d(&tA, A, K, M, lda * 2, 64, 64, 2)) return -1; }
    if (!b_mn) { if (tma::make_2d(&tB, B, N, K, ldb * 2, 64, BN / 2, 2)) return -1; }
    else       { if (tma::make_2d(&tB, B, K, N, ldb * 2, 64, 64, 2)) return -1; }
    if (!out_f32) { if (tma::make_2d(&tC, Cout, M, N, ldc * 2, 64, 128, 2)) return -1; }

    const int nunits = p.num_m * p.num_n * p.splits;
    const int grid = 2 * min(nunits, sms / 2);
    int rc;
    if (colsum != nullptr) {
        // the fused column sums exist for the one shape that needs them: a bf16 dgrad (A K-major, B MN-major), 256-wide tiles
        B200_REQUIRE(!out_f32 && !a_mn && b_mn && BN == 256, "gemm_2cta_colsum: only the bf16 dgrad form (A K-major, B [K,N], bn 256)");
        rc = launch2<256, false, true, false, false, true>(tA, tB, tC, p, grid, st);
        if (rc) return rc;
        B200_RETURN_LAUNCH_STATUS("gemm_2cta_colsum");
    }
    if (out_f32) rc = BN == 256 ? launch2_major<256, true>(a_mn, b_mn, tA, tB, tC, p, grid, st)
                                : launch2_major<128, true>(a_mn, b_mn, tA, tB, tC, p, grid, st);
    else         rc = BN == 256 ? launch2_major<256, false>(a_mn, b_mn, tA, tB, tC, p, grid, st)
                                : launch2_major<128, false>(a_mn, b_mn, tA, tB, tC, p, grid, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("gemm_2cta");
}

// All-gather fused into a GEMM (sharded data parallelism).  C[M,N] = A[M,K] * opB^T (+bias, ReLU / ReLU sign bits),
// bf16 out, where B lives in a LOCAL window of gathered layer weights (b_mn = 0: [N,K] forward; 1: [K,N] dgrad).
// The launch carries up to two gather jobs: four extra warps per CTA copy a layer's chunks from their owners'
// shards (own chunk first, then the peers over NVLink) into a local window and raise one flag per chunk.
//   * prefetch: the job belongs to the NEXT layer - its transfer hides under this GEMM's math, and it is complete
//     when the kernel is, so the consumers need no waiting at all;
//   * self (wait_flags != null, forward only): the job is this GEMM's own B operand - the tiles whose weight rows
//     are already in the window run first (n-tile order rotated to the own chunk), the others wait per chunk.
extern "C" int b200_gemm_ag(const void* A, const void* Bw, void* Cout, int M, int N, int K, long long lda, long long ldb,
                            long long ldc, int b_mn, const float* bias, int relu, void* relu_bits_out, const void* relu_bits,
                            const int* wait_flags, long long wait_chunk, long long w_offset, int world, int rank,
                            const int* gen_dev, int* err, int njobs,
                            void* const* h_src0, void* dst0, long long chunk0, long long total0, int* flags0, void* counters0,
                            void* const* h_src1, void* dst1, long long chunk1, long long total1, int* flags1, void* counters1,
                            cudaStream_t st) {
    B200_REQUIRE(world >= 1 && world <= 8, "gemm_ag: world size %d not supported", world);
    B200_REQUIRE(M > 0 && N > 0 && K > 0 && lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0, "gemm_ag: bad shape / alignment");
    B200_REQUIRE(njobs >= 0 && njobs <= 2, "gemm_ag: at most two gather jobs per launch");
    B200_REQUIRE(wait_flags == nullptr || (!b_mn && wait_chunk > 0 && wait_chunk % 8 == 0 && w_offset % 8 == 0),
                 "gemm_ag: in-kernel waiting needs a K-major B operand and 16-byte aligned chunks");
    constexpr int BN = 256;
    Params p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N; p.K = K;
    p.num_m = (M + 255) / 256;
    p.num_n = (N + BN - 1) / BN;
    p.num_kb = (K + BK - 1) / BK;
    p.splits = 1; p.kb_per_split = p.num_kb;
    p.alpha = 1.f; p.bias = bias; p.relu = relu; p.relu_bits_out = (unsigned long long*)relu_bits_out;
    p.relu_bits = (const unsigned long long*)relu_bits;
    AgParams ag;
    memset(&ag, 0, sizeof(ag));
    ag.wait_flags = wait_flags; ag.gen = gen_dev; ag.chunk = wait_chunk; ag.w_offset = w_offset;
    ag.world = world; ag.rank = rank; ag.njobs = njobs; ag.err = err;
    if (wait_flags != nullptr) {
        // start with the n-tiles whose weight rows this rank owns (they are copied first, from local memory)
        const long long own = (long long)rank * wait_chunk - w_offset;
        if (own > 0 && own < (long long)N * K) p.n_shift = (int)(((own + K - 1) / K + BN - 1) / BN) % p.num_n;
    }
    void* const* srcs[2] = {h_src0, h_src1};
    void* dsts[2] = {dst0, dst1};
    const long long chunks[2] = {chunk0, chunk1}, totals[2] = {total0, total1};
    int* flags[2] = {flags0, flags1};
    void* counters[2] = {counters0, counters1};
    for (int j = 0; j < njobs; ++j) {
        B200_REQUIRE(srcs[j] && dsts[j] && flags[j] && counters[j] && chunks[j] % 8 == 0 && totals[j] % 8 == 0,
                     "gemm_ag: incomplete or misaligned gather job %d", j);
        for (int r = 0; r < world; ++r) ag.job[j].src[r] = (const bf16*)srcs[j][r];
        ag.job[j].dst = (bf16*)dsts[j]; ag.job[j].chunk = chunks[j]; ag.job[j].total = totals[j];
        ag.job[j].flags = flags[j]; ag.job[j].counters = (unsigned int*)counters[j];
    }
    CUtensorMap tA, tB, tC;
    if (tma::make_2d(&tA, A, M, K, lda * 2, 64, 128, 2)) return -1;
    if (!b_mn) { if (tma::make_2d(&tB, Bw, N, K, ldb * 2, 64, BN / 2, 2)) return -1; }
    else       { if (tma::make_2d(&tB, Bw, K, N, ldb * 2, 64, 64, 2)) return -1; }
    if (tma::make_2d(&tC, Cout, M, N, ldc * 2, 64, 128, 2)) return -1;
    const int sms = b200_num_sms();
    const int nunits = p.num_m * p.num_n;
    const int grid = 2 * min(nunits, sms / 2);
    if (!b_mn) { if (launch2<BN, false, false, false, true>(tA, tB, tC, p, grid, st, &ag)) return -1; }
    else       { if (launch2<BN, false, true, false, true>(tA, tB, tC, p, grid, st, &ag)) return -1; }
    B200_RETURN_LAUNCH_STATUS("gemm_ag");
}
