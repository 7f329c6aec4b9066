// Persistent warp-specialised tcgen05 GEMM for sm_100a:  C[M,N] = alpha * op(A) * op(B)^T  (+ epilogue)
//
//   * operands bf16, accumulation fp32 in TMEM (two accumulator stages of BN columns, so the
//     epilogue of tile i overlaps the MMAs of tile i+1);
//   * A/B tiles staged in shared memory by TMA (128-byte swizzle) through a STAGES-deep mbarrier
//     ring; the MMA warp runs converged and one elected lane issues tcgen05.mma (M=128, N=BN, K=16 per
//     instruction) with warp-uniform operands (see ptx.cuh, `*_w` wrappers);
//   * either operand may be K-major ([rows, K] row-major - the forward "x @ W^T" layout) or
//     MN-major ([K, rows] row-major), which is what dgrad (dY @ W) and wgrad (dY^T @ X) need
//     without any transposed copies of activations or weights;
//   * warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 and 6-9 two epilogue groups (one TMEM
//     lane quarter per warp, the groups split a tile's 64-column chunks; BN = 64 uses one group);
//   * epilogues: bf16 out through swizzled smem + TMA store, with optional per-column bias, ReLU,
//     ReLU sign bits out / sign-bit mask in (ReLU backward), and per-(row, n-tile, group) softmax statistics for
//     the fused LM-head/CE path;
//     or fp32 out (plain / accumulate / atomic for split-K) for weight gradients.
//
// Replaces every cuBLAS/cuBLASLt call the reference makes through nn.Linear and autograd
// (SURVEY K3, K9-K11, K13): reference src/models/attention.py:47-55, mlp.py:53,66,
// transformer.py:71.
#include <cuda.h>

#include "b200llm.h"
#include "common.cuh"
#include "ptx.cuh"
#include "tma.cuh"
#include "gemm_epilogue.cuh"

namespace {

using gemm::Params;
using gemm::EPI_BUF_BYTES;
constexpr int BM = 128;
constexpr int BK = 64;               // 64 bf16 = 128 bytes = one swizzle row
constexpr int A_BYTES = BM * BK * 2; // 16 KiB
constexpr int NUM_THREADS = 320;   // TMA warp, MMA warp, 2 groups of 4 epilogue warps (BN = 64 uses one group)

template <int BN> struct Cfg {
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
    static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * EPI_BUF_BYTES + 256 /*barriers*/ + 1024 /*bias tile*/ + 1024 /*align*/;
};


template <int BN, bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const Params p) {
    using C = Cfg<BN>;
    constexpr int NG = BN >= 128 ? 2 : 1;   // epilogue warp groups
    extern __shared__ uint8_t smem_raw[];
    // 128B-swizzled tiles need 1024-byte alignment
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t epi_base = smem_base + C::STAGES * C::STAGE_BYTES;
    const uint32_t bar_base = epi_base + 2 * EPI_BUF_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * C::STAGES + 2 + s); };
    const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
    uint8_t* smem_gen = smem_raw + (smem_base - ptx::smem_u32(smem_raw));  // generic pointer to aligned base
    float* s_bias = reinterpret_cast<float*>(smem_gen + (bar_base + 256 - smem_base));   // [256] bias of the current tile

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        if (!OUT_F32) ptx::prefetch_tmap(&tmC);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < C::STAGES; ++s) {
            ptx::mbar_init(full_bar(s), 1);
            ptx::mbar_init(empty_bar(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(tfull_bar(s), 1);
            ptx::mbar_init(tempty_bar(s), 4 * NG);  // one arrive per epilogue warp
        }
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    pdl_sync();   // prologue done: from here on we touch memory owned by earlier kernels
    const uint32_t tmem_base = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base)), 0);

    const int ntiles = p.num_m * p.num_n;
    const int nunits = ntiles * p.splits;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        // whole warp converged; one elected lane executes each wrapper (uniform operands, see ptx.cuh)
        {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const int tile = u % ntiles, split = u / ntiles;
                const int m0 = (tile % p.num_m) * BM, n0 = (tile / p.num_m) * BN;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(empty_bar(stage), phase ^ 1);
                    ptx::mbar_arrive_expect_tx_w(full_bar(stage), C::STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sb = sa + A_BYTES;
                    if (!A_MN) {
                        ptx::tma_load_2d_w(sa, &tmA, full_bar(stage), kb * BK, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            ptx::tma_load_2d_w(sa + j * 8192, &tmA, full_bar(stage), m0 + j * 64, kb * BK);
                    }
                    if (!B_MN) {
                        ptx::tma_load_2d_w(sb, &tmB, full_bar(stage), kb * BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            ptx::tma_load_2d_w(sb + j * 8192, &tmB, full_bar(stage), n0 + j * 64, kb * BK);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        // whole warp converged, one elected lane issues inside the wrappers (uniform operands, see ptx.cuh)
        {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
            // K-major, 128B swizzle: 8-row groups 1024 B apart (SBO); LBO unused.
            // MN-major, 128B swizzle: 64-element MN chunks 8192 B apart (LBO), 8-k-row groups 1024 B apart (SBO).
            constexpr uint32_t A_LBO = A_MN ? 8192 : 16, B_LBO = B_MN ? 8192 : 16;
            constexpr uint32_t A_KSTEP = A_MN ? 2048 : 32, B_KSTEP = B_MN ? 2048 : 32;  // bytes per K=16
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
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
                        ptx::mma_bf16_ss_w(tmem_d, adesc + (uint64_t)((k * A_KSTEP) >> 4),
                                         bdesc + (uint64_t)((k * B_KSTEP) >> 4), idesc,
                                         (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    ptx::mma_commit_w(empty_bar(stage));  // frees the smem slot when these MMAs retire
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit_w(tfull_bar(acc));        // accumulator complete -> epilogue
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp < 2 + 4 * NG) {
        // ------------------------------------------------------------------ epilogue (warps 2..5 and, for BN >= 128, 6..9)
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;          // row inside the 128-row tile
        const int grp = (warp - 2) >> 2;        // epilogue group: takes the 64-column chunks grp, grp + NG, ...
        const bool store_thread = (threadIdx.x == 64 + 128 * grp);
        int acc = 0;
        uint32_t acc_phase = 0;
        int buf = NG == 2 ? grp : 0;
        const int et = (int)threadIdx.x - 64;
        auto n0_of = [&](int u) { return ((u % ntiles) / p.num_m) * BN; };
        float bias_pf[(BN + 127) / 128];
#pragma unroll
        for (int i = 0; i < (BN + 127) / 128; ++i) {
            const int j = et + 128 * NG * i;
            bias_pf[i] = (!OUT_F32 && p.bias != nullptr && (int)blockIdx.x < nunits && j < BN && n0_of(blockIdx.x) + j < p.N)
                             ? __ldg(p.bias + n0_of(blockIdx.x) + j) : 0.f;
        }
        for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
            const int tile = u % ntiles;
            const int n_blk = tile / p.num_m;
            const int m0 = (tile % p.num_m) * BM, n0 = n_blk * BN;
            const int n0_next = (u + (int)gridDim.x < nunits) ? n0_of(u + gridDim.x) : -1;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
            const uint32_t te = tempty_bar(acc);
            gemm::epilogue_tile<BN, OUT_F32, NG>(p, &tmC, tfull_bar(acc), acc_phase, taddr, m0, n0, n_blk, row, lane, epi_base,
                                                 buf, store_thread, s_bias, et, bias_pf, n0_next, grp,
                                                 [te]() { ptx::mbar_arrive(te); });
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (!OUT_F32 && store_thread) ptx::tma_store_wait_all<0>();
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
    }
}

template <int BN, bool A_MN, bool B_MN, bool OUT_F32>
int launch(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC, const Params& p, int grid,
           cudaStream_t st) {
    auto kern = gemm_kernel<BN, A_MN, B_MN, OUT_F32>;
    static bool configured = false;
    if (!configured) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM_BYTES));
        configured = true;
    }
    B200_CHECK_CUDA(b200_launch(kern, dim3(grid), dim3(NUM_THREADS), Cfg<BN>::SMEM_BYTES, st, tA, tB, tC, p));
    return 0;
}

template <int BN, bool OUT_F32>
int launch_major(int a_mn, int b_mn, const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tC,
                 const Params& p, int grid, cudaStream_t st) {
    if (!a_mn && !b_mn) return launch<BN, false, false, OUT_F32>(tA, tB, tC, p, grid, st);
    if (!a_mn && b_mn) return launch<BN, false, true, OUT_F32>(tA, tB, tC, p, grid, st);
    if (a_mn && b_mn) return launch<BN, true, true, OUT_F32>(tA, tB, tC, p, grid, st);
    return launch<BN, true, false, OUT_F32>(tA, tB, tC, p, grid, st);
}

}  // namespace

// Tile-shape heuristic shared with the Python side (for sizing the CE statistics buffer).
extern "C" int b200_gemm_pick_bn(int M, int N, int K, int out_f32) {
    // score = wave efficiency x per-tile efficiency (wider tiles amortise A traffic and the epilogue)
    const int sms = b200_num_sms();
    const int num_m = (M + BM - 1) / BM;
    const int cand[3] = {256, 128, 64};
    const float tile_eff[3] = {1.0f, 0.88f, 0.70f};
    int best = 256;
    float best_score = -1.f;
    for (int i = 0; i < 3; ++i) {
        const int bn = cand[i];
        if (bn > 64 && bn / 2 >= N) continue;  // tile mostly padding
        const int tiles = num_m * ((N + bn - 1) / bn);
        const int waves = (tiles + sms - 1) / sms;
        const float used = (float)N / (float)(((N + bn - 1) / bn) * bn);
        const float score = (float)tiles / (float)(waves * sms) * tile_eff[i] * used;
        if (score > best_score) { best_score = score; best = bn; }
    }
    (void)K; (void)out_f32;
    return best;
}

extern "C" int b200_gemm(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                         long long ldc, int a_mn, int b_mn, int out_f32, int f32_mode, float alpha,
                         const float* bias, int relu, const void* mask, int ld_mask, void* stats, void* relu_bits_out,
                         const void* relu_bits, int bn_override, int split_k, cudaStream_t st) {
    B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem");
    B200_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (16 bytes)");
    B200_REQUIRE(out_f32 ? (ldc % 4 == 0 && N % 4 == 0) : (ldc % 8 == 0), "gemm: bad ldc / N alignment");
    B200_REQUIRE(mask == nullptr || (N % 8 == 0 && ld_mask % 8 == 0), "gemm: mask needs N, ld_mask multiples of 8");
    B200_REQUIRE(bias == nullptr || N % 4 == 0, "gemm: bias needs N multiple of 4");
    int BN = bn_override > 0 ? bn_override : b200_gemm_pick_bn(M, N, K, out_f32);
    B200_REQUIRE(BN == 256 || BN == 128 || BN == 64, "gemm: BN must be 64, 128 or 256");

    Params p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N; p.K = K;
    p.num_m = (M + BM - 1) / BM;
    p.num_n = (N + BN - 1) / BN;
    p.num_kb = (K + BK - 1) / BK;
    const int sms = b200_num_sms();
    int splits = split_k > 0 ? split_k : 1;
    if (split_k == 0 && out_f32 && f32_mode == 2) {
        // auto split-K for atomic fp32 outputs (weight gradients): aim at ~3 waves of work units
        const int tiles = p.num_m * p.num_n;
        splits = max(1, min(p.num_kb / 4, (3 * sms) / max(tiles, 1)));
    }
    splits = max(1, min(splits, p.num_kb));
    p.kb_per_split = (p.num_kb + splits - 1) / splits;
    splits = (p.num_kb + p.kb_per_split - 1) / p.kb_per_split;  // no empty split
    p.splits = splits;
    B200_REQUIRE(splits == 1 || (out_f32 && f32_mode == 2), "gemm: split-K needs the atomic fp32 epilogue");
    p.alpha = alpha;
    p.bias = bias; p.relu = relu; p.mask = (const bf16*)mask; p.ld_mask = ld_mask; p.stats = (float2*)stats;
    p.relu_bits_out = (unsigned long long*)relu_bits_out; p.relu_bits = (const unsigned long long*)relu_bits;
    p.out_f32 = (float*)C; p.ldc = ldc; p.f32_mode = f32_mode;

    CUtensorMap tA, tB, tC;
    memset(&tC, 0, sizeof(tC));
    // A: K-major [M, K] box {64, 128};  MN-major [K, M] box {64, 64}
    if (!a_mn) { if (tma::make_2d(&tA, A, M, K, lda * 2, 64, BM, 2)) return -1; }
    else       { if (tma::make_2d(&tA, A, K, M, lda * 2, 64, 64, 2)) return -1; }
    if (!b_mn) { if (tma::make_2d(&tB, B, N, K, ldb * 2, 64, BN, 2)) return -1; }
    else       { if (tma::make_2d(&tB, B, K, N, ldb * 2, 64, 64, 2)) return -1; }
    if (!out_f32) { if (tma::make_2d(&tC, C, M, N, ldc * 2, 64, BM, 2)) return -1; }

    const int nunits = p.num_m * p.num_n * p.splits;
    const int grid = min(nunits, sms);
    int rc;
    if (out_f32) {
        rc = BN == 256 ? launch_major<256, true>(a_mn, b_mn, tA, tB, tC, p, grid, st)
           : BN == 128 ? launch_major<128, true>(a_mn, b_mn, tA, tB, tC, p, grid, st)
                       : launch_major<64, true>(a_mn, b_mn, tA, tB, tC, p, grid, st);
    } else {
        rc = BN == 256 ? launch_major<256, false>(a_mn, b_mn, tA, tB, tC, p, grid, st)
           : BN == 128 ? launch_major<128, false>(a_mn, b_mn, tA, tB, tC, p, grid, st)
                       : launch_major<64, false>(a_mn, b_mn, tA, tB, tC, p, grid, st);
    }
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("gemm");
}
