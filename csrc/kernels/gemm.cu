// Persistent warp-specialised tcgen05 GEMM for sm_100a:  C[M,N] = alpha * op(A) * op(B)^T  (+ epilogue)
//
//   * operands bf16, accumulation fp32 in TMEM (two accumulator stages of BN columns, so the
//     epilogue of tile i overlaps the MMAs of tile i+1);
//   * A/B tiles staged in shared memory by TMA (128-byte swizzle) through a STAGES-deep mbarrier
//     ring; one elected thread issues tcgen05.mma (M=128, N=BN, K=16 per instruction);
//   * either operand may be K-major ([rows, K] row-major - the forward "x @ W^T" layout) or
//     MN-major ([K, rows] row-major), which is what dgrad (dY @ W) and wgrad (dY^T @ X) need
//     without any transposed copies of activations or weights;
//   * warp roles: warp 0 TMA producer, warp 1 MMA issuer, warps 2-5 epilogue (one TMEM lane
//     quarter each);
//   * epilogues: bf16 out through swizzled smem + TMA store, with optional per-column bias, ReLU,
//     ReLU-backward mask and per-(row, n-tile) softmax statistics for the fused LM-head/CE path;
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

namespace {

constexpr int BM = 128;
constexpr int BK = 64;               // 64 bf16 = 128 bytes = one swizzle row
constexpr int A_BYTES = BM * BK * 2; // 16 KiB
constexpr int EPI_BUF_BYTES = BM * 64 * 2;  // one 128 x 64 bf16 staging tile
constexpr int NUM_THREADS = 192;

template <int BN> struct Cfg {
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = BN == 256 ? 4 : (BN == 128 ? 6 : 8);
    static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 2 * EPI_BUF_BYTES + 256 /*barriers*/ + 1024 /*bias tile*/ + 1024 /*align*/;
};

struct Params {
    int M, N, K;
    int num_m, num_n, num_kb, splits, kb_per_split;
    float alpha;
    // bf16 epilogue
    const float* bias;      // [N] or null
    int relu;               // apply max(0, .)
    const bf16* mask;       // [M, ld_mask] post-ReLU activations: out = mask > 0 ? v : 0
    int ld_mask;
    float2* stats;          // [num_n][M] (max, sum exp) of the bf16-rounded outputs, or null
    // fp32 epilogue
    float* out_f32;
    long long ldc;
    int f32_mode;           // 0 store, 1 accumulate (+=), 2 atomic add
};

template <int BN, bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const Params p) {
    using C = Cfg<BN>;
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

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

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
            ptx::mbar_init(tempty_bar(s), 4);  // one arrive per epilogue warp
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
    const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

    const int ntiles = p.num_m * p.num_n;
    const int nunits = ntiles * p.splits;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
                const int tile = u % ntiles, split = u / ntiles;
                const int m0 = (tile % p.num_m) * BM, n0 = (tile / p.num_m) * BN;
                const int kb0 = split * p.kb_per_split;
                const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);
                for (int kb = kb0; kb < kb1; ++kb) {
                    ptx::mbar_wait(empty_bar(stage), phase ^ 1);
                    ptx::mbar_arrive_expect_tx(full_bar(stage), C::STAGE_BYTES);
                    const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                    const uint32_t sb = sa + A_BYTES;
                    if (!A_MN) {
                        ptx::tma_load_2d(sa, &tmA, full_bar(stage), kb * BK, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)
                            ptx::tma_load_2d(sa + j * 8192, &tmA, full_bar(stage), m0 + j * 64, kb * BK);
                    }
                    if (!B_MN) {
                        ptx::tma_load_2d(sb, &tmB, full_bar(stage), kb * BK, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            ptx::tma_load_2d(sb + j * 8192, &tmB, full_bar(stage), n0 + j * 64, kb * BK);
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
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
                        ptx::mma_bf16_ss(tmem_d, adesc + (uint64_t)((k * A_KSTEP) >> 4),
                                         bdesc + (uint64_t)((k * B_KSTEP) >> 4), idesc,
                                         (kb > kb0 || k > 0) ? 1u : 0u);
                    }
                    ptx::mma_commit(empty_bar(stage));  // frees the smem slot when these MMAs retire
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit(tfull_bar(acc));        // accumulator complete -> epilogue
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------------------------ epilogue (warps 2..5)
        const int q = warp & 3;                 // TMEM lane quarter this warp may access
        const int row = q * 32 + lane;          // row inside the 128-row tile
        const bool store_thread = (threadIdx.x == 64);
        int acc = 0;
        uint32_t acc_phase = 0;
        int buf = 0;
        for (int u = blockIdx.x; u < nunits; u += gridDim.x) {
            const int tile = u % ntiles;
            const int n_blk = tile / p.num_m;
            const int m0 = (tile % p.num_m) * BM, n0 = n_blk * BN;
            const int m = m0 + row;
            if (!OUT_F32 && p.bias != nullptr) {
                // stage this tile's bias slice in smem while the MMAs are still running
                asm volatile("bar.sync 1, 128;" ::: "memory");          // previous tile's readers are done
                const int et = threadIdx.x - 64;
                for (int j = et; j < BN; j += 128) s_bias[j] = (n0 + j < p.N) ? __ldg(p.bias + n0 + j) : 0.f;
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            ptx::mbar_wait(tfull_bar(acc), acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
            float run_max = -INFINITY, run_sum = 0.f;
#pragma unroll 1
            for (int c = 0; c < BN / 64; ++c) {
                const int nc0 = n0 + c * 64;
                uint4 mk4[8];
                if (!OUT_F32 && p.mask != nullptr && m < p.M) {
                    const bf16* mk = p.mask + (size_t)m * p.ld_mask + nc0;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        mk4[j] = (nc0 + j * 8 < p.N) ? ldg_stream_u4(reinterpret_cast<const uint4*>(mk + j * 8))
                                                     : make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u);
                }
                uint32_t r[64];
                {
                    uint32_t(&r0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[0]);
                    uint32_t(&r1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&r[32]);
                    ptx::tmem_ld_32x32(taddr + c * 64, r0);
                    ptx::tmem_ld_32x32(taddr + c * 64 + 32, r1);
                    ptx::tmem_ld_wait();
                }
                if (c == BN / 64 - 1) {
                    // all TMEM reads of this accumulator stage are done: hand it back to the MMA warp
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(tempty_bar(acc));
                }
                if (OUT_F32) {
                    if (m < p.M) {
                        float* o = p.out_f32 + (size_t)m * p.ldc + nc0;
#pragma unroll
                        for (int j = 0; j < 64; j += 4) {
                            if (nc0 + j < p.N) {  // N % 4 == 0 is required by the host wrapper
                                float4 v = make_float4(__uint_as_float(r[j]) * p.alpha, __uint_as_float(r[j + 1]) * p.alpha,
                                                       __uint_as_float(r[j + 2]) * p.alpha,
                                                       __uint_as_float(r[j + 3]) * p.alpha);
                                if (p.f32_mode == 2) {
                                    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(o + j), "f"(v.x),
                                                 "f"(v.y), "f"(v.z), "f"(v.w)
                                                 : "memory");
                                } else {
                                    if (p.f32_mode == 1) {
                                        float4 old = *reinterpret_cast<float4*>(o + j);
                                        v.x += old.x; v.y += old.y; v.z += old.z; v.w += old.w;
                                    }
                                    *reinterpret_cast<float4*>(o + j) = v;
                                }
                            }
                        }
                    }
                } else {
                    // ---- elementwise epilogue in fp32 registers
                    if (p.alpha != 1.f) {
#pragma unroll
                        for (int j = 0; j < 64; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) * p.alpha);
                    }
                    if (p.bias != nullptr) {
#pragma unroll
                        for (int j = 0; j < 64; j += 4) {
                            const float4 b = *reinterpret_cast<const float4*>(s_bias + c * 64 + j);   // smem broadcast
                            r[j] = __float_as_uint(__uint_as_float(r[j]) + b.x);
                            r[j + 1] = __float_as_uint(__uint_as_float(r[j + 1]) + b.y);
                            r[j + 2] = __float_as_uint(__uint_as_float(r[j + 2]) + b.z);
                            r[j + 3] = __float_as_uint(__uint_as_float(r[j + 3]) + b.w);
                        }
                    }
                    if (p.relu) {
#pragma unroll
                        for (int j = 0; j < 64; ++j) r[j] = __float_as_uint(fmaxf(__uint_as_float(r[j]), 0.f));
                    }
                    if (p.mask != nullptr && m < p.M) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint32_t w[4] = {mk4[j].x, mk4[j].y, mk4[j].z, mk4[j].w};
#pragma unroll
                            for (int t = 0; t < 4; ++t) {
                                float2 f = unpack_bf16x2(w[t]);
                                if (!(f.x > 0.f)) r[j * 8 + 2 * t] = 0u;
                                if (!(f.y > 0.f)) r[j * 8 + 2 * t + 1] = 0u;
                            }
                        }
                    }
                    // ---- pack to bf16
                    uint32_t pk[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        pk[j] = pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
                    if (p.stats != nullptr) {
                        // online softmax statistics over the bf16-rounded values (what CE will see)
                        float cmax = -INFINITY;
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            float2 f = unpack_bf16x2(pk[j]);
                            if (nc0 + 2 * j < p.N) cmax = fmaxf(cmax, f.x);
                            if (nc0 + 2 * j + 1 < p.N) cmax = fmaxf(cmax, f.y);
                        }
                        const float new_max = fmaxf(run_max, cmax);
                        if (new_max > -INFINITY) {
                            float s = 0.f;
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                float2 f = unpack_bf16x2(pk[j]);
                                if (nc0 + 2 * j < p.N) s += __expf(f.x - new_max);
                                if (nc0 + 2 * j + 1 < p.N) s += __expf(f.y - new_max);
                            }
                            run_sum = run_sum * __expf(run_max - new_max) + s;
                            run_max = new_max;
                        }
                    }
                    // ---- stage in swizzled smem, then one thread issues the TMA store
                    if (store_thread) ptx::tma_store_wait_read<1>();  // buffer `buf` (2 stores ago) is drained
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    const uint32_t sdst = epi_base + buf * EPI_BUF_BYTES + row * 128;
#pragma unroll
                    for (int ch = 0; ch < 8; ++ch) {
                        const uint32_t addr = sdst + (((uint32_t)ch ^ ((uint32_t)row & 7u)) << 4);
                        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(pk[4 * ch]),
                                     "r"(pk[4 * ch + 1]), "r"(pk[4 * ch + 2]), "r"(pk[4 * ch + 3])
                                     : "memory");
                    }
                    ptx::fence_proxy_async_smem();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                    if (store_thread) {
                        if (nc0 < p.N) ptx::tma_store_2d(&tmC, epi_base + buf * EPI_BUF_BYTES, nc0, m0);
                        ptx::tma_store_commit();
                    }
                    buf ^= 1;
                }
            }
            if (!OUT_F32 && p.stats != nullptr && m < p.M)
                p.stats[(size_t)n_blk * p.M + m] = make_float2(run_max, run_sum);
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
    kern<<<grid, NUM_THREADS, Cfg<BN>::SMEM_BYTES, st>>>(tA, tB, tC, p);
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
                         const float* bias, int relu, const void* mask, int ld_mask, void* stats, int bn_override,
                         int split_k, cudaStream_t st) {
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
