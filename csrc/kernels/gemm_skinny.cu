// Weight-streaming GEMM for decode-sized batches:  out[M, N] = act(x[M, K] @ W[N, K]^T + bias),  M <= 128.
//
// At M = 32 the problem is a stream of the weight matrix through the SM (2 bytes per 64 MACs): the 128 x 256 tile
// kernel wastes 3/4 of every MMA on padding and, worse, has too few CTAs and too few bytes in flight to pull
// HBM bandwidth.  Here the operands are SWAPPED: the weight tile is the M = 128 operand of tcgen05.mma and the
// batch is the N operand (N = 32/64/128), so a CTA's smem ring holds 10 x 16 KiB of weights in flight, the
// accumulator is a 128 x M tile in TMEM (<= 128 columns), and K is split across CTAs so every GEMM of a decode
// step has ~one CTA per SM.  The CTAs that share a weight-row tile form a THREAD-BLOCK CLUSTER (2, 4 or 8 CTAs
// along K): each leaves its fp32 partial tile in its own shared memory, and after a cluster barrier every CTA
// reduces a slice of the batch rows by reading its peers' tiles through distributed shared memory - no global
// scratch, no atomics, no second kernel.
//
// Programmatic dependent launch: the kernel is launched with programmatic stream serialisation, lets ITS dependents
// go at once, and its TMA producer fills the whole smem ring with WEIGHT tiles before griddepcontrol.wait - weights do
// not depend on the previous kernel, so up to STAGES x 16 KiB x 148 CTAs (24 MB, most of a layer matrix) are already in
// flight while the LayerNorm / attention kernel that produces the activations is still running.  Only the activation
// loads (and everything downstream of them through the mbarrier chain) come after the wait.
// Replaces the per-token full-context fp32 forward of the reference's generate (src/models/transformer.py:107-113).
#include <cuda.h>

#include "b200llm.h"
#include "common.cuh"
#include "ptx.cuh"
#include "tma.cuh"

namespace {

constexpr int BK = 64;
constexpr int A_BYTES = 128 * BK * 2;
constexpr int NUM_THREADS = 192;

template <int NB> struct SCfg {
    static constexpr int B_BYTES = NB * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES = NB == 32 ? 10 : (NB == 64 ? 8 : 6);
    static constexpr int TMEM_COLS = NB < 32 ? 32 : NB;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256 + 1024;
};

struct SParams {
    int M, N, K, num_kb, splits, kb_per_split;
    const float* bias;
    int relu;
    bf16* out;
    long long ldc;
};

template <int NB>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const SParams p) {
    using C = SCfg<NB>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar_base = smem_base + C::STAGES * C::STAGE_BYTES;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
    const uint32_t done_bar = bar_base + 8u * (2 * C::STAGES);
    const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 1);
    uint8_t* smem_gen = smem_raw + (smem_base - ptx::smem_u32(smem_raw));

    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int n_tile = blockIdx.x, split = blockIdx.y;
    const int n0 = n_tile * 128;
    const int kb0 = split * p.kb_per_split;
    const int kb1 = min(p.num_kb, kb0 + p.kb_per_split);

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmW);
        ptx::prefetch_tmap(&tmX);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < C::STAGES; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
        ptx::mbar_init(done_bar, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base)), 0);

    if (warp == 0) {
        // whole warp converged; one elected lane executes each wrapper (uniform operands, see ptx.cuh)
        {
            // 1. weights of the first ring pass: independent of every earlier kernel, issued before the grid dependency
            const int npre = min(kb1 - kb0, C::STAGES);
            for (int i = 0; i < npre; ++i) {
                ptx::mbar_arrive_expect_tx_w(full_bar(i), C::STAGE_BYTES);
                ptx::tma_load_2d_w(smem_base + i * C::STAGE_BYTES, &tmW, full_bar(i), (kb0 + i) * BK, n0);
            }
            // 2. the activations are the previous kernel's output: wait for it (and, transitively, all earlier kernels)
            asm volatile("griddepcontrol.wait;" ::: "memory");
            for (int i = 0; i < npre; ++i)
                ptx::tma_load_2d_w(smem_base + i * C::STAGE_BYTES + A_BYTES, &tmX, full_bar(i), (kb0 + i) * BK, 0);
            // 3. steady state
            int stage = npre == C::STAGES ? 0 : npre;
            uint32_t phase = npre == C::STAGES ? 1 : 0;
            for (int kb = kb0 + npre; kb < kb1; ++kb) {
                ptx::mbar_wait(empty_bar(stage), phase ^ 1);
                ptx::mbar_arrive_expect_tx_w(full_bar(stage), C::STAGE_BYTES);
                const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                ptx::tma_load_2d_w(sa, &tmW, full_bar(stage), kb * BK, n0);            // 128 weight rows
                ptx::tma_load_2d_w(sa + A_BYTES, &tmX, full_bar(stage), kb * BK, 0);   // the whole (padded) batch
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 1) {
        // whole warp converged, one elected lane issues inside the wrappers (uniform operands, see ptx.cuh)
        {
            constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, NB, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            for (int kb = kb0; kb < kb1; ++kb) {
                ptx::mbar_wait(full_bar(stage), phase);
                ptx::tc_fence_after();
                const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
                const uint64_t adesc = ptx::umma_desc(sa, 16, 1024);
                const uint64_t bdesc = ptx::umma_desc(sa + A_BYTES, 16, 1024);
#pragma unroll
                for (int k = 0; k < BK / 16; ++k)
                    ptx::mma_bf16_ss_w(tmem, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb > kb0 || k > 0) ? 1u : 0u);
                ptx::mma_commit_w(empty_bar(stage));
                if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
            }
            ptx::mma_commit_w(done_bar);
        }
    } else {
        // epilogue: thread = weight row n (TMEM lane), NB accumulator columns = batch rows
        const int q = warp & 3;
        const int n = n0 + q * 32 + lane;
        ptx::mbar_wait(done_bar, 0);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16);
        const float b = (p.bias != nullptr && n < p.N) ? __ldg(p.bias + n) : 0.f;
        if (p.splits == 1) {
#pragma unroll 1
            for (int c = 0; c < NB / 32; ++c) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
                if (n < p.N) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int m = c * 32 + j;
                        if (m < p.M) {
                            float v = __uint_as_float(r[j]) + b;
                            if (p.relu) v = fmaxf(v, 0.f);
                            p.out[(size_t)m * p.ldc + n] = __float2bfloat16_rn(v);
                        }
                    }
                }
            }
        } else {
            // partial tile -> own smem (stage 0 is free: every MMA has retired), layout [m][128] fp32
            float* part = reinterpret_cast<float*>(smem_gen);
            const int nl = q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < NB / 32; ++c) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(taddr + c * 32, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) part[(c * 32 + j) * 128 + nl] = (kb1 > kb0) ? __uint_as_float(r[j]) : 0.f;  // empty K slice
            }
        }
    }
    if (p.splits > 1) {
        // all CTAs of the cluster have published their partial tiles
        ptx::cluster_sync_all();
        if (warp >= 2) {
            const int q = warp & 3;
            const int nl = q * 32 + lane;
            const int n = n0 + nl;
            const float b = (p.bias != nullptr && n < p.N) ? __ldg(p.bias + n) : 0.f;
            const int rows = (p.M + p.splits - 1) / p.splits;          // batch rows reduced by this CTA
            const int m_lo = split * rows, m_hi = min(p.M, m_lo + rows);
            const uint32_t local = smem_base;                          // same offset in every CTA of the cluster
            // 4 rows x up to 8 peers = 32 independent distributed-shared-memory loads in flight before the first add
            // (a remote load is ~215 clk away: the dependent one-at-a-time loop this replaces cost 3-4 us per GEMM)
            for (int m0 = m_lo; m0 < m_hi; m0 += 4) {
                float t[4][8];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int s2 = 0; s2 < 8; ++s2) {
                        t[r][s2] = 0.f;
                        if (s2 < p.splits && m0 + r < m_hi) {
                            const uint32_t ra = ptx::mapa(local + (uint32_t)((m0 + r) * 128 + nl) * 4u, (uint32_t)s2);
                            asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(t[r][s2]) : "r"(ra) : "memory");
                        }
                    }
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    if (m0 + r >= m_hi) break;
                    float v = b;
#pragma unroll
                    for (int s2 = 0; s2 < 8; ++s2) v += t[r][s2];
                    if (p.relu) v = fmaxf(v, 0.f);
                    if (n < p.N) p.out[(size_t)(m0 + r) * p.ldc + n] = __float2bfloat16_rn(v);
                }
            }
        }
        ptx::cluster_sync_all();                                       // nobody leaves while its tile may still be read
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

template <int NB>
int launch_skinny(const CUtensorMap& tW, const CUtensorMap& tX, const SParams& p, dim3 grid, cudaStream_t st) {
    auto kern = gemm_skinny_kernel<NB>;
    static bool configured = false;
    if (!configured) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SCfg<NB>::SMEM_BYTES));
        configured = true;
    }
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = SCfg<NB>::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;      // the K splits of one weight tile share a cluster
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = grid.y;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    static int pdl = -1;
    if (pdl < 0) {
        const char* e = getenv("B200_PDL_SKINNY");                // on by default for THIS kernel (weight prefetch under
        pdl = (e != nullptr && e[0] == '0') ? 0 : 1;              // the previous kernel); B200_PDL_SKINNY=0 for A/B runs
    }
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 2 : 1;
    B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tW, tX, p));
    return 0;
}

}  // namespace

// K split (= cluster size) chosen for (M, N, K): a power of two <= 8 giving about one CTA per SM
extern "C" int b200_gemm_skinny_splits(int M, int N, int K) {
    const int sms = b200_num_sms();
    const int n_tiles = (N + 127) / 128, num_kb = (K + BK - 1) / BK;
    (void)M;
    static int any = -1;
    if (any < 0) {
        const char* e = getenv("B200_SKINNY_ANY_SPLIT");       // cluster sizes 3, 5, 6, 7 as well (one CTA per SM for more shapes)
        any = (e != nullptr && e[0] == '1') ? 1 : 0;
    }
    if (any) {
        int s = sms / n_tiles;
        s = s < 1 ? 1 : (s > 8 ? 8 : s);
        while (s > 1 && num_kb / s < 4) --s;
        return s;
    }
    // Measured on B200 in graph replay (tools/bench_skinny.py, M = 32, 1.3B shapes): one CTA streams ~70 GB/s whatever
    // the split, a cluster costs ~1 us (co-scheduling, two cluster barriers, the DSMEM reduction), so splitting K only
    // pays when there are few weight-row tiles: QKV (48 tiles) 7.5 us unsplit vs 8.6 with 2 CTAs per tile, MLP-up (64
    // tiles) 8.2 vs 9.3, MLP-down (16 tiles) 16.1 unsplit, 9.3 with 4, 13.3 with 8.  Rule: about 72 CTAs in all.
    int splits = (72 + n_tiles / 2) / n_tiles;
    splits = splits < 1 ? 1 : (splits > 8 ? 8 : splits);
    if (splits == 3) splits = 4;                       // odd cluster sizes schedule badly (12.5 us vs 9.8 for QKV)
    if (splits == 5 || splits == 7) splits -= 1;
    while (splits > 1 && num_kb / splits < 4) --splits;
    return splits;
}

extern "C" int b200_gemm_skinny(const void* x, const void* W, void* out, int M, int N, int K, long long ldx, long long ldw,
                                long long ldc, const float* bias, int relu, int split_k, cudaStream_t st) {
    B200_REQUIRE(M >= 1 && M <= 128, "gemm_skinny: M must be in [1, 128] (got %d)", M);
    B200_REQUIRE(ldx % 8 == 0 && ldw % 8 == 0 && K % 8 == 0, "gemm_skinny: K / leading dimensions must be multiples of 8");
    const int NB = M <= 32 ? 32 : (M <= 64 ? 64 : 128);
    SParams p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N; p.K = K;
    p.num_kb = (K + BK - 1) / BK;
    p.splits = split_k > 0 ? split_k : b200_gemm_skinny_splits(M, N, K);
    B200_REQUIRE(p.splits >= 1 && p.splits <= 8, "gemm_skinny: split_k must be in [1, 8] (a thread-block cluster along K)");
    while (p.splits > 1 && p.num_kb < p.splits) --p.splits;
    p.kb_per_split = (p.num_kb + p.splits - 1) / p.splits;
    p.bias = bias; p.relu = relu; p.out = (bf16*)out; p.ldc = ldc;
    CUtensorMap tW, tX;
    if (tma::make_2d(&tW, W, N, K, ldw * 2, 64, 128, 2)) return -1;
    if (tma::make_2d(&tX, x, M, K, ldx * 2, 64, NB, 2)) return -1;
    dim3 grid((N + 127) / 128, p.splits);
    int rc = NB == 32 ? launch_skinny<32>(tW, tX, p, grid, st)
           : NB == 64 ? launch_skinny<64>(tW, tX, p, grid, st) : launch_skinny<128>(tW, tX, p, grid, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("gemm_skinny");
}
