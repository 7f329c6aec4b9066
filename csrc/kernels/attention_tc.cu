// Causal flash attention FORWARD on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
// One CTA = one 128-query tile of one (batch, head).  Per 128-key tile j <= diagonal:
//     S  = Q K_j^T          tcgen05.mma  M=128 N=128 K=hs   (Q, K tiles in smem via 3-D TMA)  -> TMEM
//     P  = exp2(S*c - m)    128 softmax threads, one query row each.  The row's 128 scores are read from
//                           TMEM ONCE (tcgen05.ld) and kept in registers; P is rounded to bf16 (like the
//                           reference, which casts probabilities to bf16 before attn @ v) and written back
//                           OVER the scores in TMEM, two values per 32-bit column
//     O += P V_j            tcgen05.mma  M=128 N=hs  K=128  A = P from TENSOR MEMORY (TS form), B = V read
//                           MN-major from smem                                                  -> TMEM
// O stays in TMEM for the whole row of tiles.  The running max is LAZY: it only moves (and O is
// only rescaled, with tcgen05.ld/st) when the tile maximum exceeds it by more than 2^8, so most
// tiles skip the correction entirely; the final O / l uses the same max, so the result is exact.
// Warp roles: warp 0 TMA producer (Q once, K/V double-buffered), warp 1 MMA issuer (whole warp converged,
// one elected lane per instruction, see ptx.cuh), warps 2-5 softmax + epilogue.  hs=64 runs two CTAs per SM
// (256 TMEM columns each) so one CTA's exponentials overlap the other's MMAs; hs=128 double-buffers S instead.
// CTAs are ordered heaviest query tile first (1-D grid, all (b, h) of the last tile, then the one before, ...).
//
// Replaces reference src/models/attention.py:45-58,95 (see attention.cu for the layout contract).
#include <cuda.h>

#include "b200llm.h"
#include "common.cuh"
#include "ptx.cuh"
#include "tma.cuh"

namespace {

constexpr int BM = 128, BN = 128;
constexpr int SUB = 128 * 64 * 2;   // one [128 rows x 64 cols] bf16 sub-tile = 16 KiB
constexpr int NTHREADS = 192;
constexpr int BWD_EW_WARPS = 16;
struct MaskOn { static constexpr bool value = true; };
struct MaskOff { static constexpr bool value = false; };
constexpr int NTHREADS_BWD = 64 + 32 * BWD_EW_WARPS;   // TMA warp + MMA warp + elementwise warps

// Optional forward timeline (tools/attn_trace.cu, -DATTN_TRACE): CTA 0, clock64() at the hand-offs.  Compiled out otherwise.
#ifdef ATTN_TRACE
__device__ long long g_attn_trace_fwd[3][64][8];   // [role: 0 MMA warp, 1 / 2 softmax group A / B (one warp)][tile][event]
#define TRACEF(role, n, k) do { if (blockIdx.x == 0 && (n) < 64) g_attn_trace_fwd[role][n][k] = clock64(); } while (0)
#else
#define TRACEF(role, n, k) do { } while (0)
#endif

template <int HD> struct ACfg {
    static constexpr int NSUB = HD / 64;                 // sub-tiles per Q / K / V tile
    static constexpr int TILE = NSUB * SUB;
    static constexpr int S_BUFS = HD == 64 ? 1 : 2;
    static constexpr int KV = HD == 64 ? 2 : 3;          // K / V ring depth (a cold K/V tile takes ~2500 clk to arrive)
    static constexpr int TMEM_COLS = HD == 64 ? 256 : 512;
    static constexpr int O_COL = 128 * S_BUFS;
    // smem: Q | K[2] | V[2] | barriers
    // + barriers (256 B) + alignment slack (768 B: lets two hs=64 CTAs share one SM's 227 KB)
    static constexpr int USED = TILE + KV * TILE + KV * TILE + 256;
    static constexpr int SMEM = USED + 768;
};

template <int HD>
__global__ void __launch_bounds__(NTHREADS, HD == 64 ? 2 : 1)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO,
                   float* __restrict__ lse_out, int T, int H, int HB, int nqt, float scale_log2) {
    using C = ACfg<HD>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sQ = base;
    const uint32_t sK = sQ + C::TILE;
    const uint32_t sV = sK + C::KV * C::TILE;
    const uint32_t bars = sV + C::KV * C::TILE;
    const uint32_t q_full = bars;
    auto k_full = [&](int s) { return bars + 8u * (1 + s); };            // KV (<= 3) each
    auto k_empty = [&](int s) { return bars + 8u * (4 + s); };
    auto v_full = [&](int s) { return bars + 8u * (7 + s); };
    auto v_empty = [&](int s) { return bars + 8u * (10 + s); };
    auto s_full = [&](int s) { return bars + 8u * (13 + s); };
    auto p_full = [&](int s) { return bars + 8u * (15 + s); };           // P_j ready: p_full(j & 1), phase j >> 1
    auto pv_done = [&](int s) { return bars + 8u * (17 + s); };          // PV_j commits to pv_done(j & 1): phase j >> 1
    const uint32_t tmem_slot = bars + 8u * 19;
    uint8_t* gen = smem_raw + (base - ptx::smem_u32(smem_raw));
    if (base - ptx::smem_u32(smem_raw) > 768u) __trap();   // dynamic smem window less aligned than assumed

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    // 1-D grid in longest-processing-time order: ALL (b, h) of the heaviest (last) query tile first
    const int rank = (int)blockIdx.x / HB, hb = (int)blockIdx.x - rank * HB;
    const int qt = nqt - 1 - rank;
    const int h = hb % H, b = hb / H;
    const int q0 = qt * BM;
    const int d_model = H * HD;
    const int ntiles = qt + 1;                     // causal: key tiles 0..qt

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::prefetch_tmap(&tmO);
    }
    if (warp == 1 && lane == 0) {
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < C::KV; ++s) {
            ptx::mbar_init(k_full(s), 1); ptx::mbar_init(k_empty(s), 1);
            ptx::mbar_init(v_full(s), 1); ptx::mbar_init(v_empty(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(s_full(s), 1);
        }
        ptx::mbar_init(p_full(0), 4);
        ptx::mbar_init(p_full(1), 4);
        ptx::mbar_init(pv_done(0), 1);
        ptx::mbar_init(pv_done(1), 1);
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
    const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base)), 0);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        // whole warp converged; one elected lane executes each wrapper (uniform operands, see ptx.cuh)
        {
            ptx::mbar_arrive_expect_tx_w(q_full, C::TILE);
#pragma unroll
            for (int c = 0; c < C::NSUB; ++c) ptx::tma_load_3d_w(sQ + c * SUB, &tmQKV, q_full, h * HD + c * 64, q0, b);
            for (int j = 0; j < ntiles; ++j) {
                const int st = j % C::KV;
                const uint32_t ph = ((j / C::KV) & 1) ^ 1;
                ptx::mbar_wait(k_empty(st), ph);
                ptx::mbar_arrive_expect_tx_w(k_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d_w(sK + st * C::TILE + c * SUB, &tmQKV, k_full(st), d_model + h * HD + c * 64, j * BN, b);
                ptx::mbar_wait(v_empty(st), ph);
                ptx::mbar_arrive_expect_tx_w(v_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d_w(sV + st * C::TILE + c * SUB, &tmQKV, v_full(st), 2 * d_model + h * HD + c * 64, j * BN, b);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        // whole warp converged, one elected lane issues inside the wrappers (uniform operands, see ptx.cuh)
        {
            constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(128, 128, 0, 0);
            constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, HD, 0, 1);
            // S buffer sb is rewritten by S_{j} only after PV_{j - S_BUFS} (which read the bf16 P aliased onto it) was
            // issued by this same warp: tcgen05.mma executes in issue order, so no barrier is needed for that.
            auto issue_s = [&](int j) {
                const int st = j % C::KV, sb = j % C::S_BUFS;
                ptx::mbar_wait(k_full(st), (j / C::KV) & 1);
                ptx::tc_fence_after();
                TRACEF(0, j, 0);
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk) {
                    const uint64_t ad = ptx::umma_desc(sQ + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                    const uint64_t bd = ptx::umma_desc(sK + st * C::TILE + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                    ptx::mma_bf16_ss_w(tmem + sb * 128, ad, bd, idesc_s, kk > 0 ? 1u : 0u);
                }
                ptx::mma_commit_w(s_full(sb));
                ptx::mma_commit_w(k_empty(st));
                TRACEF(0, j, 1);
            };
            ptx::mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < ntiles; ++j) {
                const int st = j % C::KV, sb = j % C::S_BUFS;
                if (C::S_BUFS == 2 && j + 1 < ntiles) issue_s(j + 1);      // other buffer: overlaps softmax_j
                ptx::mbar_wait(p_full(j & 1), (j >> 1) & 1);
                TRACEF(0, j, 2);
                ptx::mbar_wait(v_full(st), (j / C::KV) & 1);
                ptx::tc_fence_after();
                TRACEF(0, j, 3);
#pragma unroll
                for (int kk = 0; kk < BN / 16; ++kk) {
                    // A = P [128 q x 128 kv] bf16 in TMEM (8 columns per K=16 slice, aliased onto S); B = V [kv x hs], MN-major
                    const uint64_t bd = ptx::umma_desc(sV + st * C::TILE + kk * 2048, SUB, 1024);
                    ptx::mma_bf16_ts_w(tmem + C::O_COL, tmem + sb * 128 + 8 * kk, bd, idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
                }
                ptx::mma_commit_w(pv_done(j & 1));
                TRACEF(0, j, 4);
                ptx::mma_commit_w(v_empty(st));
                if (C::S_BUFS == 1 && j + 1 < ntiles) issue_s(j + 1);      // same buffer: in order behind PV_j
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax + epilogue (warps 2..5)
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int qrow = q0 + row;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < ntiles; ++j) {
            const int sb = j % C::S_BUFS;
            if (threadIdx.x == 64) TRACEF(1, j, 0);
            ptx::mbar_wait(s_full(sb), (j / C::S_BUFS) & 1);
            ptx::tc_fence_after();
            if (threadIdx.x == 64) TRACEF(1, j, 1);
            const uint32_t ts = tmem + lane_off + sb * 128;
            const bool diag = (j == qt);
            const int k0 = j * BN;
            // ---- the whole score row (128 fp32) is read from TMEM ONCE and kept in registers
            uint32_t r[4][32];
#pragma unroll
            for (int c = 0; c < 4; ++c) ptx::tmem_ld_32x32(ts + c * 32, r[c]);
            ptx::tmem_ld_wait();
            float mx = -INFINITY;
            if (diag) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (k0 + c * 32 + i > qrow) r[c][i] = 0xff800000u;      // -inf: exp2 gives exactly 0
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(r[c][i]));
            if (threadIdx.x == 64) TRACEF(1, j, 2);
            const float xmax = mx * scale_log2;
            const bool need = xmax > m_used + 8.f;          // also true on the first tile (m_used = -inf)
            const float m_new = need ? xmax : m_used;
            const float alpha = need ? exp2f(m_used - m_new) : 1.f;
            // O belongs to PV_{j-1} until it has completed - but only a RESCALE touches O, and the lazy maximum makes
            // that rare: without one the exponentials below overlap PV_{j-1}.  (P_j itself goes into S buffer j % S_BUFS,
            // whose last reader PV_{j-S_BUFS} is complete once s_full(j) is.)  Two barriers alternate so that a wait
            // skipped on earlier tiles can never alias a later phase of the same parity.
            if (j > 0 && __any_sync(0xffffffffu, need)) {
                ptx::mbar_wait(pv_done((j - 1) & 1), ((j - 1) >> 1) & 1);
                ptx::tc_fence_after();
                {
#pragma unroll 1
                    for (int c = 0; c < HD / 32; ++c) {
                        uint32_t o[32];
                        ptx::tmem_ld_32x32(tmem + lane_off + C::O_COL + c * 32, o);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        ptx::tmem_st_32x32(tmem + lane_off + C::O_COL + c * 32, o);
                    }
                }
            }
            if (threadIdx.x == 64) TRACEF(1, j, 3);
            l *= alpha;
            m_used = m_new;
            // ---- probabilities -> bf16 pairs -> back into the first 64 columns of the S buffer (A operand of PV)
            float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float p0 = exp2f(__uint_as_float(r[c][i]) * scale_log2 - m_used);
                    const float p1 = exp2f(__uint_as_float(r[c][i + 1]) * scale_log2 - m_used);
                    rs0 += p0;
                    rs1 += p1;
                    pk[i >> 1] = pack_bf16x2(p0, p1);
                }
                ptx::tmem_st_32x16(ts + c * 16, pk);
            }
            l += rs0 + rs1;
            if (threadIdx.x == 64) TRACEF(1, j, 4);
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(p_full(j & 1));
            if (threadIdx.x == 64) TRACEF(1, j, 5);
        }
        // ---- epilogue: O / l -> bf16 -> smem (Q's tile is free now) -> TMA store; LSE -> global
        ptx::mbar_wait(pv_done((ntiles - 1) & 1), ((ntiles - 1) >> 1) & 1);
        ptx::tc_fence_after();
        const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll 1
        for (int c = 0; c < HD / 32; ++c) {
            uint32_t r[32];
            ptx::tmem_ld_32x32(tmem + lane_off + C::O_COL + c * 32, r);
            ptx::tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2)
                pk[i >> 1] = pack_bf16x2(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv);
            const uint32_t dst = sQ + (c >> 1) * SUB + row * 128;
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                const uint32_t chunk = (uint32_t)((c & 1) * 4 + ch);
                const uint32_t addr = dst + ((chunk ^ ((uint32_t)row & 7u)) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(pk[4 * ch]), "r"(pk[4 * ch + 1]),
                             "r"(pk[4 * ch + 2]), "r"(pk[4 * ch + 3]) : "memory");
            }
        }
        if (qrow < T) lse_out[((size_t)b * H + h) * T + qrow] = (m_used + log2f(l)) * 0.6931471805599453f;
        ptx::tc_fence_before();
        ptx::fence_proxy_async_smem();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (threadIdx.x == 64) {
#pragma unroll
            for (int c = 0; c < C::NSUB; ++c) ptx::tma_store_3d(&tmO, sQ + c * SUB, h * HD + c * 64, q0, b);
            ptx::tma_store_commit();
            ptx::tma_store_wait_all<0>();
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

template <int HD>
int launch_fwd_1tile(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t st) {
    using C = ACfg<HD>;
    const int d = H * HD;
    CUtensorMap tq, to;
    if (tma::make_3d(&tq, qkv, 3 * d, T, B, (uint64_t)3 * d * 2, (uint64_t)T * 3 * d * 2, 64, 128, 2)) return -1;
    if (tma::make_3d(&to, out, d, T, B, (uint64_t)d * 2, (uint64_t)T * d * 2, 64, 128, 2)) return -1;
    auto kern = attn_fwd_tc_kernel<HD>;
    static bool cfg = false;
    if (!cfg) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        cfg = true;
    }
    const int nqt = (T + BM - 1) / BM;
    B200_CHECK_CUDA(b200_launch(kern, dim3(nqt * H * B), dim3(NTHREADS), C::SMEM, st, tq, to, lse, T, H, H * B, nqt,
                                scale * 1.4426950408889634f));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Forward, two query tiles per CTA ("ping-pong"): the kernel the library launches.
//
// One CTA (one per SM) = 256 consecutive queries of one (batch, head) = tiles A and B, which share every K/V tile.
// Each tile has its own score buffer and O accumulator in TMEM and its own group of 4 softmax warps.  The MMA warp
// alternates  PV_A(j), S_A(j+1), PV_B(j), S_B(j+1),  so group B runs about half a tile behind group A: while A reads
// its 64 KB of scores out of TMEM (16 B/clk per lane quarter: 1024 clk) B is in its MUFU.EX2 phase (another 1024 clk),
// and vice versa - the two phases that bound a tile (profiles/attn_bwd_timeline.md) overlap instead of adding up.
// The CTAs are persistent (static heaviest-first work list): TMEM, the K/V ring and the barriers stay live across items
// and each thread stores its own output row, so nothing of an item's prologue or epilogue holds up the next one.
// ---------------------------------------------------------------------------------------------
template <int HD> struct A2Cfg {
    static constexpr int NSUB = HD / 64;
    static constexpr int TILE = NSUB * SUB;
    static constexpr int KV = 2;                           // K / V ring depth
    // TMEM columns.  hs = 128: S_A 0 | S_B 128 | O_A 256 | O_B 384, the bf16 P of a tile written over its scores.
    // hs = 64 has room to keep P apart (PSEP): S_A 0 | S_B 128 | P_A 256 | P_B 320 | O_A 384 | O_B 448.  Then S_g(j+1) can be
    // issued as soon as the softmax group has READ S_g(j), i.e. it runs under the group's exponentials, and the group
    // never waits for scores: per tile and group the chain was softmax -> PV -> S -> softmax.
    static constexpr bool PSEP = HD == 64;
    static constexpr int TMEM_COLS = 512;
    static constexpr int P_COL = PSEP ? 256 : 0, P_STRIDE = PSEP ? 64 : 128;
    static constexpr int O_COL = PSEP ? 384 : 256;
    static constexpr int USED = 2 * TILE + 2 * KV * TILE + 256;   // Q_A Q_B | K ring | V ring | barriers
    static constexpr int SMEM = USED + 1024;
    static_assert(O_COL + 2 * HD <= 512 && (!PSEP || P_COL + 2 * P_STRIDE <= O_COL), "TMEM budget");
    static_assert(SMEM <= 232448, "smem budget");
};
constexpr int NTHREADS_FWD2 = 320;     // TMA warp, MMA warp, 4 softmax warps per query tile

template <int HD>
__global__ void __launch_bounds__(NTHREADS_FWD2, 1)
attn_fwd2_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, bf16* __restrict__ out,
                    float* __restrict__ lse_out, int T, int H, int HB, int npairs, float scale_log2) {
    using C = A2Cfg<HD>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    auto sQ = [&](int g) { return base + (uint32_t)g * C::TILE; };
    const uint32_t sK = base + 2 * C::TILE;
    const uint32_t sV = sK + C::KV * C::TILE;
    const uint32_t bars = sV + C::KV * C::TILE;
    const uint32_t q_full = bars;
    auto k_full = [&](int s) { return bars + 8u * (1 + s); };
    auto k_empty = [&](int s) { return bars + 8u * (3 + s); };
    auto v_full = [&](int s) { return bars + 8u * (5 + s); };
    auto v_empty = [&](int s) { return bars + 8u * (7 + s); };
    auto s_full = [&](int g) { return bars + 8u * (9 + g); };      // per query tile, phase = key tile j
    auto p_full = [&](int g) { return bars + 8u * (11 + g); };
    auto pv_done = [&](int g) { return bars + 8u * (13 + g); };
    auto o_free = [&](int g) { return bars + 8u * (15 + g); };    // epilogue of the previous item has read O_g
    const uint32_t q_empty = bars + 8u * 17;                      // this item's last score MMAs have read Q
    auto s_read = [&](int g) { return bars + 8u * (18 + g); };    // PSEP: the softmax group holds S_g in registers
    const uint32_t tmem_slot = bars + 8u * 20;
    uint8_t* gen = smem_raw + (base - ptx::smem_u32(smem_raw));

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    // PERSISTENT: work item i (0 = heaviest) = pair of query tiles npairs-1 - i / HB of (b, h) = i % HB; CTA c takes items
    // c, c + grid, ...  The K/V ring, the score / P / PV barriers and TMEM stay live across items, so an item's loads and
    // first score MMAs run under the previous item's last tiles and epilogue.
    const int d_model = H * HD;
    const int nitems = npairs * HB;
    struct Item { int h, b, pr; };
    auto decode = [&](int it) {
        Item w;
        const int rank = it / HB, hb = it - rank * HB;
        w.pr = npairs - 1 - rank; w.h = hb % H; w.b = hb / H;
        return w;
    };

    if (warp == 0 && lane == 0) ptx::prefetch_tmap(&tmQKV);
    if (warp == 1 && lane == 0) {
        ptx::mbar_init(q_full, 1);
        ptx::mbar_init(q_empty, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(k_full(s), 1); ptx::mbar_init(k_empty(s), 1);
            ptx::mbar_init(v_full(s), 1); ptx::mbar_init(v_empty(s), 1);
            ptx::mbar_init(s_full(s), 1); ptx::mbar_init(p_full(s), 4); ptx::mbar_init(pv_done(s), 1);
            ptx::mbar_init(o_free(s), 4); ptx::mbar_init(s_read(s), 4);
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
    const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base)), 0);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (one elected lane per wrapper)
        int gk = 0, it_k = 0;                                  // key tiles / items issued so far
        for (int it = blockIdx.x; it < nitems; it += gridDim.x, ++it_k) {
            const Item w = decode(it);
            const int nt_b = 2 * w.pr + 2;
            ptx::mbar_wait(q_empty, (it_k & 1) ^ 1);
            ptx::mbar_arrive_expect_tx_w(q_full, 2 * C::TILE);
#pragma unroll
            for (int g = 0; g < 2; ++g)
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d_w(sQ(g) + c * SUB, &tmQKV, q_full, w.h * HD + c * 64, (2 * w.pr + g) * BM, w.b);
            for (int j = 0; j < nt_b; ++j, ++gk) {
                const int st = gk & 1;
                const uint32_t ph = ((gk >> 1) & 1) ^ 1;
                ptx::mbar_wait(k_empty(st), ph);
                ptx::mbar_arrive_expect_tx_w(k_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d_w(sK + st * C::TILE + c * SUB, &tmQKV, k_full(st), d_model + w.h * HD + c * 64, j * BN, w.b);
                ptx::mbar_wait(v_empty(st), ph);
                ptx::mbar_arrive_expect_tx_w(v_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d_w(sV + st * C::TILE + c * SUB, &tmQKV, v_full(st), 2 * d_model + w.h * HD + c * 64, j * BN, w.b);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (whole warp converged, see ptx.cuh)
        constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(128, 128, 0, 0);
        constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, HD, 0, 1);
        // S_g(j) = Q_g K_j^T into score buffer g.  The buffer's last reader, the previous PV_g (bf16 P aliased onto it), was
        // issued earlier by this warp: tcgen05.mma executes in issue order, so no barrier is needed for that.
        int gk = 0, it_k = 0;                                  // key tiles / items started so far
        int cnt[2] = {0, 0};                                   // PVs issued so far per query-tile group (barrier phases)
        int scnt[2] = {0, 0};                                  // score tiles issued so far per group
        // gkj: ring index of the key tile
        auto issue_s = [&](int g, int gkj) {
            const int st = gkj & 1;
            if (C::PSEP && scnt[g] > 0) {                      // the previous scores of this group are in registers
                ptx::mbar_wait(s_read(g), (scnt[g] - 1) & 1);
                ptx::tc_fence_after();
            }
            ++scnt[g];
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
                const uint64_t ad = ptx::umma_desc(sQ(g) + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                const uint64_t bd = ptx::umma_desc(sK + st * C::TILE + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                ptx::mma_bf16_ss_w(tmem + g * 128, ad, bd, idesc_s, kk > 0 ? 1u : 0u);
            }
            ptx::mma_commit_w(s_full(g));
            if (it_k == 0) TRACEF(0, gkj, g == 0 ? 0 : 2);
        };
        // O_g += P_g(j) V_j : A = bf16 P in TMEM (8 columns per K = 16 slice, aliased onto S_g), B = V_j read MN-major
        auto issue_pv = [&](int g, int gkj, bool first) {
            const int st = gkj & 1;
            if (first) ptx::mbar_wait(o_free(g), (it_k & 1) ^ 1);     // the previous item's epilogue has read O_g
            ptx::mbar_wait(p_full(g), cnt[g] & 1);
            ptx::mbar_wait(v_full(st), (gkj >> 1) & 1);
            ptx::tc_fence_after();
#pragma unroll
            for (int kk = 0; kk < BN / 16; ++kk) {
                const uint64_t bd = ptx::umma_desc(sV + st * C::TILE + kk * 2048, SUB, 1024);
                ptx::mma_bf16_ts_w(tmem + C::O_COL + g * HD, tmem + C::P_COL + g * C::P_STRIDE + 8 * kk, bd, idesc_pv,
                                   (!first || kk > 0) ? 1u : 0u);
            }
            ptx::mma_commit_w(pv_done(g));
            if (it_k == 0) TRACEF(0, gkj, g == 0 ? 1 : 3);
            ++cnt[g];
        };
        for (int it = blockIdx.x; it < nitems; it += gridDim.x, ++it_k) {
            const Item w = decode(it);
            const int nt_a = 2 * w.pr + 1, nt_b = 2 * w.pr + 2;   // tile g = 2 pr + g needs key tiles 0 .. 2 pr + g
            ptx::mbar_wait(q_full, it_k & 1);
            ptx::mbar_wait(k_full(gk & 1), (gk >> 1) & 1);
            ptx::tc_fence_after();
            issue_s(0, gk);
            issue_s(1, gk);
            ptx::mma_commit_w(k_empty(gk & 1));
            for (int j = 0; j < nt_b; ++j, ++gk) {
                const bool more_a = j + 1 < nt_a, more_b = j + 1 < nt_b;
                if (more_b) {                                    // K_{j+1} is needed by S_A(j+1) and / or S_B(j+1)
                    ptx::mbar_wait(k_full((gk + 1) & 1), ((gk + 1) >> 1) & 1);
                    ptx::tc_fence_after();
                }
                if (C::PSEP) {
                    // P has its own columns: the next scores go out first (each waits only for its group to have READ the
                    // current ones), the PVs follow when the probabilities are written
                    if (more_a) issue_s(0, gk + 1);
                    if (more_b) {
                        issue_s(1, gk + 1);
                        ptx::mma_commit_w(k_empty((gk + 1) & 1));
                        if (j + 2 == nt_b) ptx::mma_commit_w(q_empty);
                    }
                    if (j < nt_a) issue_pv(0, gk, j == 0);
                    issue_pv(1, gk, j == 0);
                    ptx::mma_commit_w(v_empty(gk & 1));
                } else {
                    if (j < nt_a) issue_pv(0, gk, j == 0);
                    if (more_a) issue_s(0, gk + 1);
                    issue_pv(1, gk, j == 0);
                    ptx::mma_commit_w(v_empty(gk & 1));              // both PVs of key tile j are issued
                    if (more_b) {
                        issue_s(1, gk + 1);
                        ptx::mma_commit_w(k_empty((gk + 1) & 1));    // both score MMAs of key tile j+1 are issued
                        if (j + 2 == nt_b) ptx::mma_commit_w(q_empty);   // ... and they were the item's last readers of Q
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax + epilogue: group g = query tile 2 pr + g
        const int g = (warp - 2) >> 2;
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        const uint32_t ts = tmem + lane_off + g * 128;       // this group's score buffer
        const uint32_t tp = tmem + lane_off + C::P_COL + g * C::P_STRIDE;   // ... and where its bf16 P goes
        const uint32_t to = tmem + lane_off + C::O_COL + g * HD;
        int cnt = 0;                                         // tiles processed so far by this group (barrier phases)
        int ocnt = 0;                                        // ... and by the OTHER group, up to the start of this item
        for (int it = blockIdx.x; it < nitems; it += gridDim.x) {
        const Item w = decode(it);
        const int h = w.h, b = w.b;
        const int qt = 2 * w.pr + g;                         // this group's query tile = its diagonal key tile
        const int q0 = qt * BM;
        const int qrow = q0 + row;
        const int ntiles = qt + 1;
        float m_used = -INFINITY, l = 0.f;
        for (int j = 0; j < ntiles; ++j, ++cnt) {
            const bool tr = (it == (int)blockIdx.x) && lane == 0 && q == 2;
            if (tr) TRACEF(1 + g, j, 0);
            ptx::mbar_wait(s_full(g), cnt & 1);
            ptx::tc_fence_after();
            if (tr) TRACEF(1 + g, j, 1);
            // PSEP only (hs = 64, where nothing else orders the groups): the TMEM-read phase is a token the two groups pass
            // back and forth (A_0 B_0 A_1 B_1 ...) so that they do not read tensor memory at the same time.  Measured
            // neutral: the timeline shows the tile is bound by MUFU.EX2 (two groups x 128 exp per thread = 2048 clk per
            // pair of tiles on each scheduler) once the reads are out of the way.
            if (C::PSEP && (g == 0 ? (j > 0) : (j < ntiles - 1))) {
                // A_j waits for B_{j-1}; B_j waits for A_j (B's last tile has no A counterpart)
                ptx::mbar_wait(s_read(g ^ 1), (ocnt + (g == 0 ? j - 1 : j)) & 1);
            }
            const bool diag = (j == qt);
            const int k0 = j * BN;
            // ---- the whole score row (128 fp32) is read from TMEM ONCE and kept in registers
            uint32_t r[4][32];
#pragma unroll
            for (int c = 0; c < 4; ++c) ptx::tmem_ld_32x32(ts + c * 32, r[c]);
            ptx::tmem_ld_wait();
            {
                // scores are in registers: the other group may start its TMEM reads, and (PSEP) the MMA warp may overwrite
                // the buffer with the next tile's
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(s_read(g));
            }
            if (C::PSEP) {
                // P_g and O_g belong to PV_g of the previous tile until it has completed (with P aliased onto S this is
                // implied by s_full; here the scores were issued BEFORE that PV)
                if (j > 0) {
                    ptx::mbar_wait(pv_done(g), (cnt - 1) & 1);
                    ptx::tc_fence_after();
                }
            }
            float mx = -INFINITY;
            if (diag) {
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (k0 + c * 32 + i > qrow) r[c][i] = 0xff800000u;      // -inf: exp2 gives exactly 0
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int i = 0; i < 32; ++i) mx = fmaxf(mx, __uint_as_float(r[c][i]));
            if (tr) TRACEF(1 + g, j, 2);
            const float xmax = mx * scale_log2;
            const bool need = xmax > m_used + 8.f;          // also true on the first tile (m_used = -inf)
            const float m_new = need ? xmax : m_used;
            const float alpha = need ? exp2f(m_used - m_new) : 1.f;
            // O may be rescaled right away: PV_g(j-1) has completed (aliased P: S_g(j) was issued behind it, so s_full says so;
            // PSEP: the pv_done wait above), and PV_g(j) is not issued before this group's p_full arrive below.
            if (j > 0 && __any_sync(0xffffffffu, need)) {
#pragma unroll 1
                for (int c = 0; c < HD / 32; ++c) {
                    uint32_t o[32];
                    ptx::tmem_ld_32x32(to + c * 32, o);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                    ptx::tmem_st_32x32(to + c * 32, o);
                }
            }
            l *= alpha;
            m_used = m_new;
            // ---- probabilities -> bf16 pairs -> back into the first 64 columns of the score buffer (A operand of PV)
            float rs0 = 0.f, rs1 = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float p0 = exp2f(__uint_as_float(r[c][i]) * scale_log2 - m_used);
                    const float p1 = exp2f(__uint_as_float(r[c][i + 1]) * scale_log2 - m_used);
                    rs0 += p0;
                    rs1 += p1;
                    pk[i >> 1] = pack_bf16x2(p0, p1);
                }
                ptx::tmem_st_32x16(tp + c * 16, pk);
            }
            l += rs0 + rs1;
            if (tr) TRACEF(1 + g, j, 3);
            ptx::tmem_st_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(p_full(g));
            if (tr) TRACEF(1 + g, j, 4);
        }
        // ---- epilogue: O / l -> bf16 -> global (each thread writes its own row); LSE -> global
        ptx::mbar_wait(pv_done(g), (cnt - 1) & 1);
        ptx::tc_fence_after();
        const float inv = l > 0.f ? 1.f / l : 0.f;
        bf16* orow = out + ((size_t)b * T + qrow) * d_model + h * HD;
#pragma unroll 1
        for (int c = 0; c < HD / 32; ++c) {
            uint32_t o[32];
            ptx::tmem_ld_32x32(to + c * 32, o);
            ptx::tmem_ld_wait();
            if (qrow < T) {
#pragma unroll
                for (int i = 0; i < 32; i += 8)
                    *reinterpret_cast<uint4*>(orow + c * 32 + i) =
                        make_uint4(pack_bf16x2(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv),
                                   pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv),
                                   pack_bf16x2(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv),
                                   pack_bf16x2(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv));
            }
        }
        if (qrow < T) lse_out[((size_t)b * H + h) * T + qrow] = (m_used + log2f(l)) * 0.6931471805599453f;
        // O_g may be overwritten by the next item's first PV
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(o_free(g));
        ocnt += 2 * w.pr + (g ^ 1) + 1;
        }   // items
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

template <int HD>
int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t st) {
    using C = A2Cfg<HD>;
    const int d = H * HD;
    CUtensorMap tq;
    if (tma::make_3d(&tq, qkv, 3 * d, T, B, (uint64_t)3 * d * 2, (uint64_t)T * 3 * d * 2, 64, 128, 2)) return -1;
    auto kern = attn_fwd2_tc_kernel<HD>;
    static bool cfg = false;
    if (!cfg) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        cfg = true;
    }
    const int nqt = (T + BM - 1) / BM, npairs = (nqt + 1) / 2;
    const int nitems = npairs * H * B;
    const int grid = nitems < b200_num_sms() ? nitems : b200_num_sms();
    B200_CHECK_CUDA(b200_launch(kern, dim3(grid), dim3(NTHREADS_FWD2), C::SMEM, st, tq, (bf16*)out, lse, T, H, H * B, npairs,
                                scale * 1.4426950408889634f));
    return 0;
}



// =============================================================================================
// Backward on tcgen05: two passes of ONE persistent kernel template, no atomics, no fp32 scratch.
//
//   pass DKV (key-stationary):   work item = 128 keys of one (b, h); streams 64-query tiles i >= diagonal
//        S^T = K Q_i^T , dP^T = V dO_i^T                    2 MMAs (M=128 keys, N=64 queries) -> TMEM
//        P^T = exp2(S^T c - lse[q]) , dS^T = P^T (dP^T - delta[q])            thread = key row
//        dV += P^T dO_i , dK += dS^T Q_i                    2 MMAs accumulate in TMEM over all i
//   pass DQ (query-stationary):  work item = 128 queries; streams 64-key tiles j <= diagonal
//        S = Q K_j^T , dP = dO V_j^T                        2 MMAs (M=128 queries, N=64 keys)
//        dS = P (dP - delta[q])                             thread = query row
//        dQ += dS K_j                                       1 MMA accumulates in TMEM over all j
//   (the 1/sqrt(hs) factor of dS is applied once per output element in the epilogue)
//
// Both passes have the same shape: two "score" MMAs [128 x 64] with K = hs from K-major smem tiles, an elementwise
// stage, and "accumulate" MMAs [128 x hs] with K = 64 whose A operand is the bf16 result of the elementwise stage
// read straight from TENSOR MEMORY (it is written back over the fp32 scores it came from, 2 bf16 per column) and
// whose B operand is the streamed smem tile read MN-major.  Recomputing S in the second pass costs 2 extra MMAs per
// tile pair but removes the cross-CTA dQ reduction entirely (fp32 RED traffic + a conversion pass were the
// bottleneck of a single-pass version).
//
// Pipeline (what the in-kernel timeline, tools/attn_trace.cu, and ncu showed along the way):
//   * one PERSISTENT CTA per SM walks a static list of work items (heaviest stationary tiles first, round-robin
//     over CTAs): the next item's stationary tiles, its first streamed tiles and its first score MMAs are in flight
//     while the current item's accumulators are drained, so the ~5000 clk of per-item prologue / cold TMA /
//     epilogue that a CTA-per-item launch exposed (40% of the pass at T = 1024) are hidden;
//   * score accumulators are double-buffered in TMEM: the MMA warp issues scores(n+1) BEFORE the accumulate MMAs
//     of tile n, so the tensor pipe works while the elementwise warps are on tile n;
//   * 16 elementwise warps in two groups that take alternate tiles (group g owns score buffer g): the exp2-heavy
//     middle of one tile overlaps the TMEM loads / stores / barrier hand-offs of the next;
//   * streamed tiles sit in a 6-deep (hs=64) / 4-deep (hs=128) TMA ring: a cold L2 answers in ~2000 clk;
//   * every tcgen05.mma / TMA is issued with warp-uniform operands by one elected lane (see ptx.cuh): the
//     single-lane form cost 70-100 clk per MMA, three times the 32 clk a [128x64x16] MMA takes.
// Warp roles: 0 TMA (+ per-tile lse/delta staging in the key-stationary pass), 1 MMA, 2-17 elementwise + epilogue.
// =============================================================================================
constexpr int SUB64 = 64 * 64 * 2;     // [64 rows x 64 cols] bf16 sub-tile = 8 KiB

// Optional in-kernel timeline (tools/attn_trace.cu builds this file with -DATTN_TRACE): CTA 0 records clock64()
// at the pipeline hand-offs of its first work item.  Compiled out of the library.
#ifdef ATTN_TRACE
__device__ long long g_attn_trace[2][3][64][8];   // [pass][role][tile][event]
#define TRACE(role, n, k) do { if (blockIdx.x == 0 && (n) < 64) g_attn_trace[DKV ? 1 : 0][role][n][k] = clock64(); } while (0)
#else
#define TRACE(role, n, k) do { } while (0)
#endif

template <int HD, bool DKV> struct BCfg {
    static constexpr int NSUB = HD / 64;
    static constexpr int STAT = NSUB * SUB;                 // stationary tile [128 x HD]
    static constexpr int STR = NSUB * SUB64;                // streamed tile  [64 x HD]
    static constexpr int NST = HD == 64 ? 6 : 4;            // streamed stages
    static constexpr int STAT_BUFS = HD == 64 ? 2 : 1;      // stationary operand pairs resident at once
    static constexpr int NACC = DKV ? 2 : 1;                // accumulators per item (dV, dK | dQ)
    static constexpr int ACC_BUFS = (256 + 2 * NACC * HD <= 512) ? 2 : 1;
    static constexpr int ACC = 256;                         // TMEM columns: scores stage s at 128*s (S | dP), then accumulators
    static constexpr int TMEM_COLS = 512;
    // stationary x2 tensors x STAT_BUFS | streamed x2 tensors x NST stages | barriers (256) + lse/delta staging (512 per stage)
    static constexpr int USED = 2 * STAT * STAT_BUFS + 2 * NST * STR + 256 + NST * 512;
    static constexpr int SMEM = USED + 1024;
    static_assert(ACC + ACC_BUFS * NACC * HD <= 512, "TMEM budget");
    static_assert(SMEM <= 232448, "smem budget");
};

// delta[b][h][t] = sum_c out[b][t][h][c] * dout[b][t][h][c].  16-byte loads: HD/8 lanes per (row, head).
template <int HD>
__global__ void __launch_bounds__(256) attn_bwd_delta_tc_kernel(const bf16* __restrict__ out, const bf16* __restrict__ dout,
                                                               float* __restrict__ delta, int B, int T, int H) {
    pdl_sync();
    constexpr int LPR = HD / 8;                               // lanes per (row, head)
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long item = gid / LPR;                         // (m, h) flattened: m * H + h
    const int sub = (int)(gid % LPR);
    const long long nitem = (long long)B * T * H;
    float s = 0.f;
    if (item < nitem) {
        const size_t off = (size_t)item * HD + sub * 8;       // [B*T][H][HD] is contiguous in (m, h, c)
        const uint4 ua = ldg_stream_u4(reinterpret_cast<const uint4*>(out + off));
        const uint4 ud = ldg_stream_u4(reinterpret_cast<const uint4*>(dout + off));
        const float2 a0 = unpack_bf16x2(ua.x), a1 = unpack_bf16x2(ua.y), a2 = unpack_bf16x2(ua.z), a3 = unpack_bf16x2(ua.w);
        const float2 d0 = unpack_bf16x2(ud.x), d1 = unpack_bf16x2(ud.y), d2 = unpack_bf16x2(ud.z), d3 = unpack_bf16x2(ud.w);
        s = a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y + a2.x * d2.x + a2.y * d2.y + a3.x * d3.x + a3.y * d3.y;
    }
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (item < nitem && sub == 0) {
        const int h = (int)(item % H);
        const long long m = item / H;
        delta[((size_t)(m / T) * H + h) * T + (m % T)] = s;
    }
}

// tm128 / tm64: qkv viewed [B][T][3d] with 128- / 64-row boxes; tdo128 / tdo64: dO viewed [B][T][d].
// Work item i (0 = heaviest): stationary tile rank i / HB of (b, h) = i % HB; CTA c takes items c, c + grid, ...
template <int HD, bool DKV>
__global__ void __launch_bounds__(NTHREADS_BWD, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tm128, const __grid_constant__ CUtensorMap tm64,
                   const __grid_constant__ CUtensorMap tdo128, const __grid_constant__ CUtensorMap tdo64,
                   const float* __restrict__ lse, const float* __restrict__ delta, bf16* __restrict__ dqkv, int T, int H,
                   int HB, int ntile, float scale, float scale_log2) {
    using C = BCfg<HD, DKV>;
    constexpr int NST = C::NST, SB = C::STAT_BUFS, AB = C::ACC_BUFS;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    auto sA0 = [&](int p) { return base + (uint32_t)p * 2 * C::STAT; };             // stationary operand of score MMA 1 (K | Q)
    auto sA1 = [&](int p) { return base + (uint32_t)p * 2 * C::STAT + C::STAT; };   // stationary operand of score MMA 2 (V | dO)
    const uint32_t sB0 = base + SB * 2 * C::STAT;       // streamed, NST stages: (Q_i | K_j)
    const uint32_t sB1 = sB0 + NST * C::STR;            // streamed, NST stages: (dO_i | V_j)
    const uint32_t bars = sB1 + NST * C::STR;
    auto str_full = [&](int s) { return bars + 8u * s; };                         // NST each
    auto str_empty = [&](int s) { return bars + 8u * (NST + s); };
    auto vec_full = [&](int s) { return bars + 8u * (2 * NST + s); };
    auto sc_full = [&](int s) { return bars + 8u * (3 * NST + s); };              // 2 each
    auto out_full = [&](int s) { return bars + 8u * (3 * NST + 2 + s); };
    auto stat_full = [&](int s) { return bars + 8u * (3 * NST + 4 + s); };
    auto stat_empty = [&](int s) { return bars + 8u * (3 * NST + 6 + s); };
    auto acc_full = [&](int s) { return bars + 8u * (3 * NST + 8 + s); };
    auto acc_empty = [&](int s) { return bars + 8u * (3 * NST + 10 + s); };
    const uint32_t tmem_slot = bars + 8u * (3 * NST + 12);
    static_assert(8 * (3 * NST + 13) <= 256, "barrier block");
    uint8_t* gen = smem_raw + (base - ptx::smem_u32(smem_raw));
    float* s_vec = reinterpret_cast<float*>(gen + (bars + 256 - base));   // [NST stages][lse2(64) | delta(64)]

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // provably warp-uniform
    const int d_model = H * HD;
    const int n64 = (T + 63) / 64;
    const int nitems = ntile * HB;

    // decode of work item `it`: stationary tile, (b, h) and its range of streamed 64-row tiles
    struct Item { int h, b, r0, it0, niter; };
    auto decode = [&](int it) {
        Item w;
        const int rank = it / HB, hb = it - rank * HB;
        w.h = hb % H; w.b = hb / H;
        const int ts = DKV ? rank : ntile - 1 - rank;     // rank 0 = most streamed tiles
        w.r0 = ts * 128;                                    // first key (DKV) / query (DQ) row of the stationary tile
        w.it0 = DKV ? 2 * ts : 0;
        w.niter = (DKV ? n64 : min(n64, 2 * ts + 2)) - w.it0;
        return w;
    };

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tm128); ptx::prefetch_tmap(&tm64);
        ptx::prefetch_tmap(&tdo128); ptx::prefetch_tmap(&tdo64);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < NST; ++s) {
            ptx::mbar_init(str_full(s), 1); ptx::mbar_init(str_empty(s), 1); ptx::mbar_init(vec_full(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(sc_full(s), 1);
            ptx::mbar_init(out_full(s), BWD_EW_WARPS / 2);
            ptx::mbar_init(stat_full(s), 1);  ptx::mbar_init(stat_empty(s), 1);
            ptx::mbar_init(acc_full(s), 1);   ptx::mbar_init(acc_empty(s), BWD_EW_WARPS);
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
    const uint32_t tmem = __shfl_sync(0xffffffffu, *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base)), 0);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (one elected lane per wrapper) + statistics staging
        int g = 0;                                           // streamed tiles issued so far (all items)
        int k = 0;                                           // items started so far
        for (int it = blockIdx.x; it < nitems; it += gridDim.x, ++k) {
            const Item w = decode(it);
            const int sp = k % SB;
            ptx::mbar_wait(stat_empty(sp), ((k / SB) & 1) ^ 1);
            ptx::mbar_arrive_expect_tx_w(stat_full(sp), 2 * C::STAT);
#pragma unroll
            for (int c = 0; c < C::NSUB; ++c) {
                if (DKV) {
                    ptx::tma_load_3d_w(sA0(sp) + c * SUB, &tm128, stat_full(sp), d_model + w.h * HD + c * 64, w.r0, w.b);        // K
                    ptx::tma_load_3d_w(sA1(sp) + c * SUB, &tm128, stat_full(sp), 2 * d_model + w.h * HD + c * 64, w.r0, w.b);    // V
                } else {
                    ptx::tma_load_3d_w(sA0(sp) + c * SUB, &tm128, stat_full(sp), w.h * HD + c * 64, w.r0, w.b);                  // Q
                    ptx::tma_load_3d_w(sA1(sp) + c * SUB, &tdo128, stat_full(sp), w.h * HD + c * 64, w.r0, w.b);                 // dO
                }
            }
            const float* lse_g = lse + ((size_t)w.b * H + w.h) * T;
            const float* del_g = delta + ((size_t)w.b * H + w.h) * T;
            for (int n = 0; n < w.niter; ++n, ++g) {
                const int st = g % NST;
                const int row0 = (w.it0 + n) * 64;
                ptx::mbar_wait(str_empty(st), ((g / NST) & 1) ^ 1);
                if (k == 0) TRACE(0, n, 0);
                ptx::mbar_arrive_expect_tx_w(str_full(st), 2 * C::STR);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c) {
                    if (DKV) {
                        ptx::tma_load_3d_w(sB0 + st * C::STR + c * SUB64, &tm64, str_full(st), w.h * HD + c * 64, row0, w.b);    // Q_i
                        ptx::tma_load_3d_w(sB1 + st * C::STR + c * SUB64, &tdo64, str_full(st), w.h * HD + c * 64, row0, w.b);   // dO_i
                    } else {
                        ptx::tma_load_3d_w(sB0 + st * C::STR + c * SUB64, &tm64, str_full(st), d_model + w.h * HD + c * 64, row0, w.b);      // K_j
                        ptx::tma_load_3d_w(sB1 + st * C::STR + c * SUB64, &tm64, str_full(st), 2 * d_model + w.h * HD + c * 64, row0, w.b);  // V_j
                    }
                }
                if (DKV) {
                    // per-column (query) statistics of this streamed tile -> the same ring slot as the tile itself
                    // (str_empty(st), waited on above, also says the elementwise warps are done with the slot)
                    float* v = s_vec + st * 128;
                    const int q0 = row0 + lane, q1 = row0 + 32 + lane;
                    v[lane] = q0 < T ? lse_g[q0] * 1.4426950408889634f : INFINITY;
                    v[32 + lane] = q1 < T ? lse_g[q1] * 1.4426950408889634f : INFINITY;
                    v[64 + lane] = q0 < T ? del_g[q0] : 0.f;
                    v[96 + lane] = q1 < T ? del_g[q1] : 0.f;
                    __syncwarp();
                    if (lane == 0) { ptx::mbar_arrive(vec_full(st)); if (k == 0) TRACE(0, n, 1); }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        // The whole warp runs this loop converged (uniform operands, see ptx.cuh); one elected lane issues inside each
        // wrapper.  Tiles of ALL items form one sequence g = 0, 1, ...: scores(g) are issued before the accumulate
        // MMAs of tile g-1 even when g-1 belongs to the previous item.
        constexpr uint32_t id_sc = ptx::umma_idesc_bf16(128, 64, 0, 0);     // scores: both operands K-major
        constexpr uint32_t id_acc = ptx::umma_idesc_bf16(128, HD, 0, 1);    // accumulate: A from TMEM (K-major), B MN-major
        int g = 0, k = 0;
        // the tile whose accumulate MMAs are still to be issued
        bool pend = false;
        int p_g = 0, p_k = 0;
        bool p_first = false, p_last = false;
        auto accumulate = [&]() {
            const int sm = p_g & 1, stm = p_g % NST, ap = p_k % AB;
            const uint32_t b0 = sB0 + stm * C::STR, b1 = sB1 + stm * C::STR;
            const uint32_t ta = tmem + sm * 128;             // bf16 P^T at +16*kk, dS^T (dS) at +64+16*kk
            const uint32_t td = tmem + C::ACC + ap * (C::NACC * HD);
            if (p_first) ptx::mbar_wait(acc_empty(ap), ((p_k / AB) & 1) ^ 1);     // drained by the epilogue of item p_k - AB
            ptx::mbar_wait(out_full(sm), (p_g >> 1) & 1);
            ptx::tc_fence_after();
            if (p_k == 0) TRACE(1, p_g, 3);
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {                 // contraction over the 64 streamed rows, 16 per MMA
                const uint32_t acc = (!p_first || kk > 0) ? 1u : 0u;
                if (DKV) {
                    // dV += P^T dO_i ; dK += dS^T Q_i      (A from TMEM, B = streamed tile read MN-major)
                    ptx::mma_bf16_ts_w(td, ta + 16 * kk, ptx::umma_desc(b1 + kk * 2048, SUB64, 1024), id_acc, acc);
                    ptx::mma_bf16_ts_w(td + HD, ta + 64 + 16 * kk, ptx::umma_desc(b0 + kk * 2048, SUB64, 1024), id_acc, acc);
                } else {
                    // dQ += dS K_j
                    ptx::mma_bf16_ts_w(td, ta + 64 + 16 * kk, ptx::umma_desc(b0 + kk * 2048, SUB64, 1024), id_acc, acc);
                }
            }
            ptx::mma_commit_w(str_empty(stm));
            if (p_last) ptx::mma_commit_w(acc_full(ap));
            if (p_k == 0) TRACE(1, p_g, 4);
            pend = false;
        };
        for (int it = blockIdx.x; it < nitems; it += gridDim.x, ++k) {
            const Item w = decode(it);
            const int sp = k % SB;
            if (SB == 1 && pend) accumulate();               // do not hold the last tile back behind a stationary reload
            ptx::mbar_wait(stat_full(sp), (k / SB) & 1);
            for (int n = 0; n < w.niter; ++n, ++g) {
                const int st = g % NST, s = g & 1;
                const uint32_t b0 = sB0 + st * C::STR, b1 = sB1 + st * C::STR;
                ptx::mbar_wait(str_full(st), (g / NST) & 1);
                if (k == 0) TRACE(1, n, 0);
                // scores stage s is free: its last readers were the accumulate MMAs of tile g-2, issued earlier by
                // this same warp, and tcgen05.mma executes in issue order
                ptx::tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk) {
                    const uint32_t oa = (kk >> 2) * SUB + (kk & 3) * 32, ob = (kk >> 2) * SUB64 + (kk & 3) * 32;
                    ptx::mma_bf16_ss_w(tmem + s * 128, ptx::umma_desc(sA0(sp) + oa, 16, 1024), ptx::umma_desc(b0 + ob, 16, 1024),
                                       id_sc, kk > 0 ? 1u : 0u);
                }
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk) {
                    const uint32_t oa = (kk >> 2) * SUB + (kk & 3) * 32, ob = (kk >> 2) * SUB64 + (kk & 3) * 32;
                    ptx::mma_bf16_ss_w(tmem + s * 128 + 64, ptx::umma_desc(sA1(sp) + oa, 16, 1024), ptx::umma_desc(b1 + ob, 16, 1024),
                                       id_sc, kk > 0 ? 1u : 0u);
                }
                ptx::mma_commit_w(sc_full(s));
                if (n == w.niter - 1) ptx::mma_commit_w(stat_empty(sp));   // last reader of this item's stationary tiles
                if (k == 0) TRACE(1, n, 2);
                if (pend) accumulate();
                pend = true; p_g = g; p_k = k; p_first = (n == 0); p_last = (n == w.niter - 1);
            }
        }
        if (pend) accumulate();
    } else {
        // ------------------------------------------------------------------ elementwise warps (2..17)
        // Four warps share each TMEM lane quarter (tcgen05.ld reaches lanes 32*(warp%4)..+31): thread = stationary row.
        // Two groups of 8 warps take alternate tiles of the global tile sequence (group = score buffer); inside a
        // group a warp owns 32 of the 64 score columns.  The backward needs no row reductions (lse and delta are
        // inputs), so the column split is free.
        constexpr int CPW = 32;
        const int q = warp & 3;
        const int grp = (warp - 2) >> 3;
        const int part = ((warp - 2) >> 2) & 1;
        const int row = q * 32 + lane;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        int g = 0, k = 0;
        for (int it = blockIdx.x; it < nitems; it += gridDim.x, ++k) {
            const Item w = decode(it);
            const int srow = w.r0 + row;                                 // key row (DKV) / query row (DQ)
            const float* lse_g = lse + ((size_t)w.b * H + w.h) * T;
            const float* del_g = delta + ((size_t)w.b * H + w.h) * T;
            float my_l2 = INFINITY, my_dl = 0.f;
            if (!DKV && srow < T) { my_l2 = lse_g[srow] * 1.4426950408889634f; my_dl = del_g[srow]; }
            for (int n = 0; n < w.niter; ++n, ++g) {
                if ((g & 1) != grp) continue;
                const int s = g & 1, ph = (g >> 1) & 1;
                const int c0 = (w.it0 + n) * 64;                         // first streamed row = first score column
                const float* vl = s_vec + (g % NST) * 128;
                // causal mask is needed only where the streamed tile overlaps the diagonal
                const bool need_mask = DKV ? (c0 < w.r0 + 128) : (c0 + 63 > w.r0);
                if (k == 0 && warp == 2 && lane == 0) TRACE(2, n, 0);
                if (DKV) ptx::mbar_wait(vec_full(g % NST), (g / NST) & 1);
                ptx::mbar_wait(sc_full(s), ph);
                ptx::tc_fence_after();
                if (k == 0 && warp == 2 && lane == 0) TRACE(2, n, 2);
                // MASKED is a compile-time flag so the 7-of-9 off-diagonal tiles carry no compare/select instructions
                auto tile = [&](auto masked_tag) {
                    constexpr bool MASKED = decltype(masked_tag)::value;
#pragma unroll
                    for (int c = 0; c < CPW / 16; ++c) {
                        const int cb = part * CPW + c * 16;             // first score column of this 16-wide piece
                        uint32_t rs[16], rp[16];
                        ptx::tmem_ld_32x16(tmem + lane_off + s * 128 + cb, rs);
                        ptx::tmem_ld_32x16(tmem + lane_off + s * 128 + 64 + cb, rp);
                        ptx::tmem_ld_wait();
                        uint32_t pk[8], dk[8];
#pragma unroll
                        for (int e = 0; e < 16; e += 2) {
                            const int col = cb + e;
                            const float l2a = DKV ? vl[col] : my_l2, l2b = DKV ? vl[col + 1] : my_l2;
                            const float dla = DKV ? vl[64 + col] : my_dl, dlb = DKV ? vl[64 + col + 1] : my_dl;
                            // Rows / columns past T need no extra test: TMA zero-fills them, out-of-range QUERIES carry
                            // lse = +inf (p = 0), and out-of-range KEYS are either masked by causality or only feed
                            // output rows that are never stored.
                            float p0 = exp2f(__uint_as_float(rs[e]) * scale_log2 - l2a);
                            float p1 = exp2f(__uint_as_float(rs[e + 1]) * scale_log2 - l2b);
                            if (MASKED) {
                                // key index must be <= query index
                                const int key0 = DKV ? srow : c0 + col, qry0 = DKV ? c0 + col : srow;
                                if (key0 > qry0) p0 = 0.f;
                                if (DKV ? (key0 > qry0 + 1) : (key0 + 1 > qry0)) p1 = 0.f;
                            }
                            const float s0 = p0 * (__uint_as_float(rp[e]) - dla);
                            const float s1 = p1 * (__uint_as_float(rp[e + 1]) - dlb);
                            if (DKV) pk[e >> 1] = pack_bf16x2(p0, p1);
                            dk[e >> 1] = pack_bf16x2(s0, s1);
                        }
                        // bf16 results go back into the first 8 of the 16 TMEM columns this warp just read: they are the
                        // K=16 slice number cb/16 of the A operand of the accumulate MMAs (2 bf16 per 32-bit column)
                        if (DKV) ptx::tmem_st_32x8(tmem + lane_off + s * 128 + cb, pk);
                        ptx::tmem_st_32x8(tmem + lane_off + s * 128 + 64 + cb, dk);
                    }
                };
                if (need_mask) tile(MaskOn{}); else tile(MaskOff{});
                if (k == 0 && warp == 2 && lane == 0) TRACE(2, n, 4);
                ptx::tmem_st_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(out_full(s));
                if (k == 0 && warp == 2 && lane == 0) TRACE(2, n, 5);
            }
            // ---- epilogue of this item: accumulator rows -> bf16 -> dqkv (thread = stationary row; the four warps of a
            //      lane quarter take every fourth 16-column piece)
            const int ap = k % AB;
            ptx::mbar_wait(acc_full(ap), (k / AB) & 1);
            ptx::tc_fence_after();
            const uint32_t td = tmem + lane_off + C::ACC + ap * (C::NACC * HD);
            bf16* g0 = dqkv + ((size_t)w.b * T + srow) * (3 * d_model) + w.h * HD + (DKV ? 2 * d_model : 0);   // dV | dQ
            bf16* g1 = dqkv + ((size_t)w.b * T + srow) * (3 * d_model) + w.h * HD + d_model;                    // dK
            const float f0 = DKV ? 1.f : scale;
#pragma unroll 1
            for (int c = part + 2 * grp; c < HD / 16; c += 4) {
                uint32_t r0v[16], r1v[16];
                ptx::tmem_ld_32x16(td + c * 16, r0v);
                if (DKV) ptx::tmem_ld_32x16(td + HD + c * 16, r1v);
                ptx::tmem_ld_wait();
                if (srow < T) {
#pragma unroll
                    for (int e = 0; e < 16; e += 8) {
                        *reinterpret_cast<uint4*>(g0 + c * 16 + e) =
                            make_uint4(pack_bf16x2(f0 * __uint_as_float(r0v[e]), f0 * __uint_as_float(r0v[e + 1])),
                                       pack_bf16x2(f0 * __uint_as_float(r0v[e + 2]), f0 * __uint_as_float(r0v[e + 3])),
                                       pack_bf16x2(f0 * __uint_as_float(r0v[e + 4]), f0 * __uint_as_float(r0v[e + 5])),
                                       pack_bf16x2(f0 * __uint_as_float(r0v[e + 6]), f0 * __uint_as_float(r0v[e + 7])));
                        if (DKV)
                            *reinterpret_cast<uint4*>(g1 + c * 16 + e) =
                                make_uint4(pack_bf16x2(scale * __uint_as_float(r1v[e]), scale * __uint_as_float(r1v[e + 1])),
                                           pack_bf16x2(scale * __uint_as_float(r1v[e + 2]), scale * __uint_as_float(r1v[e + 3])),
                                           pack_bf16x2(scale * __uint_as_float(r1v[e + 4]), scale * __uint_as_float(r1v[e + 5])),
                                           pack_bf16x2(scale * __uint_as_float(r1v[e + 6]), scale * __uint_as_float(r1v[e + 7])));
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(acc_empty(ap));
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, C::TMEM_COLS);
    }
}

template <int HD, bool DKV>
int launch_bwd_pass(const CUtensorMap& t128, const CUtensorMap& t64, const CUtensorMap& d128, const CUtensorMap& d64,
                    const float* lse, const float* delta, void* dqkv, int B, int T, int H, float scale, cudaStream_t st) {
    using C = BCfg<HD, DKV>;
    auto kern = attn_bwd_tc_kernel<HD, DKV>;
    static bool cfg = false;
    if (!cfg) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM));
        cfg = true;
    }
    const int ntile = (T + 127) / 128;
    const int nitems = ntile * H * B;
    const int grid = nitems < b200_num_sms() ? nitems : b200_num_sms();
    B200_CHECK_CUDA(b200_launch(kern, dim3(grid), dim3(NTHREADS_BWD), C::SMEM, st, t128, t64, d128, d64, lse, delta,
                                (bf16*)dqkv, T, H, H * B, ntile, scale, scale * 1.4426950408889634f));
    return 0;
}

template <int HD>
int launch_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta, void* dqkv, int B, int T,
               int H, float scale, cudaStream_t st) {
    const int d = H * HD;
    CUtensorMap t128, t64, d128, d64;
    if (tma::make_3d(&t128, qkv, 3 * d, T, B, (uint64_t)3 * d * 2, (uint64_t)T * 3 * d * 2, 64, 128, 2)) return -1;
    if (tma::make_3d(&t64, qkv, 3 * d, T, B, (uint64_t)3 * d * 2, (uint64_t)T * 3 * d * 2, 64, 64, 2)) return -1;
    if (tma::make_3d(&d128, dout, d, T, B, (uint64_t)d * 2, (uint64_t)T * d * 2, 64, 128, 2)) return -1;
    if (tma::make_3d(&d64, dout, d, T, B, (uint64_t)d * 2, (uint64_t)T * d * 2, 64, 64, 2)) return -1;
    const long long delta_threads = (long long)B * T * H * (HD / 8);
    B200_CHECK_CUDA(b200_launch(attn_bwd_delta_tc_kernel<HD>, dim3((unsigned)((delta_threads + 255) / 256)), dim3(256), 0, st,
                                (const bf16*)out, (const bf16*)dout, delta, B, T, H));
    if (launch_bwd_pass<HD, true>(t128, t64, d128, d64, lse, delta, dqkv, B, T, H, scale, st)) return -1;
    if (launch_bwd_pass<HD, false>(t128, t64, d128, d64, lse, delta, dqkv, B, T, H, scale, st)) return -1;
    return 0;
}

}  // namespace

extern "C" int b200_attn_fwd_tc_tiles(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale,
                                      int tiles, cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_fwd_tc: head size must be 64 or 128 (got %d)", hs);
    B200_REQUIRE((H * hs) % 8 == 0, "attn_fwd_tc: model width must be a multiple of 8");
    B200_REQUIRE(tiles >= 0 && tiles <= 2, "attn_fwd_tc: tiles per CTA must be 0 (default), 1 or 2");
    // Query tiles per CTA.  hs = 128: two (one persistent CTA per SM, the two softmax groups run half a tile apart:
    // 0.197 -> 0.159 ms on the 1.3B shape).  hs = 64: one, two CTAs per SM (0.060 ms; the two-tile kernel measures
    // 0.061 ms there).  B200_ATTN_FWD_TILES=1|2 overrides the default.
    static int env_tiles = -1;
    if (env_tiles < 0) {
        const char* e = getenv("B200_ATTN_FWD_TILES");
        env_tiles = (e != nullptr && (e[0] == '1' || e[0] == '2')) ? e[0] - '0' : 0;
    }
    const int use = tiles ? tiles : (env_tiles ? env_tiles : (hs == 128 ? 2 : 1));
    int rc;
    if (use == 1) rc = hs == 64 ? launch_fwd_1tile<64>(qkv, out, lse, B, T, H, scale, st) : launch_fwd_1tile<128>(qkv, out, lse, B, T, H, scale, st);
    else rc = hs == 64 ? launch_fwd<64>(qkv, out, lse, B, T, H, scale, st) : launch_fwd<128>(qkv, out, lse, B, T, H, scale, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("attn_fwd_tc");
}

extern "C" int b200_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale,
                                cudaStream_t st) {
    return b200_attn_fwd_tc_tiles(qkv, out, lse, B, T, H, hs, scale, 0, st);
}

// delta: [B, H, T] fp32 scratch (row sums of dO * O); nothing else is needed.
extern "C" int b200_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                                void* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_bwd_tc: head size must be 64 or 128 (got %d)", hs);
    int rc = hs == 64 ? launch_bwd<64>(qkv, out, dout, lse, delta, dqkv, B, T, H, scale, st)
                      : launch_bwd<128>(qkv, out, dout, lse, delta, dqkv, B, T, H, scale, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("attn_bwd_tc");
}
