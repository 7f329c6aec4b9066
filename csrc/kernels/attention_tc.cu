// Causal flash attention FORWARD on the 5th-generation tensor cores (tcgen05 + TMEM + TMA).
//
// One CTA = one 128-query tile of one (batch, head).  Per 128-key tile j <= diagonal:
//     S  = Q K_j^T          tcgen05.mma  M=128 N=128 K=hs   (Q, K tiles in smem via 3-D TMA)  -> TMEM
//     P  = exp2(S*c - m)    128 softmax threads, one query row each, S read with tcgen05.ld;
//                           P is rounded to bf16 and written to swizzled smem (like the reference,
//                           which casts probabilities to bf16 before attn @ v)
//     O += P V_j            tcgen05.mma  M=128 N=hs  K=128  (V is the MN-major B operand)     -> TMEM
// O stays in TMEM for the whole row of tiles.  The running max is LAZY: it only moves (and O is
// only rescaled, with tcgen05.ld/st) when the tile maximum exceeds it by more than 2^8, so most
// tiles skip the correction entirely; the final O / l uses the same max, so the result is exact.
// Warp roles: warp 0 TMA producer (Q once, K/V double-buffered), warp 1 MMA issuer, warps 2-5
// softmax + epilogue.  hs=64 runs two CTAs per SM (256 TMEM columns each) so one CTA's
// exponentials overlap the other's MMAs; hs=128 double-buffers S instead.
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

template <int HD> struct ACfg {
    static constexpr int NSUB = HD / 64;                 // sub-tiles per Q / K / V tile
    static constexpr int TILE = NSUB * SUB;
    static constexpr int S_BUFS = HD == 64 ? 1 : 2;
    static constexpr int TMEM_COLS = HD == 64 ? 256 : 512;
    static constexpr int O_COL = 128 * S_BUFS;
    // smem: Q | K[2] | V[2] | P (2 sub-tiles) | barriers
    // + barriers (256 B) + alignment slack (768 B: lets two hs=64 CTAs share one SM's 227 KB)
    static constexpr int USED = TILE + 2 * TILE + 2 * TILE + 2 * SUB + 256;
    static constexpr int SMEM = USED + 768;
};

template <int HD>
__global__ void __launch_bounds__(NTHREADS, HD == 64 ? 2 : 1)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO,
                   float* __restrict__ lse_out, int T, int H, float scale_log2) {
    using C = ACfg<HD>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sQ = base;
    const uint32_t sK = sQ + C::TILE;
    const uint32_t sV = sK + 2 * C::TILE;
    const uint32_t sP = sV + 2 * C::TILE;
    const uint32_t bars = sP + 2 * SUB;
    const uint32_t q_full = bars;
    auto k_full = [&](int s) { return bars + 8u * (1 + s); };
    auto k_empty = [&](int s) { return bars + 8u * (3 + s); };
    auto v_full = [&](int s) { return bars + 8u * (5 + s); };
    auto v_empty = [&](int s) { return bars + 8u * (7 + s); };
    auto s_full = [&](int s) { return bars + 8u * (9 + s); };
    auto s_empty = [&](int s) { return bars + 8u * (11 + s); };
    const uint32_t p_full = bars + 8u * 13;
    const uint32_t pv_done = bars + 8u * 14;
    const uint32_t tmem_slot = bars + 8u * 15;
    uint8_t* gen = smem_raw + (base - ptx::smem_u32(smem_raw));
    if (base - ptx::smem_u32(smem_raw) > 768u) __trap();   // dynamic smem window less aligned than assumed

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = gridDim.x - 1 - blockIdx.x;     // heavy (late) query tiles first
    const int h = blockIdx.y, b = blockIdx.z;
    const int q0 = qt * BM;
    const int d_model = H * HD;
    const int ntiles = qt + 1;                     // causal: key tiles 0..qt

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmQKV);
        ptx::prefetch_tmap(&tmO);
    }
    if (warp == 1 && lane == 0) {
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(k_full(s), 1); ptx::mbar_init(k_empty(s), 1);
            ptx::mbar_init(v_full(s), 1); ptx::mbar_init(v_empty(s), 1);
            ptx::mbar_init(s_full(s), 1); ptx::mbar_init(s_empty(s), 4);
        }
        ptx::mbar_init(p_full, 4);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(gen + (tmem_slot - base));

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            ptx::mbar_arrive_expect_tx(q_full, C::TILE);
#pragma unroll
            for (int c = 0; c < C::NSUB; ++c) ptx::tma_load_3d(sQ + c * SUB, &tmQKV, q_full, h * HD + c * 64, q0, b);
            for (int j = 0; j < ntiles; ++j) {
                const int st = j & 1;
                const uint32_t ph = ((j >> 1) & 1) ^ 1;
                ptx::mbar_wait(k_empty(st), ph);
                ptx::mbar_arrive_expect_tx(k_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d(sK + st * C::TILE + c * SUB, &tmQKV, k_full(st), d_model + h * HD + c * 64, j * BN, b);
                ptx::mbar_wait(v_empty(st), ph);
                ptx::mbar_arrive_expect_tx(v_full(st), C::TILE);
#pragma unroll
                for (int c = 0; c < C::NSUB; ++c)
                    ptx::tma_load_3d(sV + st * C::TILE + c * SUB, &tmQKV, v_full(st), 2 * d_model + h * HD + c * 64, j * BN, b);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc_s = ptx::umma_idesc_bf16(128, 128, 0, 0);
            constexpr uint32_t idesc_pv = ptx::umma_idesc_bf16(128, HD, 0, 1);
            auto issue_s = [&](int j) {
                const int st = j & 1, sb = j % C::S_BUFS;
                ptx::mbar_wait(k_full(st), (j >> 1) & 1);
                ptx::mbar_wait(s_empty(sb), ((j / C::S_BUFS) & 1) ^ 1);
                ptx::tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < HD / 16; ++kk) {
                    const uint64_t ad = ptx::umma_desc(sQ + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                    const uint64_t bd = ptx::umma_desc(sK + st * C::TILE + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                    ptx::mma_bf16_ss(tmem + sb * 128, ad, bd, idesc_s, kk > 0 ? 1u : 0u);
                }
                ptx::mma_commit(s_full(sb));
                ptx::mma_commit(k_empty(st));
            };
            ptx::mbar_wait(q_full, 0);
            issue_s(0);
            for (int j = 0; j < ntiles; ++j) {
                const int st = j & 1;
                if (j + 1 < ntiles) issue_s(j + 1);
                ptx::mbar_wait(p_full, j & 1);
                ptx::mbar_wait(v_full(st), (j >> 1) & 1);
                ptx::tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < BN / 16; ++kk) {
                    // A = P [128 q x 128 kv], K-major, two 64-wide sub-tiles; B = V [kv x hs], MN-major
                    const uint64_t ad = ptx::umma_desc(sP + (kk >> 2) * SUB + (kk & 3) * 32, 16, 1024);
                    const uint64_t bd = ptx::umma_desc(sV + st * C::TILE + kk * 2048, SUB, 1024);
                    ptx::mma_bf16_ss(tmem + C::O_COL, ad, bd, idesc_pv, (j > 0 || kk > 0) ? 1u : 0u);
                }
                ptx::mma_commit(pv_done);
                ptx::mma_commit(v_empty(st));
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
            ptx::mbar_wait(s_full(sb), (j / C::S_BUFS) & 1);
            ptx::tc_fence_after();
            const uint32_t ts = tmem + lane_off + sb * 128;
            const bool diag = (j == qt);
            const int k0 = j * BN;
            // ---- pass 1: row maximum
            float mx = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(ts + c * 32, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    float v = __uint_as_float(r[i]);
                    if (diag && (k0 + c * 32 + i > qrow)) v = -INFINITY;
                    mx = fmaxf(mx, v);
                }
            }
            const float xmax = mx * scale_log2;
            const bool need = xmax > m_used + 8.f;          // also true on the first tile (m_used = -inf)
            const float m_new = need ? xmax : m_used;
            const float alpha = need ? exp2f(m_used - m_new) : 1.f;
            // O and the P buffer belong to PV_{j-1} until it has completed
            if (j > 0) {
                ptx::mbar_wait(pv_done, (j - 1) & 1);
                ptx::tc_fence_after();
                if (__any_sync(0xffffffffu, need)) {
#pragma unroll 1
                    for (int c = 0; c < HD / 32; ++c) {
                        uint32_t r[32];
                        ptx::tmem_ld_32x32(tmem + lane_off + C::O_COL + c * 32, r);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
                        ptx::tmem_st_32x32(tmem + lane_off + C::O_COL + c * 32, r);
                    }
                    ptx::tmem_st_wait();
                }
            }
            l *= alpha;
            m_used = m_new;
            // ---- pass 2: probabilities -> bf16 -> swizzled smem (K-major A operand of the PV MMA)
            float rs = 0.f;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t r[32];
                ptx::tmem_ld_32x32(ts + c * 32, r);
                ptx::tmem_ld_wait();
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float p0 = exp2f(__uint_as_float(r[i]) * scale_log2 - m_used);
                    float p1 = exp2f(__uint_as_float(r[i + 1]) * scale_log2 - m_used);
                    if (diag) {
                        if (k0 + c * 32 + i > qrow) p0 = 0.f;
                        if (k0 + c * 32 + i + 1 > qrow) p1 = 0.f;
                    }
                    rs += p0 + p1;
                    pk[i >> 1] = pack_bf16x2(p0, p1);
                }
                // 32 columns = 4 chunks of 16 B; column block c -> sub-tile c/2, chunks (c%2)*4 .. +3
                const uint32_t dst = sP + (c >> 1) * SUB + row * 128;
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    const uint32_t chunk = (uint32_t)((c & 1) * 4 + ch);
                    const uint32_t addr = dst + ((chunk ^ ((uint32_t)row & 7u)) << 4);
                    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(pk[4 * ch]), "r"(pk[4 * ch + 1]),
                                 "r"(pk[4 * ch + 2]), "r"(pk[4 * ch + 3]) : "memory");
                }
            }
            l += rs;
            ptx::tc_fence_before();
            ptx::fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                ptx::mbar_arrive(p_full);
                ptx::mbar_arrive(s_empty(sb));
            }
        }
        // ---- epilogue: O / l -> bf16 -> smem (Q's tile is free now) -> TMA store; LSE -> global
        ptx::mbar_wait(pv_done, (ntiles - 1) & 1);
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
int launch_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, float scale, cudaStream_t st) {
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
    dim3 grid((T + BM - 1) / BM, H, B);
    kern<<<grid, NTHREADS, C::SMEM, st>>>(tq, to, lse, T, H, scale * 1.4426950408889634f);
    return 0;
}

}  // namespace

extern "C" int b200_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale,
                                cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_fwd_tc: head size must be 64 or 128 (got %d)", hs);
    B200_REQUIRE((H * hs) % 8 == 0, "attn_fwd_tc: model width must be a multiple of 8");
    int rc = hs == 64 ? launch_fwd<64>(qkv, out, lse, B, T, H, scale, st) : launch_fwd<128>(qkv, out, lse, B, T, H, scale, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("attn_fwd_tc");
}
