// Causal flash attention, forward and backward, on the packed QKV activation.
//
// Replaces, per layer of the reference, H x (3 small GEMMs are already packed elsewhere) the
// q@k^T bmm, the tril compare + masked_fill + fp32 softmax over a materialised (B,T,T) matrix,
// the attn@v bmm and the torch.cat of heads (reference src/models/attention.py:45-58,95), and
// their autograd.  Nothing of size T x T ever reaches HBM; the causal mask is implicit
// (upper-triangular kv tiles are skipped), the log-sum-exp of every row is saved for backward.
//
// Layout: qkv [B*T, 3*H*hs] bf16 = [q | k | v] thirds with head h at columns h*hs (exactly what the
// packed QKV GEMM writes); out [B*T, H*hs] head-major = the reference's concatenation of heads.
//
// This file is the warp-level tensor-core implementation (mma.sync m16n8k16 bf16, ldmatrix,
// cp.async double buffering).  Scores and probabilities are kept in fp32 registers (the
// reference rounds scores to bf16 before the softmax under autocast - this is strictly more
// accurate), P is rounded to bf16 for the PV product exactly like the reference.
#include "b200llm.h"
#include "common.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t saddr, const void* gptr, bool pred) {
    const int sz = pred ? 16 : 0;  // src-size 0 => zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(saddr), "l"(gptr), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3, uint32_t addr) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Swizzled [rows][HD] bf16 tile: 16-byte chunk c of row r lives at chunk (c ^ (r & 7)).
template <int HD>
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
    return base + row * (HD * 2) + ((chunk ^ (row & 7)) << 4);
}

// Cooperative async load of `rows` x HD bf16 from global (row pitch ld elements) into a swizzled tile.
// Rows >= valid_rows are zero-filled.
template <int HD, int THREADS>
__device__ __forceinline__ void load_tile_async(uint32_t sbase, const bf16* g, long long ld, int rows, int valid_rows,
                                                int tid) {
    constexpr int CH = HD / 8;  // chunks per row
    for (int i = tid; i < rows * CH; i += THREADS) {
        const int r = i / CH, c = i % CH;
        const bool ok = r < valid_rows;
        cp_async16(tile_addr<HD>(sbase, r, c), g + (ok ? (size_t)r * ld + c * 8 : 0), ok);
    }
}

// ldmatrix address helpers (see the fragment-layout notes in DESIGN.md)
//  A operand, storage [m][k]:            rows m0.., chunk kc
template <int HD> __device__ __forceinline__ uint32_t addr_A(uint32_t base, int m0, int kchunk, int lane) {
    const int j = lane >> 3, i = lane & 7;
    return tile_addr<HD>(base, m0 + (j & 1) * 8 + i, kchunk + (j >> 1));
}
//  B operand, storage [n][k] (non-transposed), two n-blocks (n0..n0+15)
template <int HD> __device__ __forceinline__ uint32_t addr_B_nk(uint32_t base, int n0, int kchunk, int lane) {
    const int j = lane >> 3, i = lane & 7;
    return tile_addr<HD>(base, n0 + (j >> 1) * 8 + i, kchunk + (j & 1));
}
//  B operand, storage [k][n] (use .trans), two n-blocks (nchunk, nchunk+1)
template <int HD> __device__ __forceinline__ uint32_t addr_B_kn(uint32_t base, int k0, int nchunk, int lane) {
    const int j = lane >> 3, i = lane & 7;
    return tile_addr<HD>(base, k0 + (j & 1) * 8 + i, nchunk + (j >> 1));
}
//  A operand, storage [k][m] (use .trans): rows k0.., m chunk
template <int HD> __device__ __forceinline__ uint32_t addr_A_km(uint32_t base, int k0, int mchunk, int lane) {
    const int j = lane >> 3, i = lane & 7;
    return tile_addr<HD>(base, k0 + (j >> 1) * 8 + i, mchunk + (j & 1));
}

// =============================================================================================
// Forward
// =============================================================================================
constexpr int FWD_BM = 128, FWD_BN = 64, FWD_THREADS = 256;

template <int HD>
__global__ void __launch_bounds__(FWD_THREADS) attn_fwd_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ out,
                                                              float* __restrict__ lse_out, int T, int H,
                                                              float scale_log2) {
    pdl_sync();
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sQ = smem_u32(smem);
    const uint32_t sK = sQ + FWD_BM * HD * 2;            // 2 stages
    const uint32_t sV = sK + 2 * FWD_BN * HD * 2;        // 2 stages
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t4 = lane & 3;
    const int qblk = gridDim.x - 1 - blockIdx.x;         // heavy (late) query blocks first
    const int h = blockIdx.y, b = blockIdx.z;
    const int q0 = qblk * FWD_BM;
    const long long ld = 3LL * H * HD;
    const bf16* qg = qkv + ((size_t)b * T + q0) * ld + h * HD;
    const bf16* kg = qkv + (size_t)b * T * ld + (size_t)H * HD + h * HD;
    const bf16* vg = kg + (size_t)H * HD;

    const int q_rows = min(FWD_BM, T - q0);
    const int kv_end = min(T, q0 + FWD_BM);              // causal: keys < q0 + BM
    const int nkv = (kv_end + FWD_BN - 1) / FWD_BN;

    load_tile_async<HD, FWD_THREADS>(sQ, qg, ld, FWD_BM, q_rows, tid);
    load_tile_async<HD, FWD_THREADS>(sK, kg, ld, FWD_BN, min(FWD_BN, T), tid);
    load_tile_async<HD, FWD_THREADS>(sV, vg, ld, FWD_BN, min(FWD_BN, T), tid);
    cp_async_commit();

    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    uint32_t qf[HD / 16][4];
    const int wrow0 = warp * 16;                         // this warp's query rows inside the tile

    for (int j = 0; j < nkv; ++j) {
        const int st = j & 1;
        if (j + 1 < nkv) {
            const int k0n = (j + 1) * FWD_BN;
            load_tile_async<HD, FWD_THREADS>(sK + (st ^ 1) * FWD_BN * HD * 2, kg + (size_t)k0n * ld, ld, FWD_BN,
                                             min(FWD_BN, T - k0n), tid);
            load_tile_async<HD, FWD_THREADS>(sV + (st ^ 1) * FWD_BN * HD * 2, vg + (size_t)k0n * ld, ld, FWD_BN,
                                             min(FWD_BN, T - k0n), tid);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (j == 0) {
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk)
                ldsm_x4(qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3], addr_A<HD>(sQ, wrow0, kk * 2, lane));
        }
        const int k0 = j * FWD_BN;
        const uint32_t sKs = sK + st * FWD_BN * HD * 2, sVs = sV + st * FWD_BN * HD * 2;
        // warp-uniform skip: this warp's rows are all before this kv tile
        const bool warp_active = (q0 + wrow0 + 15) >= k0;
        if (warp_active) {
            float s[FWD_BN / 8][4];
#pragma unroll
            for (int i = 0; i < FWD_BN / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
            for (int kk = 0; kk < HD / 16; ++kk) {
#pragma unroll
                for (int nb = 0; nb < FWD_BN / 16; ++nb) {
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4(b0, b1, b2, b3, addr_B_nk<HD>(sKs, nb * 16, kk * 2, lane));
                    mma16816(s[2 * nb], qf[kk], b0, b1);
                    mma16816(s[2 * nb + 1], qf[kk], b2, b3);
                }
            }
            // causal / tail mask (only tiles that touch the diagonal or the sequence end)
            const int qr0 = q0 + wrow0 + g, qr1 = qr0 + 8;
            if (k0 + FWD_BN - 1 > q0 + wrow0 || k0 + FWD_BN > T) {
#pragma unroll
                for (int i = 0; i < FWD_BN / 8; ++i) {
                    const int c0 = k0 + i * 8 + 2 * t4, c1 = c0 + 1;
                    if (c0 > qr0 || c0 >= T) s[i][0] = -INFINITY;
                    if (c1 > qr0 || c1 >= T) s[i][1] = -INFINITY;
                    if (c0 > qr1 || c0 >= T) s[i][2] = -INFINITY;
                    if (c1 > qr1 || c1 >= T) s[i][3] = -INFINITY;
                }
            }
            // online softmax (rows g and g+8 of the warp tile; 4 lanes share a row)
            float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
            for (int i = 0; i < FWD_BN / 8; ++i) {
                mx[0] = fmaxf(mx[0], fmaxf(s[i][0], s[i][1]));
                mx[1] = fmaxf(mx[1], fmaxf(s[i][2], s[i][3]));
            }
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
                mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
            }
            float corr[2], msc[2];
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const float mnew = fmaxf(m_run[r], mx[r]);
                msc[r] = mnew == -INFINITY ? 0.f : mnew * scale_log2;
                corr[r] = exp2f(m_run[r] == -INFINITY ? -INFINITY : m_run[r] * scale_log2 - msc[r]);
                m_run[r] = mnew;
            }
            float rs[2] = {0.f, 0.f};
            uint32_t pf[FWD_BN / 16][4];
#pragma unroll
            for (int i = 0; i < FWD_BN / 8; ++i) {
                const float p0 = exp2f(s[i][0] * scale_log2 - msc[0]);
                const float p1 = exp2f(s[i][1] * scale_log2 - msc[0]);
                const float p2 = exp2f(s[i][2] * scale_log2 - msc[1]);
                const float p3 = exp2f(s[i][3] * scale_log2 - msc[1]);
                rs[0] += p0 + p1;
                rs[1] += p2 + p3;
                pf[i >> 1][(i & 1) * 2] = pack_bf16x2(p0, p1);
                pf[i >> 1][(i & 1) * 2 + 1] = pack_bf16x2(p2, p3);
            }
            l_run[0] = l_run[0] * corr[0] + rs[0];
            l_run[1] = l_run[1] * corr[1] + rs[1];
#pragma unroll
            for (int i = 0; i < HD / 8; ++i) {
                o[i][0] *= corr[0]; o[i][1] *= corr[0];
                o[i][2] *= corr[1]; o[i][3] *= corr[1];
            }
            // O += P V
#pragma unroll
            for (int kk = 0; kk < FWD_BN / 16; ++kk) {
#pragma unroll
                for (int nb = 0; nb < HD / 16; ++nb) {
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4_t(b0, b1, b2, b3, addr_B_kn<HD>(sVs, kk * 16, nb * 2, lane));
                    mma16816(o[2 * nb], pf[kk], b0, b1);
                    mma16816(o[2 * nb + 1], pf[kk], b2, b3);
                }
            }
        }
        __syncthreads();  // everyone done with stage st before it is refilled
    }
    // finalize: normalise, write O and LSE
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
        l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
    const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f, inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
    const int r0 = q0 + wrow0 + g, r1 = r0 + 8;
    bf16* og = out + ((size_t)b * T) * (H * HD) + h * HD;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
        const int c = i * 8 + 2 * t4;
        if (r0 < T) *reinterpret_cast<uint32_t*>(og + (size_t)r0 * (H * HD) + c) = pack_bf16x2(o[i][0] * inv0, o[i][1] * inv0);
        if (r1 < T) *reinterpret_cast<uint32_t*>(og + (size_t)r1 * (H * HD) + c) = pack_bf16x2(o[i][2] * inv1, o[i][3] * inv1);
    }
    if (t4 == 0) {
        const float ln2 = 0.6931471805599453f;
        float* lg = lse_out + ((size_t)b * H + h) * T;
        if (r0 < T) lg[r0] = (m_run[0] * scale_log2 + log2f(l_run[0])) * ln2;
        if (r1 < T) lg[r1] = (m_run[1] * scale_log2 + log2f(l_run[1])) * ln2;
    }
}

// =============================================================================================
// Backward
// =============================================================================================
// delta[b,h,q] = sum_c dO[q,c] * O[q,c]
template <int HD>
__global__ void __launch_bounds__(256) attn_bwd_delta_kernel(const bf16* __restrict__ out, const bf16* __restrict__ dout,
                                                            float* __restrict__ delta, int B, int T, int H) {
    pdl_sync();
    // one warp per (row m, head h); lane handles HD/32 elements
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int total = B * T * H;
    if (gw >= total) return;
    const int h = gw % H;
    const int m = gw / H;  // b*T + t
    const size_t off = (size_t)m * (H * HD) + h * HD + lane * (HD / 32);
    float s = 0.f;
    if (HD == 64) {
        float2 a = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(out + off));
        float2 d = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(dout + off));
        s = a.x * d.x + a.y * d.y;
    } else {
        uint2 ua = *reinterpret_cast<const uint2*>(out + off), ud = *reinterpret_cast<const uint2*>(dout + off);
        float2 a0 = unpack_bf16x2(ua.x), a1 = unpack_bf16x2(ua.y), d0 = unpack_bf16x2(ud.x), d1 = unpack_bf16x2(ud.y);
        s = a0.x * d0.x + a0.y * d0.y + a1.x * d1.x + a1.y * d1.y;
    }
    s = warp_sum(s);
    if (lane == 0) {
        const int b = m / T, t = m % T;
        delta[((size_t)b * H + h) * T + t] = s;
    }
}

constexpr int BWD_BN = 64, BWD_BM = 64, BWD_THREADS = 128;

template <int HD>
__global__ void __launch_bounds__(BWD_THREADS) attn_bwd_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ dout,
                                                              const float* __restrict__ lse,
                                                              const float* __restrict__ delta,
                                                              float* __restrict__ dq_acc, bf16* __restrict__ dqkv, int T,
                                                              int H, float scale, float scale_log2) {
    pdl_sync();
    extern __shared__ __align__(1024) uint8_t smem[];
    constexpr int TILE = 64 * HD * 2;
    const uint32_t sK = smem_u32(smem);
    const uint32_t sV = sK + TILE;
    const uint32_t sQ = sV + TILE;        // 2 stages
    const uint32_t sdO = sQ + 2 * TILE;   // 2 stages
    const uint32_t sdS = sdO + 2 * TILE;  // [64 kv][64 q] bf16, swizzled as a HD=64 tile
    float* s_lse = reinterpret_cast<float*>(smem + 6 * TILE + 64 * 64 * 2);  // 2 stages x 64
    float* s_delta = s_lse + 128;                                            // 2 stages x 64

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t4 = lane & 3;
    const int kvblk = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int k0 = kvblk * BWD_BN;
    const long long ld = 3LL * H * HD;
    const long long ldo = (long long)H * HD;
    const bf16* qg = qkv + (size_t)b * T * ld + h * HD;
    const bf16* kg = qg + (size_t)H * HD;
    const bf16* vg = kg + (size_t)H * HD;
    const bf16* dog = dout + (size_t)b * T * ldo + h * HD;
    const float* lse_g = lse + ((size_t)b * H + h) * T;
    const float* delta_g = delta + ((size_t)b * H + h) * T;

    const int kv_rows = min(BWD_BN, T - k0);
    const int qb0 = k0 / BWD_BM;                       // first q block that can see these keys
    const int nqb = (T + BWD_BM - 1) / BWD_BM;

    auto load_q_stage = [&](int qb, int st) {
        const int q0 = qb * BWD_BM;
        const int rows = min(BWD_BM, T - q0);
        load_tile_async<HD, BWD_THREADS>(sQ + st * TILE, qg + (size_t)q0 * ld, ld, BWD_BM, rows, tid);
        load_tile_async<HD, BWD_THREADS>(sdO + st * TILE, dog + (size_t)q0 * ldo, ldo, BWD_BM, rows, tid);
        if (tid < 64) {
            const int q = q0 + tid;
            s_lse[st * 64 + tid] = q < T ? lse_g[q] * 1.4426950408889634f : INFINITY;  // log2 units
            s_delta[st * 64 + tid] = q < T ? delta_g[q] : 0.f;
        }
    };

    load_tile_async<HD, BWD_THREADS>(sK, kg + (size_t)k0 * ld, ld, BWD_BN, kv_rows, tid);
    load_tile_async<HD, BWD_THREADS>(sV, vg + (size_t)k0 * ld, ld, BWD_BN, kv_rows, tid);
    load_q_stage(qb0, 0);
    cp_async_commit();

    float dk[HD / 8][4], dv[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
        dk[i][0] = dk[i][1] = dk[i][2] = dk[i][3] = 0.f;
        dv[i][0] = dv[i][1] = dv[i][2] = dv[i][3] = 0.f;
    }
    const int wkv0 = warp * 16;  // this warp's kv rows inside the tile

    for (int qb = qb0; qb < nqb; ++qb) {
        const int st = (qb - qb0) & 1;
        if (qb + 1 < nqb) {
            load_q_stage(qb + 1, st ^ 1);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const int q0 = qb * BWD_BM;
        const uint32_t sQs = sQ + st * TILE, sdOs = sdO + st * TILE;
        const float* lse_s = s_lse + st * 64;
        const float* del_s = s_delta + st * 64;

        // S^T = K Q^T  and  dP^T = V dO^T   (rows: this warp's 16 kv, cols: 64 q)
        float sT[BWD_BM / 8][4], dpT[BWD_BM / 8][4];
#pragma unroll
        for (int i = 0; i < BWD_BM / 8; ++i) {
            sT[i][0] = sT[i][1] = sT[i][2] = sT[i][3] = 0.f;
            dpT[i][0] = dpT[i][1] = dpT[i][2] = dpT[i][3] = 0.f;
        }
#pragma unroll
        for (int kk = 0; kk < HD / 16; ++kk) {
            uint32_t ka[4], va[4];
            ldsm_x4(ka[0], ka[1], ka[2], ka[3], addr_A<HD>(sK, wkv0, kk * 2, lane));
            ldsm_x4(va[0], va[1], va[2], va[3], addr_A<HD>(sV, wkv0, kk * 2, lane));
#pragma unroll
            for (int nb = 0; nb < BWD_BM / 16; ++nb) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4(b0, b1, b2, b3, addr_B_nk<HD>(sQs, nb * 16, kk * 2, lane));
                mma16816(sT[2 * nb], ka, b0, b1);
                mma16816(sT[2 * nb + 1], ka, b2, b3);
                ldsm_x4(b0, b1, b2, b3, addr_B_nk<HD>(sdOs, nb * 16, kk * 2, lane));
                mma16816(dpT[2 * nb], va, b0, b1);
                mma16816(dpT[2 * nb + 1], va, b2, b3);
            }
        }
        // P^T = exp2(S^T * scale_log2 - lse2[q]) with the causal mask (q >= kv) and tails;
        // dS^T = P^T * (dP^T - delta[q]) * scale
        const int kvr0 = k0 + wkv0 + g, kvr1 = kvr0 + 8;
        const bool diag = (q0 < k0 + BWD_BN);
        uint32_t pT[BWD_BM / 16][4], dsT[BWD_BM / 16][4];
#pragma unroll
        for (int i = 0; i < BWD_BM / 8; ++i) {
            const int qc0 = i * 8 + 2 * t4, qc1 = qc0 + 1;
            const float l0 = lse_s[qc0], l1 = lse_s[qc1];
            const float d0 = del_s[qc0], d1 = del_s[qc1];
            float p00 = exp2f(sT[i][0] * scale_log2 - l0), p01 = exp2f(sT[i][1] * scale_log2 - l1);
            float p10 = exp2f(sT[i][2] * scale_log2 - l0), p11 = exp2f(sT[i][3] * scale_log2 - l1);
            if (diag || kvr1 >= T) {
                const int qa0 = q0 + qc0, qa1 = q0 + qc1;
                if (qa0 < kvr0 || kvr0 >= T) p00 = 0.f;
                if (qa1 < kvr0 || kvr0 >= T) p01 = 0.f;
                if (qa0 < kvr1 || kvr1 >= T) p10 = 0.f;
                if (qa1 < kvr1 || kvr1 >= T) p11 = 0.f;
            }
            const float s00 = p00 * (dpT[i][0] - d0) * scale, s01 = p01 * (dpT[i][1] - d1) * scale;
            const float s10 = p10 * (dpT[i][2] - d0) * scale, s11 = p11 * (dpT[i][3] - d1) * scale;
            pT[i >> 1][(i & 1) * 2] = pack_bf16x2(p00, p01);
            pT[i >> 1][(i & 1) * 2 + 1] = pack_bf16x2(p10, p11);
            const uint32_t ds0 = pack_bf16x2(s00, s01), ds1 = pack_bf16x2(s10, s11);
            dsT[i >> 1][(i & 1) * 2] = ds0;
            dsT[i >> 1][(i & 1) * 2 + 1] = ds1;
            // stash dS^T [kv][q] in smem for the dQ product (row = kv, 16-byte chunk = q/8)
            const int r0 = wkv0 + g, r1 = r0 + 8;
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_addr<64>(sdS, r0, i) + t4 * 4), "r"(ds0) : "memory");
            asm volatile("st.shared.b32 [%0], %1;" ::"r"(tile_addr<64>(sdS, r1, i) + t4 * 4), "r"(ds1) : "memory");
        }
        // dV += P^T dO ; dK += dS^T Q      (k = q index, B operand stored [q][hd] -> .trans)
#pragma unroll
        for (int kk = 0; kk < BWD_BM / 16; ++kk) {
#pragma unroll
            for (int nb = 0; nb < HD / 16; ++nb) {
                uint32_t b0, b1, b2, b3;
                ldsm_x4_t(b0, b1, b2, b3, addr_B_kn<HD>(sdOs, kk * 16, nb * 2, lane));
                mma16816(dv[2 * nb], pT[kk], b0, b1);
                mma16816(dv[2 * nb + 1], pT[kk], b2, b3);
                ldsm_x4_t(b0, b1, b2, b3, addr_B_kn<HD>(sQs, kk * 16, nb * 2, lane));
                mma16816(dk[2 * nb], dsT[kk], b0, b1);
                mma16816(dk[2 * nb + 1], dsT[kk], b2, b3);
            }
        }
        __syncthreads();  // dS^T tile complete
        // dQ[q, :] += dS[q, kv] K[kv, :]   (warp handles 16 q rows; A = dS from sdS [kv][q] via .trans)
        {
            float dq[HD / 8][4];
#pragma unroll
            for (int i = 0; i < HD / 8; ++i) dq[i][0] = dq[i][1] = dq[i][2] = dq[i][3] = 0.f;
            const int wq0 = warp * 16;
#pragma unroll
            for (int kk = 0; kk < BWD_BN / 16; ++kk) {
                uint32_t a[4];
                ldsm_x4_t(a[0], a[1], a[2], a[3], addr_A_km<64>(sdS, kk * 16, wq0 / 8, lane));
#pragma unroll
                for (int nb = 0; nb < HD / 16; ++nb) {
                    uint32_t b0, b1, b2, b3;
                    ldsm_x4_t(b0, b1, b2, b3, addr_B_kn<HD>(sK, kk * 16, nb * 2, lane));
                    mma16816(dq[2 * nb], a, b0, b1);
                    mma16816(dq[2 * nb + 1], a, b2, b3);
                }
            }
            const int qr0 = q0 + wq0 + g, qr1 = qr0 + 8;
            float* dqg = dq_acc + ((size_t)b * T) * ldo + h * HD;
#pragma unroll
            for (int i = 0; i < HD / 8; ++i) {
                const int c = i * 8 + 2 * t4;
                if (qr0 < T)
                    asm volatile("red.global.add.v2.f32 [%0], {%1,%2};" ::"l"(dqg + (size_t)qr0 * ldo + c), "f"(dq[i][0]),
                                 "f"(dq[i][1]) : "memory");
                if (qr1 < T)
                    asm volatile("red.global.add.v2.f32 [%0], {%1,%2};" ::"l"(dqg + (size_t)qr1 * ldo + c), "f"(dq[i][2]),
                                 "f"(dq[i][3]) : "memory");
            }
        }
        __syncthreads();  // stage st and sdS free for reuse
    }
    // write dK, dV (bf16) into the k / v thirds of dqkv
    bf16* dkg = dqkv + (size_t)b * T * ld + (size_t)H * HD + h * HD;
    bf16* dvg = dkg + (size_t)H * HD;
    const int r0 = k0 + wkv0 + g, r1 = r0 + 8;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
        const int c = i * 8 + 2 * t4;
        if (r0 < T) {
            *reinterpret_cast<uint32_t*>(dkg + (size_t)r0 * ld + c) = pack_bf16x2(dk[i][0], dk[i][1]);
            *reinterpret_cast<uint32_t*>(dvg + (size_t)r0 * ld + c) = pack_bf16x2(dv[i][0], dv[i][1]);
        }
        if (r1 < T) {
            *reinterpret_cast<uint32_t*>(dkg + (size_t)r1 * ld + c) = pack_bf16x2(dk[i][2], dk[i][3]);
            *reinterpret_cast<uint32_t*>(dvg + (size_t)r1 * ld + c) = pack_bf16x2(dv[i][2], dv[i][3]);
        }
    }
}

// dq_acc fp32 [M, H*hs] -> q third of dqkv (bf16), and re-zero dq_acc for the next call
__global__ void __launch_bounds__(256) attn_bwd_dq_store_kernel(float* __restrict__ dq_acc, bf16* __restrict__ dqkv,
                                                               long long M, int HHD) {
    pdl_sync();
    const long long n4 = M * HHD / 4;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 v = reinterpret_cast<float4*>(dq_acc)[i];
        reinterpret_cast<float4*>(dq_acc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        const long long e = i * 4;
        const long long m = e / HHD;
        const int c = (int)(e % HHD);
        *reinterpret_cast<uint2*>(dqkv + m * 3 * HHD + c) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    }
}

template <int HD>
int attn_fwd_launch(const bf16* qkv, bf16* out, float* lse, int B, int T, int H, float scale, cudaStream_t st) {
    constexpr int SMEM = FWD_BM * HD * 2 + 4 * FWD_BN * HD * 2;
    static bool cfg = false;
    if (!cfg) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        cfg = true;
    }
    dim3 grid((T + FWD_BM - 1) / FWD_BM, H, B);
    B200_CHECK_CUDA(b200_launch(attn_fwd_kernel<HD>, dim3(grid), dim3(FWD_THREADS), SMEM, st, qkv, out, lse, T, H, scale * 1.4426950408889634f));
    return 0;
}

template <int HD>
int attn_bwd_launch(const bf16* qkv, const bf16* out, const bf16* dout, const float* lse, float* delta, float* dq_acc,
                    bf16* dqkv, int B, int T, int H, float scale, cudaStream_t st) {
    constexpr int TILE = 64 * HD * 2;
    constexpr int SMEM = 6 * TILE + 64 * 64 * 2 + 4 * 64 * 4;
    static bool cfg = false;
    if (!cfg) {
        B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
        cfg = true;
    }
    const int total_warps = B * T * H;
    B200_CHECK_CUDA(b200_launch(attn_bwd_delta_kernel<HD>, dim3((total_warps + 7) / 8), dim3(256), 0, st, out, dout, delta, B, T, H));
    dim3 grid((T + BWD_BN - 1) / BWD_BN, H, B);
    B200_CHECK_CUDA(b200_launch(attn_bwd_kernel<HD>, dim3(grid), dim3(BWD_THREADS), SMEM, st, qkv, dout, lse, delta, dq_acc, dqkv, T, H, scale,
                                                         scale * 1.4426950408889634f));
    const long long M = (long long)B * T;
    B200_CHECK_CUDA(b200_launch(attn_bwd_dq_store_kernel, dim3(b200_num_sms() * 8), dim3(256), 0, st, dq_acc, dqkv, M, H * HD));
    return 0;
}

}  // namespace

extern "C" int b200_attn_fwd(const void* qkv, void* out, float* lse, int B, int T, int H, int hs, float scale,
                             cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_fwd: head size must be 64 or 128 (got %d)", hs);
    int rc = hs == 64 ? attn_fwd_launch<64>((const bf16*)qkv, (bf16*)out, lse, B, T, H, scale, st)
                      : attn_fwd_launch<128>((const bf16*)qkv, (bf16*)out, lse, B, T, H, scale, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("attn_fwd");
}

// dq_acc must be all zeros on entry (it is left zeroed on exit).
extern "C" int b200_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, float* delta,
                             float* dq_acc, void* dqkv, int B, int T, int H, int hs, float scale, cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_bwd: head size must be 64 or 128 (got %d)", hs);
    int rc = hs == 64 ? attn_bwd_launch<64>((const bf16*)qkv, (const bf16*)out, (const bf16*)dout, lse, delta, dq_acc,
                                            (bf16*)dqkv, B, T, H, scale, st)
                      : attn_bwd_launch<128>((const bf16*)qkv, (const bf16*)out, (const bf16*)dout, lse, delta, dq_acc,
                                             (bf16*)dqkv, B, T, H, scale, st);
    if (rc) return rc;
    B200_RETURN_LAUNCH_STATUS("attn_bwd");
}
