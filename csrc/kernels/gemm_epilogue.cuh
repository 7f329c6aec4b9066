// Shared epilogue of the tcgen05 GEMM kernels (1-CTA gemm.cu and 2-CTA gemm2.cu): one 128-row x BN-column
// accumulator tile, TMEM -> registers -> (bf16: bias / ReLU / ReLU-mask / softmax statistics -> swizzled smem ->
// TMA store) or (fp32: store / accumulate / atomic add).  Called by NG groups of 4 epilogue warps (128 threads
// per group, thread <-> accumulator row); group g takes the 64-column chunks g, g + NG, ... of the tile, so with
// NG = 2 the epilogue of a tile takes half as long (one warp per scheduler cannot hide its own instruction
// latencies: with 4 warps the bias/ReLU/sign-bit and CE-statistics epilogues of the K = 768 GEMMs were longer than
// their main loops).  `arrive_tempty()` hands the TMEM stage back to the MMA issuer (once per warp).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace gemm {

constexpr int EPI_BUF_BYTES = 128 * 64 * 2;  // one 128 x 64 bf16 staging tile

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
    // ReLU sign bits, one 64-bit word per (row, 64-column chunk): [M][ceil(N/64)].  The forward GEMM that applies the
    // ReLU writes them (bit j = bf16 output column 64*w + j is > 0); the matching dgrad GEMM reads them back instead of
    // the activations: 32 B per row and 256-column tile instead of 512 B, and no uncoalesced 16-byte loads (a dgrad with
    // the activation mask ran 2.5x slower than without: 32 LSU wavefronts per mask load instruction).
    unsigned long long* relu_bits_out;
    const unsigned long long* relu_bits;
    // fp32 epilogue
    float* out_f32;
    long long ldc;
    int f32_mode;           // 0 store, 1 accumulate (+=), 2 atomic add
    int n_shift;            // rotate the n-tile order (all-gather fused GEMM: start with the locally owned rows)
    float* colsum;          // [N] fp32 or null: += column sums of the bf16 outputs (bias gradient of the producing Linear),
                            // summed from the staged smem tile, one fp32 atomic per column and 128-row tile
};

// All-gather fused into the GEMM (sharded data parallelism): the B operand (a weight matrix) lives in a flat
// bf16 buffer whose layer range [layer_start, layer_end) is split in `world` contiguous chunks, chunk c owned
// (kept up to date) by rank c.  flags[c] >= *epoch means chunk c of this layer is valid in the LOCAL buffer.
// All-gather jobs a GEMM launch can carry (sharded data parallelism, csrc/comm/zero.cu + gemm2.cu).  A job copies one
// layer's bf16 weights, chunk by chunk, from the ranks that own them (peer-mapped shard memory, generic loads over
// NVLink) into a local window; the GEMM's own B operand is always a TMA tensor map over LOCAL memory (the window).
struct AgJob {
    const bf16* src[8];       // src[r]: rank r's chunk of the layer (own rank: the local shard)
    bf16* dst;                // local window that receives the whole layer
    long long chunk;          // elements per rank
    long long total;          // elements of the layer (the last chunk may be short)
    int* flags;               // [world] raised to *gen once chunk c sits in the window
    unsigned int* counters;   // [world] zero-initialised arrival counters
};
struct AgParams {
    const int* wait_flags;    // flags of the window that holds THIS GEMM's B operand (null: B is complete, no waiting)
    const int* gen;           // device-side gather generation (incremented once per forward / backward pass)
    long long chunk;          // elements per rank of the B operand's layer
    long long w_offset;       // first element of the B operand inside its window
    int world, rank;
    int njobs;                // gather jobs carried by this launch (0..2); the first may be the B operand's own layer
    int* err;                 // raised when a wait times out
    AgJob job[2];
};

// tfull_bar: local mbarrier (shared::cta address) the MMA issuer commits to; taddr: TMEM address of this
// thread's lane quarter at the accumulator stage's first column; m0: first global row of this CTA's 128 rows.
// bias_pf: this thread's share of the CURRENT tile's bias slice, loaded one tile ahead (n0_next: first column of the
// next tile of this CTA, < 0 if none) so the global-load latency never sits in front of a tile's epilogue.
template <int BN, bool OUT_F32, int NG, bool COLSUM = false, class ArriveFn>
__device__ __forceinline__ void epilogue_tile(const Params& p, const CUtensorMap* tmC, uint32_t tfull_bar, uint32_t acc_phase,
                                              uint32_t taddr, int m0, int n0, int n_blk, int row, int lane,
                                              uint32_t epi_base, int& buf, bool store_thread, float* s_bias, int et,
                                              float (&bias_pf)[(BN + 127) / 128], int n0_next, int grp,
                                              ArriveFn arrive_tempty) {
    // et: thread index among ALL epilogue threads (0 .. 128*NG-1); grp = et / 128; buf: staging buffer (NG = 1: the two
    // buffers alternate; NG = 2: group g owns buffer g)
    constexpr int NCH = BN / 64;
    constexpr int NTE = 128 * NG;                        // epilogue threads
    constexpr int NPF = (BN + 127) / 128;
    const int m = m0 + row;
    const uint32_t bar_id = 1 + grp;
    if (!OUT_F32 && p.bias != nullptr) {
        // stage this tile's bias slice in smem while the MMAs are still running
        asm volatile("bar.sync 3, %0;" ::"n"(NTE) : "memory");          // previous tile's readers are done
#pragma unroll
        for (int i = 0; i < NPF; ++i) {
            const int j = et + NTE * i;
            if (i * NTE < BN) {
                if (j < BN) s_bias[j] = bias_pf[i];
                bias_pf[i] = (n0_next >= 0 && j < BN && n0_next + j < p.N) ? __ldg(p.bias + n0_next + j) : 0.f;
            }
        }
        asm volatile("bar.sync 3, %0;" ::"n"(NTE) : "memory");
    }
    // ReLU sign bits of this row for the whole tile: loaded while the MMAs are still running
    const int ld_bits = (p.N + 63) >> 6;
    unsigned long long mbits[NCH];
    if (!OUT_F32 && p.relu_bits != nullptr) {
#pragma unroll
        for (int c = 0; c < NCH; ++c)
            mbits[c] = (m < p.M && (n0 >> 6) + c < ld_bits) ? __ldg(p.relu_bits + (size_t)m * ld_bits + (n0 >> 6) + c) : 0ull;
    }
    ptx::mbar_wait(tfull_bar, acc_phase);
    ptx::tc_fence_after();
    float run_max = -INFINITY, run_sum = 0.f;
#pragma unroll 1
    for (int c = grp; c < NCH; c += NG) {
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
        if (c + NG >= NCH) {
            // all TMEM reads of this warp for this accumulator stage are done: hand it back to the MMA warp
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) arrive_tempty();
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
            if (p.relu_bits != nullptr) {
                unsigned long long mb = 0ull;          // register select: a dynamic index would put mbits[] in local memory
#pragma unroll
                for (int i = 0; i < NCH; ++i)
                    if (i == c) mb = mbits[i];
                const uint32_t lo = (uint32_t)mb, hi = (uint32_t)(mb >> 32);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if (!(lo & (1u << j))) r[j] = 0u;
                    if (!(hi & (1u << j))) r[32 + j] = 0u;
                }
            }
            // ---- pack to bf16
            uint32_t pk[32];
#pragma unroll
            for (int j = 0; j < 32; ++j)
                pk[j] = pack_bf16x2(__uint_as_float(r[2 * j]), __uint_as_float(r[2 * j + 1]));
            if (p.relu_bits_out != nullptr) {
                // sign bits of the bf16-ROUNDED outputs (what the activation tensor holds); after the ReLU they are >= 0
                // four independent partial words per half: one OR chain of 32 would serialise the lone epilogue warp
                uint32_t lw[4] = {0u, 0u, 0u, 0u}, hw[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    if (pk[j] & 0x00007fffu) lw[j & 3] |= 1u << (2 * j);
                    if (pk[j] & 0x7fff0000u) lw[j & 3] |= 1u << (2 * j + 1);
                    if (pk[16 + j] & 0x00007fffu) hw[j & 3] |= 1u << (2 * j);
                    if (pk[16 + j] & 0x7fff0000u) hw[j & 3] |= 1u << (2 * j + 1);
                }
                const uint32_t lo = (lw[0] | lw[1]) | (lw[2] | lw[3]), hi = (hw[0] | hw[1]) | (hw[2] | hw[3]);
                if (m < p.M && nc0 < p.N)
                    p.relu_bits_out[(size_t)m * ld_bits + (nc0 >> 6)] = ((unsigned long long)hi << 32) | lo;
            }
            if (p.stats != nullptr) {
                // online softmax statistics over the bf16-rounded values (what CE will see)
                if (nc0 + 64 <= p.N) {
                    // full chunk (all but the last of the vocabulary): packed bf16x2 maximum, then exp2 with four
                    // independent accumulators.  (The per-element bounds tests and the serial sum of the general
                    // path made the LM-head epilogue as long as its main loop.)
                    bf162 mx2 = *reinterpret_cast<const bf162*>(&pk[0]);
#pragma unroll
                    for (int j = 1; j < 32; ++j) mx2 = __hmax2(mx2, *reinterpret_cast<const bf162*>(&pk[j]));
                    const float2 mxf = __bfloat1622float2(mx2);
                    const float new_max = fmaxf(run_max, fmaxf(mxf.x, mxf.y));
                    const float nm2 = new_max * 1.4426950408889634f;
                    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        const float2 f = unpack_bf16x2(pk[j]), g = unpack_bf16x2(pk[j + 1]);
                        s0 += exp2f(fmaf(f.x, 1.4426950408889634f, -nm2));
                        s1 += exp2f(fmaf(f.y, 1.4426950408889634f, -nm2));
                        s2 += exp2f(fmaf(g.x, 1.4426950408889634f, -nm2));
                        s3 += exp2f(fmaf(g.y, 1.4426950408889634f, -nm2));
                    }
                    run_sum = run_sum * exp2f(fmaf(run_max, 1.4426950408889634f, -nm2)) + ((s0 + s1) + (s2 + s3));
                    run_max = new_max;
                } else {
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
            }
            // ---- stage in swizzled smem, then one thread issues the TMA store.  Buffer `buf` is free: the store that
            //      read it (two chunks ago) was waited for before the PREVIOUS chunk's barrier, so one barrier per
            //      chunk is enough.
            if (NG == 2) {
                // this group's single buffer: its previous store must have left it
                if (store_thread) ptx::tma_store_wait_read<0>();
                asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            }
            const uint32_t sdst = epi_base + buf * EPI_BUF_BYTES + row * 128;
#pragma unroll
            for (int ch = 0; ch < 8; ++ch) {
                const uint32_t addr = sdst + (((uint32_t)ch ^ ((uint32_t)row & 7u)) << 4);
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(pk[4 * ch]),
                             "r"(pk[4 * ch + 1]), "r"(pk[4 * ch + 2]), "r"(pk[4 * ch + 3])
                             : "memory");
            }
            ptx::fence_proxy_async_smem();
            if (NG == 1 && store_thread) ptx::tma_store_wait_read<0>();  // the previous store has left the OTHER buffer
            asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
            if (store_thread) {
                if (nc0 < p.N) ptx::tma_store_2d(tmC, epi_base + buf * EPI_BUF_BYTES, nc0, m0);
                ptx::tma_store_commit();
            }
            if (COLSUM) {
                // Column sums of the staged 128 x 64 bf16 chunk (the values the consumer will read): thread t sums column
                // t % 64 over rows [64 * (t / 64), +64) of the swizzled tile - the 32 lanes of a warp read 64 consecutive
                // bytes of one row per step (conflict-free) - and adds its half to colsum[n] with one fp32 atomic.  The
                // buffer is not overwritten before the next chunk's barrier, which every thread passes after these reads.
                const int tg = et - 128 * grp;               // 0..127 inside this epilogue group
                const int cj = tg & 63, r_lo = (tg >> 6) * 64;
                const uint32_t cbase = epi_base + buf * EPI_BUF_BYTES + (uint32_t)((cj & 7) * 2);
                float cs = 0.f;
#pragma unroll 4
                for (int rr = 0; rr < 64; ++rr) {
                    const int rw = r_lo + rr;
                    const uint32_t addr = cbase + (uint32_t)rw * 128u + ((((uint32_t)cj >> 3) ^ ((uint32_t)rw & 7u)) << 4);
                    unsigned short hv;
                    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(addr));
                    if (m0 + rw < p.M) cs += __bfloat162float(__ushort_as_bfloat16(hv));
                }
                if (nc0 + cj < p.N) atomicAdd(p.colsum + nc0 + cj, cs);
            }
            if (NG == 1) buf ^= 1;
        }
    }
    // one (max, sum exp) entry per row, n-tile and epilogue group: the CE kernels merge any number of entries per row
    if (!OUT_F32 && p.stats != nullptr && m < p.M && grp < NCH)
        p.stats[((size_t)n_blk * NG + grp) * p.M + m] = make_float2(run_max, run_sum);
}

}  // namespace gemm
