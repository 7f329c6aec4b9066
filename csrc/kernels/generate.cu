// Generation kernels: KV-cache append, single-position decode attention, fused softmax sampler.
//
// The reference re-runs the whole model over the whole context for every new token and samples
// with softmax + torch.multinomial on fp32 logits (reference src/models/transformer.py:96-114).
// Here each step touches one position: its K/V rows are appended to a [B, H, Tmax, hs] cache, the
// query attends to the cache with an online softmax, and the next token is drawn from the
// last-position logits by inverse-CDF sampling inside one kernel (no host sync per token).
#include "b200llm.h"
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256) kv_append_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ kc,
                                                        bf16* __restrict__ vc, int B, int t, int H, int hs, int Tmax,
                                                        int pos0, const int* __restrict__ pos_dev) {
    pdl_sync();
    if (pos_dev != nullptr) pos0 = *pos_dev;
    const int cpr = hs / 8;  // 16-byte chunks per (row, head)
    const long long total = (long long)B * t * H * cpr;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cpr);
        const int h = (int)((i / cpr) % H);
        const int tt = (int)((i / ((long long)cpr * H)) % t);
        const int b = (int)(i / ((long long)cpr * H * t));
        const size_t src = ((size_t)b * t + tt) * (3 * H * hs) + (size_t)H * hs + h * hs + c * 8;
        const size_t dst = (((size_t)b * H + h) * Tmax + pos0 + tt) * hs + c * 8;
        *reinterpret_cast<uint4*>(kc + dst) = *reinterpret_cast<const uint4*>(qkv + src);
        *reinterpret_cast<uint4*>(vc + dst) = *reinterpret_cast<const uint4*>(qkv + src + (size_t)H * hs);
    }
}

constexpr int DEC_WARPS = 8;

template <int HD>
__global__ void __launch_bounds__(DEC_WARPS * 32) attn_decode_kernel(const bf16* __restrict__ qkv,
                                                                     const bf16* __restrict__ kc,
                                                                     const bf16* __restrict__ vc, bf16* __restrict__ out,
                                                                     int H, int Tmax, int len, float scale,
                                                                     const int* __restrict__ pos_dev, int append) {
    pdl_sync();
    if (pos_dev != nullptr) len = *pos_dev + 1;
    constexpr int VPL = HD / 32;  // output columns per lane
    __shared__ float s_q[HD];
    __shared__ float s_m[DEC_WARPS], s_l[DEC_WARPS];
    __shared__ float s_acc[DEC_WARPS][HD];
    const int h = blockIdx.x, b = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bf16* kb = kc + ((size_t)b * H + h) * Tmax * HD;
    const bf16* vb = vc + ((size_t)b * H + h) * Tmax * HD;
    if (threadIdx.x < HD) {
        const size_t qoff = (size_t)b * 3 * H * HD + h * HD + threadIdx.x;
        s_q[threadIdx.x] = __bfloat162float(qkv[qoff]) * scale;
        if (append) {
            // fused KV-cache append: this position's key / value rows go into the cache before they are attended to
            const_cast<bf16*>(kb)[(size_t)(len - 1) * HD + threadIdx.x] = qkv[qoff + (size_t)H * HD];
            const_cast<bf16*>(vb)[(size_t)(len - 1) * HD + threadIdx.x] = qkv[qoff + 2 * (size_t)H * HD];
        }
    }
    __syncthreads();

    float m = -INFINITY, l = 0.f, acc[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] = 0.f;

    for (int j0 = warp * 32; j0 < len; j0 += DEC_WARPS * 32) {
        // phase A: lane <-> key j0 + lane, full dot product
        const int j = j0 + lane;
        float s = -INFINITY;
        if (j < len) {
            const uint4* kr = reinterpret_cast<const uint4*>(kb + (size_t)j * HD);
            float d = 0.f;
#pragma unroll
            for (int c = 0; c < HD / 8; ++c) {
                const uint4 u = append ? kr[c] : __ldg(kr + c);   // L1-allocating: a lane walks its key row 16 B at a time
                const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y), f2 = unpack_bf16x2(u.z),
                             f3 = unpack_bf16x2(u.w);
                d += f0.x * s_q[c * 8] + f0.y * s_q[c * 8 + 1] + f1.x * s_q[c * 8 + 2] + f1.y * s_q[c * 8 + 3] +
                     f2.x * s_q[c * 8 + 4] + f2.y * s_q[c * 8 + 5] + f3.x * s_q[c * 8 + 6] + f3.y * s_q[c * 8 + 7];
            }
            s = d;
        }
        // phase B: online softmax over the 32 keys
        const float mnew = fmaxf(m, warp_max(s));
        const float corr = __expf(m - mnew);
        const float p = j < len ? __expf(s - mnew) : 0.f;
        l = l * corr + warp_sum(p);
        m = mnew;
#pragma unroll
        for (int i = 0; i < VPL; ++i) acc[i] *= corr;
        // phase C: acc += sum_j p_j V[j, lane cols]
        const int nk = min(32, len - j0);
#pragma unroll 8
        for (int jj = 0; jj < nk; ++jj) {
            const float pj = __shfl_sync(0xffffffffu, p, jj);
            const bf16* vr = vb + (size_t)(j0 + jj) * HD + lane * VPL;
            if (VPL == 2) {
                const float2 f = unpack_bf16x2(*reinterpret_cast<const uint32_t*>(vr));
                acc[0] += pj * f.x; acc[1] += pj * f.y;
            } else {
                const uint2 u = *reinterpret_cast<const uint2*>(vr);
                const float2 f0 = unpack_bf16x2(u.x), f1 = unpack_bf16x2(u.y);
                acc[0] += pj * f0.x; acc[1] += pj * f0.y; acc[2] += pj * f1.x; acc[3] += pj * f1.y;
            }
        }
    }
    if (lane == 0) { s_m[warp] = m; s_l[warp] = l; }
#pragma unroll
    for (int i = 0; i < VPL; ++i) s_acc[warp][lane * VPL + i] = acc[i];
    __syncthreads();
    if (threadIdx.x < HD) {
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < DEC_WARPS; ++w) M = fmaxf(M, s_m[w]);
        float L = 0.f, o = 0.f;
#pragma unroll
        for (int w = 0; w < DEC_WARPS; ++w) {
            const float f = s_m[w] == -INFINITY ? 0.f : __expf(s_m[w] - M);
            L += s_l[w] * f;
            o += s_acc[w][threadIdx.x] * f;
        }
        out[(size_t)b * H * HD + h * HD + threadIdx.x] = __float2bfloat16_rn(L > 0.f ? o / L : 0.f);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// Split-KV decode attention.  Grid (H, B, S): the S CTAs of a (batch row, head) pair each take 1/S of the cached keys,
// so the grid has B*H*S CTAs of 4 warps however long the context is and every SM streams the cache.  A lane owns
// HD/32 contiguous dimensions of every key / value row: all loads are coalesced (a warp reads a whole row per
// instruction), 8 key rows and 8 value rows are in flight per warp before the first use (4 KB per warp, ~128 KB per SM
// at 8 CTAs per SM), the dot products are completed with xor-shuffles, and the online softmax runs 8 keys at a time.
// Partials (m, l, acc) of the splits go to `scratch`; the LAST CTA of a (b, h) pair to arrive (atomic ticket) merges
// them and writes the output row - one launch, no second kernel.  With append != 0 the warp that meets position
// len-1 takes that key / value row from qkv and stores it into the cache (fused KV append).
// ---------------------------------------------------------------------------------------------------------------
constexpr int DS_WARPS = 4;
constexpr int DS_KB = 8;

template <int HD>
__global__ void __launch_bounds__(DS_WARPS * 32) attn_decode_split_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ kc,
                                                                          bf16* __restrict__ vc, bf16* __restrict__ out,
                                                                          float* __restrict__ scratch, unsigned int* tickets,
                                                                          int H, int Tmax, int len, float scale_log2,
                                                                          const int* __restrict__ pos_dev, int append) {
    pdl_sync();
    if (pos_dev != nullptr) len = *pos_dev + 1;
    constexpr int VPL = HD / 32;                 // dimensions per lane: 2 (hs 64) or 4 (hs 128)
    __shared__ float s_m[DS_WARPS], s_l[DS_WARPS];
    __shared__ float s_acc[DS_WARPS][HD];
    __shared__ unsigned int s_ticket;
    const int h = blockIdx.x, b = blockIdx.y, sp = blockIdx.z, S = gridDim.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    bf16* kb = kc + ((size_t)b * H + h) * Tmax * HD + lane * VPL;
    bf16* vb = vc + ((size_t)b * H + h) * Tmax * HD + lane * VPL;
    const bf16* qrow = qkv + (size_t)b * 3 * H * HD + h * HD + lane * VPL;
    float q[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) q[i] = __bfloat162float(qrow[i]) * scale_log2;

    const int per = ((len + S - 1) / S + DS_KB - 1) / DS_KB * DS_KB;      // keys per split, a multiple of the block
    const int lo = sp * per, hi = min(len, lo + per);
    float m = -INFINITY, l = 0.f, acc[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] = 0.f;

    for (int j0 = lo + warp * DS_KB; j0 < hi; j0 += DS_WARPS * DS_KB) {
        float kf[DS_KB][VPL], vf[DS_KB][VPL];
        // ---- all 16 row loads of the block are issued before anything is used
#pragma unroll
        for (int i = 0; i < DS_KB; ++i) {
            const int j = j0 + i;
            const bool fresh = append && j == len - 1;           // this step's own key / value: from qkv, into the cache
            const bf16* kr = fresh ? qrow + (size_t)H * HD : kb + (size_t)j * HD;
            const bf16* vr = fresh ? qrow + 2 * (size_t)H * HD : vb + (size_t)j * HD;
            if (j < hi) {
                if (VPL == 2) {
                    const uint32_t ku = *reinterpret_cast<const uint32_t*>(kr), vu = *reinterpret_cast<const uint32_t*>(vr);
                    const float2 a = unpack_bf16x2(ku), c = unpack_bf16x2(vu);
                    kf[i][0] = a.x; kf[i][1] = a.y; vf[i][0] = c.x; vf[i][1] = c.y;
                    if (fresh) {
                        *reinterpret_cast<uint32_t*>(kb + (size_t)j * HD) = ku;
                        *reinterpret_cast<uint32_t*>(vb + (size_t)j * HD) = vu;
                    }
                } else {
                    const uint2 ku = *reinterpret_cast<const uint2*>(kr), vu = *reinterpret_cast<const uint2*>(vr);
                    const float2 a0 = unpack_bf16x2(ku.x), a1 = unpack_bf16x2(ku.y), c0 = unpack_bf16x2(vu.x), c1 = unpack_bf16x2(vu.y);
                    kf[i][0] = a0.x; kf[i][1] = a0.y; kf[i][VPL - 2] = a1.x; kf[i][VPL - 1] = a1.y;
                    vf[i][0] = c0.x; vf[i][1] = c0.y; vf[i][VPL - 2] = c1.x; vf[i][VPL - 1] = c1.y;
                    if (fresh) {
                        *reinterpret_cast<uint2*>(kb + (size_t)j * HD) = ku;
                        *reinterpret_cast<uint2*>(vb + (size_t)j * HD) = vu;
                    }
                }
            } else {
#pragma unroll
                for (int e = 0; e < VPL; ++e) { kf[i][e] = 0.f; vf[i][e] = 0.f; }
            }
        }
        // ---- 8 dot products: per-lane partials, then xor-shuffle butterflies (every lane ends with every score)
        float sc[DS_KB];
#pragma unroll
        for (int i = 0; i < DS_KB; ++i) {
            float d = 0.f;
#pragma unroll
            for (int e = 0; e < VPL; ++e) d = fmaf(kf[i][e], q[e], d);
            sc[i] = d;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int i = 0; i < DS_KB; ++i) sc[i] += __shfl_xor_sync(0xffffffffu, sc[i], o);
        float mx = m;
#pragma unroll
        for (int i = 0; i < DS_KB; ++i) {
            if (j0 + i >= hi) sc[i] = -INFINITY;
            mx = fmaxf(mx, sc[i]);
        }
        const float corr = exp2f(m - mx);                     // m = -inf on the first block: exp2(-inf) = 0
        m = mx;
        l *= corr;
#pragma unroll
        for (int e = 0; e < VPL; ++e) acc[e] *= corr;
#pragma unroll
        for (int i = 0; i < DS_KB; ++i) {
            const float p = exp2f(sc[i] - mx);                // masked keys: exp2(-inf) = 0
            l += p;
#pragma unroll
            for (int e = 0; e < VPL; ++e) acc[e] = fmaf(p, vf[i][e], acc[e]);
        }
    }
    // ---- merge the warps of this CTA
    if (lane == 0) { s_m[warp] = m; s_l[warp] = l; }
#pragma unroll
    for (int e = 0; e < VPL; ++e) s_acc[warp][lane * VPL + e] = acc[e];
    __syncthreads();
    float M = -INFINITY, L = 0.f, o = 0.f;
    if (threadIdx.x < HD) {
#pragma unroll
        for (int w = 0; w < DS_WARPS; ++w) M = fmaxf(M, s_m[w]);
#pragma unroll
        for (int w = 0; w < DS_WARPS; ++w) {
            const float f = s_m[w] == -INFINITY ? 0.f : exp2f(s_m[w] - M);
            L += s_l[w] * f;
            o += s_acc[w][threadIdx.x] * f;
        }
    }
    bf16* orow = out + (size_t)b * H * HD + h * HD;
    if (S == 1) {
        if (threadIdx.x < HD) orow[threadIdx.x] = __float2bfloat16_rn(L > 0.f ? o / L : 0.f);
        return;
    }
    // ---- publish this split's partial; the last split to arrive merges all of them
    float* mine = scratch + (((size_t)b * H + h) * S + sp) * (HD + 2);
    if (threadIdx.x < HD) mine[2 + threadIdx.x] = o;
    if (threadIdx.x == 0) { mine[0] = M; mine[1] = L; }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_ticket = atomicAdd(tickets + b * H + h, 1u);
    __syncthreads();
    if (s_ticket != (unsigned int)(S - 1)) return;
    __threadfence();
    if (threadIdx.x < HD) {
        const float* base = scratch + ((size_t)b * H + h) * S * (HD + 2);
        float MM = -INFINITY;
        for (int s2 = 0; s2 < S; ++s2) MM = fmaxf(MM, __ldcg(base + s2 * (HD + 2)));
        float LL = 0.f, oo = 0.f;
        for (int s2 = 0; s2 < S; ++s2) {
            const float ms = __ldcg(base + s2 * (HD + 2));
            const float f = ms == -INFINITY ? 0.f : exp2f(ms - MM);
            LL += __ldcg(base + s2 * (HD + 2) + 1) * f;
            oo += __ldcg(base + s2 * (HD + 2) + 2 + threadIdx.x) * f;
        }
        orow[threadIdx.x] = __float2bfloat16_rn(LL > 0.f ? oo / LL : 0.f);
    }
    if (threadIdx.x == 0) tickets[b * H + h] = 0;             // ready for the next launch
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

constexpr int SAMP_THREADS = 1024;

// One CTA per row: max -> sum exp -> inverse CDF at u * sum.
__global__ void __launch_bounds__(SAMP_THREADS) sample_kernel(const bf16* __restrict__ logits, long long ld,
                                                              long long* __restrict__ next, int V, float inv_temp,
                                                              unsigned long long seed, unsigned long long step,
                                                              const int* __restrict__ pos_dev) {
    pdl_sync();
    if (pos_dev != nullptr) step += (unsigned long long)*pos_dev;
    __shared__ float s_red[32];
    __shared__ int s_pick;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bf16* row = logits + (size_t)b * ld;
    const int per = (V + SAMP_THREADS - 1) / SAMP_THREADS;
    const int i0 = tid * per, i1 = min(V, i0 + per);
    float mx = -INFINITY;
    for (int i = i0; i < i1; ++i) mx = fmaxf(mx, __bfloat162float(row[i]) * inv_temp);
    mx = warp_max(mx);
    if (lane == 0) s_red[warp] = mx;
    __syncthreads();
    mx = s_red[lane];
    mx = warp_max(mx);
    __syncthreads();
    float part = 0.f;
    for (int i = i0; i < i1; ++i) part += __expf(__bfloat162float(row[i]) * inv_temp - mx);
    // inclusive scan of per-thread partial sums
    float v = part;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    if (lane == 31) s_red[warp] = v;
    __syncthreads();
    if (warp == 0) {
        float w = s_red[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float n = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += n;
        }
        s_red[lane] = w;
    }
    if (tid == 0) s_pick = -1;
    __syncthreads();
    const float incl = v + (warp > 0 ? s_red[warp - 1] : 0.f);
    const float total = s_red[31];
    const unsigned long long r = mix64(mix64(seed ^ (step * 0xD1B54A32D192ED03ull)) + (unsigned long long)b);
    const float u = (float)((r >> 40) + 0.5) * (1.0f / 16777216.0f);  // (0,1)
    const float target = u * total;
    const float excl = incl - part;
    if (target >= excl && target < incl && i0 < i1) {
        float c = excl;
        int pick = i1 - 1;
        for (int i = i0; i < i1; ++i) {
            c += __expf(__bfloat162float(row[i]) * inv_temp - mx);
            if (target < c) { pick = i; break; }
        }
        atomicMax(&s_pick, pick);
    }
    __syncthreads();
    if (tid == 0) {
        int pick = s_pick;
        if (pick < 0) {  // rounding fell off the end: take the last token with non-zero mass
            pick = V - 1;
        }
        next[b] = pick;
    }
}

}  // namespace

extern "C" int b200_kv_append(const void* qkv, void* kcache, void* vcache, int B, int t, int H, int hs, int Tmax, int pos0,
                              const int* pos_dev, cudaStream_t st) {
    B200_REQUIRE(hs % 8 == 0 && pos0 >= 0 && pos0 + t <= Tmax, "kv_append: bad shape (hs=%d pos0=%d t=%d Tmax=%d)", hs,
                 pos0, t, Tmax);
    const long long total = (long long)B * t * H * (hs / 8);
    long long nb = (total + 255) / 256;
    const long long cap = (long long)b200_num_sms() * 8;
    const int blocks = (int)(nb < cap ? nb : cap);
    B200_CHECK_CUDA(b200_launch(kv_append_kernel, dim3(max(blocks, 1)), dim3(256), 0, st, (const bf16*)qkv, (bf16*)kcache, (bf16*)vcache, B, t, H, hs, Tmax,
                                                      pos0, pos_dev));
    B200_RETURN_LAUNCH_STATUS("kv_append");
}

extern "C" int b200_attn_decode(const void* qkv, void* kcache, void* vcache, void* out, int B, int H, int hs,
                                int Tmax, int len, float scale, const int* pos_dev, int append, cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_decode: head size must be 64 or 128");
    B200_REQUIRE(len >= 1 && len <= Tmax, "attn_decode: bad len");
    dim3 grid(H, B);
    if (hs == 64)
        B200_CHECK_CUDA(b200_launch(attn_decode_kernel<64>, dim3(grid), dim3(DEC_WARPS * 32), 0, st, (const bf16*)qkv, (const bf16*)kcache,
                                                                (const bf16*)vcache, (bf16*)out, H, Tmax, len, scale, pos_dev, append));
    else
        B200_CHECK_CUDA(b200_launch(attn_decode_kernel<128>, dim3(grid), dim3(DEC_WARPS * 32), 0, st, (const bf16*)qkv, (const bf16*)kcache,
                                                                 (const bf16*)vcache, (bf16*)out, H, Tmax, len, scale, pos_dev, append));
    B200_RETURN_LAUNCH_STATUS("attn_decode");
}

extern "C" int b200_sample(const void* logits, long long ld, long long* next, int B, int V, float temperature, int top_k,
                           unsigned long long seed, unsigned long long step, const int* pos_dev, cudaStream_t st) {
    B200_REQUIRE(top_k <= 0, "sample: top_k filtering is applied by the caller (mask logits), pass top_k=0");
    B200_REQUIRE(temperature > 0.f, "sample: temperature must be > 0");
    B200_CHECK_CUDA(b200_launch(sample_kernel, dim3(B), dim3(SAMP_THREADS), 0, st, (const bf16*)logits, ld, next, V, 1.f / temperature, seed, step, pos_dev));
    B200_RETURN_LAUNCH_STATUS("sample");
}

__global__ void incr_kernel(int* p, int delta) {
    pdl_sync(); *p += delta; }
extern "C" int b200_incr_int(int* p, int delta, cudaStream_t st) {
    B200_CHECK_CUDA(b200_launch(incr_kernel, dim3(1), dim3(1), 0, st, p, delta));
    B200_RETURN_LAUNCH_STATUS("incr_int");
}

// seq[b, col] = next[b]   (col = *pos_dev + 1 when pos_dev is given)
__global__ void scatter_tokens_kernel(const long long* next, long long* seq, int B, long long ld, int col,
                                      const int* pos_dev) {
    pdl_sync();
    if (pos_dev != nullptr) col = *pos_dev + 1;
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b < B) seq[(size_t)b * ld + col] = next[b];
}
extern "C" int b200_scatter_tokens(const long long* next, long long* seq, int B, long long ld, int col,
                                   const int* pos_dev, cudaStream_t st) {
    B200_CHECK_CUDA(b200_launch(scatter_tokens_kernel, dim3((B + 127) / 128), dim3(128), 0, st, next, seq, B, ld, col, pos_dev));
    B200_RETURN_LAUNCH_STATUS("scatter_tokens");
}

// out[b, :] = in[(b * T + T - 1), :]  - last-position rows for the last-token-only LM head
__global__ void gather_last_rows_kernel(const uint4* in, uint4* out, int B, int T, int row_chunks) {
    pdl_sync();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B * row_chunks) {
        const int b = i / row_chunks, c = i % row_chunks;
        out[(size_t)b * row_chunks + c] = in[((size_t)b * T + T - 1) * row_chunks + c];
    }
}
extern "C" int b200_gather_last_rows(const void* in, void* out, int B, int T, int row_bytes, cudaStream_t st) {
    B200_REQUIRE(row_bytes % 16 == 0, "gather_last_rows: row_bytes must be a multiple of 16");
    const int rc = row_bytes / 16;
    B200_CHECK_CUDA(b200_launch(gather_last_rows_kernel, dim3((B * rc + 255) / 256), dim3(256), 0, st, (const uint4*)in, (uint4*)out, B, T, rc));
    B200_RETURN_LAUNCH_STATUS("gather_last_rows");
}


// Split-KV form of b200_attn_decode (same contract).  scratch: B * H * splits * (hs + 2) floats; tickets: B * H zeroed
// unsigned ints (left zero).  splits >= 1; with splits == 1 scratch / tickets may be null.
extern "C" int b200_attn_decode_split(const void* qkv, void* kcache, void* vcache, void* out, float* scratch, void* tickets,
                                      int B, int H, int hs, int Tmax, int len, float scale, const int* pos_dev, int append,
                                      int splits, cudaStream_t st) {
    B200_REQUIRE(hs == 64 || hs == 128, "attn_decode_split: head size must be 64 or 128");
    B200_REQUIRE(len >= 1 && len <= Tmax, "attn_decode_split: bad len");
    B200_REQUIRE(splits >= 1 && splits <= 64 && (splits == 1 || (scratch != nullptr && tickets != nullptr)),
                 "attn_decode_split: splits > 1 needs scratch and tickets");
    const float scale_log2 = scale * 1.4426950408889634f;
    dim3 grid(H, B, splits);
    if (hs == 64)
        B200_CHECK_CUDA(b200_launch(attn_decode_split_kernel<64>, grid, dim3(DS_WARPS * 32), 0, st, (const bf16*)qkv, (bf16*)kcache,
                                    (bf16*)vcache, (bf16*)out, scratch, (unsigned int*)tickets, H, Tmax, len, scale_log2, pos_dev, append));
    else
        B200_CHECK_CUDA(b200_launch(attn_decode_split_kernel<128>, grid, dim3(DS_WARPS * 32), 0, st, (const bf16*)qkv, (bf16*)kcache,
                                    (bf16*)vcache, (bf16*)out, scratch, (unsigned int*)tickets, H, Tmax, len, scale_log2, pos_dev, append));
    B200_RETURN_LAUNCH_STATUS("attn_decode_split");
}
