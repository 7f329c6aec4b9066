// In-kernel timeline of the tcgen05 attention backward (one CTA's pipeline hand-offs, clock64 deltas).
// Build + run:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --use_fast_math --expt-relaxed-constexpr \
//               -DATTN_TRACE -Iinclude -Icsrc tools/attn_trace.cu -o build/attn_trace -lcudart && build/attn_trace 64
#include "../csrc/kernels/misc.cu"
#include "../csrc/kernels/attention_tc.cu"
#include <vector>

int main(int argc, char** argv) {
    const int hs = argc > 1 ? atoi(argv[1]) : 64;
    const int H = hs == 64 ? 12 : 16, B = hs == 64 ? 16 : 8, T = hs == 64 ? 1024 : 2048, d = H * hs;
    const size_t nq = (size_t)B * T * 3 * d, no = (size_t)B * T * d;
    std::vector<bf16> hq(nq), ho(no);
    unsigned s = 1234567u;
    auto rnd = [&]() { s = s * 1664525u + 1013904223u; return ((s >> 8) & 0xffff) / 65536.f - 0.5f; };
    for (auto& v : hq) v = __float2bfloat16(rnd());
    for (auto& v : ho) v = __float2bfloat16(rnd());
    bf16 *qkv, *out, *dout, *dqkv; float *lse, *delta;
    cudaMalloc(&qkv, nq * 2); cudaMalloc(&dqkv, nq * 2); cudaMalloc(&out, no * 2); cudaMalloc(&dout, no * 2);
    cudaMalloc(&lse, (size_t)B * H * T * 4); cudaMalloc(&delta, (size_t)B * H * T * 4);
    cudaMemcpy(qkv, hq.data(), nq * 2, cudaMemcpyHostToDevice);
    cudaMemcpy(dout, ho.data(), no * 2, cudaMemcpyHostToDevice);
    const float scale = 1.f / sqrtf((float)hs);
    if (b200_attn_fwd_tc(qkv, out, lse, B, T, H, hs, scale, 0)) { printf("fwd: %s\n", b200_last_error()); return 1; }
    for (int it = 0; it < 3; ++it)
        if (b200_attn_bwd_tc(qkv, out, dout, lse, delta, dqkv, B, T, H, hs, scale, 0)) { printf("bwd: %s\n", b200_last_error()); return 1; }
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("sync failed\n"); return 1; }
    {
        static long long f[3][64][8];
        cudaMemcpyFromSymbol(f, g_attn_trace_fwd, sizeof(f));
        const long long t0 = f[0][0][0];
        printf("pass=FWD hs=%d, CTA 0, first work item (last pair of query tiles); clk relative to the first S_A issue\n", hs);
        printf("%3s | %35s | %44s | %44s\n", "j", "mma: S_A PV_A S_B PV_B issued", "group A: top s_full loaded exp_done arrived",
               "group B: top s_full loaded exp_done arrived");
        for (int n = 0; n < 16; ++n) {
            printf("%3d |", n);
            for (int k = 0; k < 4; ++k) printf(" %8lld", f[0][n][k] ? f[0][n][k] - t0 : 0);
            printf(" |");
            for (int k = 0; k < 5; ++k) printf(" %8lld", f[1][n][k] ? f[1][n][k] - t0 : 0);
            printf(" |");
            for (int k = 0; k < 5; ++k) printf(" %8lld", f[2][n][k] ? f[2][n][k] - t0 : 0);
            printf("\n");
        }
    }
    static long long all[2][3][64][8];
    cudaMemcpyFromSymbol(all, g_attn_trace, sizeof(all));
    for (int pass = 1; pass >= 0; --pass) {
        auto& tr = all[pass];
        const long long t0 = tr[1][0][0];
        printf("pass=%s hs=%d, CTA 0 (heaviest tile); clk relative to the MMA warp's first str_full\n", pass ? "DKV" : "DQ", hs);
        printf("%3s | %17s | %44s | %53s\n", "n", "str_empty vec_done", "str_full sc_empty sc_issued out_full acc_issued",
               "top vec_full sc_full out_empty computed arrived");
        for (int n = 0; n < 16; ++n) {
            printf("%3d |", n);
            for (int k = 0; k < 2; ++k) printf(" %8lld", tr[0][n][k] ? tr[0][n][k] - t0 : 0);
            printf(" |");
            for (int k = 0; k < 5; ++k) printf(" %8lld", tr[1][n][k] ? tr[1][n][k] - t0 : 0);
            printf(" |");
            for (int k = 0; k < 6; ++k) printf(" %8lld", tr[2][n][k] ? tr[2][n][k] - t0 : 0);
            printf("\n");
        }
    }
    return 0;
}
