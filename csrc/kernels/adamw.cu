// Fused AdamW over a flat fp32 parameter buffer, emitting the bf16 shadow weights the GEMMs read.
//
// Replaces torch.optim.AdamW's foreach multi-tensor path over 534..3590 tensors (reference
// scripts/train_transformer.py:92,126) with ONE launch.  The math follows torch's
// operation order (decoupled weight decay on every parameter, lerp for exp_avg, bias
// corrections applied as step_size = lr/bc1 and denom = sqrt(v)/sqrt(bc2) + eps) so that it
// tracks torch.optim.AdamW to rounding.  Hyper-parameters come from an 8-float device array so
// a captured CUDA graph picks up the new learning rate / bias corrections on every replay.  The
// derived scalars are computed on the host in double and rounded once, like torch does:
//   hyper = { 1 - lr*wd, beta2, 1 - beta1, 1 - beta2, lr / bias_correction1, sqrt(bias_correction2),
//             eps, grad_scale }
#include "b200llm.h"
#include "common.cuh"

__device__ __forceinline__ void adamw_elem(float& p, float g, float& m, float& v, float decay, float b2, float w1,
                                           float w2, float step_size, float bc2s, float eps, float gs) {
    g *= gs;
    p = p * decay;                         // param.mul_(1 - lr * weight_decay)
    m = fmaf(w1, g - m, m);                // exp_avg.lerp_(grad, 1 - beta1)
    v = fmaf(w2 * g, g, v * b2);           // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    const float denom = __fdiv_rn(__fsqrt_rn(v), bc2s) + eps;
    p = fmaf(-step_size, __fdiv_rn(m, denom), p);   // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                    float* __restrict__ m, float* __restrict__ v,
                                                    bf16* __restrict__ pb, const float* __restrict__ hyper,
                                                    long long n) {
    pdl_sync();
    const float decay = hyper[0], b2 = hyper[1], w1 = hyper[2], w2 = hyper[3], step_size = hyper[4],
                bc2s = hyper[5], eps = hyper[6], gs = hyper[7];
    const long long n4 = n >> 2;
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < n4; i += stride) {
        float4 pv = reinterpret_cast<float4*>(p)[i];
        float4 gv = ldg_stream_f4(reinterpret_cast<const float4*>(g) + i);
        float4 mv = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        adamw_elem(pv.x, gv.x, mv.x, vv.x, decay, b2, w1, w2, step_size, bc2s, eps, gs);
        adamw_elem(pv.y, gv.y, mv.y, vv.y, decay, b2, w1, w2, step_size, bc2s, eps, gs);
        adamw_elem(pv.z, gv.z, mv.z, vv.z, decay, b2, w1, w2, step_size, bc2s, eps, gs);
        adamw_elem(pv.w, gv.w, mv.w, vv.w, decay, b2, w1, w2, step_size, bc2s, eps, gs);
        reinterpret_cast<float4*>(p)[i] = pv;
        reinterpret_cast<float4*>(m)[i] = mv;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (pb != nullptr)
            reinterpret_cast<uint2*>(pb)[i] = make_uint2(pack_bf16x2(pv.x, pv.y), pack_bf16x2(pv.z, pv.w));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (long long j = n4 << 2; j < n; ++j) {
            float pj = p[j], mj = m[j], vj = v[j];
            adamw_elem(pj, g[j], mj, vj, decay, b2, w1, w2, step_size, bc2s, eps, gs);
            p[j] = pj; m[j] = mj; v[j] = vj;
            if (pb != nullptr) pb[j] = __float2bfloat16_rn(pj);
        }
    }
}

extern "C" int b200_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, const float* hyper,
                          long long n, cudaStream_t st) {
    if (n <= 0) return 0;
    B200_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0 &&
                     (uintptr_t)p_bf16 % 8 == 0,
                 "adamw: buffers must be 16-byte aligned");
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = (long long)b200_num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    B200_CHECK_CUDA(b200_launch(adamw_kernel, dim3((int)blocks), dim3(256), 0, st, p, g, m, v, (bf16*)p_bf16, hyper, n));
    B200_RETURN_LAUNCH_STATUS("adamw");
}
