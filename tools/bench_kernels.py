"""Per-kernel micro-benchmarks (CUDA events, L2 flushed between iterations) against the measured
peaks of MEASURED_PEAKS.json and, for GEMMs, against torch.matmul (cuBLAS) on the same shapes."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pretraining_llm_b200 import ops  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = {"hbm_gbs": 6538.3, "bf16_tflops": 1669.2, "bf16_tflops_sustained": 1396.0}
try:
    peaks.update(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))))
except Exception:
    pass
flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        ops.l2_flush(flush_buf)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


rows = []
BF16 = torch.bfloat16


def gemm_case(name, M, N, K, a_mn=False, b_mn=False, f32=False, **kw):
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(BF16)
    b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(BF16)
    out = torch.empty(M, N, device="cuda", dtype=torch.float32 if f32 else BF16)
    ms = timeit(lambda: ops.gemm(a, b, out=out, a_mn=a_mn, b_mn=b_mn, two_cta=False, **kw))
    kw2 = dict(kw); kw2.pop("bn", None)
    ms2 = timeit(lambda: ops.gemm(a, b, out=out, a_mn=a_mn, b_mn=b_mn, two_cta=True, **kw2))
    A = a.t() if a_mn else a
    Bt = b if b_mn else b.t()
    ref = torch.empty(M, N, device="cuda", dtype=BF16)
    ms_ref = timeit(lambda: torch.matmul(A, Bt, out=ref))
    tf = 2.0 * M * N * K / ms / 1e9
    rows.append(dict(kernel=name, shape=f"{M}x{N}x{K}", ms=round(ms, 4), tflops=round(tf, 1),
                     frac_of_measured_burst=round(tf / peaks["bf16_tflops"], 3), ms_2cta=round(ms2, 4),
                     tflops_2cta=round(2.0 * M * N * K / ms2 / 1e9, 1), cublas_ms=round(ms_ref, 4),
                     cublas_tflops=round(2.0 * M * N * K / ms_ref / 1e9, 1)))
    print(rows[-1], flush=True)


M = 16384
d = 768
gemm_case("qkv fwd", M, 3 * d, d)
gemm_case("mlp up fwd (+bias,relu)", M, 4 * d, d, bias=torch.zeros(4 * d, device="cuda"), relu=True)
gemm_case("mlp down fwd", M, d, 4 * d)
gemm_case("lm head fwd", M, 50304, d, bn=256)
gemm_case("mlp down dgrad", M, 4 * d, d, b_mn=True)
gemm_case("mlp up dgrad", M, d, 4 * d, b_mn=True)
gemm_case("qkv dgrad", M, d, 3 * d, b_mn=True)
gemm_case("lm head dgrad", M, d, 50304, b_mn=True)
gemm_case("qkv wgrad", 3 * d, d, M, a_mn=True, b_mn=True, f32=True)
gemm_case("mlp up wgrad", 4 * d, d, M, a_mn=True, b_mn=True, f32=True)
gemm_case("mlp down wgrad", d, 4 * d, M, a_mn=True, b_mn=True, f32=True)
gemm_case("lm head wgrad", 50304, d, M, a_mn=True, b_mn=True, f32=True)
gemm_case("square 8192", 8192, 8192, 8192)
gemm_case("1.3b mlp up", 16384, 8192, 2048)

# attention
for hs, H in ((64, 12), (128, 16)):
    B, T = (16, 1024) if hs == 64 else (8, 2048)
    qkv = torch.randn(B * T, 3 * H * hs, device="cuda").to(BF16)
    out, lse = ops.attn_fwd(qkv, B, T, H, hs)
    ms = timeit(lambda: ops.attn_fwd(qkv, B, T, H, hs, out=out, lse=lse))
    fl = 4.0 * B * H * T * T * hs / 2
    rows.append(dict(kernel=f"attn fwd hs{hs}", shape=f"B{B} T{T} H{H}", ms=round(ms, 4), tflops=round(fl / ms / 1e9, 1)))
    print(rows[-1], flush=True)
    dout = torch.randn_like(out)
    dqkv = torch.empty_like(qkv)
    delta = torch.empty(B, H, T, device="cuda")
    dq = torch.zeros(B * T, H * hs, device="cuda")
    ms = timeit(lambda: ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=dqkv, delta=delta, dq_acc=dq))
    rows.append(dict(kernel=f"attn bwd hs{hs}", shape=f"B{B} T{T} H{H}", ms=round(ms, 4), tflops=round(2.5 * fl / ms / 1e9, 1)))
    print(rows[-1], flush=True)

# memory-bound kernels
x = torch.randn(M, d, device="cuda")
br = torch.randn(M, d, device="cuda").to(BF16)
g = torch.ones(d, device="cuda"); b = torch.zeros(d, device="cuda")
xo = torch.empty_like(x); h = torch.empty(M, d, device="cuda", dtype=BF16)
mean = torch.empty(M, device="cuda"); rstd = torch.empty(M, device="cuda")
ms = timeit(lambda: ops.ln_fwd(x, g, b, branch=br, x_out=xo, out=h, mean=mean, rstd=rstd))
by = M * d * (4 + 2 + 4 + 2)
rows.append(dict(kernel="ln_fwd(+residual)", shape=f"{M}x{d}", ms=round(ms, 4), gbs=round(by / ms / 1e6, 1),
                 frac_of_measured_hbm=round(by / ms / 1e6 / peaks["hbm_gbs"], 3)))
print(rows[-1], flush=True)
dy = torch.randn(M, d, device="cuda").to(BF16)
dx = torch.randn(M, d, device="cuda"); dxb = torch.empty(M, d, device="cuda", dtype=BF16)
dg = torch.zeros(d, device="cuda"); db = torch.zeros(d, device="cuda"); cs = torch.zeros(d, device="cuda")
part = torch.empty(ops.ln_bwd_nparts(M, d) * 3 * d, device="cuda")
ms = timeit(lambda: ops.ln_bwd(dy, x, mean, rstd, g, dg, db, dx_in=dx, dx_out=dx, dx_bf16=dxb, colsum_out=cs, partial=part))
by = M * d * (2 + 4 + 4 + 4 + 2)
rows.append(dict(kernel="ln_bwd(+residual,+colsum)", shape=f"{M}x{d}", ms=round(ms, 4), gbs=round(by / ms / 1e6, 1),
                 frac_of_measured_hbm=round(by / ms / 1e6 / peaks["hbm_gbs"], 3)))
print(rows[-1], flush=True)
n = 156_000_000 // 4 * 4
p = torch.randn(n, device="cuda"); gr = torch.randn(n, device="cuda"); m1 = torch.zeros(n, device="cuda"); v1 = torch.zeros(n, device="cuda")
pb = torch.empty(n, device="cuda", dtype=BF16); hy = torch.tensor(ops.adamw_hyper(5e-4, 1), device="cuda")
ms = timeit(lambda: ops.adamw(p, gr, m1, v1, pb, hy))
by = n * (4 * 4 + 3 * 4 + 2)
rows.append(dict(kernel="adamw", shape=f"{n}", ms=round(ms, 4), gbs=round(by / ms / 1e6, 1),
                 frac_of_measured_hbm=round(by / ms / 1e6 / peaks["hbm_gbs"], 3)))
print(rows[-1], flush=True)
del p, gr, m1, v1, pb
logits = torch.randn(M, 50304, device="cuda").to(BF16)
lse = torch.logsumexp(logits[:2048].float(), -1).repeat(8)
tgt = torch.randint(0, 50304, (M,), device="cuda"); dbias = torch.zeros(50304, device="cuda")
ms = timeit(lambda: ops.ce_grad(logits, lse, tgt, dbias, 50304, 1.0 / M), iters=5)
by = M * 50304 * 4
rows.append(dict(kernel="ce_grad (in place)", shape=f"{M}x50304", ms=round(ms, 4), gbs=round(by / ms / 1e6, 1),
                 frac_of_measured_hbm=round(by / ms / 1e6 / peaks["hbm_gbs"], 3)))
print(rows[-1], flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump({"peaks": peaks, "rows": rows}, open(os.path.join(ROOT, "gpurun_out", "kernel_bench.json"), "w"), indent=1)
