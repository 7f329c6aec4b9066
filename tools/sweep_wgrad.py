import os, sys, itertools
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=8, warm=2):
    for _ in range(warm): fn()
    ts = []
    for _ in range(iters):
        ops.l2_flush(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort(); return ts[len(ts)//2]
K = 16384
for (M, N) in ((2304, 768), (3072, 768), (768, 3072)):
    a = torch.randn(K, M, device="cuda").to(BF16); b = torch.randn(K, N, device="cuda").to(BF16)
    out = torch.zeros(M, N, device="cuda")
    res = []
    for two, bn, sk in itertools.product((False, True), (128, 256), (1, 2, 3, 4, 6, 8)):
        mode = 2 if sk > 1 else 0
        try:
            ms = timeit(lambda: ops.gemm(a, b, out=out, a_mn=True, b_mn=True, f32_mode=mode, bn=bn, split_k=sk, two_cta=two))
        except Exception as e:
            continue
        res.append((ms, two, bn, sk))
    res.sort()
    ref = timeit(lambda: torch.matmul(a.t(), b))
    print(f"wgrad {M}x{N}x{K}: cublas {ref:.4f} ms; best:", [(round(m, 4), '2cta' if t else '1cta', bn, sk) for m, t, bn, sk in res[:6]], flush=True)
# same for small-N dgrads / fwd with N=768
M = 16384
for (N, Kk, bmn) in ((768, 3072, False), (768, 3072, True), (768, 2304, True), (2304, 768, False), (3072, 768, False)):
    a = torch.randn(M, Kk, device="cuda").to(BF16)
    b = torch.randn((Kk, N) if bmn else (N, Kk), device="cuda").to(BF16)
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    res = []
    for two, bn in itertools.product((False, True), (64, 128, 256)):
        if two and bn == 64: continue
        ms = timeit(lambda: ops.gemm(a, b, out=out, b_mn=bmn, bn=bn, two_cta=two))
        res.append((round(ms, 4), '2cta' if two else '1cta', bn))
    res.sort()
    ref = timeit(lambda: torch.matmul(a, b if bmn else b.t()))
    print(f"gemm {M}x{N}x{Kk} b_mn={bmn}: cublas {ref:.4f};", res, flush=True)
