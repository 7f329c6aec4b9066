import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pretraining_llm_b200 import ops
BF16 = torch.bfloat16
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
def timeit(fn, iters=10, warm=3):
    for _ in range(warm): fn()
    ts = []
    for _ in range(iters):
        ops.l2_flush(flush)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort(); return ts[len(ts)//2]
for hs, H, B, T in ((64, 12, 16, 1024), (128, 16, 8, 2048), (64, 12, 4, 4096)):
    qkv = torch.randn(B*T, 3*H*hs, device="cuda").to(BF16)
    fl = 4.0*B*H*T*T*hs/2
    for impl in ("mma", "tc"):
        out, lse = ops.attn_fwd(qkv, B, T, H, hs, impl=impl)
        ms = timeit(lambda: ops.attn_fwd(qkv, B, T, H, hs, out=out, lse=lse, impl=impl))
        print(f"attn fwd {impl} hs{hs} B{B} T{T} H{H}: {ms:.4f} ms  {fl/ms/1e9:.1f} TFLOP/s", flush=True)
        dout = torch.randn_like(out); dqkv = torch.empty_like(qkv); delta = torch.empty(B, H, T, device="cuda")
        dq = torch.zeros(B*T, H*hs, device="cuda"); 
        ms = timeit(lambda: ops.attn_bwd(qkv, out, dout, lse, B, T, H, hs, dqkv=dqkv, delta=delta, dq_acc=dq, impl=impl))
        print(f"attn bwd {impl} hs{hs} B{B} T{T} H{H}: {ms:.4f} ms  {2.5*fl/ms/1e9:.1f} TFLOP/s", flush=True)
