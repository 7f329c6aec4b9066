"""Roofline of the two fused compute + collective paths, in isolation, device-timed, max over ranks (run under torchrun).

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_dp_kernel.py [--model gpt2-small|1.3b]

1. reduce path - gradient reduce(-scatter) + AdamW (+ weight push), ALL buckets back to back on the communication
   stream with nothing else running: `dp_reduce_adamw` (replicated mode) and `zero_reduce_adamw` (sharded mode).
   Per rank and step (P parameters, W ranks), bytes that must cross NVLink:
       multimem:  in  = 4P/W (the switch returns one reduced copy of the owned slice)      out = 4P (W-1)/W is read OUT of
                  this GPU by the switch for the peers' slices; pushes: 2P/W bf16 out once (replicated), nothing (sharded)
       per-peer:  in  = 4P (W-1)/W,  out = 2P (W-1)/W (replicated) | 0 (sharded)
   and local HBM traffic 16P/W (master, m, v read + write) + 4P (own gradients read by somebody) + pushes landing.
   Roofline = max(NVLink bytes per direction / 770 GB/s measured peer bandwidth, HBM bytes / measured copy bandwidth).
2. all-gather path - the layer's QKV GEMM carrying the all-gather of the next layer's weights (`b200_gemm_ag`):
   roofline = max(GEMM FLOPs / measured sustained cuBLAS bf16 rate, (W-1)/W x layer bytes / 770 GB/s).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from bench import MODELS  # noqa: E402
from pretraining_llm_b200 import ops  # noqa: E402
from pretraining_llm_b200.models import Transformer  # noqa: E402
from pretraining_llm_b200.parallel import make_data_parallel  # noqa: E402
from pretraining_llm_b200.parallel.bootstrap import init_distributed  # noqa: E402
from pretraining_llm_b200.utils.flops import measured_peaks  # noqa: E402

NVLINK = 770e9      # measured peer-copy bandwidth per direction per GPU (profiling recipe); nominal 900e9

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="gpt2-small")
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--tokens", type=int, default=16384)
args = ap.parse_args()
ctx = init_distributed({"device": "cuda", "ddp_backend": "nccl"})
m = MODELS[args.model]
W = ctx.world
peaks = measured_peaks()
HBM = float(peaks.get("hbm_gbs", 6538.3)) * 1e9
TF = float(peaks.get("bf16_tflops_sustained", 1396.0)) * 1e12


def timed(fn, iters):
    ts = []
    for it in range(iters + 3):
        if W > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            ts.append(e0.elapsed_time(e1))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda")
    if W > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(d):
    if ctx.rank == 0:
        print(json.dumps(d), flush=True)


def build(mode):
    torch.manual_seed(0)
    model = Transformer(m["n_head"], m["n_embed"], m["context_length"], m["vocab_size"], N_BLOCKS=m["n_blocks"]).to(ctx.device)
    eng = model.engine
    P = eng.layout.total
    dp = make_data_parallel(eng, ctx, mode)
    eng.set_hyper(5e-4, 1)
    return model, eng, dp, P


# ---------------------------------------------------------------------------------------- 1. reduce path, both modes
for mode in ("ddp", "zero"):
    model, eng, dp, P = build(mode)
    if mode == "ddp":
        eng.ensure_grad()
        eng.grad.normal_(0, 1e-3)
    else:
        for b in (*dp.sym_gwin, dp.sym_head_g, dp.sym_embed_g):
            b.view(torch.float32).normal_(0, 1e-3)

    def step():
        dp.begin_step()
        dp.reduce_and_step()

    ms = timed(step, args.iters)
    dp.check_errors()
    mc = bool(dp.multicast)
    if mode == "zero":      # the block buckets alias two windows: every bucket is reduced, so P parameters move all the same
        pass
    nv_in = 4 * P / W if mc else 4 * P * (W - 1) / W
    push = 2 * P / W if mode == "ddp" else 0.0
    nv_out = (4 * P * (W - 1) / W if mc else 0.0) + (push if mc else push * (W - 1))
    hbm_bytes = 16 * P / W + 4 * P + 2 * P / W + (2 * P if mode == "ddp" else 0)
    t_nv = max(nv_in, nv_out) / NVLINK * 1e3
    t_hbm = hbm_bytes / HBM * 1e3
    roof = max(t_nv, t_hbm)
    emit({"path": "reduce(-scatter) + AdamW" + (" + weight push" if mode == "ddp" else ""), "kernel": f"{mode}: all {dp.nb} buckets",
          "model": args.model, "world": W, "params": P, "multimem": mc, "ms": round(ms, 4),
          "nvlink_in_MB": round(nv_in / 1e6, 1), "nvlink_out_MB": round(nv_out / 1e6, 1), "t_nvlink_ms": round(t_nv, 4),
          "t_hbm_ms": round(t_hbm, 4), "roofline_ms": round(roof, 4), "fraction_of_roofline": round(roof / ms, 3)})
    if mode == "ddp":
        del model, eng, dp
        torch.cuda.empty_cache()

# ---------------------------------------------------------------------------------------- 2. all-gather + GEMM (sharded engine from above)
lay = eng.layout
M, d = args.tokens, lay.d
a = (torch.randn(M, d, device=ctx.device) * 0.5).to(torch.bfloat16)
out = torch.empty(M, 3 * d, device=ctx.device, dtype=torch.bfloat16)
layer = 1 if lay.L > 2 else 0


def ag():
    dp.begin_pass()
    dp.gemm_fwd(layer, a, out)           # QKV GEMM of `layer` + all-gather of layer+1 (and of itself for layer 0)


def plain():
    ops.gemm(a, dp.W(f"attn_blocks.{layer}.attn.wqkv"), out=out)


dp.begin_pass()
dp.gemm_fwd(0, a, out)                   # fills window 0 / 1 once so that every operand is valid
ms_ag = timed(ag, args.iters)
ms_plain = timed(plain, args.iters)
dp.check_errors()
flops = 2.0 * M * 3 * d * d
layer_bytes = 2 * dp.plan.size(dp.plan.block_bucket(layer + 1))
t_c = flops / TF * 1e3
t_nv = layer_bytes * (W - 1) / W / NVLINK * 1e3
roof = max(t_c, t_nv)
emit({"path": "all-gather + GEMM", "kernel": "gemm2<256,AG>: QKV GEMM carrying the next layer's all-gather", "model": args.model,
      "world": W, "M": M, "N": 3 * d, "K": d, "gathered_MB": round(layer_bytes / 1e6, 1), "ms": round(ms_ag, 4),
      "ms_same_gemm_without_gather": round(ms_plain, 4), "t_compute_ms": round(t_c, 4), "t_nvlink_ms": round(t_nv, 4),
      "roofline_ms": round(roof, 4), "fraction_of_roofline": round(roof / ms_ag, 3)})
ctx.shutdown()
