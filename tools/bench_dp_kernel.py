"""Roofline of the fused gradient path (P2P reduce-scatter + AdamW + all-gather kernel, csrc/comm/dp.cu) in isolation:
all buckets of the model, nothing else running, device-timed, max over ranks.

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_dp_kernel.py [--model gpt2-small|1.3b] [--dp-mode ddp|zero]

Bytes per rank and step (P parameters, W ranks):   NVLink in  = (W-1)/W * 4P   (fp32 gradient slices pulled from peers)
                                                    NVLink out = (W-1)/W * 2P   (bf16 weights pushed; "zero" mode keeps matrices local)
                                                    HBM        = 4P (own grads) ... + 16P/W (m, v, master read+write) + pushes landing
Roofline = max(NVLink bytes / 900 GB/s, HBM bytes / measured copy bandwidth)."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from bench import MODELS
from pretraining_llm_b200.models import Transformer
from pretraining_llm_b200.parallel.bootstrap import init_distributed
from pretraining_llm_b200.parallel.dp import DataParallel

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="gpt2-small")
ap.add_argument("--dp-mode", default="ddp")
ap.add_argument("--iters", type=int, default=20)
args = ap.parse_args()
ctx = init_distributed({"device": "cuda", "ddp_backend": "nccl"})
m = MODELS[args.model]
torch.manual_seed(0)
model = Transformer(m["n_head"], m["n_embed"], m["context_length"], m["vocab_size"], N_BLOCKS=m["n_blocks"]).to(ctx.device)
eng = model.engine
eng.dp = DataParallel(eng, ctx, mode=args.dp_mode)
dp = eng.dp
eng.set_hyper(5e-4, 1)
eng.grad.normal_(0, 1e-3)
W, P = ctx.world, eng.flat.numel()
ts = []
for it in range(args.iters + 3):
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    dp.begin_step()
    dp.reduce_and_step()
    e1.record()
    torch.cuda.synchronize()
    if it >= 3:
        ts.append(e0.elapsed_time(e1))
dp.check_errors()
t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
ms = float(t.item())
if ctx.rank == 0:
    peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else {}
    hbm = float(peaks.get("hbm_gbs", peaks.get("copy_gbs", 6538.0))) if isinstance(peaks, dict) else 6538.0
    nv_in = (W - 1) / W * 4 * P
    nv_out = (W - 1) / W * 2 * P if args.dp_mode == "ddp" else 0.0
    hbm_bytes = 4 * P + 16 * P / W + 2 * P + 2 * P / W      # own-slice grads of all ranks' buffers are remote; local: full grads read by peers + state + weights
    t_nv = max(nv_in, nv_out) / 900e9 * 1e3
    t_hbm = hbm_bytes / (hbm * 1e9) * 1e3
    roof = max(t_nv, t_hbm)
    print(json.dumps({"kernel": "dp_reduce_adamw (all buckets)", "model": args.model, "mode": args.dp_mode, "world": W, "params": P,
                      "ms": round(ms, 4), "nvlink_in_MB": round(nv_in / 1e6, 1), "nvlink_out_MB": round(nv_out / 1e6, 1),
                      "t_nvlink_ms": round(t_nv, 4), "t_hbm_ms": round(t_hbm, 4), "roofline_ms": round(roof, 4),
                      "fraction_of_roofline": round(roof / ms, 3)}))
ctx.shutdown()
