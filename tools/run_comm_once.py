"""One eager step of the data-parallel machinery with ONE rank, bracketed by cudaProfilerStart/Stop, so that ncu can
capture the communication kernels on a single GPU (ncu must never wrap a multi-rank command): the fused reduce+AdamW
kernels (replicated and sharded), the all-gather GEMMs (forward K-major and backward MN-major variants), the
owner-read embedding forward, the bucket wait.  With world = 1 the peer pointers alias the local buffers: same code,
same instruction mix, no NVLink hop.

    python tools/run_comm_once.py --mode zero|ddp [--model gpt2-small]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import MODELS  # noqa: E402
from pretraining_llm_b200.models import Transformer  # noqa: E402
from pretraining_llm_b200.parallel import make_data_parallel  # noqa: E402
from pretraining_llm_b200.parallel.bootstrap import DistContext  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mode", default="zero")
ap.add_argument("--model", default="gpt2-small")
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--layers", type=int, default=4)
args = ap.parse_args()
m = MODELS[args.model]
T, V = m["context_length"], m["vocab_size"]
torch.manual_seed(0)
model = Transformer(m["n_head"], m["n_embed"], T, V, N_BLOCKS=args.layers).cuda()
eng = model.engine
make_data_parallel(eng, DistContext(rank=0, local_rank=0, world=1, device="cuda:0"), args.mode)
x = torch.randint(0, V, (args.batch, T), device="cuda")
y = torch.randint(0, V, (args.batch, T), device="cuda")
for _ in range(2):
    eng.train_step(x, y, 5e-4, use_graph=False)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
loss = eng.train_step(x, y, 5e-4, use_graph=False)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
eng.dp.check_errors()
print(f"{args.mode} step ok, loss {loss.item():.4f}")
