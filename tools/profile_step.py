"""Run a few eager training steps of the flagship config and bracket ONE step with
cudaProfilerStart/Stop so `ncu --profile-from-start off` captures exactly its kernels.
Also prints a CUDA-event time per op family (eager, serialised) when run without ncu."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import MODELS  # noqa: E402
from pretraining_llm_b200.models import Transformer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="gpt2-small")
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--steps", type=int, default=3)
args = ap.parse_args()
m = MODELS[args.model]
T, V = m["context_length"], m["vocab_size"]
torch.manual_seed(0)
model = Transformer(m["n_head"], m["n_embed"], T, V, N_BLOCKS=m["n_blocks"]).cuda()
eng = model.engine
x = torch.randint(0, V, (args.batch, T), device="cuda")
y = torch.randint(0, V, (args.batch, T), device="cuda")
for _ in range(args.steps):
    eng.train_step(x, y, 5e-4, use_graph=False)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
loss = eng.train_step(x, y, 5e-4, use_graph=False)
e1.record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(f"eager step {e0.elapsed_time(e1):.3f} ms, loss {loss.item():.4f}")
