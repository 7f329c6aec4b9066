"""Bracket two eager decode steps of the 1.3B-shaped model with cudaProfilerStart/Stop (for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import MODELS
from pretraining_llm_b200.models import Transformer
from pretraining_llm_b200.runtime.generation import Generator
m = MODELS["1.3b"]
torch.manual_seed(0)
model = Transformer(m["n_head"], m["n_embed"], m["context_length"], m["vocab_size"], N_BLOCKS=m["n_blocks"]).cuda().eval()
eng = model.engine
gen = Generator(eng)
prompt = torch.randint(0, m["vocab_size"], (32, 256), device="cuda")
with torch.no_grad():
    out = gen.generate(prompt, 4, use_graph=False)
    db = gen._buffers(32)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    gen._decode_step(db, 1.0, 1)
    e1.record()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
print(f"eager decode step {e0.elapsed_time(e1):.3f} ms")
