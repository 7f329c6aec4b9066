"""Generation benchmark (BASELINE.json config 5): autoregressive sampling, batch 32, 512 new tokens,
~1.3B-shaped checkpoint (random init), tokens/s = batch * new_tokens / device time.

    python tools/bench_generate.py [--impl ours|reference] [--model 1.3b] [--batch 32] [--new 512] [--prompt 16]

ours      : Transformer.generate on the native kernels (flash prefill + KV cache + graph-replayed decode step)
reference : the unmodified reference Transformer.generate (fp32, full re-forward per token, torch.multinomial)
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import MODELS  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--impl", default="ours")
ap.add_argument("--model", default="1.3b")
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--new", type=int, default=512)
ap.add_argument("--prompt", type=int, default=16)
ap.add_argument("--ref-new", type=int, default=0, help="reference only: time this many tokens and extrapolate nothing")
ap.add_argument("--ref-autocast", action="store_true",
                help="reference only, for information: run its generate under bf16 autocast (its script runs fp32)")
args = ap.parse_args()
m = MODELS[args.model]
T, V = m["context_length"], m["vocab_size"]
torch.manual_seed(0)
prompt = torch.randint(0, V, (args.batch, args.prompt), device="cuda")

if args.impl == "ours":
    from pretraining_llm_b200.models import Transformer
    model = Transformer(m["n_head"], m["n_embed"], T, V, N_BLOCKS=m["n_blocks"]).cuda().eval()
    gen = lambda n: model.generate(prompt, n, generator=torch.Generator().manual_seed(1))
    new = args.new
else:
    from baseline.ref_import import import_reference_models
    R = import_reference_models()
    model = R.Transformer(m["n_head"], m["n_embed"], T, V, m["n_blocks"]).cuda().eval()
    def gen(n):
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16, enabled=args.ref_autocast):
            return model.generate(prompt, n)
    new = args.ref_new or args.new

gen(4)                     # warm-up (allocations, graph capture, cuBLAS handles)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
out = gen(new)
e1.record()
torch.cuda.synchronize()
wall = time.perf_counter() - t0
ms = e0.elapsed_time(e1)
assert out.shape == (args.batch, args.prompt + new)
print(json.dumps({"metric": "generation tokens/sec", "impl": args.impl, "value": args.batch * new / (ms / 1e3),
                  "unit": "tokens/s", "batch": args.batch, "new_tokens": new, "prompt": args.prompt,
                  "ms_total": ms, "ms_per_token_step": ms / new, "wall_s": wall,
                  "config": {"model": f"{args.model}-shaped decoder {m['n_blocks']}L d={m['n_embed']} H={m['n_head']} ctx={T}",
                             "dtype": "bf16 weights/KV cache, fp32 softmax" if args.impl == "ours" else
                             ("bf16 autocast (information only)" if args.ref_autocast else "fp32 (reference default)"),
                             "data": "random prompt, random-init weights"}}))
