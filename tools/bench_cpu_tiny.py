"""BASELINE.json config 1: the tiny decoder (4 layers, d=256, 4 heads, context 256, vocab 50304, batch 8) trained on
synthetic tokens on the CPU in a single process - the reference's own model and recipe against this framework's CPU
path (plain PyTorch over the packed parameter store), same seed, same batches, bf16 autocast, AdamW(lr=5e-4)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from baseline.ref_import import import_reference_models  # noqa: E402
from pretraining_llm_b200.config import default_config  # noqa: E402
from pretraining_llm_b200.models import Transformer  # noqa: E402
from pretraining_llm_b200.runtime.trainer import Trainer  # noqa: E402

H, d, ctx, V, L, B = 4, 256, 256, 50304, 4, 8
STEPS, WARM = 6, 2
torch.manual_seed(0)
g = torch.Generator().manual_seed(1)
data = []
for _ in range(STEPS + WARM):
    xy = torch.randint(0, V, (B, ctx + 1), generator=g)
    data.append((xy[:, :-1].contiguous(), xy[:, 1:].contiguous()))

R = import_reference_models()
torch.manual_seed(1234)
ref = R.Transformer(H, d, ctx, V, L)
opt = torch.optim.AdamW(ref.parameters(), lr=5e-4)


def ref_step(x, y):
    with torch.autocast("cpu", dtype=torch.bfloat16):
        _, loss = ref(x, y)
    loss.backward()
    opt.step()
    opt.zero_grad(set_to_none=True)
    return float(loss)


torch.manual_seed(1234)
ours = Transformer(H, d, ctx, V, N_BLOCKS=L)
ours.load_state_dict(ref.state_dict())
tr = Trainer(ours, iter(()), None, config=dict(default_config, dtype="bfloat16", t_lr=5e-4, grad_accum_steps=1))


def our_step(x, y):
    return float(tr._train_step(x, y, 5e-4))


out = {}
for name, fn in (("reference", ref_step), ("ours", our_step)):
    losses = []
    for i, (x, y) in enumerate(data):
        if i == WARM:
            t0 = time.perf_counter()
        losses.append(fn(x, y))
    dt = (time.perf_counter() - t0) / STEPS
    out[name] = {"ms_per_step": round(dt * 1e3, 1), "tokens_per_s": round(B * ctx / dt, 1), "losses": [round(v, 4) for v in losses]}
out["max_loss_diff"] = max(abs(a - b) for a, b in zip(out["reference"]["losses"], out["ours"]["losses"]))
out["config"] = f"tiny {L}L d={d} H={H} ctx={ctx} V={V} batch {B}, CPU, {torch.get_num_threads()} threads, bf16 autocast, eager"
print(json.dumps(out))
