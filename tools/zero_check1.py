"""Single-GPU check of the sharded ("zero") data-parallel machinery with world = 1: the all-gather jobs copy the own
shard into the rolling windows, the reduce kernel runs with one rank, the embedding rows come out of the master
shard.  Everything the multi-GPU path does except the NVLink hops, against the plain single-GPU engine."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from pretraining_llm_b200.models import Transformer  # noqa: E402
from pretraining_llm_b200.parallel import make_data_parallel  # noqa: E402
from pretraining_llm_b200.parallel.bootstrap import DistContext  # noqa: E402


def run(cfg, L, B, T, steps, lr, graph):
    torch.manual_seed(0)
    a = Transformer(**cfg, N_BLOCKS=L).to("cuda")
    torch.manual_seed(0)
    b = Transformer(**cfg, N_BLOCKS=L).to("cuda")
    make_data_parallel(a.engine, DistContext(rank=0, local_rank=0, world=1, device="cuda:0"), "zero")
    worst = 0.0
    for s in range(steps):
        g = torch.Generator().manual_seed(s)
        xy = torch.randint(0, cfg["vocab_size"], (B, T + 1), generator=g)
        x, y = xy[:, :-1].contiguous().cuda(), xy[:, 1:].contiguous().cuda()
        la = float(a.engine.train_step(x, y, lr, use_graph=graph).item())
        lb = float(b.engine.train_step(x, y, lr, use_graph=graph).item())
        worst = max(worst, abs(la - lb))
    a.engine.dp.check_errors()
    a.engine.dp.gather_full_state()
    dw = (a.engine.flat - b.engine.flat).abs()
    dm = (a.engine.exp_avg - b.engine.exp_avg).abs().max().item()
    print(f"L={L} d={cfg['n_embed']} graph={graph}: max|dloss| {worst:.2e}  mean|dW| {dw.mean().item():.2e} "
          f"max|dW| {dw.max().item():.2e}  max|dm| {dm:.2e}")
    assert worst < 2e-3 and dm < 1e-4 and dw.mean().item() < 5e-3 * lr * steps, "sharded path diverges from the oracle"
    a.engine.dp.release_full()
    # evaluation forward runs sharded too
    _, _, le = a.engine.forward(x, y, train=False)
    _, _, lo = b.engine.forward(x, y, train=False)
    assert abs(float(le) - float(lo)) < 2e-3, (float(le), float(lo))


def run_accum(cfg, L, B, T, steps, lr, accum):
    """Gradient accumulation in sharded mode (each micro-step reduce-scattered into the owners' fp32 accumulators,
    AdamW on the last) against the plain engine accumulating in its flat gradient buffer."""
    torch.manual_seed(0)
    a = Transformer(**cfg, N_BLOCKS=L).to("cuda")
    torch.manual_seed(0)
    b = Transformer(**cfg, N_BLOCKS=L).to("cuda")
    make_data_parallel(a.engine, DistContext(rank=0, local_rank=0, world=1, device="cuda:0"), "zero")
    worst = 0.0
    for s in range(steps):
        for eng in (a.engine, b.engine):
            eng.set_hyper(lr)
            if eng.dp is not None:
                eng.dp.begin_step()
            tot = 0.0
            for k in range(accum):
                g = torch.Generator().manual_seed(100 * s + k)
                xy = torch.randint(0, cfg["vocab_size"], (B, T + 1), generator=g)
                x, y = xy[:, :-1].contiguous().cuda(), xy[:, 1:].contiguous().cuda()
                w, _, loss = eng.forward(x, y, train=True)
                eng.backward(w, grad_scale=1.0 / accum, accumulate=k > 0, dp_sync=(k == accum - 1))
                tot += float(loss)
            eng.optimizer_step()
            if eng is a.engine:
                la = tot
            else:
                worst = max(worst, abs(la - tot) / accum)
    a.engine.dp.check_errors()
    a.engine.dp.gather_full_state()
    dw = (a.engine.flat - b.engine.flat).abs()
    dm = (a.engine.exp_avg - b.engine.exp_avg).abs().max().item()
    print(f"accum={accum} L={L}: max|dloss| {worst:.2e}  mean|dW| {dw.mean().item():.2e}  max|dm| {dm:.2e}")
    assert worst < 2e-3 and dm < 1e-4 and dw.mean().item() < 5e-3 * lr * steps, "sharded accumulation diverges from the oracle"
    a.engine.dp.release_full()


if __name__ == "__main__":
    run_accum(dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000), 3, 4, 128, 3, 1e-3, 3)
    run(dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000), 3, 4, 128, 5, 1e-3, False)
    run(dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000), 4, 4, 128, 6, 1e-3, True)
    run(dict(n_head=12, n_embed=768, context_length=1024, vocab_size=50304), 2, 4, 1024, 3, 5e-4, True)
    print("ZERO_CHECK1_OK")
