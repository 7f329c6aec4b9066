"""Multi-GPU check THROUGH the Trainer (run under torchrun): the paths tools/dp_check.py does not touch.

  1. gradient accumulation with data parallelism (`grad_accum_steps=2`, replicated mode): the fused reduce+AdamW
     kernels of the last micro-step must see the step's learning rate / bias corrections (staged before the loop);
  2. `Trainer.save_checkpoint` with the masters / moments spread over the ranks (replicated, vocab-parallel and sharded
     modes): the file rank 0 writes is reloaded into a plain single-GPU model and compared with a single-GPU oracle
     that saw the concatenation of all ranks' batches;
  3. resume: a second Trainer loads that file (`resume_from`), trains one more step, and still matches the oracle.

DP_MODE=ddp|zero, DP_VP=0|1 select the mode; prints DP_TRAINER_CHECK_OK on success.
"""
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pretraining_llm_b200.config import default_config  # noqa: E402
from pretraining_llm_b200.models import Transformer  # noqa: E402
from pretraining_llm_b200.parallel.bootstrap import init_distributed  # noqa: E402
from pretraining_llm_b200.runtime.trainer import Trainer  # noqa: E402

CFG = dict(n_head=4, n_embed=256, context_length=128, vocab_size=1000)
L, B, T, LR = 3, 4, 128, 1e-3
MODE = os.environ.get("DP_MODE", "ddp")
VP = os.environ.get("DP_VP", "0") == "1"
ACCUM = 2                                      # gradient accumulation in every mode (sharded: owner-side accumulators)
STEPS = 3


def batch(rank, step, micro):
    g = torch.Generator().manual_seed(100000 * step + 100 * micro + rank)
    xy = torch.randint(0, CFG["vocab_size"], (B, T + 1), generator=g)
    return xy[:, :-1].contiguous().cuda(), xy[:, 1:].contiguous().cuda()


def stream(rank, first_step=0):
    step = first_step
    while True:
        for micro in range(ACCUM):
            yield batch(rank, step, micro)
        step += 1


def oracle_step(oracle, W, step):
    """One optimizer step of the single-GPU oracle on everything the W ranks saw in `step`."""
    eng = oracle.engine
    eng.ensure_grad()
    eng.sync_weights()
    k = 0
    for micro in range(ACCUM):
        for r in range(W):
            x, y = batch(r, step, micro)
            w, _, _ = eng.forward(x, y, train=True)
            eng.backward(w, grad_scale=1.0 / (W * ACCUM), accumulate=k > 0)
            k += 1
    eng.set_hyper(LR)
    eng.optimizer_step()


def main():
    ctx = init_distributed({"device": "cuda", "ddp_backend": "nccl"})
    W, r = ctx.world, ctx.rank
    tmp = tempfile.gettempdir()
    path = os.path.join(tmp, f"b200_dp_trainer_check_{MODE}_{int(VP)}.pt")
    conf = dict(default_config, dp_mode=MODE, vocab_parallel=VP, grad_accum_steps=ACCUM, t_train_steps=STEPS, t_lr=LR,
                t_eval_steps=10 ** 9, t_out_path=path, dtype="bfloat16", use_cuda_graph=True, lr_schedule="reference")
    torch.manual_seed(0)
    model = Transformer(**CFG, N_BLOCKS=L).to(ctx.device)
    oracle = None
    if r == 0:
        torch.manual_seed(0)
        oracle = Transformer(**CFG, N_BLOCKS=L).to(ctx.device)
    it = stream(r)
    tr = Trainer(model, it, None, config=conf, dist_ctx=ctx)
    for step in range(STEPS):
        xb, yb = next(it)
        tr._train_step(xb, yb, LR)
        tr.step += 1
        if oracle is not None:
            oracle_step(oracle, W, step)
    torch.cuda.synchronize()
    model.engine.dp.check_errors()
    tr.save_checkpoint(path)                       # collective
    dist.barrier()
    ok = True
    if r == 0:
        ck = torch.load(path, map_location="cuda", weights_only=False)
        plain = Transformer(**CFG, N_BLOCKS=L).to(ctx.device)
        plain.load_state_dict(ck["model_state_dict"])
        dw = (plain.flat_params - oracle.engine.flat).abs()
        lay = plain.layout
        m = torch.zeros(lay.total)
        v = torch.zeros(lay.total)
        step_saved, _ = lay.from_reference_optimizer_state(ck["optimizer_state_dict"], m, v)
        dm = (m.cuda() - oracle.engine.exp_avg).abs().max().item()
        print(f"[{MODE}{'+vp' if VP else ''} W={W} accum={ACCUM}] checkpoint vs oracle: mean|dW| {dw.mean().item():.3e} "
              f"({dw.mean().item() / (LR * STEPS):.2e} of lr*steps)  max|dm| {dm:.3e}  saved step {step_saved}")
        ok = dw.mean().item() < 5e-3 * LR * STEPS and dm < 1e-4 and step_saved == STEPS
    # ---- resume in a fresh trainer, one more step
    torch.manual_seed(1)
    model2 = Transformer(**CFG, N_BLOCKS=L).to(ctx.device)
    conf2 = dict(conf, resume_from=path, t_train_steps=STEPS + 1)
    it2 = stream(r, first_step=STEPS)
    tr2 = Trainer(model2, it2, None, config=conf2, dist_ctx=ctx)
    tr2.load_checkpoint(path)
    assert tr2.step == STEPS
    xb, yb = next(it2)
    tr2._train_step(xb, yb, LR)
    torch.cuda.synchronize()
    model2.engine.dp.check_errors()
    model2.engine.dp.gather_full_state()
    if r == 0:
        oracle_step(oracle, W, STEPS)
        dw2 = (model2.engine.flat - oracle.engine.flat).abs().mean().item()
        print(f"after resume + 1 step: mean|dW| {dw2:.3e} ({dw2 / (LR * (STEPS + 1)):.2e} of lr*steps)")
        ok = ok and dw2 < 5e-3 * LR * (STEPS + 1)
        assert ok, "Trainer-level data-parallel check failed"
        print("DP_TRAINER_CHECK_OK")
        os.remove(path)
    ctx.shutdown()


if __name__ == "__main__":
    main()
