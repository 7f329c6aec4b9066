"""Reference arm of bench.py: the UNMODIFIED reference tree (baseline/_ref) on the same config.

Everything on the measured path is the reference's own code: ``src.models.transformer.Transformer``,
the module-level DDP bootstrap of ``scripts/train_transformer.py`` (``init_process_group`` from the
RANK/LOCAL_RANK/WORLD_SIZE environment), its ``Trainer`` (``_get_lr``, ``_train_step`` = bf16 autocast
forward + backward) followed by the three optimizer lines of ``Trainer.train``
(``scaler.step; scaler.update; zero_grad(set_to_none=True)``), ``torch.optim.AdamW`` defaults,
``torch.compile`` when ``TORCH_COMPILE`` is "1" (the reference default) and
``data_loader.get_batch_iterator`` for the end-to-end number.

The reference cannot run as shipped (missing config keys ``ddp_backend``/``dtype``/
``val_path``, ``n_blocks`` vs ``N_BLOCKS``).  The repairs are applied to the IN-MEMORY config dict and
by building the model with the right keyword; no reference file is edited.  Gradient sync runs on
every step (the as-written alternate-step sync makes replicas diverge, D8) and the model is built on
every rank from the same seed.  None of this repo's models, kernels or engine are imported here.
"""
from __future__ import annotations

import importlib
import json
import os
import sys
import tempfile
import time


def _import_reference(root: str):
    """Put the reference tree first on sys.path and drop this repo's same-named facades."""
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for k in [k for k in sys.modules if k.split(".")[0] in ("src", "config", "data_loader", "scripts")]:
        del sys.modules[k]
    sys.path[:] = [root] + [p for p in sys.path if os.path.abspath(p or ".") != repo]
    cfg_mod = importlib.import_module("config.config")
    return cfg_mod


def run_reference_bench(args, root: str, m: dict, lr: float) -> None:
    import numpy as np
    import torch

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    B = args.batch or (8 if args.model == "1.3b" else 16)
    T, V = m["context_length"], m["vocab_size"]

    cfg_mod = _import_reference(root)
    config = cfg_mod.default_config
    tmpdir = tempfile.mkdtemp(prefix="refbench_")
    data_path = os.path.join(tmpdir, f"synthetic_{rank}.bin")
    rng = np.random.default_rng(100 + rank)
    rng.integers(0, V, size=4 * B * (T + 1) * max(world, 1) + 4096, dtype=np.uint16).tofile(data_path)
    config.update({
        # model / training shape of this benchmark
        "vocab_size": V, "context_length": T, "n_embed": m["n_embed"], "n_head": m["n_head"],
        "n_blocks": m["n_blocks"], "t_batch_size": B, "t_lr": lr, "t_train_steps": 10 ** 9,
        "train_path": data_path, "dev_path": data_path,
        # keys the reference's own script reads but its config never defines
        "ddp_backend": "nccl", "dtype": "bfloat16", "val_path": data_path,
        "t_out_path": os.path.join(tmpdir, "ref.pt"),
    })
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if world > 1:
        torch.cuda.set_device(local)
    # importing the script runs the reference's own DDP bootstrap (train_transformer.py:14-33)
    ref = importlib.import_module("scripts.train_transformer")
    from src.models.transformer import Transformer
    from data_loader.data_loader import get_batch_iterator

    device = ref.device if world > 1 else "cuda"
    torch.manual_seed(1234)
    model = Transformer(n_head=m["n_head"], n_embed=m["n_embed"], context_length=T, vocab_size=V,
                        N_BLOCKS=m["n_blocks"]).to(device)
    compiled = False
    if ref.compile_model:
        try:
            model = torch.compile(model)
            compiled = True
        except Exception as e:  # pragma: no cover
            print(f"[reference] torch.compile unavailable: {e}", file=sys.stderr)
    if world > 1:
        from torch.nn.parallel import DistributedDataParallel as DDP
        # stock path: DDP(model, device_ids=[...]) exactly as the reference writes it.  REF_BROADCAST_BUFFERS=0 is an
        # information-only variant: it stops DDP from re-broadcasting the 576 MiB of tril mask buffers
        # every step; the driver's arm never sets it.
        ddp_kw = {} if os.environ.get("REF_BROADCAST_BUFFERS", "1") == "1" else {"broadcast_buffers": False}
        model = DDP(model, device_ids=[ref.ddp_local_rank], **ddp_kw)
    optimizer = torch.optim.AdamW(model.parameters(), lr=config["t_lr"])
    train_it = get_batch_iterator(data_path, B, T, device, world > 1, rank, world)
    trainer = ref.Trainer(model, train_it, train_it, optimizer)

    g = torch.Generator().manual_seed(100 + rank)
    staged = []
    for _ in range(4):
        xy = torch.randint(0, V, (B, T + 1), generator=g)
        staged.append((xy[:, :-1].contiguous().to(device), xy[:, 1:].contiguous().to(device)))

    def one_step(xb, yb):
        for group in optimizer.param_groups:                  # train_transformer.py:79-81 (constant lr here)
            group["lr"] = lr
        if world > 1:
            model.require_backward_grad_sync = True           # repaired: sync every step (D8)
        loss = trainer._train_step(xb, yb)                    # :64-70
        trainer.scaler.step(optimizer)                        # :92-94
        trainer.scaler.update()
        optimizer.zero_grad(set_to_none=True)
        return loss

    import torch.distributed as dist

    def barrier_sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.append(repo)
    from pretraining_llm_b200.utils.clocks import ClockSampler  # measurement utility only (nvidia-smi reader)
    clocks = ClockSampler(local, period_ms=100).__enter__()

    warm = max(args.warmup, 3)
    try:
        for i in range(warm):
            one_step(*staged[i % 4])
    except Exception as e:
        if not compiled:
            raise
        # Inductor/Triton could not build on this box: fall back to the reference's eager mode
        print(f"[reference] compiled step failed ({type(e).__name__}); falling back to TORCH_COMPILE=0",
              file=sys.stderr)
        compiled = False
        base = model.module if world > 1 else model
        base = getattr(base, "_orig_mod", base)
        model = base
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            model = DDP(base, device_ids=[ref.ddp_local_rank])
        optimizer = torch.optim.AdamW(model.parameters(), lr=config["t_lr"])
        trainer = ref.Trainer(model, train_it, train_it, optimizer)
        for i in range(warm):
            one_step(*staged[i % 4])

    barrier_sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks.mark_start()
    e0.record()
    for i in range(args.steps):
        loss = one_step(*staged[i % 4])
    e1.record()
    barrier_sync()
    clocks.mark_end()
    time.sleep(0.15)
    clocks.__exit__(None, None, None)
    ms = max_over_ranks(e0.elapsed_time(e1))
    final_loss = float(loss.item())
    tokens = B * T * world * args.steps

    e2e = None
    if not args.skip_e2e:
        for _ in range(2):
            xb, yb = next(train_it)
            one_step(xb, yb).item()
        barrier_sync()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            xb, yb = next(train_it)          # the reference's loader: memmap -> stack -> (pinned) H2D
            _ = one_step(xb, yb).item()      # D2H read of the loss
        barrier_sync()
        dt = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": tokens / dt, "unit": "tokens/s", "h2d_bytes_per_step": 2 * B * T * 8 * world,
               "d2h_bytes_per_step": 4 * world, "ms_per_step": dt * 1e3 / args.steps}
    if rank == 0:
        print(json.dumps({
            "metric": "train tokens/sec (device-timed, max over ranks)", "value": tokens / (ms / 1e3),
            "unit": "tokens/s", "impl": "reference", "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 autocast", "data": "synthetic random tokens, random-init weights",
            "config": {"model": f"{args.model}-shaped decoder {m['n_blocks']}L d={m['n_embed']} H={m['n_head']} "
                                f"seq={T} vocab={V}", "global_batch": B * world, "seq_len": T,
                       "parallelism": f"dp{world}",
                       "optimizer": "AdamW (torch defaults), every step",
                       "l2": "no flush: per-step working set (activations + weights + optimizer state, >4 GB) "
                             "is far larger than the 126 MB L2"},
            "arm_details": {"torch_compile": compiled,
                            "ddp": ("torch DDP/NCCL, sync every step" + ("" if os.environ.get("REF_BROADCAST_BUFFERS", "1") == "1"
                                                                             else ", broadcast_buffers=False (variant)"))
                                   if world > 1 else "none",
                            "reference_root": root},
            "final_loss": final_loss, "clocks": clocks.summary(), "e2e": e2e, "gpu_launches": 0,
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()
