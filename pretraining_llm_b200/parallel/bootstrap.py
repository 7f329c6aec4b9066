"""Process-group bootstrap: one process per GPU, ``torch.distributed`` for the plumbing.

Replaces the module-import side effects of reference ``scripts/train_transformer.py:14-33``
(``init_process_group(backend=config['ddp_backend'])``, RANK/LOCAL_RANK/WORLD_SIZE from the
environment, ``cuda:{local_rank}``).  NCCL (or gloo on CPU) is used only for rendezvous, handle
exchange, tiny scalar reductions and as the measured baseline; gradient traffic goes through the
kernels in ``csrc/comm`` over peer-mapped memory (``parallel/dp.py``).
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    local_rank: int = 0
    world: int = 1
    device: str = "cpu"
    initialized_here: bool = False

    @property
    def master(self) -> bool:
        return self.rank == 0

    @property
    def ddp(self) -> bool:
        return self.world > 1

    # tiny host-driven collectives (metrics only - never on the gradient path)
    def all_reduce_mean(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            t = t.clone()
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            t /= self.world
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            t = t.clone()
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    def all_reduce_grads(self, model) -> None:
        """CPU / plain-PyTorch path only: average ``.grad`` across ranks (gloo or NCCL)."""
        if self.world <= 1:
            return
        for p in model.parameters():
            if p.grad is not None:
                dist.all_reduce(p.grad, op=dist.ReduceOp.SUM)
                p.grad /= self.world

    def barrier(self) -> None:
        if self.world > 1:
            dist.barrier()

    def shutdown(self) -> None:
        if self.initialized_here and dist.is_initialized():
            dist.destroy_process_group()


def init_distributed(config: Optional[dict] = None, device: Optional[str] = None,
                     timeout_s: int = 600) -> DistContext:
    """Join the job described by the environment (torchrun / ``python -m torch.distributed.run``)."""
    cfg = config or {}
    dev_pref = device or cfg.get("device", "cuda")
    use_cuda = dev_pref.startswith("cuda") and torch.cuda.is_available()
    if int(os.environ.get("RANK", -1)) == -1:
        return DistContext(device=dev_pref if (use_cuda or not dev_pref.startswith("cuda")) else "cpu")
    rank = int(os.environ["RANK"])
    local_rank = int(os.environ.get("LOCAL_RANK", rank))
    world = int(os.environ["WORLD_SIZE"])
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    backend = cfg.get("ddp_backend", "nccl") if use_cuda else "gloo"
    initialized_here = False
    if use_cuda:
        torch.cuda.set_device(local_rank)
    if not dist.is_initialized():
        kw = {}
        if use_cuda:
            kw["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend=backend, timeout=datetime.timedelta(seconds=timeout_s), **kw)
        initialized_here = True
    return DistContext(rank=rank, local_rank=local_rank, world=world,
                       device=f"cuda:{local_rank}" if use_cuda else "cpu", initialized_here=initialized_here)
