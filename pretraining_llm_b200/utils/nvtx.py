"""NVTX ranges per kernel family, off unless ``B200_NVTX=1``.

With the switch on, every ``ops`` wrapper is bracketed by an NVTX range named after its family (gemm, attn_fwd,
ln_bwd, dp_reduce_adamw, ...) and the engine marks forward / backward / optimizer, so an Nsight Systems or
``ncu --nvtx`` capture groups the launches the way ``profiles/step_launches.md`` does.  Off, ``range`` is a
no-op context manager that costs one attribute test per call (the ranges are host-side only: nothing is
captured into CUDA graphs either way).
"""
from __future__ import annotations

import contextlib
import os

ENABLED = os.environ.get("B200_NVTX", "0") == "1"
_null = contextlib.nullcontext()


def range(name: str):  # noqa: A001 - mirrors torch.cuda.nvtx.range
    if not ENABLED:
        return _null
    import torch
    return torch.cuda.nvtx.range(name)


def push(name: str) -> None:
    if ENABLED:
        import torch
        torch.cuda.nvtx.range_push(name)


def pop() -> None:
    if ENABLED:
        import torch
        torch.cuda.nvtx.range_pop()
