"""Hyper-parameter surface of the framework.

Same 17 keys as the reference's ``default_config`` (reference ``config/config.py:29-47``)
with the same default values, plus the keys the reference's scripts read but never define
(``val_path``, ``dtype``, ``ddp_backend``, ``dataset_name``, ``tokenizer_name`` -- SURVEY
defects D2-D5) and the knobs this framework adds (``seed``, ``grad_accum_steps``,
``dp_mode``, ...).

Overrides without editing the file: ``B200_CFG_<KEY>=value`` environment variables, or
``apply_overrides(cfg, ["key=value", ...])`` from a CLI.  Named presets for the
benchmark shapes of BASELINE.json live in ``PRESETS``.
"""
from __future__ import annotations

import ast
import os
from typing import Iterable

# --- model (reference config/config.py:4-8) ---------------------------------------------
VOCAB_SIZE = 50304
CONTEXT_LENGTH = 512
N_EMBED = 2048
N_HEAD = 16
N_BLOCKS = 64

# --- data (reference :11-12; ".h5" there, but the loader memmaps raw uint16 -> ".bin") ----
TRAIN_PATH = "data/train/pile_train.bin"
DEV_PATH = "data/val/pile_dev.bin"

# --- training (reference :15-23) -----------------------------------------------------------
T_BATCH_SIZE = 32
T_CONTEXT_LENGTH = 16
T_TRAIN_STEPS = 200000
T_EVAL_STEPS = 1000
T_EVAL_ITERS = 250
T_LR_DECAY_STEP = 50000
T_LR = 5e-4
T_LR_DECAYED = 5e-5
T_OUT_PATH = "models/transformer_B.pt"

DEVICE = "cuda"

default_config = {
    # the reference's 17 keys, same names, same order
    "vocab_size": VOCAB_SIZE,
    "context_length": CONTEXT_LENGTH,
    "n_embed": N_EMBED,
    "n_head": N_HEAD,
    "n_blocks": N_BLOCKS,
    "train_path": TRAIN_PATH,
    "dev_path": DEV_PATH,
    "t_batch_size": T_BATCH_SIZE,
    "t_context_length": T_CONTEXT_LENGTH,
    "t_train_steps": T_TRAIN_STEPS,
    "t_eval_steps": T_EVAL_STEPS,
    "t_eval_iters": T_EVAL_ITERS,
    "t_lr_decay_step": T_LR_DECAY_STEP,
    "t_lr": T_LR,
    "t_lr_decayed": T_LR_DECAYED,
    "t_out_path": T_OUT_PATH,
    "device": DEVICE,
    # keys the reference scripts read but its config lacks (train_transformer.py:17,41,134;
    # data_preprocess.py:12-13)
    "val_path": DEV_PATH,
    "dtype": "bfloat16",
    "ddp_backend": "nccl",
    "dataset_name": "openwebtext",
    "tokenizer_name": "gpt2",
    # additions of this framework
    "seed": 1337,
    "grad_accum_steps": 1,
    "dp_mode": "ddp",            # "ddp" (replicated bf16 weights) | "zero" (sharded)
    "vocab_parallel": False,     # multi-GPU: shard the LM head + cross-entropy over the vocabulary
    "lr_schedule": "reference",  # "reference" = warm-up then constant (train_transformer.py:43-49)
                                 # "decay"     = additionally honour t_lr_decay_step / t_lr_decayed
    "synthetic_data": False,     # True: random tokens of the right shape, no files needed
    "save_every": 0,             # >0: periodic checkpoint every N steps
    "resume_from": "",           # path of a checkpoint to resume from
    "log_jsonl": "",             # optional JSONL metrics file
    "use_cuda_graph": True,
}

# Benchmark shapes named by BASELINE.json.  Head counts follow SURVEY section 6.
PRESETS = {
    "tiny": dict(n_blocks=4, n_embed=256, n_head=4, context_length=256, t_batch_size=8),
    "gpt2-small": dict(n_blocks=12, n_embed=768, n_head=12, context_length=1024, t_batch_size=16),
    "1.3b": dict(n_blocks=24, n_embed=2048, n_head=16, context_length=2048, t_batch_size=8),
    "reference-default": dict(n_blocks=N_BLOCKS, n_embed=N_EMBED, n_head=N_HEAD,
                              context_length=CONTEXT_LENGTH, t_batch_size=T_BATCH_SIZE),
}


def _coerce(old, text: str):
    if isinstance(old, bool):
        return text.lower() in ("1", "true", "yes", "on")
    if isinstance(old, int):
        return int(text)
    if isinstance(old, float):
        return float(text)
    if isinstance(old, str):
        return text
    try:
        return ast.literal_eval(text)
    except Exception:
        return text


def apply_overrides(cfg: dict, items: Iterable[str]) -> dict:
    """Apply ``key=value`` strings (``preset=<name>`` expands a PRESETS entry)."""
    for item in items:
        if "=" not in item:
            raise ValueError(f"override must be key=value, got {item!r}")
        k, v = item.split("=", 1)
        k = k.strip()
        if k == "preset":
            cfg.update(PRESETS[v])
            continue
        if k not in cfg:
            raise KeyError(f"unknown config key {k!r}")
        cfg[k] = _coerce(cfg[k], v)
    return cfg


def apply_env_overrides(cfg: dict, environ=None) -> dict:
    environ = os.environ if environ is None else environ
    if "B200_CFG_PRESET" in environ:
        cfg.update(PRESETS[environ["B200_CFG_PRESET"]])
    for k in list(cfg):
        name = "B200_CFG_" + k.upper()
        if name in environ:
            cfg[k] = _coerce(cfg[k], environ[name])
    return cfg


def validate(cfg: dict) -> dict:
    """Fail early on the classes of mistakes the reference lets through."""
    if cfg["n_embed"] % cfg["n_head"]:
        raise ValueError("n_embed must be divisible by n_head")
    if cfg["dtype"] not in ("bfloat16", "float16", "float32"):
        raise ValueError(f"unsupported dtype {cfg['dtype']}")
    if cfg["dp_mode"] not in ("ddp", "zero"):
        raise ValueError("dp_mode must be 'ddp' or 'zero'")
    if cfg["grad_accum_steps"] < 1:
        raise ValueError("grad_accum_steps must be >= 1")
    return cfg


def model_args(cfg: dict) -> dict:
    """The constructor kwargs the reference builds (train_transformer.py:114)."""
    return {k: cfg[k] for k in ("vocab_size", "n_embed", "n_head", "n_blocks", "context_length")}


apply_env_overrides(default_config)
