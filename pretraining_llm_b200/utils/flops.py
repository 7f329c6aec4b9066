"""Model FLOPs of one training token and model-FLOP utilisation (MFU) against the MEASURED tensor peak.

Count (the usual 6 x matmul parameters, causal attention at half the square):
  * every weight matrix (QKV, MLP up / down, LM head - the reference model has no attention output projection,
    src/models/attention.py:95): 2 FLOPs per parameter forward, 4 backward;
  * attention scores and P.V: 2 x 2 x T x d per token forward for a full square, halved by causality, x 3 with
    the backward = 6 T d per layer (the recomputation inside the flash backward is NOT counted: model FLOPs).
The denominator is ``bf16_tflops_sustained`` of MEASURED_PEAKS.json (cuBLAS bf16 under the power cap on this
pool's B200s); 1396 TFLOP/s if the file is absent.
"""
from __future__ import annotations

import json
import os
from functools import lru_cache

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def matmul_params(n_blocks: int, d: int, vocab: int) -> int:
    return n_blocks * (3 * d * d + 8 * d * d) + vocab * d


def model_flops_per_token(n_blocks: int, d: int, vocab: int, T: int) -> float:
    return 6.0 * matmul_params(n_blocks, d, vocab) + 6.0 * n_blocks * T * d


@lru_cache(maxsize=1)
def measured_peaks() -> dict:
    try:
        with open(os.path.join(_ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {"bf16_tflops_sustained": 1396.0, "bf16_tflops": 1669.2, "hbm_gbs": 6538.3, "fallback": True}


def mfu(tokens_per_s: float, n_blocks: int, d: int, vocab: int, T: int, n_gpus: int = 1) -> float:
    peak = float(measured_peaks().get("bf16_tflops_sustained", 1396.0)) * 1e12 * max(n_gpus, 1)
    return tokens_per_s * model_flops_per_token(n_blocks, d, vocab, T) / peak
