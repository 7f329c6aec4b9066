"""pretraining_llm_b200 - a Blackwell (B200, sm_100a) native decoder-only LM pretraining and
generation framework with the capabilities of Flink-ddd/pretraining-llm.

Layers:
  config            hyper-parameter dict (reference surface + additions)
  data.loader       token-batch iterator (reference signature, pinned prefetch)
  models            Transformer facade over a flat packed parameter store, checkpoint translator
  ops               ctypes bindings of libb200llm.so (csrc/: tcgen05 GEMM, flash attention, ...)
  runtime           engine (static fwd/bwd kernel schedule, CUDA graphs), trainer, generation
  parallel          symmetric-memory bootstrap, fused gradient all-reduce + AdamW over NVLink
  utils             device timing, clocks sampling, logging
"""
from .config import default_config  # noqa: F401

__version__ = "0.1.0"


def __getattr__(name):
    if name == "Transformer":
        from .models.transformer import Transformer
        return Transformer
    raise AttributeError(name)
