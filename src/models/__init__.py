"""Reference-compatible import path ``src.models`` (reference ``src/models/__init__.py:2-5``)."""
from pretraining_llm_b200.models import MLP, Head, MultiHeadAttention, Block, Transformer  # noqa: F401
