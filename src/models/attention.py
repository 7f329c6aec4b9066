"""Reference-compatible module path ``src.models.attention`` (reference ``src/models/attention.py``);
the classes are implemented in ``pretraining_llm_b200/models``."""
from pretraining_llm_b200.models import Head, MultiHeadAttention  # noqa: F401
