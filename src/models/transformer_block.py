"""Reference-compatible module path ``src.models.transformer_block`` (reference ``src/models/transformer_block.py``);
the classes are implemented in ``pretraining_llm_b200/models``."""
from pretraining_llm_b200.models import Block  # noqa: F401
