"""Reference-compatible module path ``src.models.transformer`` (reference ``src/models/transformer.py``);
the classes are implemented in ``pretraining_llm_b200/models``."""
from pretraining_llm_b200.models import Transformer  # noqa: F401
