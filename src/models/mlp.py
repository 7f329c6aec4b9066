"""Reference-compatible module path ``src.models.mlp`` (reference ``src/models/mlp.py``);
the classes are implemented in ``pretraining_llm_b200/models``."""
from pretraining_llm_b200.models import MLP  # noqa: F401
