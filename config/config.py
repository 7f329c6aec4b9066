"""Reference-compatible import path: ``from config.config import default_config``.

Same constants and the same 17 dict keys as the reference (``config/config.py:4-47``) plus the
keys its scripts read but never defined; the implementation lives in
``pretraining_llm_b200/config.py``.
"""
from pretraining_llm_b200.config import *  # noqa: F401,F403
from pretraining_llm_b200.config import (PRESETS, apply_env_overrides, apply_overrides,  # noqa: F401
                                         default_config, model_args, validate)
