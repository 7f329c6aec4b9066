"""Reference-compatible import path: ``from data_loader.data_loader import get_batch_iterator``
(reference ``data_loader/data_loader.py:7-51``); implementation in
``pretraining_llm_b200/data/loader.py``."""
from pretraining_llm_b200.data.loader import get_batch_iterator  # noqa: F401
