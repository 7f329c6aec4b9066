"""Import the reference's modules without colliding with this repo's same-named facades.

The reference is not an installable package (no setup.py / pyproject - `pip install` refuses
it); "installing" it means putting its tree on sys.path.  This repo ships facades with the same
top-level names (`src`, `config`, `data_loader`), so the reference tree is imported under a
private alias by temporarily swapping sys.path / sys.modules.
"""
from __future__ import annotations

import importlib
import os
import sys
from types import ModuleType
from typing import Optional

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_CACHE: dict = {}


def reference_root() -> Optional[str]:
    for p in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(p, "src", "models")):
            return p
    return None


def _import_isolated(names) -> Optional[dict]:
    root = reference_root()
    if root is None:
        return None
    key = tuple(names)
    if key in _CACHE:
        return _CACHE[key]
    shadowed = ("src", "config", "data_loader", "scripts")
    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in shadowed}
    for k in saved:
        del sys.modules[k]
    saved_path = list(sys.path)
    sys.path[:] = [root] + [p for p in sys.path if os.path.abspath(p or ".") != ROOT]
    try:
        out = {n: importlib.import_module(n) for n in names}
    finally:
        ref_loaded = {k: v for k, v in sys.modules.items() if k.split(".")[0] in shadowed}
        for k in ref_loaded:
            del sys.modules[k]
        sys.modules.update(saved)
        sys.path[:] = saved_path
    _CACHE[key] = out
    return out


def import_reference_models() -> Optional[ModuleType]:
    """The reference's ``src.models`` (Transformer, Block, MLP, Head, MultiHeadAttention)."""
    mods = _import_isolated(["src.models"])
    return None if mods is None else mods["src.models"]


def import_reference_data_loader() -> Optional[ModuleType]:
    mods = _import_isolated(["data_loader.data_loader"])
    return None if mods is None else mods["data_loader.data_loader"]


def import_reference_config() -> Optional[ModuleType]:
    mods = _import_isolated(["config.config"])
    return None if mods is None else mods["config.config"]
