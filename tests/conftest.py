"""pytest configuration: the `gpu` marker and shared helpers.

CPU tier (``pytest -m "not gpu"``) runs here without a GPU; GPU tier (``pytest -m gpu``) runs on a
B200 and exercises the native library - it fails, never skips, if the library is missing there.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200) and the native sm_100a library")
    config.addinivalue_line("markers", "slow: long-running")


def reference_root():
    """Where the unmodified reference lives: baseline/_ref (travels to the GPU box) or /root/reference."""
    for p in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        if os.path.isdir(os.path.join(p, "src", "models")):
            return p
    return None


@pytest.fixture(scope="session")
def ref_models():
    """The reference's own ``src.models`` package, imported from an isolated sys.path entry."""
    from baseline.ref_import import import_reference_models
    mod = import_reference_models()
    if mod is None:
        pytest.skip("reference tree not available")
    return mod
