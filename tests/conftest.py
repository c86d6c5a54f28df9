import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")
    config.addinivalue_line("markers", "slow: long-running")


@pytest.fixture(autouse=True)
def _seed():
    np.random.seed(0)


@pytest.fixture(autouse=True)
def _restore_trace_printers():
    """Several test modules silence the movers' trace lines by replacing the modules' ``_print``; tests that read the
    lines through ``capsys`` must not depend on the order the files run in."""
    import importlib

    mods = [importlib.import_module("movement_b200." + name) for name in ("mover", "monge_ampere", "laplacian_smoothing", "spring")]
    saved = [(m, m._print) for m in mods]
    yield
    for m, fn in saved:
        m._print = fn


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
