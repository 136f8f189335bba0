import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def ffk():
    import ffk_import

    return ffk_import.load()


@pytest.fixture(scope="session")
def K(ffk):
    """The host-side mirror of Kernels::<Op> (flexflow-train_b200/kernels.py)."""
    import importlib

    return importlib.import_module("flexflow_train_b200.kernels")


@pytest.fixture(scope="session")
def oracle():
    import oracle as o

    return o.load()
