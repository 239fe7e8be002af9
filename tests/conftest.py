import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE_DIRS = [os.path.join(ROOT, "baseline", "_ref"), "/root/reference"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 with `pytest -m gpu`)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def reference_module(name: str):
    """Import a module of the UNMODIFIED reference (oracle only; never copied into product code)."""
    import importlib.util
    for d in REFERENCE_DIRS:
        path = os.path.join(d, name + ".py")
        if os.path.exists(path):
            spec = importlib.util.spec_from_file_location("_reference_" + name, path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return mod
    pytest.skip("reference sources not available")


@pytest.fixture
def ref_model_module():
    return reference_module("model")
