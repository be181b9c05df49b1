import importlib.util
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _reference_harness_path():
    for base in (os.path.join(ROOT, "baseline", "_ref"), "/root/reference"):
        p = os.path.join(base, "benchmarking", "train_harness.py")
        if os.path.exists(p):
            return p
    return None


@pytest.fixture(scope="session")
def reference_module():
    """The unmodified reference harness, imported read-only as the oracle."""
    p = _reference_harness_path()
    if p is None:
        pytest.skip("reference harness not available")
    spec = importlib.util.spec_from_file_location("reference_train_harness", p)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def native():
    """libb200bench.so front-end; GPU tests must fail, not skip, when it is missing on a GPU box."""
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from b200bench import ops

    return ops.load()
