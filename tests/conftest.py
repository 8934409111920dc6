import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def reference():
    """The unmodified reference, importable only in the authoring container (skips elsewhere)."""
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference tree not present (expected on the GPU box)")
    return ref_import.load()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from b_pinn_kalman_filter_b200 import _lib
    _lib.lib()  # fail loudly if the extension is missing on a GPU box
    return torch.device("cuda:0")


def golden_path(name):
    return os.path.join(GOLDEN, name)
