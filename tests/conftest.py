"""pytest configuration: registers the `gpu` marker and puts the repo root on sys.path.

`-m "not gpu"` covers the oracle against the golden vectors, the host logic and the C-ABI
export table (no compute calls); `-m gpu` holds the parity tests proper, which call the CUDA
kernels through the C ABI and compare against the oracle / golden fixtures."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU")


@pytest.fixture(scope="session")
def golden_outputs():
    import torch
    return torch.load(os.path.join(GOLDEN, "ngswin_outputs.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden_maps():
    import torch
    return torch.load(os.path.join(GOLDEN, "index_maps.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden_contract():
    import torch
    return torch.load(os.path.join(GOLDEN, "state_dict_contract.pt"), weights_only=False)


@pytest.fixture(scope="session")
def golden_w3():
    """W3 = trained-like weights + the unmodified reference's outputs on them (oracle/make_w3.py)."""
    import torch
    return torch.load(os.path.join(GOLDEN, "w3_trained.pt"), weights_only=False)
