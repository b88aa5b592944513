import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def product():
    """The product library, built in-tree if needed.  Never falls back to anything else."""
    import nntoolkitcore_b200 as nn
    if not nn.LIB_PATH.exists():
        nn.build()
    return nn.lib()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled by oracle/Makefile (checker only)."""
    import helpers
    return helpers.reference_lib()


@pytest.fixture(scope="session")
def port():
    import helpers
    return helpers.oracle_port()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible")
    import nntoolkitcore_b200 as nn
    lib = nn.lib()
    assert lib.NNTKInit(0) == 0, lib.NNTKLastError()
    return torch.device("cuda:0")
