import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def golden():
    import json
    import numpy as np
    g = json.loads((ROOT / "tests" / "golden" / "golden.json").read_text())
    g["arrays"] = dict(np.load(ROOT / "tests" / "golden" / "golden_small.npz"))
    return g


@pytest.fixture(scope="session")
def ph():
    """The product package with libheat_b200.so built (nvcc cross-compiles without a GPU)."""
    from parallel_heat_b200 import build
    build.build()
    import parallel_heat_b200
    parallel_heat_b200.load()
    return parallel_heat_b200


def have_reference_binaries():
    return (ROOT / "oracle" / "_ref" / "mpi_heat_20x20_O3").exists()
