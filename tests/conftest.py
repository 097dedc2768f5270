"""pytest configuration: the `gpu` marker, import paths, shared fixtures.

`-m "not gpu"` tests run on the CPU-only build container (oracle vs goldens, host logic, C-ABI exports);
`-m gpu` tests are the parity tests proper: they call the CUDA engine through the C ABI on a B200 and
compare with the oracle (`oracle/nerf_oracle.py`) and the committed goldens (`tests/golden/*.npz`).
Nothing here reads /root/reference.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


def load_golden(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def oracle():
    from oracle import nerf_oracle
    return nerf_oracle


@pytest.fixture(scope="session")
def built_lib():
    """Builds libnerf_b200.so if it is stale (nvcc cross-compiles without a GPU)."""
    import nerf_replication_b200.build as b
    return b.build()
