"""pytest setup: registers the `gpu` marker, puts the repo root on sys.path, builds the oracle
(test infrastructure) and makes sure the product library exists."""
import os
import sys

import pytest

# Tests run several shards of one scene on ONE GPU (pmb_group): every stream of every shard gets its own hardware queue,
# so that a stream waiting for another shard's flag never sits in front of that shard's work (read at CUDA start-up).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as orc
    orc.build()
    return orc


@pytest.fixture(scope="session")
def lib():
    """the product library; built here if the .so is absent (nvcc cross-compiles without a GPU)"""
    from pmtech_b200 import build, capi
    if not os.path.exists(capi.SO_PATH):
        build.build()
    return capi.load()


@pytest.fixture(scope="session")
def gpu(lib):
    from pmtech_b200 import capi
    if capi.device_count() < 1:
        pytest.fail("no CUDA device: -m gpu tests must run on the GPU box (there is no CPU fallback)")
    return lib
