import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def cport():
    from oracle import cport as cp
    cp.build()
    return cp


@pytest.fixture(scope="session")
def emu():
    from tests import emu_util
    emu_util.build_emu()
    return emu_util


@pytest.fixture(scope="session")
def gpu_engine():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from lidar_slam_b200 import batch
    return batch.Engine(0)        # raises (does not skip) if liblslam.so is missing
