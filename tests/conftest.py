import sys
from pathlib import Path

import pytest

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def tiny_cfg():
    from gru_mpi_cuda_b200 import GRUConfig
    return GRUConfig.tiny()


@pytest.fixture(scope="session")
def tiny_params(tiny_cfg):
    from gru_mpi_cuda_b200.utils import weights as W
    return W.random_init(tiny_cfg, seed=11)
