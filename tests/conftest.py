import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def cuda_ops():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sheeprl_prey_b200._lib import CudaOps

    return CudaOps("cuda:0")


@pytest.fixture(scope="session")
def torch_ops():
    from oracle.torch_ops import TorchOps

    return TorchOps()


@pytest.fixture()
def cuda_ops_bf16():
    """A fresh bf16-precision CudaOps (own mirror registry, so tests cannot leak mirrors of freed tensors)."""
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from sheeprl_prey_b200._lib import CudaOps

    return CudaOps("cuda:0", scratch_bytes=4 << 20, precision="bf16")
