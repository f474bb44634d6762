import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU; run with -m gpu")


# GPU tests never run silently on a machine without a GPU: they are deselected by `-m "not gpu"`, and if selected without
# a device the fixture below fails them loudly (no CPU fallback exists).
@pytest.fixture(scope="session")
def cuda_device():
    import torch

    if not torch.cuda.is_available():
        pytest.fail("GPU test selected but no CUDA device is visible (stablediffusion_b200 has no CPU fallback)")
    return torch.device("cuda", 0)
