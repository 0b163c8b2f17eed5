import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "reference: needs the read-only reference tree (build container only)")


@pytest.fixture(scope="session")
def gold():
    import numpy as np

    return {n: dict(np.load(os.path.join(GOLD, n + ".npz"))) for n in ("fk_user05", "geometry", "warp", "model", "crop_cameras", "eval_user05", "crop_matrices", "torchdata")}


@pytest.fixture(scope="session")
def cuda_lib():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from absolutetrack_b200 import _cabi

    return _cabi.lib()
