import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "mdct4_ref.npz"))


@pytest.fixture(scope="session")
def golden_config1():
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", "mdct4_config1.npz"))


@pytest.fixture(scope="session", autouse=True)
def built_library():
    from pix2pixhdaudiosr_b200 import build
    return build.build()
