import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle"), os.path.join(ROOT, "deep-learning-air-hockey-robot_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure), built on demand."""
    import orc as _orc
    _orc.build()
    _orc.lib()
    return _orc


@pytest.fixture(scope="session")
def table():
    from gym_air_hockey_b200.assets import make_table_image
    return make_table_image()


@pytest.fixture(scope="session")
def noise_table():
    """A table image of pure noise: any indexing slip in the rasteriser shows up as a pixel mismatch."""
    return np.random.RandomState(7).randint(0, 256, size=(700, 450, 3)).astype(np.uint8)


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch
