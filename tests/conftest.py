import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def _ensure_library():
    """The C-ABI tests load safe_exploration_b200/libsecem.so; build it if this checkout has not been built yet
    (nvcc cross-compiles without a GPU, ~2 min).  The library is never replaced by a CPU path: a failed build fails
    the tests that need it."""
    lib = os.path.join(ROOT, 'safe_exploration_b200', 'libsecem.so')
    if not os.path.exists(lib):
        import subprocess
        subprocess.run(['make', '-C', os.path.join(ROOT, 'safe_exploration_b200', 'csrc'), '-j',
                        str(max(1, min(8, os.cpu_count() or 1)))], capture_output=True)


def pytest_configure(config):
    _ensure_library()
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (B200); run with -m gpu')
    torch.set_default_dtype(torch.double)  # like the reference's hot-path test modules


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason='no CUDA device')
    for item in items:
        if 'gpu' in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    with np.load(os.path.join(GOLDEN, name), allow_pickle=False) as f:
        return {k: f[k] for k in f.files}


@pytest.fixture(scope='session')
def golden_algebra():
    return load_golden('ellipsoid_algebra.npz')


@pytest.fixture(scope='session')
def golden_onestep():
    return load_golden('onestep_invpend.npz')


@pytest.fixture(scope='session')
def golden_multistep():
    return load_golden('multistep_invpend.npz')


@pytest.fixture(scope='session')
def golden_cem():
    return load_golden('cem_pendulum.npz')
