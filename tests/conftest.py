import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


_CASES = {}


def get_case(name):
    """Seeded synthetic registration problems shared by the tests (built once per session)."""
    from xchu_slam_b200 import synth
    if name in _CASES:
        return _CASES[name]
    if name == "small":      # 5k-pt scan vs 60k-pt localmap
        c = synth.make_case(n_source=5000, n_target=60000, seed=synth.BASE_SEED + 1)
    elif name == "c1":       # BASELINE configs[0]: 20k vs 300k, res 1.0
        c = synth.make_case(n_source=20000, n_target=300000, seed=synth.BASE_SEED)
    elif name == "tiny":
        c = synth.make_case(n_source=1500, n_target=20000, seed=synth.BASE_SEED + 2)
    elif name == "c3":       # BASELINE configs[2]: 20k-pt scan vs 5M-pt localmap, res 0.5
        c = synth.make_case(n_source=20000, n_target=5000000, seed=synth.BASE_SEED + 3, keyframe_leaf=0.25)
    elif name == "c4":       # BASELINE configs[3]: undownsampled 120k-pt scan vs 1M-pt map, res 2.0
        c = synth.make_case(n_source=120000, n_target=1000000, seed=synth.BASE_SEED + 4, dense=True)
    else:
        raise KeyError(name)
    _CASES[name] = c
    return c


@pytest.fixture(scope="session")
def case_small():
    return get_case("small")


@pytest.fixture(scope="session")
def case_c1():
    return get_case("c1")


@pytest.fixture(scope="session")
def case_c3():
    return get_case("c3")


@pytest.fixture(scope="session")
def case_c4():
    return get_case("c4")


@pytest.fixture(scope="session")
def case_tiny():
    return get_case("tiny")


def rel_err(a, b, scale=None):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    s = np.max(np.abs(b)) if scale is None else scale
    return float(np.max(np.abs(a - b)) / (s if s > 0 else 1.0))
