import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Builds (if stale) and loads the product package; the oracle is built on first use by oracle.lib()."""
    import __graft_entry__ as ge
    ge.build()
    import rtb200_loader
    return rtb200_loader.load()


@pytest.fixture(scope="session")
def R(built):
    return built


@pytest.fixture(scope="session")
def O():
    import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def earth(R):
    return R.load_earth()


EARTH_SCENES = ("earth", "random", "next_week_final")


@pytest.fixture(scope="session")
def oracle_scene(O, earth):
    cache = {}

    def get(name, lit=False, synth_n=30):
        key = (name, lit, synth_n)
        if key not in cache:
            cache[key] = O.OracleScene(name, earth=earth if name in EARTH_SCENES else None, lights_mode=int(lit), synth_n=synth_n)
        return cache[key]
    return get


@pytest.fixture(scope="session")
def gpu(R):
    if R.core().rt_device_count() < 1:
        pytest.fail("gpu-marked test on a box where librt_b200.so sees no CUDA device")
    return R


def rel_close(a, b, tol):
    return np.abs(a - b) <= tol * (np.abs(b) + 1e-3)
