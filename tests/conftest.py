import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# k_trace_pool (trace_pool.cu) is opt-in and the library reads the switch once per process: the GPU suite runs with it on, so
# every FULL-lighting handle WITHOUT hit buffers takes the pool and every handle with them the sequential form - both are compared
os.environ.setdefault("VXRT_POOL", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as o
    o.build()
    return o


@pytest.fixture(scope="session")
def noise(orc):
    return orc.blue_noise()


_world_cache = {}


@pytest.fixture(scope="session")
def world(orc):
    """world(name) -> (materials, minefield) of a seeded scene, built by the oracle and cached."""
    from raytrace_b200 import scenes

    def get(name):
        if name not in _world_cache:
            dims, kind, seed = scenes.WORLDS[name]
            m = orc.generate_world(*dims, kind, seed)
            _world_cache[name] = orc.build_minefield(m)
        return _world_cache[name]
    return get
