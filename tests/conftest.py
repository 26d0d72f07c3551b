import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


# The suite runs the library's defaults (the sequential form of the trace pass); the pool form (trace_pool.cu) and the wavefront form
# (trace_wave.cu) are selected per handle by the tests that cover them (vxrt_set_trace_form).


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
