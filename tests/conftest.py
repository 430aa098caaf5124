import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU restatement of the reference shaders (test infrastructure)."""
    from oracle import binding
    binding.lib()
    return binding


@pytest.fixture(scope="session")
def scenes():
    """Small procedural scenes, built once per session."""
    from path_tracer_b200 import scene as S
    S.ensure_spectrum_table()
    cache = {}

    def get(name, *params):
        key = (name,) + params
        if key not in cache:
            cache[key] = S.Scene.make(name, *params)
        return cache[key]
    return get


@pytest.fixture(scope="session")
def gpu():
    """Context on cuda:0 through the C ABI; fails loudly when the CUDA path is unusable."""
    from path_tracer_b200 import integrator as I
    return I.Context(0)
