import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

MODELS = os.path.join(ROOT, "myosuite_mjx_b200", "models")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def model_path(name):
    return os.path.join(MODELS, name + ".npz")


@pytest.fixture(scope="session")
def models():
    from myosuite_mjx_b200 import mjcf
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = mjcf.Model.load(model_path(name))
        return cache[name]
    return get
