import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def srt():
    return importlib.import_module("simd-ray-tracer_b200")


@pytest.fixture(scope="session")
def oracle():
    import orclib
    return orclib.Oracle()


@pytest.fixture(scope="session")
def backend(srt):
    b = srt.Backend(0)
    yield b
    b.close()
