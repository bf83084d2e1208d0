import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """librtcore.so and liboracle.so must exist; build them if a fresh checkout has neither."""
    import subprocess
    if not os.path.exists(os.path.join(ROOT, "rustraytracer_b200", "librtcore.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "rustraytracer_b200", "csrc"), "-s"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
