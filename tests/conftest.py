import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_checkers():
    """The oracle is test infrastructure: (re)build it when its source is newer than the .so."""
    so = os.path.join(ROOT, "oracle", "libkmi_oracle.so")
    src = os.path.join(ROOT, "oracle", "kmi_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(src) > os.path.getmtime(so):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "oracle"), so])
    yield
