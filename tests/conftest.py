import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def gm_ls_blob():
    return open(os.path.join(GOLDEN, "gm_ls.params.bin"), "rb").read()


@pytest.fixture(scope="session")
def gm_ls_ir():
    return np.load(os.path.join(GOLDEN, "gm_ls.ir.npy"))


@pytest.fixture(scope="session")
def subaru_blob():
    return open(os.path.join(GOLDEN, "subaru_ej25.params.bin"), "rb").read()


@pytest.fixture(scope="session")
def subaru_ir():
    return np.load(os.path.join(GOLDEN, "subaru_ej25.ir.npy"))


@pytest.fixture(scope="session")
def honda_blob():
    return open(os.path.join(GOLDEN, "honda_trx520.params.bin"), "rb").read()


@pytest.fixture(scope="session")
def honda_ir():
    return np.load(os.path.join(GOLDEN, "honda_trx520.ir.npy"))
