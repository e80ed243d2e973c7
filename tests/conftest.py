import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")


def _ensure_built():
    from libpoporon_b200.build import build_native, lib_dir

    have = all(os.path.exists(os.path.join(lib_dir(), f)) for f in ("libpoporon_cuda.so", "libpoporon.so"))
    if not have:
        build_native()
    from oraclelib import ORACLE_DIR, build_oracle

    if not os.path.exists(os.path.join(ORACLE_DIR, "liboracle.so")):
        build_oracle()


_ensure_built()


@pytest.fixture(scope="session")
def orc():
    from oraclelib import Oracle

    return Oracle("port")


@pytest.fixture(scope="session")
def ref():
    from oraclelib import Oracle, have_reference

    if not have_reference():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return Oracle("reference")


@pytest.fixture(scope="session")
def gpu_device():
    import libpoporon_b200.api as api

    if api.cuda.poporon_cuda_device_count() <= 0:
        pytest.fail("no CUDA device visible: -m gpu tests must run on a GPU box (no CPU fallback exists)")
    return 0


def golden(name):
    import numpy as np

    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
