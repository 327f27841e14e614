"""pytest configuration.

Markers: ``gpu`` = needs a B200 and the built CUDA library (run with ``-m gpu`` on the GPU box).  Everything else
runs on CPU (``-m "not gpu"``): the oracle against the reference's golden vectors, host logic, ABI checks.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from quest_b200 import capi  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device and quest_b200/lib/libQuEST.so")


def _ensure_oracle():
    if not os.path.isfile(capi.ORACLE_LIB):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "oracle"], check=True,
                       stdout=subprocess.DEVNULL)
    return capi.load(capi.ORACLE_LIB)


@pytest.fixture(scope="session")
def oracle():
    """Our CPU restatement (oracle/quest_oracle.c) under the product's host front-end."""
    return _ensure_oracle()


@pytest.fixture(scope="session")
def oracle_env(oracle):
    return oracle.createQuESTEnv()


@pytest.fixture(scope="session")
def reference():
    """The reference's own CPU/OpenMP build (oracle/_ref), when it has been built (it travels to the GPU box)."""
    if not os.path.isfile(capi.REFERENCE_LIB):
        pytest.skip("oracle/_ref/libQuEST.so not built (needs /root/reference at build time)")
    return capi.load(capi.REFERENCE_LIB)


@pytest.fixture(scope="session")
def reference_env(reference):
    return reference.createQuESTEnv()


@pytest.fixture(scope="session")
def product():
    """The CUDA library.  Fails (does not skip) when missing: the product has no CPU stand-in."""
    return capi.load_product()


@pytest.fixture(scope="session")
def product_env(product):
    return product.createQuESTEnv()
