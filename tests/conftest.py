import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")
    config.addinivalue_line("markers", "slow: larger CPU case")


@pytest.fixture(scope="session")
def O():
    import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def R(O):
    r = O.ref()
    if r is None:
        pytest.skip("oracle/_ref/libref.so not built (no /root/reference)")
    return r


@pytest.fixture(scope="session")
def sb():
    """the product package (CUDA path through the C-ABI)"""
    import sb200

    return sb200.load()


@pytest.fixture(scope="session")
def RS(O):
    r = O.ref_scalar()
    if r is None:
        pytest.skip("oracle/_ref/libref_scalar.so not built (no /root/reference)")
    return r


@pytest.fixture(scope="session")
def RDB():
    """the reference's DuckDB-linked code (oracle/_ref/libref_duckdb.so, `make -C oracle ref_duckdb`)"""
    from oracle import refdb

    if not refdb.available():
        pytest.skip("oracle/_ref/libref_duckdb.so not built")
    return refdb
