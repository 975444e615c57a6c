"""Shared fixtures.  `-m "not gpu"` runs here (no GPU): oracle vs golden vectors, host logic,
C-ABI symbol check, gloo multi-rank logic.  `-m gpu` runs on a B200: the parity tests proper,
all through the C ABI of libbricks_b200.so."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
STENCILS = (0, 7, 13, 25, 125)
#: the project's bar for stencil output (BASELINE.json north_star): max relative error in fp64
REL_TOL = 1e-12


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built():
    """Build (or find) libbricks_b200.so and the oracle; returns the capi module."""
    import __graft_entry__ as ge
    ge.build()
    from bricklib_b200 import capi
    capi.lib()
    return capi


@pytest.fixture(scope="session")
def oracle(built):
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def golden_decomp():
    with open(os.path.join(GOLDEN, "decomp.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def ctx(built):
    c = built.Context(0)
    yield c
    c.close()


def relerr(got, exp):
    got, exp = np.asarray(got), np.asarray(exp)
    return float(np.max(np.abs(got - exp) / np.maximum(np.abs(exp), 1e-300)))


def fnv1a64(values):
    h = 0xCBF29CE484222325
    for v in np.asarray(values).reshape(-1).tolist():
        h ^= int(v)
        h = (h * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % h
