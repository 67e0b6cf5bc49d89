import importlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def monitor_short(v):
    """KSPMonitorDefaultShort formatting (src/ksp/ksp/interface/iterativ.c:495-501): %G above 1e-9."""
    if v > 1e-9:
        return "%G" % v
    if v > 1e-11:
        return "%5.3e" % v
    return "< 1.e-11"


@pytest.fixture(scope="session")
def P():
    return importlib.import_module("workloads.problems")


@pytest.fixture(scope="session")
def pb():
    return importlib.import_module("petsc-dev_b200")


@pytest.fixture(scope="session")
def gpu(pb):
    """The initialised library on cuda:0; fails loudly (no CPU fallback) if there is no device."""
    pb.init(0)
    return pb


@pytest.fixture(scope="session")
def R():
    from oracle import refpetsc
    if not refpetsc.available():
        pytest.skip("oracle/_ref/libpetsc_ref.so not built (run oracle/build_ref.py where /root/reference exists)")
    refpetsc.lib()
    return refpetsc


@pytest.fixture(scope="session")
def goldens():
    return json.load(open(os.path.join(GOLDEN, "reference_goldens.json")))


@pytest.fixture(scope="session")
def fixtures():
    return np.load(os.path.join(GOLDEN, "oracle_fixtures.npz"))


def ex5_matrix():
    """src/mat/examples/tests/ex5.c -rectA: m = 8, n = 10, C_ij = 0.1*(10(i+1)+j+1) (MatScale by alpha = 0.1)."""
    m, n = 8, 10
    ip = np.arange(0, (m + 1) * n, n, dtype=np.int32)
    j = np.tile(np.arange(n, dtype=np.int32), m)
    a = np.array([(10.0 * (i + 1) + jj + 1) for i in range(m) for jj in range(n)])
    a = a * 0.1        # MatScale_SeqAIJ: BLASscal_ on the value array
    return m, n, ip, j, a


def tridiag(n=10):
    """src/ksp/pc/examples/tests/ex2.c: tridiag(-1, 2, -1)."""
    ip, j, a = [0], [], []
    for i in range(n):
        for c, v in ((i - 1, -1.0), (i, 2.0), (i + 1, -1.0)):
            if 0 <= c < n:
                j.append(c); a.append(v)
        ip.append(len(j))
    return np.array(ip, np.int32), np.array(j, np.int32), np.array(a)
