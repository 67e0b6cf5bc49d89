"""MatSetValuesBatch oracle (CPU): the reference's host path (oracle/_ref) against the matrix the reference's own golden
output prints for ex4 (src/ksp/ksp/examples/tutorials/output/ex4_0.out, first MatView) and against the committed
fixture; the numpy restatement against the reference library on meshes and on hostile inputs."""
import importlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = importlib.import_module("petsc-dev_b200.problems")
from oracle import assembly_restate as AR          # noqa: E402
from oracle import refpetsc as R                   # noqa: E402

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "reference_goldens.json")))
FX = np.load(os.path.join(ROOT, "tests", "golden", "oracle_fixtures.npz"))
needs_ref = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built")


def golden_ex4():
    rows = GOLD["ksp_ex4_0"]["rows"]
    i = np.cumsum([0] + [len(r) for r in rows]).astype(np.int32)
    j = np.array([c for r in rows for c, _ in r], dtype=np.int32)
    a = np.array([v for r in rows for _, v in r])
    return i, j, a


def hostile_batch(seed, m=40, ne=300, nl=4):
    """Repeated ids inside an element, negative ids, a hub row that collects thousands of contributions."""
    rng = np.random.RandomState(seed)
    rows = rng.randint(0, m, size=(ne, nl)).astype(np.int32)
    rows[rng.rand(ne, nl) < 0.15] = 0                      # hub
    rows[rng.rand(ne, nl) < 0.05] = -1                     # skipped
    rows[::7, 1] = rows[::7, 0]                            # the same id twice in one element
    vals = rng.standard_normal(ne * nl * nl)
    vals[rng.rand(vals.size) < 0.1] = 0.0                  # explicit zeros stay in the pattern
    return m, nl, rows.reshape(-1), vals


def test_restatement_matches_ex4_golden():
    N, nl, er, em = P.elements_tri2d(3, 3)
    i, j, a = AR.setvaluesbatch(N, N, er, em, nl)
    gi, gj, ga = golden_ex4()
    assert np.array_equal(i, gi) and np.array_equal(j, gj) and np.array_equal(a, ga)


@needs_ref
def test_reference_matches_ex4_golden():
    N, nl, er, em = P.elements_tri2d(3, 3)
    i, j, a = R.mat_setvaluesbatch(N, N, er, em, nl).csr()
    gi, gj, ga = golden_ex4()
    assert np.array_equal(i, gi) and np.array_equal(j, gj) and np.array_equal(a, ga)


@needs_ref
@pytest.mark.parametrize("case", ["tri", "hex", "hostile0", "hostile1", "empty"])
def test_restatement_matches_reference(case):
    if case == "tri":
        N, nl, er, em = P.elements_tri2d(17, 9)
    elif case == "hex":
        N, nl, er, em = P.elements_hex3d(7)
    elif case == "empty":
        N, nl, er, em = 5, 3, np.zeros(0, np.int32), np.zeros(0)
    else:
        N, nl, er, em = hostile_batch(int(case[-1]))
    i, j, a = AR.setvaluesbatch(N, N, er, em, nl)
    ri, rj, ra = R.mat_setvaluesbatch(N, N, er, em, nl).csr()
    assert np.array_equal(i, ri) and np.array_equal(j, rj)
    assert np.array_equal(a.view(np.int64), ra.view(np.int64))           # bit-exact, signed zeros included


def test_restatement_matches_fixture():
    N, nl, er, em = P.elements_hex3d(6)
    i, j, a = AR.setvaluesbatch(N, N, er, em, nl)
    assert np.array_equal(i, FX["hex6_batch_i"]) and np.array_equal(j, FX["hex6_batch_j"]) and np.array_equal(a, FX["hex6_batch_a"])


def test_out_of_range_is_an_error():
    # the reference checks the range only in debug builds (aij.c:313,327 under PETSC_USE_DEBUG); the optimised oracle
    # library would write out of bounds, so only the restatement (and the device path, tests/test_gpu_assembly.py) is held to it
    er = np.array([0, 1, 9], np.int32)
    with pytest.raises(IndexError):
        AR.setvaluesbatch(5, 5, er, np.ones(9), 3)
