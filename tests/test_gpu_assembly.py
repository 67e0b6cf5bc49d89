"""Device assembly (b200_mat_setvaluesbatch = MatSetValuesBatch, matrix.c:1698) on the B200 against the reference's host
path run by the oracle library (one MatSetValues/ADD_VALUES per element), the matrix printed in the reference's golden
output ex4_0.out, and the committed fixture.  Integer arrays and values are BIT-EXACT: every nonzero is the
left-to-right sum of its contributions in element order."""
import importlib
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from test_assembly import golden_ex4, hostile_batch     # noqa: E402

pytestmark = pytest.mark.gpu


def same(got, want):
    gi, gj, ga = got
    wi, wj, wa = want
    return (np.array_equal(gi, wi) and np.array_equal(gj, wj) and
            np.array_equal(np.asarray(ga).view(np.int64), np.asarray(wa).view(np.int64)))


def test_ex4_golden(gpu, P):
    N, nl, er, em = P.elements_tri2d(3, 3)
    A = gpu.Mat.set_values_batch(N, N, nl, er, em)
    assert A.nnz() == 41
    assert same(A.csr(), golden_ex4())


def test_hex_fixture(gpu, P, fixtures):
    N, nl, er, em = P.elements_hex3d(6)
    A = gpu.Mat.set_values_batch(N, N, nl, er, em)
    assert same(A.csr(), (fixtures["hex6_batch_i"], fixtures["hex6_batch_j"], fixtures["hex6_batch_a"]))


@pytest.mark.parametrize("case", ["tri_129x65", "hex_24", "hostile0", "hostile1", "hostile_big", "empty", "nl1"])
def test_against_reference(gpu, P, R, case):
    if case == "tri_129x65":
        N, nl, er, em = P.elements_tri2d(129, 65)
    elif case == "hex_24":
        N, nl, er, em = P.elements_hex3d(24)
    elif case == "empty":
        N, nl, er, em = 7, 3, np.zeros(0, np.int32), np.zeros(0)
    elif case == "nl1":
        rng = np.random.RandomState(5)
        N, nl, er, em = 50, 1, rng.randint(-1, 50, size=4000).astype(np.int32), rng.standard_normal(4000)
    elif case == "hostile_big":
        # the hub row collects ~15 % of 40000 ids x nl: sorted by one CTA in global scratch; a few hundred more rows
        # exceed the warp path's 512 entries and sort in the CTA's shared memory
        N, nl, er, em = hostile_batch(3, m=300, ne=10000, nl=4)
    else:
        N, nl, er, em = hostile_batch(int(case[-1]))
    A = gpu.Mat.set_values_batch(N, N, nl, er, em)
    ref = R.mat_setvaluesbatch(N, N, er, em, nl)
    assert same(A.csr(), ref.csr())
    # the assembled matrix multiplies like any other (row binning was computed from the device offsets)
    xh = P.random_vector(N, 3)
    x, y = gpu.Vec(N, xh), gpu.Vec(N)
    A.mult(x, y)
    want = ref.mult(xh)
    ri, rj, ra = ref.csr()
    scale = np.bincount(np.repeat(np.arange(N), np.diff(ri)), weights=np.abs(ra * xh[rj]), minlength=N) if ra.size else np.zeros(N)
    assert np.all(np.abs(y.get_array() - want) <= 1e-12 * scale)


def test_same_pattern_new_values(gpu, P, R):
    N, nl, er, em = P.elements_hex3d(14)
    A = gpu.Mat.set_values_batch(N, N, nl, er, em, keep_plan=True)
    rng = np.random.RandomState(8)
    for _ in range(2):
        em2 = rng.standard_normal(em.size)
        A.batch_values(em2)
        assert same(A.csr(), R.mat_setvaluesbatch(N, N, er, em2, nl).csr())
    B = gpu.Mat.set_values_batch(N, N, nl, er, em)
    with pytest.raises(RuntimeError):
        B.batch_values(em)                                  # no plan kept


def test_device_resident_elements(gpu, P, R):
    import torch
    N, nl, er, em = P.elements_hex3d(16)
    der, dem = torch.from_numpy(er).cuda(), torch.from_numpy(em).cuda()
    torch.cuda.synchronize()
    A = gpu.Mat.set_values_batch_device(N, N, er.size // nl, nl, der.data_ptr(), dem.data_ptr(), keep_plan=True)
    ref = R.mat_setvaluesbatch(N, N, er, em, nl)
    assert same(A.csr(), ref.csr())
    dem2 = dem * 3.0
    torch.cuda.synchronize()
    A.batch_values_device(dem2.data_ptr())
    gpu.synchronize()
    assert same(A.csr(), R.mat_setvaluesbatch(N, N, er, em * 3.0, nl).csr())


def test_out_of_range(gpu):
    er = np.array([0, 1, 9], np.int32)
    with pytest.raises(RuntimeError, match="too large"):
        gpu.Mat.set_values_batch(5, 5, 3, er, np.ones(9))
    er = np.array([0, 1, 4], np.int32)
    with pytest.raises(RuntimeError, match="Column too large"):
        gpu.Mat.set_values_batch(5, 4, 3, er, np.ones(9))


def test_size_independent_properties(gpu, P):
    """At a size the host path would take minutes for: row sums of the Laplacian-like element matrices vanish, the
    pattern is the 27-point box, and assembling twice gives identical bits (no dependence on the atomics' order)."""
    nx = 96
    N, nl, er, em = P.elements_hex3d(nx)
    A = gpu.Mat.set_values_batch(N, N, nl, er, em)
    assert A.nnz() == (3 * nx - 2) ** 3
    one, y = gpu.Vec(N), gpu.Vec(N)
    one.set(1.0)
    A.mult(one, y)
    assert np.max(np.abs(y.get_array())) < 1e-13
    i1, j1, a1 = A.csr()
    i2, j2, a2 = gpu.Mat.set_values_batch(N, N, nl, er, em).csr()
    assert same((i1, j1, a1), (i2, j2, a2))
    assert np.all(np.diff(j1)[np.setdiff1d(np.arange(j1.size - 1), i1[1:-1] - 1)] > 0)      # columns ascend inside every row


@pytest.mark.parametrize("size", [2, 3])
def test_local_rows_of_a_parallel_matrix(gpu, P, R, size):
    """b200_mat_setvaluesbatch_rows on the element lists every (virtual) rank holds after the exchange: bit-identical
    to the restatement with the row window, and — for exact data — to the rows of the reference's global assembly."""
    import ctypes as C
    from oracle import assembly_restate as AR
    from test_assembly import expected_after_exchange, split_elements
    par = importlib.import_module("petsc-dev_b200.parallel")
    pb = gpu
    for exact in (False, True):
        N, nl, er, em = P.elements_hex3d(9, scale_values=not exact)
        if exact:
            em = np.round(em * 3.0)
        ranges = par.split_ownership(N, size)
        parts = split_elements(er, em, nl, size)
        ri, rj, ra = R.mat_setvaluesbatch(N, N, er, em, nl).csr()
        for r in range(size):
            rows, mats = expected_after_exchange(ranges, parts, nl, r)
            rows, mats = np.ascontiguousarray(rows, dtype=np.int32), np.ascontiguousarray(mats)
            rs, re_ = int(ranges[r]), int(ranges[r + 1])
            h = C.c_void_p()
            rc = pb.lib.b200_mat_setvaluesbatch_rows(re_ - rs, rs, N, C.c_longlong(rows.size // nl), nl,
                                                     rows.ctypes.data_as(C.c_void_p), mats.ctypes.data_as(C.c_void_p), 0, 0, C.byref(h))
            assert rc == 0, pb.lib.b200_last_error()
            L = pb.Mat(h, re_ - rs, N)
            got = L.csr()
            assert same(got, AR.setvaluesbatch(re_ - rs, N, rows, mats, nl, rstart=rs))
            assert np.array_equal(got[0], ri[rs:re_ + 1] - ri[rs]) and np.array_equal(got[1], rj[ri[rs]:ri[re_]])
            want = ra[ri[rs]:ri[re_]]
            if exact:
                assert np.array_equal(got[2], want)
            else:
                assert np.all(np.abs(got[2] - want) <= 1e-13 * np.maximum(np.abs(want), 1.0))
