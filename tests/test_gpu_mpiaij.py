"""The MPIAIJ operator's device kernels (diagonal block, compressed off-diagonal block added from the result) with
P virtual ranks emulated in one process on one GPU: the test plays the scatter (lvec = x[garray], a pure copy), the
library does diag-mult + offdiag-multadd, and the result is compared with the reference's MatMult_SeqAIJ +
MatMultAdd_SeqAIJ pair per rank (oracle/mpiaij_restate.mult) — bit-exact for the stencil cases."""
import importlib

import numpy as np
import pytest

from test_mpiaij_host import _gens, serial_plans

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("lanes", [-1, 1])
@pytest.mark.parametrize("name,size", [("lap7", 2), ("lap7", 8), ("s27", 4), ("bs3", 3), ("plaw", 4), ("lap2d", 1)])
def test_virtual_ranks(gpu, R, P, name, size, lanes):
    par = importlib.import_module("petsc-dev_b200.parallel")
    from oracle import mpiaij_restate as O
    N, bs, gen = _gens(P)[name]
    ranges, host = serial_plans(par, size, N, gen, bs)
    def gen0(r0, r1):
        ai, aj, aa = gen(r0, r1)
        return ai - ai[0], aj, aa
    _, S = O.all_structures(size, N, gen0, bs)
    xh, yh = P.random_vector(N, 2), P.random_vector(N, 3)
    want = O.mult(R, ranges, S, xh)
    want_add = O.mult(R, ranges, S, xh, add_y=yh)
    got, got_add = np.empty(N), np.empty(N)
    for r in range(size):
        rs, re_ = int(ranges[r]), int(ranges[r + 1])
        h = dict(host[r], m=re_ - rs, N=N)
        gpu.spmv_tuning(lanes_per_row=lanes)          # 1: one lane per row = the reference's left-to-right row sums
        try:
            A = par.create_mpiaij(h)
        finally:
            gpu.spmv_tuning()
        g = h["split"]["garray"]
        x, y, z = gpu.Vec(re_ - rs, xh[rs:re_]), gpu.Vec(re_ - rs, yh[rs:re_]), gpu.Vec(re_ - rs)
        lv = gpu.Vec(max(g.size, 1), xh[g] if g.size else np.zeros(1))
        gpu._chk(gpu.lib.b200_mat_mpiaij_set_lvec(A.h, lv.p), "set_lvec")
        A.mult(x, z)
        got[rs:re_] = z.get_array()
        A.mult_add(x, y, z)
        got_add[rs:re_] = z.get_array()
        if size > 1 and name != "plaw":
            assert A.nnz() == h["split"]["Aj"].size + h["split"]["Bj"].size
        del A
    if name == "plaw" or (lanes != 1 and name in ("s27", "bs3")):
        sc = np.maximum(np.abs(want), 1.0)
        assert np.all(np.abs(got - want) <= 1e-12 * sc) and np.all(np.abs(got_add - want_add) <= 1e-12 * (sc + np.abs(yh)))
    else:
        assert np.array_equal(got, want)
        assert np.array_equal(got_add, want_add)


def test_scatter_single_rank_self_part(gpu, P):
    """The device VecScatter on one rank is the self part only (vpscat.c:1890-1903): forward/reverse, INSERT/ADD, repeated
    sources, and the range checks of the constructor."""
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, ny = 50, 8
    ranges = par.split_ownership(N, 1)
    inidx = np.array([7, 3, 3, 49, 0, 12, 25, 7], np.int32)
    inidy = np.array([0, 1, 2, 3, 4, 5, 6, 7], np.int32)
    S = par.create_scatter(0, 1, ranges, inidx, inidy, ny)
    xh = np.arange(N, dtype=np.float64) + 0.5
    yh = np.linspace(-1, 1, ny)
    x, y = gpu.Vec(N, xh), gpu.Vec(ny, yh)
    S.scatter(x, y)
    assert np.array_equal(y.get_array(), xh[inidx])
    y.set_array(yh)
    S.scatter(x, y, add=True)
    assert np.array_equal(y.get_array(), yh + xh[inidx])
    y.set_array(yh)
    S.scatter(x, y, add=True, reverse=True)                # repeated targets (3 and 7 twice): segment plan, list order
    want = xh.copy()
    for i, t in enumerate(inidx):                          # the reference's loop x[idx[i]] += y[i] (vpscat.c:520-570)
        want[t] = want[t] + yh[i]
    assert np.array_equal(x.get_array(), want)             # bit-exact and reproducible: no floating-point atomics
    first = x.get_array().copy()
    for _ in range(3):
        x.set_array(xh)
        S.scatter(x, y, add=True, reverse=True)
        assert np.array_equal(x.get_array(), first)
    x.set_array(xh)
    S.scatter(x, y, add=False, reverse=True)               # INSERT with repeated targets: the last entry of the list wins
    want = xh.copy()
    for i, t in enumerate(inidx):
        want[t] = yh[i]
    assert np.array_equal(x.get_array(), want)
    with pytest.raises(gpu.PetscB200Error) as e:
        par.create_scatter(0, 1, ranges, np.array([1], np.int32), np.array([9], np.int32), ny)
    assert e.value.code == 63


def test_sequential_scatter_layouts(gpu):
    """b200_scatter_create_local: y[dst[i]] (+)= x[src[i]] and its reverse, repeated targets in list order
    (VecScatterBegin_SGToSG, vscat.c:344-376)."""
    par = importlib.import_module("petsc-dev_b200.parallel")
    rng = np.random.default_rng(5)
    nx, ny, n = 4000, 3000, 10000
    src = rng.integers(0, nx, n).astype(np.int32)
    dst = rng.integers(0, ny, n).astype(np.int32)          # many targets named several times
    S = par.create_scatter_local(nx, ny, src, dst)
    xh, yh = rng.standard_normal(nx), rng.standard_normal(ny)
    x, y = gpu.Vec(nx, xh), gpu.Vec(ny, yh)
    S.scatter(x, y, add=True)
    want = yh.copy()
    for i in range(n):
        want[dst[i]] = want[dst[i]] + xh[src[i]]
    assert np.array_equal(y.get_array(), want)
    y.set_array(yh)
    S.scatter(x, y)
    want = yh.copy()
    for i in range(n):
        want[dst[i]] = xh[src[i]]
    assert np.array_equal(y.get_array(), want)
    y.set_array(yh)
    S.scatter(x, y, add=True, reverse=True)
    want = xh.copy()
    for i in range(n):
        want[src[i]] = want[src[i]] + yh[dst[i]]
    assert np.array_equal(x.get_array(), want)
    with pytest.raises(gpu.PetscB200Error):
        par.create_scatter_local(nx, ny, np.array([nx], np.int32), np.array([0], np.int32))


@pytest.mark.parametrize("name,size", [("lap7", 2), ("lap7", 8), ("s27", 4), ("bs3", 3), ("plaw", 4), ("lap2d", 1)])
def test_device_split_of_the_local_rows(gpu, R, P, name, size):
    """b200_mat_split_device (MatSetValues_MPIAIJ routing + MatSetUpMultiply compaction on the device): garray and the two
    blocks must be what the host split builds — garray compared directly, the blocks through the operator assembled from
    them (b200_mat_create_mpiaij_split), whose products must be bit-identical to those of the host-split operator."""
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, bs, gen = _gens(P)[name]
    ranges, host = serial_plans(par, size, N, gen, bs)
    xh, yh = P.random_vector(N, 2), P.random_vector(N, 3)
    for r in range(size):
        rs, re_ = int(ranges[r]), int(ranges[r + 1])
        ai, aj, aa = gen(rs, re_)
        ai = (ai - ai[0]).astype(np.int32)
        gpu.spmv_strips(0)
        try:
            L = gpu.Mat.from_csr(re_ - rs, N, ai, aj, aa)
        finally:
            gpu.spmv_strips()
        S, nzA, nzB, g = par.split_device(L, rs, re_)
        del L
        sp = host[r]["split"]
        assert np.array_equal(g, sp["garray"]) and nzA == sp["Aj"].size and nzB == sp["Bj"].size
        h = dict(host[r], m=re_ - rs, N=N)
        gpu.spmv_tuning(lanes_per_row=1)
        try:
            A = par.create_mpiaij_from_split(S, h)
            A0 = par.create_mpiaij(h)
        finally:
            gpu.spmv_tuning()
        assert A.nnz() == A0.nnz()
        x, y = gpu.Vec(re_ - rs, xh[rs:re_]), gpu.Vec(re_ - rs, yh[rs:re_])
        z, z0 = gpu.Vec(re_ - rs), gpu.Vec(re_ - rs)
        lv = gpu.Vec(max(g.size, 1), xh[g] if g.size else np.zeros(1))
        for M in (A, A0):
            gpu._chk(gpu.lib.b200_mat_mpiaij_set_lvec(M.h, lv.p), "set_lvec")
        A.mult(x, z); A0.mult(x, z0)
        assert np.array_equal(z.get_array(), z0.get_array())
        A.mult_add(x, y, z); A0.mult_add(x, y, z0)
        assert np.array_equal(z.get_array(), z0.get_array())
        d, d0 = gpu.Vec(re_ - rs), gpu.Vec(re_ - rs)
        A.get_diagonal(d); A0.get_diagonal(d0)
        assert np.array_equal(d.get_array(), d0.get_array())
        del A, A0
