"""MatSetValuesBatch oracle (CPU): the reference's host path (oracle/_ref) against the matrix the reference's own golden
output prints for ex4 (src/ksp/ksp/examples/tutorials/output/ex4_0.out, first MatView) and against the committed
fixture; the numpy restatement against the reference library on meshes and on hostile inputs."""
import importlib
import json
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = importlib.import_module("workloads.problems")
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


# ---- parallel matrix: who gets which element (petsc-dev_b200/parallel.py::batch_exchange) ---------------------------
def expected_after_exchange(ranges, parts, nl, rank):
    """Own elements first, then — sender by sender, in the sender's order — the elements that touch this rank's rows."""
    rows = [parts[rank][0].reshape(-1, nl)]
    mats = [parts[rank][1].reshape(-1, nl * nl)]
    for s, (er, em) in enumerate(parts):
        if s == rank:
            continue
        er2 = er.reshape(-1, nl)
        hit = np.any((er2 >= ranges[rank]) & (er2 < ranges[rank + 1]), axis=1)
        rows.append(er2[hit]); mats.append(em.reshape(-1, nl * nl)[hit])
    return np.concatenate(rows).reshape(-1), np.concatenate(mats).reshape(-1)


def split_elements(er, em, nl, size):
    """Contiguous chunks of the element list, one per rank (so interface elements touch rows of two ranks)."""
    ne = er.size // nl
    cut = [ne * r // size for r in range(size + 1)]
    return [(er[cut[r] * nl:cut[r + 1] * nl], em[cut[r] * nl * nl:cut[r + 1] * nl * nl]) for r in range(size)]


def _gloo_batch_worker(rank, size, port, q):
    import sys
    sys.path.insert(0, ROOT)
    import torch.distributed as td
    td.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=size)
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, nl, er, em = P.elements_hex3d(7)
    ranges = par.split_ownership(N, size)
    mine = split_elements(er, em, nl, size)[rank]
    rows, mats = par.batch_exchange(rank, size, ranges, mine[0], mine[1], nl)
    q.put((rank, rows, mats))
    td.barrier()
    td.destroy_process_group()


def test_gloo_batch_exchange_world_size_2():
    import torch.multiprocessing as mp
    par = importlib.import_module("petsc-dev_b200.parallel")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31000 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_batch_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = {}
    for _ in range(2):
        r, rows, mats = q.get(timeout=120)
        got[r] = (rows, mats)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    N, nl, er, em = P.elements_hex3d(7)
    ranges = par.split_ownership(N, 2)
    parts = split_elements(er, em, nl, 2)
    for r in range(2):
        wr, wm = expected_after_exchange(ranges, parts, nl, r)
        assert np.array_equal(got[r][0], wr) and np.array_equal(got[r][1], wm)
        assert wr.size > parts[r][0].size                    # something did cross the interface


@needs_ref
@pytest.mark.parametrize("size", [2, 3])
def test_restated_local_rows_add_up_to_the_reference_matrix(size):
    """Exact-arithmetic data (small integers): the per-rank local-row assemblies of the exchanged element lists,
    stacked, are bit-identical to the reference's assembly of the whole element list — whatever the order in which
    the off-process contributions arrive (which the reference leaves to MPI)."""
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, nl, er, em = P.elements_hex3d(6, scale_values=False)
    em = np.round(em * 3.0)                                   # -1, 0, 4 -> integers: every sum exact
    ranges = par.split_ownership(N, size)
    parts = split_elements(er, em, nl, size)
    ri, rj, ra = R.mat_setvaluesbatch(N, N, er, em, nl).csr()
    for r in range(size):
        rows, mats = expected_after_exchange(ranges, parts, nl, r)
        rs, re_ = int(ranges[r]), int(ranges[r + 1])
        i, j, a = AR.setvaluesbatch(re_ - rs, N, rows, mats, nl, rstart=rs)
        assert np.array_equal(i, ri[rs:re_ + 1] - ri[rs])
        assert np.array_equal(j, rj[ri[rs]:ri[re_]]) and np.array_equal(a, ra[ri[rs]:ri[re_]])


@pytest.mark.parametrize("mesh", ["tri", "hostile"])
@pytest.mark.parametrize("size", [1, 2, 3, 5])
def test_exchange_plan_helpers_with_virtual_ranks(size, mesh):
    """b200_batch_send_plan / b200_batch_recv_plan (the integer logic shared by the PETSc plugin's MPI path and
    parallel.py) driven for every virtual rank in one process: publish one buffer per rank, pick the shares."""
    import ctypes as C
    pb = importlib.import_module("petsc-dev_b200")
    par = importlib.import_module("petsc-dev_b200.parallel")
    llp = C.POINTER(C.c_longlong)
    ip = C.POINTER(C.c_int)
    if mesh == "tri":
        N, nl, er, em = P.elements_tri2d(13, 9)
        er = er.copy(); er[5] = -1                              # an ignored id changes nothing about the routing of the rest
    else:                                                       # random ids: elements that touch 1..4 ranks, repeats, ignored ids
        N, nl, er, em = hostile_batch(7, m=97, ne=400, nl=4)
    ranges = np.ascontiguousarray(par.split_ownership(N, size), dtype=np.int32)
    parts = split_elements(er, em, nl, size)
    bufs, counts = [], np.zeros((size, size), np.int64)
    for r in range(size):
        rows = np.ascontiguousarray(parts[r][0], dtype=np.int32)
        ne = rows.size // nl
        cnt = np.zeros(size, np.int64)
        assert pb.lib.b200_batch_send_plan(size, r, ranges.ctypes.data_as(ip), C.c_longlong(ne), nl, rows.ctypes.data_as(ip),
                                           cnt.ctypes.data_as(llp), None) == 0
        order = np.zeros(max(int(cnt.sum()), 1), np.int64)
        assert pb.lib.b200_batch_send_plan(size, r, ranges.ctypes.data_as(ip), C.c_longlong(ne), nl, rows.ctypes.data_as(ip),
                                           cnt.ctypes.data_as(llp), order.ctypes.data_as(llp)) == 0
        order = order[:int(cnt.sum())]
        assert cnt[r] == 0
        counts[r] = cnt
        bufs.append((rows.reshape(-1, nl)[order], parts[r][1].reshape(-1, nl * nl)[order]))
    for r in range(size):
        off, num = np.zeros(size, np.int64), np.zeros(size, np.int64)
        assert pb.lib.b200_batch_recv_plan(size, r, np.ascontiguousarray(counts).ctypes.data_as(llp), off.ctypes.data_as(llp),
                                           num.ctypes.data_as(llp)) == 0
        rows = [parts[r][0].reshape(-1, nl)] + [bufs[s][0][off[s]:off[s] + num[s]] for s in range(size) if num[s]]
        mats = [parts[r][1].reshape(-1, nl * nl)] + [bufs[s][1][off[s]:off[s] + num[s]] for s in range(size) if num[s]]
        wr, wm = expected_after_exchange(ranges, parts, nl, r)
        assert np.array_equal(np.concatenate(rows).reshape(-1), wr) and np.array_equal(np.concatenate(mats).reshape(-1), wm)
    bad = np.array([0, 1, N], np.int32)                         # a row beyond the matrix is an error, not a routing decision
    cnt = np.zeros(size, np.int64)
    assert pb.lib.b200_batch_send_plan(size, 0, ranges.ctypes.data_as(ip), C.c_longlong(1), 3, bad.ctypes.data_as(ip),
                                       cnt.ctypes.data_as(llp), None) == 63
