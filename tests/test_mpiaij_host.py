"""Host-side integer structures of the row-partitioned operator (C ABI) against the independent numpy restatement in
oracle/mpiaij_restate.py — bit-exact — plus the world_size-2 gloo path of petsc-dev_b200/parallel.py."""
import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gens(P):
    return {
        "lap2d": (30 * 20, 1, lambda r0, r1: P.laplace2d_5pt(30, 20, r0, r1)),
        "lap7": (10 ** 3, 1, lambda r0, r1: P.laplace3d_7pt(10, r0=r0, r1=r1)),
        "s27": (9 ** 3, 1, lambda r0, r1: P.stencil3d_27pt(9, r0=r0, r1=r1)),
        "bs3": (3 * 6 ** 3, 3, lambda r0, r1: P.elasticity_bs3(6, r0=r0, r1=r1)),
        "plaw": (2000, 1, lambda r0, r1: P.powerlaw(2000, r0, r1, seed=4, xm=4.0, lmax=300)),
    }


def serial_plans(par, size, N, gen, bs):
    """Drive the C ABI for all virtual ranks in one process (the exchange is a dictionary lookup)."""
    ranges = par.split_ownership(N, size, bs)
    host = []
    for r in range(size):
        ai, aj, aa = gen(int(ranges[r]), int(ranges[r + 1]))
        ai = ai - ai[0]
        m = int(ranges[r + 1] - ranges[r])
        sp = par.mpiaij_split(m, int(ranges[r]), int(ranges[r + 1]), ai, aj, aa)
        g = sp["garray"]
        blk = par.use_block_is(g, bs)
        pbs = bs if blk else 1
        inidx = (g[::pbs] // pbs).astype(np.int32)
        frm = par.scatter_plan_from(inidx, np.arange(inidx.size, dtype=np.int32), pbs, ranges, r, g.size)
        host.append(dict(split=sp, from_=frm, bs=pbs, block_is=blk))
    for p in range(size):
        procs, lengths, req = [], [], []
        for r in range(size):
            f = host[r]["from_"]
            c = int(f["req_counts"][p])
            if r != p and c:
                off = int(f["req_counts"][:p].sum())
                procs.append(r); lengths.append(c); req.append(f["req_values"][off:off + c])
        req = np.concatenate(req) if req else np.zeros(0, np.int32)
        host[p]["to"] = par.scatter_plan_to(np.array(procs, np.int32), np.array(lengths, np.int32), req,
                                            int(ranges[p]), int(ranges[p + 1]))
    return ranges, host


@pytest.mark.parametrize("name", ["lap2d", "lap7", "s27", "bs3", "plaw"])
@pytest.mark.parametrize("size", [1, 2, 3, 8])
def test_structures_bit_exact(pb, P, name, size):
    par = importlib.import_module("petsc-dev_b200.parallel")
    from oracle import mpiaij_restate as O
    N, bs, gen = _gens(P)[name]
    def gen0(r0, r1):
        ai, aj, aa = gen(r0, r1)
        return ai - ai[0], aj, aa
    ranges, host = serial_plans(par, size, N, gen, bs)
    oranges, S = O.all_structures(size, N, gen0, bs)
    assert np.array_equal(ranges, oranges)
    for r in range(size):
        sp, s = host[r]["split"], S[r]
        for k in ("Ai", "Aj", "Bi", "Bj", "garray"):
            assert sp[k].dtype == np.int32 and np.array_equal(sp[k], s[k]), (name, size, r, k)
        assert np.array_equal(sp["Aa"], s["Aa"]) and np.array_equal(sp["Ba"], s["Ba"])
        assert host[r]["block_is"] == s["block_is"] and host[r]["bs"] == s["bs"]
        f, t = host[r]["from_"], host[r]["to"]
        assert np.array_equal(f["from_procs"], s["from_procs"])
        assert np.array_equal(f["from_starts"], s["from_starts"])
        assert np.array_equal(f["from_indices"], s["from_indices"])
        assert np.array_equal(t["to_procs"], s["to_procs"])
        assert np.array_equal(t["to_starts"], s["to_starts"])
        assert np.array_equal(t["to_indices"], s["to_indices"])
        assert f["nlocal"] == 0                       # garray never holds an owned column
        # colmap: colmap[garray[i]] = i + 1 (mpiaij.c:373)
        cm = par.colmap(N, sp["garray"])
        assert np.array_equal(np.nonzero(cm)[0], sp["garray"]) and np.array_equal(cm[sp["garray"]], np.arange(1, sp["garray"].size + 1))
        # reassembly: A/B/garray reproduce the original rows exactly
        ai, aj, aa = gen0(int(ranges[r]), int(ranges[r + 1]))
        m = int(ranges[r + 1] - ranges[r])
        for row in range(0, m, max(1, m // 50)):
            ca = sp["Aj"][sp["Ai"][row]:sp["Ai"][row + 1]] + int(ranges[r])
            cb = sp["garray"][sp["Bj"][sp["Bi"][row]:sp["Bi"][row + 1]]]
            assert np.array_equal(np.sort(np.concatenate([ca, cb])), aj[ai[row]:ai[row + 1]])


def test_general_scatter_with_self_part(pb):
    """VecScatterCreate_PtoS with unsorted indices and locally owned entries (the general MPI->Seq case)."""
    par = importlib.import_module("petsc-dev_b200.parallel")
    ranges = np.array([0, 4, 8, 12], np.int32)
    inidx = np.array([9, 1, 5, 4, 11, 0, 6], np.int32)
    inidy = np.arange(7, dtype=np.int32)
    f = par.scatter_plan_from(inidx, inidy, 1, ranges, 1, 7)
    assert f["nfrom"] == 2 and list(f["from_procs"]) == [0, 2]
    assert list(f["from_starts"]) == [0, 2, 4]
    assert list(f["from_indices"]) == [1, 5, 0, 4]
    assert list(f["req_counts"]) == [2, 0, 2] and list(f["req_values"]) == [1, 0, 9, 11]
    assert f["nlocal"] == 3 and list(f["to_local"]) == [1, 0, 2] and list(f["from_local"]) == [2, 3, 6]
    with pytest.raises(pb.PetscB200Error) as e:
        par.scatter_plan_from(np.array([12], np.int32), np.array([0], np.int32), 1, ranges, 0, 1)
    assert e.value.code == 77
    with pytest.raises(pb.PetscB200Error) as e:
        par.scatter_plan_to(np.array([0], np.int32), np.array([1], np.int32), np.array([3], np.int32), 4, 8)
    assert e.value.code == 63


def _gloo_worker(rank, size, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as td
    td.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=size)
    P = importlib.import_module("workloads.problems")
    par = importlib.import_module("petsc-dev_b200.parallel")
    N = 12 ** 3
    ranges = par.split_ownership(N, size)
    ai, aj, aa = P.stencil3d_27pt(12, r0=int(ranges[rank]), r1=int(ranges[rank + 1]))
    h = par.setup_mpiaij_host(rank, size, ranges, ai - ai[0], aj, aa, N)
    q.put((rank, {k: h["to"][k] for k in ("to_procs", "to_starts", "to_indices")},
           {k: h["from_"][k] for k in ("from_procs", "from_starts")}, h["split"]["garray"]))
    td.barrier()
    td.destroy_process_group()


def test_gloo_world_size_2(P):
    """The N > 1 host path (request exchange over torch.distributed) on CPU with gloo, world_size 2."""
    import torch.multiprocessing as mp
    from oracle import mpiaij_restate as O
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = {}
    for _ in range(2):
        r, to, frm, g = q.get(timeout=120)
        got[r] = (to, frm, g)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    def gen0(r0, r1):
        ai, aj, aa = P.stencil3d_27pt(12, r0=r0, r1=r1)
        return ai - ai[0], aj, aa
    _, S = O.all_structures(2, 12 ** 3, gen0)
    for r in range(2):
        to, frm, g = got[r]
        assert np.array_equal(g, S[r]["garray"])
        for k in ("to_procs", "to_starts", "to_indices"):
            assert np.array_equal(to[k], S[r][k]), k
        for k in ("from_procs", "from_starts"):
            assert np.array_equal(frm[k], S[r][k]), k


def _gloo_scatter_worker(rank, size, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as td
    td.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=size)
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, bs = 120, 3
    ranges = par.split_ownership(N, size, bs)
    rng = np.random.RandomState(7 + rank)
    inidx = rng.randint(0, N // bs, size=25).astype(np.int32)          # block ids, repeats allowed
    inidy = rng.permutation(25).astype(np.int32)
    frm = par.scatter_plan_from(inidx, inidy, bs, ranges, rank, 25 * bs)
    procs, lengths, req = par.exchange_requests(frm, rank, size)
    to = par.scatter_plan_to(procs, lengths, req, int(ranges[rank]), int(ranges[rank + 1]))
    q.put((rank, inidx, inidy, {k: np.asarray(v) for k, v in frm.items()}, {k: np.asarray(v) for k, v in to.items()}))
    td.barrier()
    td.destroy_process_group()


def test_gloo_general_scatter_plan(pb):
    """VecScatterCreate_PtoS for arbitrary (repeated, unsorted, partly local) block indices over 2 gloo ranks: what every
    owner is asked to send equals, requester by requester in ascending rank, what the requesters' receive lists expect."""
    import torch.multiprocessing as mp
    par = importlib.import_module("petsc-dev_b200.parallel")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31000 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_scatter_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got = {}
    for _ in range(2):
        r, inidx, inidy, frm, to = q.get(timeout=120)
        got[r] = (inidx, inidy, frm, to)
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    N, bs = 120, 3
    ranges = par.split_ownership(N, 2, bs)
    for r in range(2):
        inidx, inidy, frm, to = got[r]
        o = 1 - r
        owned = (bs * inidx >= ranges[r]) & (bs * inidx < ranges[r + 1])
        # self part, in list order (vpscat.c:1890-1903)
        assert np.array_equal(frm["to_local"], bs * inidx[owned] - ranges[r]) and np.array_equal(frm["from_local"], bs * inidy[owned])
        # remote part: requests in list order, receive positions alongside
        assert int(frm["nfrom"]) == (1 if (~owned).any() else 0)
        assert np.array_equal(frm["req_values"], bs * inidx[~owned]) and np.array_equal(frm["from_indices"], bs * inidy[~owned])
        # the owner's send list is exactly that request, rebased to its range
        oto = got[o][3]
        if (~owned).any():
            assert list(oto["to_procs"]) == [r]
            assert np.array_equal(oto["to_indices"], bs * inidx[~owned] - ranges[o])
            assert list(oto["to_starts"]) == [0, int((~owned).sum())]


def test_split_is_the_same_on_any_number_of_host_threads(monkeypatch):
    """b200_mpiaij_split works on row ranges in parallel for large inputs; B200_HOST_THREADS forces that path on a small
    one: diag/off-diag blocks and garray must be identical to the one-thread result."""
    import importlib
    par = importlib.import_module("petsc-dev_b200.parallel")
    P = importlib.import_module("workloads.problems")
    N = 6000
    ranges = par.split_ownership(N, 3)
    for rank in range(3):
        r0, r1 = int(ranges[rank]), int(ranges[rank + 1])
        ip, j, a = P.powerlaw(N, r0, r1)
        ip = (ip - ip[0]).astype(np.int32)
        monkeypatch.setenv("B200_HOST_THREADS", "1")
        one = par.mpiaij_split(r1 - r0, r0, r1, ip, j, a)
        for T in ("2", "7", "16"):
            monkeypatch.setenv("B200_HOST_THREADS", T)
            many = par.mpiaij_split(r1 - r0, r0, r1, ip, j, a)
            for key in one:
                assert np.array_equal(one[key], many[key]), (rank, T, key)
