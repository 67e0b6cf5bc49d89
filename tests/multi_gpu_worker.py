"""One rank of the real multi-GPU parity run (launched by tests/test_gpu_multi.py through torch.distributed.run, one
process per GPU, NCCL).  Every rank builds its row slab of the same global matrix, the library exchanges the ghost
entries GPU to GPU, and rank 0 compares the gathered result with the reference's own sequential kernels applied per
rank (oracle/mpiaij_restate.mult = MatMult_SeqAIJ on the diagonal block + MatMultAdd_SeqAIJ on the off-diagonal block,
mpiaij.c:1111-1114) and the KSP residual histories with the reference's KSPSolve on the whole matrix.
Prints one JSON line on rank 0."""
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    rank, lrank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    transport = sys.argv[1] if len(sys.argv) > 1 else "nccl"
    torch.cuda.set_device(lrank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
    pb = importlib.import_module("petsc-dev_b200")
    par = importlib.import_module("petsc-dev_b200.parallel")
    P = importlib.import_module("workloads.problems")
    pb.init(lrank)
    par.comm_init(rank, world, dist, device="cuda")
    if transport == "p2p":
        par.enable_p2p(rank, world, dist)
    cases = {
        "lap2d": (40 * 30, 1, lambda r0, r1: P.laplace2d_5pt(40, 30, r0, r1)),
        "lap7": (24 ** 3, 1, lambda r0, r1: P.laplace3d_7pt(24, r0=r0, r1=r1)),
        "s27": (16 ** 3, 1, lambda r0, r1: P.stencil3d_27pt(16, r0=r0, r1=r1)),
        "bs3": (3 * 8 ** 3, 3, lambda r0, r1: P.elasticity_bs3(8, r0=r0, r1=r1)),
        "plaw": (20000, 1, lambda r0, r1: P.powerlaw(20000, r0, r1, seed=4, xm=4.0, lmax=600)),
    }
    out = {"world": world, "transport": transport, "cases": {}}
    R = None
    if rank == 0:
        from oracle import refpetsc as R
        from oracle import mpiaij_restate as O
    for name, (N, bs, gen) in cases.items():
        ranges = par.split_ownership(N, world, bs)
        rs, re_ = int(ranges[rank]), int(ranges[rank + 1])
        m = re_ - rs
        ai, aj, aa = gen(rs, re_)
        ai = ai - ai[0]
        lanes = -1 if name == "plaw" else 1           # 1: reference left-to-right row sums -> bit-exact
        pb.spmv_tuning(lanes_per_row=lanes)
        try:
            host = par.setup_mpiaij_host(rank, world, ranges, ai, aj, aa, N, bs=bs, dist=dist, device="cuda")
            A = par.create_mpiaij(host)
        finally:
            pb.spmv_tuning()
        # same pattern, new values (b200_mat_update_values_mpiaij): away and back again; everything below then runs
        # on the refreshed operator, and the doubled one must give exactly twice the product
        sp = host["split"]
        xh0 = P.random_vector(N, 2)
        x0, z0, z1 = pb.Vec(re_ - rs, xh0[rs:re_]), pb.Vec(re_ - rs), pb.Vec(re_ - rs)
        A.mult(x0, z0)
        A.update_values_mpiaij(2.0 * np.asarray(sp["Aa"]), 2.0 * np.asarray(sp["Ba"]))
        A.mult(x0, z1)
        refresh_ok = bool(np.array_equal(z1.get_array(), 2.0 * z0.get_array()))
        A.update_values_mpiaij(sp["Aa"], sp["Ba"])
        xh, yh = P.random_vector(N, 2), P.random_vector(N, 3)
        x, y, z = pb.Vec(m, xh[rs:re_]), pb.Vec(m, yh[rs:re_]), pb.Vec(m)

        def gather(v):
            mx = int(np.max(np.diff(ranges)))
            t = torch.zeros(mx, dtype=torch.float64, device="cuda")
            t[:m] = torch.from_numpy(v.get_array()).cuda()
            parts = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(parts, t)
            return np.concatenate([parts[r][: int(ranges[r + 1] - ranges[r])].cpu().numpy() for r in range(world)])

        for _ in range(3):                            # repeated products: the double-buffered ghost vector must stay right
            A.mult(x, z)
        got = gather(z)
        A.mult_add(x, y, z)
        got_add = gather(z)
        for _ in range(3):                            # reverse exchange (ADD): repeated, both parities of the return buffers
            A.mult_transpose(x, z)
        got_t = gather(z)
        A.mult_transpose_add(x, y, z)
        got_tadd = gather(z)
        # vector reductions across ranks
        dot = x.dot(y)
        nrm = x.norm()
        # KSP CG / GMRES on the distributed operator (SPD cases only)
        ksp = {}
        if name in ("lap2d", "lap7", "s27", "bs3"):
            ones = pb.Vec(m); ones.set(1.0)
            b, sol = pb.Vec(m), pb.Vec(m)
            A.mult(ones, b)
            for kt, fused in (("cg", 0), ("cg", 1), ("gmres", 0)):
                sol.set(0.0)
                r = pb.ksp_solve(A, b, sol, ksp_type=kt, pc_type="jacobi", rtol=1e-8, fused=fused, check_every=1)
                ksp["%s_%d" % (kt, fused)] = dict(its=int(r["its"]), reason=int(r["reason"]), history=[float(v) for v in r["history"]],
                                                  sol=gather(sol).tolist() if N <= 5000 else None)
        rk = torch.tensor([1.0 if refresh_ok else 0.0], device="cuda")
        dist.all_reduce(rk, op=dist.ReduceOp.MIN)
        refresh_ok = bool(rk.item() == 1.0)
        if rank == 0:
            def gen0(r0, r1):
                a_i, a_j, a_a = gen(r0, r1)
                return a_i - a_i[0], a_j, a_a
            _, S = O.all_structures(world, N, gen0, bs)
            want = O.mult(R, ranges, S, xh)
            want_add = O.mult(R, ranges, S, xh, add_y=yh)
            res = {}
            want_t = O.mult_transpose(R, ranges, S, xh)
            want_tadd = O.mult_transpose(R, ranges, S, xh, add_y=yh)
            sct = np.maximum(np.abs(want_t), 1.0)
            if name in ("plaw", "bs3"):               # several lanes per transposed row (bs3: 81 entries) -> tolerance
                res["multtranspose_ok"] = bool(np.all(np.abs(got_t - want_t) <= 1e-12 * sct) and
                                               np.all(np.abs(got_tadd - want_tadd) <= 1e-12 * (sct + np.abs(yh))))
            else:
                res["multtranspose_ok"] = bool(np.array_equal(got_t, want_t) and np.array_equal(got_tadd, want_tadd))
            res["multtranspose_max_abs_err"] = float(np.max(np.abs(got_t - want_t)))
            if name == "plaw":
                sc = np.maximum(np.abs(want), 1.0)
                res["mult_ok"] = bool(np.all(np.abs(got - want) <= 1e-12 * sc))
                res["multadd_ok"] = bool(np.all(np.abs(got_add - want_add) <= 1e-12 * (sc + np.abs(yh))))
                res["bit_exact"] = False
            else:
                res["mult_ok"] = bool(np.array_equal(got, want))
                res["multadd_ok"] = bool(np.array_equal(got_add, want_add))
                res["bit_exact"] = True
            res["max_abs_err"] = float(np.max(np.abs(got - want)))
            res["refresh_ok"] = refresh_ok
            res["dot_rel"] = float(abs(dot - R.vec_dot(xh, yh)) / abs(R.vec_dot(xh, yh)))
            res["norm_rel"] = float(abs(nrm - R.vec_norm(xh)) / R.vec_norm(xh))
            if ksp:
                fi, fj, fa = gen0(0, N)
                ref = R.RefMat(N, N, fi, fj, fa, no_inode=True)
                bh = ref.mult(np.ones(N))
                for key, r in ksp.items():
                    kt = key.split("_")[0]
                    rr = ref.ksp_solve(bh, kt, "jacobi", rtol=1e-8)
                    k = min(len(r["history"]), len(rr["history"]))
                    h, hr = np.array(r["history"][:k]), np.array(rr["history"][:k])
                    # gate: compare while the residual is above 1e-6 of the start (rounding-amplified tail excluded)
                    gate = hr > 1e-6 * hr[0]
                    dev = float(np.max(np.abs(h[gate] - hr[gate]) / hr[gate]))
                    res[key] = dict(its=r["its"], ref_its=int(rr["its"]), reason=r["reason"], ref_reason=int(rr["reason"]), hist_dev=dev)
            out["cases"][name] = res
        del A
    # ---- general device VecScatter (parallel <-> sequential), integer-valued data so every summation order is exact ----
    sc = {}
    for bs in (1, 3):
        N = 999
        ranges = par.split_ownership(N, world, bs)
        rs, re_ = int(ranges[rank]), int(ranges[rank + 1])
        m = re_ - rs
        rng = np.random.RandomState(100 + rank)
        nblk = 60
        inidx = rng.randint(0, N // bs, size=nblk).astype(np.int32)        # block ids, repeats allowed, some owned here
        inidx[:5] = (rs // bs) + np.arange(5) % max(m // bs, 1)              # force a self part
        inidy = rng.permutation(nblk).astype(np.int32)                      # distinct positions in the sequential vector
        ny = nblk * bs
        S = par.create_scatter(rank, world, ranges, inidx, inidy, ny, bs=bs, dist=dist, device="cuda")
        xg = np.arange(N, dtype=np.float64) * 3.0 + 1.0
        y0 = np.arange(ny, dtype=np.float64) * 7.0 + rank
        gi = (bs * inidx[:, None] + np.arange(bs)[None, :]).reshape(-1)       # scalar global ids
        yi = (bs * inidy[:, None] + np.arange(bs)[None, :]).reshape(-1)
        x, y = pb.Vec(m, xg[rs:re_]), pb.Vec(ny, y0)
        ok = True
        for _ in range(2):                                                  # both buffer parities
            S.scatter(x, y)                                                  # forward INSERT
            want = y0.copy(); want[yi] = xg[gi]
            ok = ok and np.array_equal(y.get_array(), want)
        y.set_array(y0)
        S.scatter(x, y, add=True)                                            # forward ADD
        want = y0.copy(); np.add.at(want, yi, xg[gi])
        ok = ok and np.array_equal(y.get_array(), want)
        # reverse ADD: every rank's y flows back and accumulates on the owners
        y.set_array(y0)
        allgi = [torch.zeros(gi.size, dtype=torch.int64, device="cuda") for _ in range(world)]
        allyv = [torch.zeros(gi.size, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(allgi, torch.from_numpy(gi.astype(np.int64)).cuda())
        dist.all_gather(allyv, torch.from_numpy(y0[yi]).cuda())
        for rep in range(2):
            x.set_array(xg[rs:re_])
            S.scatter(x, y, add=True, reverse=True)
            want = xg.copy()
            for r in range(world):
                np.add.at(want, allgi[r].cpu().numpy(), allyv[r].cpu().numpy())
            ok = ok and np.array_equal(x.get_array(), want[rs:re_])
        sc["bs%d" % bs] = bool(ok)
        del S
    # ---- MatSetValuesBatch on the parallel type: every rank brings a chunk of the element list, interface elements
    # travel to the owners of the rows they touch, the local rows are assembled on the device ----
    N, nl, er, em = P.elements_hex3d(14)
    ne = er.size // nl
    e0, e1 = ne * rank // world, ne * (rank + 1) // world
    ranges = par.split_ownership(N, world)
    rs, re_ = int(ranges[rank]), int(ranges[rank + 1])
    Ab, (bi, bj, ba) = par.set_values_batch_mpiaij(rank, world, ranges, N, nl, er[e0 * nl:e1 * nl], em[e0 * nl * nl:e1 * nl * nl],
                                                   dist=dist, device="cuda", return_host=True)
    from oracle import assembly_restate as AR
    gi, gj, ga = AR.setvaluesbatch(N, N, er, em, nl)             # the whole matrix, every rank for itself
    lo, hi = gi[rs], gi[re_]
    batch_ok = bool(np.array_equal(bi, gi[rs:re_ + 1] - lo) and np.array_equal(bj, gj[lo:hi]) and
                    np.all(np.abs(ba - ga[lo:hi]) <= 1e-13 * np.maximum(np.abs(ga[lo:hi]), 1.0)))
    xg = P.random_vector(N, 5)
    xb, yb = pb.Vec(re_ - rs, xg[rs:re_]), pb.Vec(re_ - rs)
    Ab.mult(xb, yb)
    want = np.array([np.dot(ga[gi[r]:gi[r + 1]], xg[gj[gi[r]:gi[r + 1]]]) for r in range(rs, re_)])
    sc_rows = np.array([np.sum(np.abs(ga[gi[r]:gi[r + 1]] * xg[gj[gi[r]:gi[r + 1]]])) for r in range(rs, re_)])
    batch_ok = batch_ok and bool(np.all(np.abs(yb.get_array() - want) <= 1e-12 * np.maximum(sc_rows, 1e-300)))
    bk = torch.tensor([1.0 if batch_ok else 0.0], device="cuda")
    dist.all_reduce(bk, op=dist.ReduceOp.MIN)
    out["batch_assembly_ok"] = bool(bk.item() == 1.0)
    del Ab
    oks = torch.tensor([1.0 if all(sc.values()) else 0.0], device="cuda")
    dist.all_reduce(oks, op=dist.ReduceOp.MIN)
    out["scatter"] = dict(sc, all_ranks=bool(oks.item() == 1.0))
    if rank == 0:
        print("MULTI_GPU_RESULT " + json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
