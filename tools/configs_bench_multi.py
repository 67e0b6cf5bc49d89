#!/usr/bin/env python3
"""BASELINE.json configs[2..4] at their multi-GPU sizes (run under torch.distributed.run, one rank per GPU):
  C3  3-D 27-point stencil 512^3, MPIAIJ row slabs, KSPCG + Jacobi
  C4  irregular power-law matrix, 10 M rows (~50 nnz/row), MatMult + KSPGMRES(30) + Jacobi
  C5  bs = 3 blocked 27-point stencil, 3 x 128^3 unknowns, KSPCG + Jacobi
One JSON line per config on rank 0: MatMult ms, aggregate algorithmic GB/s, per-GPU fraction of the measured HBM peak,
solver iterations/s.  Sizes can be scaled down for a smoke run:  --scale 0.25  (grid edge / row count multiplier).

  python -m torch.distributed.run --nproc-per-node 8 ... tools/configs_bench_multi.py [c3 c4 c5] [--scale S] [--transport p2p|nccl]
"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

rank, lrank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(lrank)
dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
pb = importlib.import_module("petsc-dev_b200")
par = importlib.import_module("petsc-dev_b200.parallel")
P = importlib.import_module("workloads.problems")
pb.init(lrank)
par.comm_init(rank, world, dist, device="cuda")
argv = sys.argv[1:]
scale = float(argv[argv.index("--scale") + 1]) if "--scale" in argv else 1.0
transport = argv[argv.index("--transport") + 1] if "--transport" in argv else "p2p"
which = [a for a in argv if a in ("c3", "c4", "c5")] or ["c3", "c5", "c4"]
if transport == "p2p" and not par.enable_p2p(strict=False):
    transport = "nccl"
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0


def allsum(v):
    t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t)
    return float(t.item())


def allmax(v):
    t = torch.tensor([float(v)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run(tag, desc, N, bs, gen, solver, iters):
    t0 = time.perf_counter()
    ranges = par.split_ownership(N, world, bs)
    rs, re_ = int(ranges[rank]), int(ranges[rank + 1])
    ai, aj, aa = gen(rs, re_)
    nz_local = len(aa)
    host = par.setup_mpiaij_host(rank, world, ranges, ai - ai[0], aj, aa, N, bs=bs, dist=dist, device="cuda")
    A = par.create_mpiaij(host)
    m = host["m"]
    nghost = int(host["split"]["garray"].size)
    del ai, aj, aa, host
    t_setup = time.perf_counter() - t0
    x, y = pb.Vec(m, P.random_vector(m, 2, r0=rs)), pb.Vec(m)
    for _ in range(3):
        A.mult(x, y)
    pb.synchronize(); dist.barrier()
    pb.timer_start()
    reps = 20
    for _ in range(reps):
        A.mult(x, y)
    ms = allmax(pb.timer_stop()) / reps
    Btot = allsum(A.bytes())
    nz = allsum(nz_local)
    out = {"config": tag, "desc": desc, "n_gpus": world, "transport": transport, "rows": N, "nnz": int(nz), "nghost_rank0": nghost,
           "matmult_ms": ms, "matmult_gbs_total": Btot / ms / 1e6, "per_gpu_frac_of_measured_peak": Btot / ms / 1e6 / world / PEAK,
           "gflops_total": (2.0 * nz - N) / ms / 1e6, "strips_rank0": A.strips(), "setup_s": t_setup}
    one, b, sol = pb.Vec(m), pb.Vec(m), pb.Vec(m)
    if solver == "cg":
        one.set(1.0); A.mult(one, b)
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=3, fused=1, histmax=0)
        dist.barrier()
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=iters, fused=1, check_every=iters, histmax=0)
        t = allmax(r["solve_ms"])
        out["kspcg_fused"] = {"iter_per_s": r["its"] / t * 1e3, "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"],
                              "gbs_vs_Bcg_min": (Btot + 88.0 * N) / (t / max(r["its"], 1)) / 1e6}
    else:
        A.mult(x, b)
        pb.ksp_solve(A, b, sol, ksp_type="gmres", pc_type="jacobi", rtol=1e-50, max_it=3, histmax=0)
        dist.barrier()
        r = pb.ksp_solve(A, b, sol, ksp_type="gmres", pc_type="jacobi", rtol=1e-50, abstol=1e-300, max_it=iters, histmax=0)
        t = allmax(r["solve_ms"])
        out["kspgmres30"] = {"iter_per_s": r["its"] / t * 1e3, "ms_per_iter": t / max(r["its"], 1), "iterations": r["its"], "reason": r["reason"]}
    if rank == 0:
        print(json.dumps(out), flush=True)
    del A, x, y, one, b, sol
    dist.barrier()


if "c3" in which:
    nx = int(argv[argv.index("--c3-nx") + 1]) if "--c3-nx" in argv else max(int(round(512 * scale)), 8)
    run("C3", "3-D 27-point stencil %d^3, row slabs" % nx, nx ** 3, 1, lambda r0, r1: P.stencil3d_27pt(nx, r0=r0, r1=r1), "cg", 50)
if "c5" in which:
    nx = max(int(round(128 * scale)), 4)
    run("C5", "bs=3 blocked 27-point stencil, 3 x %d^3 unknowns" % nx, 3 * nx ** 3, 3, lambda r0, r1: P.elasticity_bs3(nx, r0=r0, r1=r1), "cg", 50)
if "c4" in which:
    rows = max(int(round(10_000_000 * scale)), 1000)
    run("C4", "irregular power-law row lengths, %d rows" % rows, rows, 1, lambda r0, r1: P.powerlaw(rows, r0, r1), "gmres", 60)
dist.barrier()
dist.destroy_process_group()
