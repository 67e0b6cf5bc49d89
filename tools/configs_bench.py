#!/usr/bin/env python3
"""All five BASELINE.json configs on ONE B200 (sizes that fit one GPU and 32-bit nnz), one JSON line each:
MatMult ms / algorithmic GB/s / fraction of the measured HBM peak, and the Krylov solver rate the config names
(KSPCG + Jacobi, or KSPGMRES(30) + Jacobi for the irregular matrix).  Development/evidence tool; the headline
measurement is bench.py.

  python tools/configs_bench.py [c1 c2 c3 c4 c5] [--plaw-rows 10000000]
"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0
args = [a for a in sys.argv[1:] if not a.startswith("--")]
which = args or ["c1", "c2", "c3", "c4", "c5"]
plaw_rows = 10_000_000
if "--plaw-rows" in sys.argv:
    plaw_rows = int(sys.argv[sys.argv.index("--plaw-rows") + 1])


def run(tag, desc, m, ip, j, a, solver, flush=False, iters=100):
    t0 = time.perf_counter()
    A = pb.Mat.from_csr(m, m, ip, j, a)
    nz = len(a)
    del ip, j, a
    t_up = time.perf_counter() - t0
    x, y = pb.Vec(m, P.random_vector(m, 2)), pb.Vec(m)
    for _ in range(3):
        A.mult(x, y)
    reps = 20
    if flush:                       # matrix smaller than L2: flush between launches, time each launch alone
        ms = 0.0
        for _ in range(reps):
            pb.flush_l2(); pb.synchronize(); pb.timer_start(); A.mult(x, y); ms += pb.timer_stop()
        ms /= reps
    else:
        pb.synchronize(); pb.timer_start()
        for _ in range(reps):
            A.mult(x, y)
        ms = pb.timer_stop() / reps
    B = A.bytes()
    out = {"config": tag, "desc": desc, "rows": m, "nnz": int(nz), "matmult_ms": ms, "matmult_gbs": B / ms / 1e6,
           "matmult_frac_of_measured_peak": B / ms / 1e6 / PEAK, "gflops": A.flops() / ms / 1e6, "algorithmic_bytes": B,
           "l2": "flushed between launches" if flush else "inputs larger than L2", "bins": A.bin_info(), "upload_s": t_up}
    one, b, sol = pb.Vec(m), pb.Vec(m), pb.Vec(m)
    one.set(1.0); A.mult(one, b)
    if solver == "cg":
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=5, fused=1, histmax=0)
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=iters, fused=1, check_every=iters, histmax=0)
        out["kspcg_fused"] = {"iter_per_s": r["its"] / r["solve_ms"] * 1e3, "ms_per_iter": r["solve_ms"] / r["its"], "iterations": r["its"],
                              "gbs_vs_Bcg_min": (B + 88.0 * m) / (r["solve_ms"] / r["its"]) / 1e6}
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=iters, fused=0, histmax=0)
        out["kspcg_op_by_op"] = {"iter_per_s": r["its"] / r["solve_ms"] * 1e3, "ms_per_iter": r["solve_ms"] / r["its"]}
    else:
        # b = A * random so GMRES does not converge at once on the diagonally dominant irregular matrix
        A.mult(x, b)
        pb.ksp_solve(A, b, sol, ksp_type="gmres", pc_type="jacobi", rtol=1e-50, max_it=5, histmax=0)
        its = 60
        r = pb.ksp_solve(A, b, sol, ksp_type="gmres", pc_type="jacobi", rtol=1e-50, abstol=1e-300, max_it=its, histmax=0)
        # bytes of one restart cycle of 30: per step k (1..30) SpMV + Jacobi 24n + MDot 8n(k+1) + MAXPY 8n(k+2) + norm 8n + scale 16n
        cyc = sum(B + 24.0 * m + 8.0 * m * (k + 1) + 8.0 * m * (k + 2) + 24.0 * m for k in range(1, 31))
        out["kspgmres30"] = {"iter_per_s": r["its"] / r["solve_ms"] * 1e3, "ms_per_iter": r["solve_ms"] / max(r["its"], 1), "iterations": r["its"],
                             "reason": r["reason"], "gbs_vs_unfused_bytes": (cyc / 30.0) / (r["solve_ms"] / max(r["its"], 1)) / 1e6}
    print(json.dumps(out), flush=True)
    del A


if "c1" in which:
    ip, j, a = P.laplace2d_5pt(1000, 1000)
    run("C1", "ex2 2-D 5-pt Laplacian m=n=1000 (matrix 80 MB < L2)", 10 ** 6, ip, j, a, "cg", flush=True)
if "c2" in which:
    ip, j, a = P.laplace3d_7pt(256)
    run("C2", "3-D 7-pt Laplacian 256^3", 256 ** 3, ip, j, a, "cg")
if "c3" in which:
    ip, j, a = P.stencil3d_27pt(256)
    run("C3@1GPU", "3-D 27-pt stencil 256^3 (512^3 needs >= 2 ranks: nnz > 2^31)", 256 ** 3, ip, j, a, "cg", iters=50)
if "c5" in which:
    ip, j, a = P.elasticity_bs3(128)
    run("C5", "bs=3 blocked 27-pt stencil, 3 x 128^3 unknowns", 3 * 128 ** 3, ip, j, a, "cg", iters=50)
if "c4" in which:
    t0 = time.perf_counter()
    ip, j, a = P.powerlaw(plaw_rows)
    print("# powerlaw generated in %.1f s: nnz %d, mean %.1f, max row %d" % (time.perf_counter() - t0, len(a), len(a) / plaw_rows, int(np.diff(ip).max())), flush=True)
    run("C4", "irregular power-law row lengths, %d rows" % plaw_rows, plaw_rows, ip, j, a, "gmres")
