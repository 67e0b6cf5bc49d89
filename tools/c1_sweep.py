#!/usr/bin/env python3
"""Small-problem sweep (development tool): ex2's 2-D 5-point Laplacian at m = n = 100 ... 1000, fused KSPCG + Jacobi as the
persistent cooperative kernel and as four launches per iteration, and the MatMult launch shapes, one line each."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
for m in (100, 200, 300, 500, 700, 1000):
    ip, j, a = P.laplace2d_5pt(m, m)
    n = m * m
    A = pb.Mat.from_csr(n, n, ip, j, a)
    one, b, sol = pb.Vec(n), pb.Vec(n), pb.Vec(n)
    one.set(1.0); A.mult(one, b)
    out = {}
    for mode, env in (("persistent", {"B200_CG_PERSISTENT": "1"}), ("launches", {"B200_CG_GRAPH": "0"}), ("graph", {})):
        os.environ.update(env)
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=20, fused=1, histmax=0)
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=1280, fused=1, check_every=256, histmax=0)
        out[mode] = 1e3 * r["solve_ms"] / r["its"]
        for k in env: del os.environ[k]
    x, y = pb.Vec(n, P.random_vector(n, 2)), pb.Vec(n)
    for _ in range(5): A.mult(x, y)
    pb.synchronize(); pb.timer_start()
    for _ in range(200): A.mult(x, y)
    mm = 1e3 * pb.timer_stop() / 200
    print("m=%4d rows %7d nnz %8d  CG us/iter: persistent %.2f  four launches %.2f  graph replay %.2f   MatMult (L2-warm) %.2f us" % (m, n, len(a), out["persistent"], out["launches"], out["graph"], mm), flush=True)
    del A
if "cfg" in sys.argv:
    ip, j, a = P.laplace2d_5pt(1000, 1000); n = 10 ** 6
    for cfg in range(8):
        pb.spmv_tuning(shape=cfg)
        try:
            A = pb.Mat.from_csr(n, n, ip, j, a)
        except Exception as e:
            print("cfg", cfg, "failed", e); continue
        x, y = pb.Vec(n, P.random_vector(n, 2)), pb.Vec(n)
        for _ in range(5): A.mult(x, y)
        pb.synchronize(); pb.timer_start()
        for _ in range(200): A.mult(x, y)
        mm = 1e3 * pb.timer_stop() / 200
        one, b, sol = pb.Vec(n), pb.Vec(n), pb.Vec(n)
        one.set(1.0); A.mult(one, b)
        pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=20, fused=1, histmax=0)
        r = pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=1280, fused=1, check_every=256, histmax=0)
        print("1000^2 shape %d: MatMult (L2-warm) %.2f us, CG (graph replay) %.2f us/iter  %s" % (cfg, mm, 1e3 * r["solve_ms"] / r["its"], A.bin_info()), flush=True)
        del A
    pb.spmv_tuning()
