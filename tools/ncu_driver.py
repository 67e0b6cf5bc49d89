#!/usr/bin/env python3
"""Tiny driver for ncu captures (development tool): builds one BASELINE matrix and runs a few MatMults / a short CG.
  python tools/ncu_driver.py lap7 256 [cg]"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
kind = sys.argv[1] if len(sys.argv) > 1 else "lap7"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 256
if kind == "lap7":
    ip, j, a = P.laplace3d_7pt(nx); m = nx ** 3
elif kind == "s27":
    ip, j, a = P.stencil3d_27pt(nx); m = nx ** 3
elif kind == "bs3":
    ip, j, a = P.elasticity_bs3(nx); m = 3 * nx ** 3
elif kind == "plaw":
    cache = os.path.join(os.environ.get("B200_CACHE_DIR", "/tmp"), "plaw_%d.npz" % nx)     # the second (ncu) run of a call reuses it
    if os.path.exists(cache):
        z = np.load(cache); ip, j, a = z["ip"], z["j"], z["a"]
    else:
        ip, j, a = P.powerlaw(nx)
        np.savez(cache, ip=ip, j=j, a=a)
    m = nx
A = pb.Mat.from_csr(m, m, ip, j, a)
x, y = pb.Vec(m, P.random_vector(m, 2)), pb.Vec(m)
for _ in range(4):
    A.mult(x, y)
pb.synchronize()
if "cg" in sys.argv:
    b, sol = pb.Vec(m), pb.Vec(m)
    x.set(1.0); A.mult(x, b)
    pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=6, fused=1, histmax=0)
    pb.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=6, fused=0, histmax=0)
if "gmres" in sys.argv:
    b, sol = pb.Vec(m), pb.Vec(m)
    x.set(1.0); A.mult(x, b)
    pb.ksp_solve(A, b, sol, ksp_type="gmres", pc_type="jacobi", rtol=1e-50, max_it=31, histmax=0)
print("done", A.bin_info())
