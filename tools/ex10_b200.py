#!/usr/bin/env python3
"""ex10-style driver (reference src/ksp/ksp/examples/tutorials/ex10.c): load a PETSc binary matrix (and right-hand side, if
the file holds one; otherwise b = A * ones), solve on the B200 with the options given, print what ex10 prints.

  python tools/ex10_b200.py -f0 matrix.bin [-ksp_type cg|gmres] [-pc_type jacobi|none] [-ksp_rtol 1e-5] [-ksp_max_it N]
                            [-ksp_gmres_restart 30] [-ksp_monitor]
"""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def opt(name, default=None, flag=False):
    if name in sys.argv:
        return True if flag else sys.argv[sys.argv.index(name) + 1]
    return default


def main():
    pb = importlib.import_module("petsc-dev_b200")
    bio = importlib.import_module("petsc-dev_b200.binaryio")
    f0 = opt("-f0")
    if not f0:
        raise SystemExit("Must indicate binary file with the -f0 option")        # ex10.c:93
    t0 = time.perf_counter()
    M, N, ai, aj, aa, bh = bio.read_system(f0)
    if M != N:
        raise SystemExit("This example is not intended for rectangular matrices (%d, %d)" % (M, N))   # ex10.c:170
    pb.init(0)
    A = pb.Mat.from_csr(M, N, ai, aj, aa)
    t_load = time.perf_counter() - t0
    x, b = pb.Vec(M), pb.Vec(M)
    if bh is None:
        ones = pb.Vec(M); ones.set(1.0); A.mult(ones, b)                        # ex10.c:131-135
    else:
        b.set_array(bh)
    r = pb.ksp_solve(A, b, x, ksp_type=opt("-ksp_type", "gmres"), pc_type=opt("-pc_type", "jacobi"),
                     rtol=float(opt("-ksp_rtol", 1e-5)), max_it=int(opt("-ksp_max_it", 10000)),
                     gmres_restart=int(opt("-ksp_gmres_restart", 30)))
    if opt("-ksp_monitor", flag=True):
        for k, v in enumerate(r["history"]):
            print("%3d KSP Residual norm %14.12e " % (k, v))
    # true residual, as ex10 reports it (ex10.c:383-388)
    u = pb.Vec(M)
    A.mult(x, u)
    u.axpy(-1.0, b)
    print("Number of iterations = %3d" % r["its"])
    print("Residual norm %G" % u.norm())
    print("# reason %d, load+upload %.3f s, KSPSolve %.3f ms (%.1f it/s), rows %d, nnz %d" %
          (r["reason"], t_load, r["solve_ms"], r["its"] / max(r["solve_ms"], 1e-9) * 1e3, M, len(aa)))
    return r


if __name__ == "__main__":
    main()
