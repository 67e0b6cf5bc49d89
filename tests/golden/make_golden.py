#!/usr/bin/env python3
"""Generates the committed fixtures of tests/golden/ (run HERE, where /root/reference exists):

  reference_goldens.json  numbers transcribed from the reference's own golden outputs for this path
      src/mat/examples/tests/output/ex5_11_A.out          MatMult / MatGetDiagonal on the 8x10 AIJ matrix of ex5.c -rectA
      src/ksp/ksp/examples/tutorials/output/ex2_1.out     GMRES residual history of ex2 -m 5 -n 5 (refine_always)
      src/ksp/ksp/examples/tutorials/output/ex5_1.out     GMRES+Jacobi history of ex5 (first solve)
      src/ksp/pc/examples/tests/output/ex2_1.out          CG + PCNONE history on the 10x10 tridiagonal matrix
      src/ksp/ksp/examples/tutorials/output/ex4_0.out     MatSetValuesBatch: the matrix assembled from ex4's 8 triangles
  oracle_fixtures.npz     outputs of the reference library itself (oracle/_ref, built by oracle/build_ref.py) on
      seeded inputs: Vec kernels, MatMult/MatMultAdd, CG and GMRES residual histories.  The GPU tests compare
      against these even where the oracle .so is absent.
"""
import importlib
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def monitor_norms(path, first_block_only=True):
    out = []
    for line in open(path):
        m = re.match(r"\s*(\d+) KSP Residual norm (.*)$", line)
        if not m:
            continue
        it, val = int(m.group(1)), m.group(2).strip()
        if it == 0 and out and first_block_only:
            break
        out.append(val)
    return out


def numbers_after(path, header, count):
    lines = open(path).read().splitlines()
    k = next(i for i, l in enumerate(lines) if l.startswith(header))
    vals = []
    for l in lines[k + 1:]:
        try:
            vals.append(float(l))
        except ValueError:
            if vals:
                break
        if len(vals) == count:
            break
    return vals


def main():
    g = {}
    ex5 = os.path.join(REF, "src/mat/examples/tests/output/ex5_11_A.out")
    g["mat_ex5_11_A"] = {"source": "src/mat/examples/tests/output/ex5_11_A.out",
                         "matmult": numbers_after(ex5, "testing MatMult()", 8)}
    lines = open(ex5).read().splitlines()
    k = max(i for i, l in enumerate(lines) if l.startswith("Vector Object"))
    g["mat_ex5_11_A"]["diagonal"] = [float(x) for x in lines[k + 2:k + 10]]
    g["ksp_ex2_1"] = {"source": "src/ksp/ksp/examples/tutorials/output/ex2_1.out",
                      "history": monitor_norms(os.path.join(REF, "src/ksp/ksp/examples/tutorials/output/ex2_1.out")),
                      "iterations": 4, "error_norm": "0.000392701"}
    g["ksp_ex5_1"] = {"source": "src/ksp/ksp/examples/tutorials/output/ex5_1.out",
                      "history": monitor_norms(os.path.join(REF, "src/ksp/ksp/examples/tutorials/output/ex5_1.out"))}
    g["pc_ex2_1"] = {"source": "src/ksp/pc/examples/tests/output/ex2_1.out",
                     "history": monitor_norms(os.path.join(REF, "src/ksp/pc/examples/tests/output/ex2_1.out"))}
    # MatSetValuesBatch: the assembled 9 x 9 matrix the reference's DEVICE assembly printed (first MatView of ex4_0.out,
    # "type: seqaijcusp"): 41 entries, explicit zeros kept
    ex4 = os.path.join(REF, "src/ksp/ksp/examples/tutorials/output/ex4_0.out")
    lines = open(ex4).read().splitlines()
    k = next(i for i, l in enumerate(lines) if l.strip() == "type: seqaijcusp")
    rows = []
    for l in lines[k + 1:]:
        m = re.match(r"row (\d+):(.*)$", l)
        if not m:
            break
        rows.append([[int(c), float(v)] for c, v in re.findall(r"\((\d+), ([-+0-9.eE]+)\)", m.group(2))])
    g["ksp_ex4_0"] = {"source": "src/ksp/ksp/examples/tutorials/output/ex4_0.out", "grid": [3, 3], "rows": rows}
    json.dump(g, open(os.path.join(HERE, "reference_goldens.json"), "w"), indent=1)

    # ---- oracle outputs on seeded inputs ----
    from oracle import refpetsc as R
    P = importlib.import_module("workloads.problems")
    fx = {}
    n = 10007
    x, y, w = (P.random_vector(n, s) for s in (11, 12, 13))
    fx["vec_n"] = np.array(n)
    fx["vec_dot"] = np.array(R.vec_dot(x, y))
    fx["vec_norm2"] = np.array(R.vec_norm(x, R.NORM_2))
    fx["vec_norm1"] = np.array(R.vec_norm(x, R.NORM_1))
    fx["vec_norminf"] = np.array(R.vec_norm(x, R.NORM_INFINITY))
    ys = [P.random_vector(n, 20 + k) for k in range(7)]
    fx["vec_mdot7"] = R.vec_mdot(x, ys)
    al = P.random_vector(7, 99)
    fx["vec_maxpy7"] = R.vec_maxpy(y, al, ys)
    yy = y.copy(); R.vec_elementwise(R.OP_AXPY, n, a=0.37, x=x, y=yy); fx["vec_axpy"] = yy
    yy = y.copy(); R.vec_elementwise(R.OP_AYPX, n, a=-0.81, x=x, y=yy); fx["vec_aypx"] = yy
    ww = w.copy(); R.vec_elementwise(R.OP_PMULT, n, x=x, y=y, w=ww); fx["vec_pmult"] = ww
    # matrices
    ip, j, a = P.laplace3d_7pt(12)
    N = 12 ** 3
    A = R.RefMat(N, N, ip, j, a)
    xv = P.random_vector(N, 2)
    fx["lap7_12_mult"] = A.mult(xv)
    fx["lap7_12_multadd"] = A.multadd(xv, P.random_vector(N, 3))
    b = A.mult(np.ones(N))
    r = A.ksp_solve(b, "cg", "jacobi", rtol=1e-8)
    fx["lap7_12_cg_hist"] = r["history"]; fx["lap7_12_cg_its"] = np.array(r["its"])
    r = A.ksp_solve(b, "gmres", "jacobi", rtol=1e-8)
    fx["lap7_12_gmres_hist"] = r["history"]; fx["lap7_12_gmres_its"] = np.array(r["its"])
    ip, j, a = P.powerlaw(3000, seed=4, xm=5.0, lmax=800)
    A = R.RefMat(3000, 3000, ip, j, a)
    xv = P.random_vector(3000, 5)
    fx["plaw3000_mult"] = A.mult(xv)
    b = A.mult(np.ones(3000))
    r = A.ksp_solve(b, "gmres", "jacobi", rtol=1e-8)
    fx["plaw3000_gmres_hist"] = r["history"]; fx["plaw3000_gmres_its"] = np.array(r["its"])
    # MatSetValuesBatch through the reference's host path (one MatSetValues/ADD_VALUES per element)
    Nh, nl, er, em = P.elements_hex3d(6)
    hi, hj, ha = R.mat_setvaluesbatch(Nh, Nh, er, em, nl).csr()
    fx["hex6_batch_i"], fx["hex6_batch_j"], fx["hex6_batch_a"] = hi, hj, ha
    np.savez_compressed(os.path.join(HERE, "oracle_fixtures.npz"), **fx)
    print("wrote", len(g), "reference goldens and", len(fx), "oracle fixtures")


if __name__ == "__main__":
    main()
