#!/usr/bin/env python3
"""Micro-benchmark of the hot kernels (development tool): MatMult GB/s on the BASELINE matrix shapes, vector kernels."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
which = sys.argv[1:] or ["lap7", "s27", "plaw", "vec"]

def bench_mat(name, m, ip, j, a, reps=20):
    A = pb.Mat.from_csr(m, m, ip, j, a)
    x, y = pb.Vec(m, P.random_vector(m, 2)), pb.Vec(m)
    for _ in range(3): A.mult(x, y)
    pb.synchronize(); pb.timer_start()
    for _ in range(reps): A.mult(x, y)
    ms = pb.timer_stop() / reps
    B = A.bytes()
    print("%-8s m=%d nz=%d  %.3f ms  %.0f GB/s (%.1f%% of 6447.5)  %s" % (name, m, len(a), ms, B / ms / 1e6, B / ms / 1e6 / 64.475, A.bin_info()), flush=True)
    return A, x, y

if "lap7" in which:
    ip, j, a = P.laplace3d_7pt(256); bench_mat("lap7", 256 ** 3, ip, j, a)
if "s27" in which:
    ip, j, a = P.stencil3d_27pt(160); bench_mat("s27", 160 ** 3, ip, j, a)
if "bs3" in which:
    ip, j, a = P.elasticity_bs3(96); bench_mat("bs3", 3 * 96 ** 3, ip, j, a)
if "plaw" in which:
    ip, j, a = P.powerlaw(2000000); bench_mat("plaw", 2000000, ip, j, a)
if "lap2d" in which:
    ip, j, a = P.laplace2d_5pt(1000, 1000); bench_mat("lap2d", 1000000, ip, j, a, reps=200)
if "vec" in which:
    n = 256 ** 3
    x, y, w = pb.Vec(n, P.random_vector(n, 1)), pb.Vec(n, P.random_vector(n, 2)), pb.Vec(n)
    vs = [pb.Vec(n, P.random_vector(n, 10 + k)) for k in range(30)]
    al = np.linspace(0.1, 1, 30)
    def t(name, f, nbytes, reps=20):
        for _ in range(3): f()
        pb.synchronize(); pb.timer_start()
        for _ in range(reps): f()
        ms = pb.timer_stop() / reps
        print("%-10s %.3f ms  %.0f GB/s (%.1f%%)" % (name, ms, nbytes / ms / 1e6, nbytes / ms / 1e6 / 64.475), flush=True)
    t("axpy", lambda: y.axpy(0.5, x), 24 * n)
    t("aypx", lambda: y.aypx(0.5, x), 24 * n)
    t("pmult", lambda: w.pointwise_mult(x, y), 24 * n)
    t("copy", lambda: x.copy(w), 16 * n)
    t("set", lambda: w.set(1.0), 8 * n)
    t("dot", lambda: x.dot(y), 16 * n)
    t("norm", lambda: x.norm(), 8 * n)
    for nv in (4, 8, 16, 30):
        t("mdot%d" % nv, lambda: x.mdot(vs[:nv]), 8 * n * (nv + 1))
        t("maxpy%d" % nv, lambda: y.maxpy(al[:nv], vs[:nv]), 8 * n * (nv + 2))
