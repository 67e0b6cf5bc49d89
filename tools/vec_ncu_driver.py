#!/usr/bin/env python3
"""Development tool for ncu: one launch of every Krylov vector kernel on 256^3-entry vectors (the C2 vector length), plus
the live CUDA-event timing of each (printed) so that profiles/ can hold achieved GB/s per kernel next to the ncu DRAM bytes."""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
n = 256 ** 3
timed = "--time" in sys.argv
x, y, w = pb.Vec(n, P.random_vector(n, 1)), pb.Vec(n, P.random_vector(n, 2)), pb.Vec(n)
vs = [pb.Vec(n, P.random_vector(n, 10 + k)) for k in range(30)]
al = np.linspace(0.1, 1, 30)
ops = [("VecAXPY", lambda: y.axpy(0.5, x), 24), ("VecAYPX", lambda: y.aypx(0.5, x), 24), ("VecWAXPY", lambda: w.waxpy(0.5, x, y), 24),
       ("VecPointwiseMult", lambda: w.pointwise_mult(x, y), 24), ("VecCopy", lambda: x.copy(w), 16), ("VecSet", lambda: w.set(1.0), 8),
       ("VecScale", lambda: w.scale(1.0001), 16), ("VecDot", lambda: x.dot(y), 16), ("VecNorm2", lambda: x.norm(), 8),
       ("VecMDot8", lambda: x.mdot(vs[:8]), 8 * 9), ("VecMDot30", lambda: x.mdot(vs[:30]), 8 * 31),
       ("VecMAXPY8", lambda: y.maxpy(al[:8] * 1e-3, vs[:8]), 8 * 10), ("VecMAXPY30", lambda: y.maxpy(al[:30] * 1e-3, vs[:30]), 8 * 32)]
out = []
for name, f, bpe in ops:
    reps = 20 if timed else 1
    if timed:
        for _ in range(3):
            f()
    pb.synchronize(); pb.timer_start()
    for _ in range(reps):
        f()
    ms = pb.timer_stop() / reps
    out.append({"op": name, "ms": ms, "algorithmic_bytes": bpe * n, "gbs": bpe * n / ms / 1e6})
if timed:
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        peak = 6650.0
    for o in out:
        o["frac_of_measured_peak"] = o["gbs"] / peak
        print(json.dumps(o))
