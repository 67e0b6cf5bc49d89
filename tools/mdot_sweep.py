#!/usr/bin/env python3
"""Development tool: VecMDot / VecMAXPY achieved GB/s for every vector count 1..32 (n = 256^3)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200"); P = importlib.import_module("workloads.problems")
pb.init(0)
n = 256 ** 3
base = P.random_vector(n, 1)                    # one host vector for every buffer: bandwidth does not depend on the values
x, y = pb.Vec(n, base), pb.Vec(n, base)
vs = [pb.Vec(n, base) for k in range(32)]
al = np.linspace(1e-4, 1e-3, 32)
for nv in range(1, 33):
    res = []
    for f, bpe in ((lambda: x.mdot(vs[:nv]), 8 * (nv + 1)), (lambda: y.maxpy(al[:nv], vs[:nv]), 8 * (nv + 2))):
        for _ in range(2): f()
        pb.synchronize(); pb.timer_start()
        for _ in range(10): f()
        ms = pb.timer_stop() / 10
        res.append((ms, bpe * n / ms / 1e6))
    print("nv %2d  mdot %.4f ms %6.0f GB/s   maxpy %.4f ms %6.0f GB/s" % (nv, res[0][0], res[0][1], res[1][0], res[1][1]), flush=True)
