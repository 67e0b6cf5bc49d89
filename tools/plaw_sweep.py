#!/usr/bin/env python3
"""Tuning sweep of the irregular-matrix SpMV (development tool): C4 (power-law, N rows) generated once, then one MatMult
timing per (lanes per row, long-row threshold, launch shape, column strips, park) combination.
  python tools/plaw_sweep.py [N]"""
import importlib, itertools, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
pb.init(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 10_000_000
cache = os.path.join(os.environ.get("B200_CACHE_DIR", "/tmp"), "plaw_%d.npz" % N)
if os.path.exists(cache):
    z = np.load(cache); ip, j, a = z["ip"], z["j"], z["a"]
else:
    t0 = time.perf_counter(); ip, j, a = P.powerlaw(N); print("# generated in %.1f s" % (time.perf_counter() - t0), flush=True)
xh = P.random_vector(N, 2)
ref = None
if "cap" in sys.argv:
    combos = [dict(cap=c, long_thresh=t) for t in (256, 512, 1024, 2048) for c in (0, 8, 16, 32, 64)]
    combos += [dict(cap=16, long_thresh=2048, strips=s) for s in (2, 3)] if N > 6_000_000 else []
else:
    combos = [dict()]                                                            # automatic everything
    combos += [dict(lanes=g) for g in (2, 4, 8)]
    combos += [dict(long_thresh=t) for t in (64, 128, 512)]
    combos += [dict(shape=s) for s in (1, 3, 5)]
    combos += [dict(strips=s) for s in (2, 3, 6, 8)] if N > 6_000_000 else []
    combos += [dict(lanes=2, long_thresh=128), dict(lanes=4, long_thresh=128), dict(lanes=2, long_thresh=64), dict(lanes=4, long_thresh=64),
               dict(lanes=2, shape=1), dict(lanes=4, shape=1), dict(park=0), dict(park=0, lanes=4), dict(park=0, lanes=8)]
for c in combos:
    pb._chk(pb.lib.b200_spmv_set_park_cap(int(c.get("cap", 32))), "park cap")
    pb.spmv_tuning(shape=c.get("shape", -1), lanes_per_row=c.get("lanes", -1), long_thresh=c.get("long_thresh", -1), park=c.get("park", -1))
    pb.spmv_strips(c.get("strips", -1))
    t0 = time.perf_counter()
    try:
        A = pb.Mat.from_csr(N, N, ip, j, a)
    except Exception as e:
        print(c, "failed:", e, flush=True); continue
    up = time.perf_counter() - t0
    x, y = pb.Vec(N, xh), pb.Vec(N)
    for _ in range(3): A.mult(x, y)
    pb.synchronize(); pb.timer_start()
    for _ in range(10): A.mult(x, y)
    ms = pb.timer_stop() / 10
    yh = y.get_array()
    if ref is None: ref = yh.copy()
    err = float(np.max(np.abs(yh - ref)) / max(np.max(np.abs(ref)), 1e-300))
    print("%-40s %.3f ms  %.0f GB/s  upload %.2f s  max dev from first %.1e  %s" % (c, ms, A.bytes() / ms / 1e6, up, err, A.bin_info()), flush=True)
    del A, x, y
pb.spmv_tuning(); pb.spmv_strips(-1)
