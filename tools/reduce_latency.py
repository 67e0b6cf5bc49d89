#!/usr/bin/env python3
"""VecDot / VecNorm / VecMDot as the caller sees them (result valid on return): time per call by the host clock and by CUDA
events, at n = 256^3 (the C2 vector), 10^6 and 10^4.  Run once with B200_REDUCE_POLL=0 (plain stream synchronise) and
once without (the host polls the mapped result slots) to see what the completion signal costs."""
import importlib, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pb = importlib.import_module("petsc-dev_b200")
pb.init(0)
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
mode = "sync" if os.environ.get("B200_REDUCE_POLL", "1") == "0" else "poll"
for n in (256 ** 3, 10 ** 6, 10 ** 4):
    x, y = pb.Vec(n), pb.Vec(n)
    x.set(0.5); y.set(-2.0)
    ys = [y, x, y, x]
    for name, f, bpe, want in (("VecDot", lambda: x.dot(y), 16, -float(n)), ("VecNorm1", lambda: x.norm(pb.NORM_1), 8, 0.5 * n),
                               ("VecMDot4", lambda: x.mdot(ys)[0], 40, -float(n))):
        reps = 200
        for _ in range(10):
            assert f() == want
        pb.synchronize(); pb.timer_start(); t0 = time.perf_counter()
        for _ in range(reps):
            f()
        wall = (time.perf_counter() - t0) / reps * 1e6
        ev = pb.timer_stop() / reps * 1e3
        print(json.dumps({"mode": mode, "op": name, "n": n, "us_per_call_host": round(wall, 2), "us_per_call_events": round(ev, 2),
                          "gbs": round(bpe * n / wall / 1e3, 1), "frac_of_measured_peak": round(bpe * n / wall / 1e3 / peak, 3)}))
