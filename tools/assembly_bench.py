#!/usr/bin/env python3
"""Device assembly (MatSetValuesBatch) on one B200: Q1 hexahedra on an nx^3 vertex grid, element arrays resident on the
device.  Reports the full assembly (symbolic + numeric, wall clock around the call: it allocates and reads the row
offsets back), the values-only refill through the kept plan (CUDA events), and the reference's host path on a bounded
sample (oracle library, one core).  One JSON line.

  python tools/assembly_bench.py [--nx 128] [--ref-nx 40] [--reps 10]"""
import importlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
pb = importlib.import_module("petsc-dev_b200")
P = importlib.import_module("workloads.problems")
argv = sys.argv[1:]
def opt(name, default):
    return int(argv[argv.index(name) + 1]) if name in argv else default
nx, ref_nx, reps = opt("--nx", 128), opt("--ref-nx", 40), opt("--reps", 10)
pb.init(0)
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    PEAK = 6650.0
N, nl, er, em = P.elements_hex3d(nx)
ne = er.size // nl
der, dem = torch.from_numpy(er).cuda(), torch.from_numpy(em).cuda()
torch.cuda.synchronize()
l0 = pb.launch_count()
full = []
NFULL = 6
for _ in range(NFULL):
    t0 = time.perf_counter()
    A = pb.Mat.set_values_batch_device(N, N, ne, nl, der.data_ptr(), dem.data_ptr(), keep_plan=True)
    pb.synchronize()
    full.append((time.perf_counter() - t0) * 1e3)
    if _ < NFULL - 1:
        del A
launches_full = (pb.launch_count() - l0) // NFULL
nz = A.nnz()
fat = em.size
for _ in range(3):
    A.batch_values_device(dem.data_ptr())
pb.synchronize()
pb.timer_start()
for _ in range(reps):
    A.batch_values_device(dem.data_ptr())
ms_vals = pb.timer_stop() / reps
bytes_vals = 8.0 * fat + 4.0 * fat + 4.0 * (nz + 1) + 8.0 * nz          # element values + plan (perm, run starts) + nonzeros
# MatMult on the assembled operator, for scale
x, y = pb.Vec(N), pb.Vec(N)
x.set(1.0)
for _ in range(3):
    A.mult(x, y)
pb.synchronize(); pb.timer_start()
for _ in range(reps):
    A.mult(x, y)
ms_mult = pb.timer_stop() / reps
out = {"workload": "MatSetValuesBatch, Q1 hexahedra on %d^3 vertices" % nx, "rows": N, "elements": ne, "nl": nl, "fat_entries": fat, "nnz": nz,
       "assemble_full_ms": min(full), "assemble_full_ms_median": float(np.median(full)), "assemble_full_ms_all": full, "assemble_full_Melem_per_s": ne / min(full) / 1e3,
       "kernels_per_full_assembly": launches_full,
       "values_only_ms": ms_vals, "values_only_Melem_per_s": ne / ms_vals / 1e3,
       "values_only_gbs": bytes_vals / ms_vals / 1e6, "values_only_frac_of_measured_peak": bytes_vals / ms_vals / 1e6 / PEAK,
       "matmult_ms": ms_mult, "peak_gbs": PEAK}
try:
    from oracle import refpetsc as R
    if R.available():
        Nr, _, rer, rem = P.elements_hex3d(ref_nx)
        t0 = time.perf_counter()
        M = R.mat_setvaluesbatch(Nr, Nr, rer, rem, nl)
        dt = time.perf_counter() - t0
        out["cpu_reference"] = {"kind": "reference", "cores": 1, "sample": "hexahedra on %d^3 vertices (%d elements), preallocated" % (ref_nx, rer.size // nl),
                                "ms": dt * 1e3, "Melem_per_s": rer.size / nl / dt / 1e6}
except Exception as e:                                   # the oracle is optional here
    out["cpu_reference"] = {"unavailable": str(e)[:200]}
print(json.dumps(out))
