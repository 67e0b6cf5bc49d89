#!/usr/bin/env python3
"""BASELINE.json configs[0] through the drop-in boundary: the reference's UNMODIFIED ksp ex2 (2-D 5-pt Laplacian m=n=1000,
KSPCG + PCJACOBI) run three ways on this box, KSPSolve wall time from the program's own -log_summary:
  reference : stock CPU types (seqaij / seq), one core (the only PETSc on this box is the 1-rank MPIUNI build)
  b200      : -dll_prepend libpetscb200plugin.so -mat_type aijb200 -vec_type b200, the reference's KSPCG op by op
  b200-fused: same + -ksp_type cgb200 (device-resident fused CG)
Prints one JSON line.  Development/evidence tool."""
import json, os, re, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "bin", "ksp_ex2")
PLUGIN = os.path.join(ROOT, "petsc-dev_b200", "lib", "libpetscb200plugin.so")
m = sys.argv[1] if len(sys.argv) > 1 else "1000"
base = ["-m", m, "-n", m, "-pc_type", "jacobi", "-log_summary"]
arms = {"reference": ["-ksp_type", "cg"],
        "b200": ["-ksp_type", "cg", "-dll_prepend", PLUGIN, "-mat_type", "aijb200", "-vec_type", "b200"],
        "b200_fused": ["-ksp_type", "cgb200", "-dll_prepend", PLUGIN, "-mat_type", "aijb200", "-vec_type", "b200"]}
out = {"workload": "ksp ex2 -m %s -n %s -ksp_type cg -pc_type jacobi (configs[0])" % (m, m)}
for name, extra in arms.items():
    t0 = time.perf_counter()
    r = subprocess.run([EXE] + base + extra, capture_output=True, text=True, timeout=1800)
    wall = time.perf_counter() - t0
    if r.returncode:
        out[name] = {"error": r.stderr[-400:]}
        continue
    txt = r.stdout
    ev = {}
    for l in txt.splitlines():
        f = l.split()
        if f and f[0] in ("KSPSolve", "MatMult", "VecDot", "VecAXPY", "VecAYPX", "VecNorm", "PCApply", "MatAssemblyEnd") and len(f) > 4:
            try:
                ev[f[0]] = {"count": int(f[1]), "seconds": float(f[3])}
            except ValueError:
                pass
    mo = re.search(r"Norm of error (\S+),? iterations (\d+)", txt)
    out[name] = {"norm_of_error": mo.group(1) if mo else None, "iterations": int(mo.group(2)) if mo else None,
                 "kspsolve_s": ev.get("KSPSolve", {}).get("seconds"), "events": ev, "process_wall_s": wall}
    if out[name]["iterations"] and out[name]["kspsolve_s"]:
        out[name]["iter_per_s"] = out[name]["iterations"] / out[name]["kspsolve_s"]
print(json.dumps(out))
