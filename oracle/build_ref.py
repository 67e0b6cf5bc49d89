#!/usr/bin/env python3
"""Build the UNMODIFIED reference (erdc/petsc-dev, PETSc 3.3-dev) CPU library from the sources where
they lie under /root/reference into oracle/_ref/ (git-ignored; travels to the GPU box as a built .so).

TEST INFRASTRUCTURE ONLY.  The result, oracle/_ref/libpetsc_ref.so, is
  * the parity oracle for tests/ (the reference's own MatMult_SeqAIJ / MatMult_MPIAIJ / Vec kernels /
    KSPSolve_CG / KSPSolve_GMRES / PCJACOBI code, compiled, not restated), and
  * the timed CPU baseline of bench.py (`cpu_baseline.kind == "reference"`, `--impl reference`).
Nothing in the product path (petsc-dev_b200/) links or loads it, with one documented exception: the
drop-in demonstration plugin (petsc-dev_b200/plugin) needs *a* PETSc to plug into, and on this box the only
PETSc is this build.

The reference's own build system cannot run here (config/BuildSystem is an un-vendored submodule and needs
Python 2), so this script does what `conf/rules` would do: it walks the per-directory `makefile`s, honours
their `#requires...` directives the way the reference's builder does, and compiles every `SOURCEC` file
with gcc against a hand-written petscconf.h (oracle/refconf/).  BLAS level-1/2 and the few LAPACK entry
points are supplied by oracle/refblas.c (textbook reference-BLAS loops; see its header).  No reference source
file is copied into the repo; objects and the .so go to oracle/_ref/ only.
"""
import concurrent.futures as cf
import hashlib
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("PETSC_REFERENCE_DIR", "/root/reference")
OUT = os.path.join(HERE, "_ref")
OBJ = os.path.join(OUT, "obj")
CONF = os.path.join(HERE, "refconf")
LIB = os.path.join(OUT, "libpetsc_ref.so")

ROOTS = ["src/sys", "src/vec", "src/mat", "src/dm", "src/ksp"]
SKIP_DIR_NAMES = {"examples", "tests", "tutorials", "benchmarks", "docs", "contrib"}
# directories whose code needs things this box does not have (X11, pthread classes, Fortran, C++ sieve ...)
SKIP_DIR_PREFIX = ("ftn-", "f90-", "f2003-")
SKIP_PATHS = {
    "src/sys/fsrc", "src/sys/ams", "src/sys/yaml", "src/sys/totalview",
    "src/sys/classes/draw/impls/x", "src/sys/classes/draw/impls/opengl", "src/sys/classes/draw/impls/win32",
    "src/sys/classes/viewer/impls/socket", "src/sys/classes/viewer/impls/mathematica",
    "src/sys/classes/viewer/impls/matlab", "src/sys/classes/viewer/impls/hdf5",
    "src/sys/classes/viewer/impls/ams", "src/sys/classes/viewer/impls/netcdf",
    "src/sys/time/nt_time", "src/sys/time/rs6000_time",
    "src/dm/impls/mesh", "src/dm/impls/cartesian", "src/dm/impls/iga", "src/dm/impls/akkt", "src/dm/ao/impls/mapping",
    "src/mat/impls/python", "src/ksp/ksp/impls/python", "src/ksp/pc/impls/python",
    "src/mat/impls/aij/seq/ftn-kernels", "src/mat/ftn-kernels",
}
# per-file exclusions (files that need generated headers or missing system pieces)
SKIP_FILES = set()

# hot-path files get -O3 (the timed CPU baseline); everything else -O1 to keep the build short
HOT = ("src/mat/impls/aij/seq/aij.c", "src/mat/impls/aij/seq/inode.c", "src/mat/impls/aij/mpi/",
       "src/vec/vec/impls/seq/", "src/vec/vec/impls/mpi/", "src/vec/vec/utils/vscat.c",
       "src/vec/vec/utils/vpscat.c", "src/ksp/ksp/impls/cg/cg.c", "src/ksp/ksp/impls/gmres/")

HAVE = {"PETSC_HAVE_MPIUNI", "PETSC_USE_LOG", "PETSC_HAVE_RAND", "PETSC_HAVE_DRAND48", "PETSC_USE_INFO"}  # packages we "have" for #requirespackage purposes
MPIINC = os.path.join(REF, "include/mpiuni")       # where <mpi.h> comes from: the reference's one-rank MPIUNI ...
MPIDEFS = []
MPILIB = []                                        # ... whose mpi.c is compiled with the rest (src/sys/mpiuni)
SHIM = os.path.join(os.path.dirname(HERE), "petsc-dev_b200", "mpishim")
SHIMLIBDIR = os.path.join(os.path.dirname(HERE), "petsc-dev_b200", "lib")


def use_mpi_shim():
    """Second flavour of the same build: the SAME unmodified sources against the N-rank b200mpi subset
    (petsc-dev_b200/mpishim) instead of MPIUNI -> oracle/_ref_mpi/libpetsc_ref_mpi.so.  Run with
    petsc-dev_b200/lib/b200mpiexec -n P; this is the P > 1 oracle (real MatSetUpMultiply_MPIAIJ / VecScatterCreate_PtoS /
    MatMult_MPIAIJ / KSPSolve on P ranks) and the all-cores CPU baseline of bench.py."""
    global OUT, OBJ, LIB, HAVE, MPIINC, MPILIB, MPIDEFS
    OUT = os.path.join(HERE, "_ref_mpi")
    OBJ = os.path.join(OUT, "obj")
    LIB = os.path.join(OUT, "libpetsc_ref_mpi.so")
    HAVE = HAVE - {"PETSC_HAVE_MPIUNI"}
    MPIINC = SHIM
    MPIDEFS = ["-DPETSC_HAVE_MPI_TYPE_DUP", "-DPETSC_HAVE_MPI_FINALIZED"]   # what configure would have found
    MPILIB = ["-L", SHIMLIBDIR, "-lb200mpi", "-Wl,-rpath,$ORIGIN/../../petsc-dev_b200/lib"]


def parse_makefile(path):
    txt = open(path, errors="replace").read()
    txt = txt.replace("\\\n", " ")
    req_ok = True
    for m in re.finditer(r"^#requires(\w+)\s+(.*)$", txt, re.M):
        kind, val = m.group(1), m.group(2).strip().strip("'\"")
        if kind in ("package", "define", "function"):
            if val not in HAVE:
                req_ok = False
        elif kind == "language":
            if val.upper() in ("CXXONLY", "CXX"):
                req_ok = False
        elif kind == "scalar":
            if val != "real":
                req_ok = False
        elif kind == "precision":
            if val != "double":
                req_ok = False
    def var(name):
        m = re.search(r"^%s\s*=(.*)$" % name, txt, re.M)
        return m.group(1).split() if m else []
    return req_ok, var("SOURCEC"), var("DIRS")


def walk(rel, out):
    mk = os.path.join(REF, rel, "makefile")
    if not os.path.exists(mk):
        return
    ok, srcs, dirs = parse_makefile(mk)
    if not ok:
        return
    for s in srcs:
        p = os.path.join(rel, s)
        if s.endswith(".c") and p not in SKIP_FILES and os.path.exists(os.path.join(REF, p)):
            out.append(p)
    for d in dirs:
        if d in SKIP_DIR_NAMES or d.startswith(SKIP_DIR_PREFIX):
            continue
        sub = os.path.join(rel, d)
        if sub in SKIP_PATHS:
            continue
        walk(sub, out)


def cflags(rel):
    opt = "-O3" if rel.startswith(HOT) else "-O1"
    # -ffp-contract=off: the reference's arithmetic is plain IEEE mul+add (no FMA contraction), so the
    # oracle is reproducible across x86 hosts with/without FMA units.
    return ["gcc", "-c", "-fPIC", opt, "-ffp-contract=off", "-fno-strict-aliasing", "-w", "-Werror=implicit-function-declaration",
            "-I", CONF, "-I", os.path.join(REF, "include"), "-I", MPIINC,
            "-I", REF, "-D__SDIR__=\"%s/\"" % os.path.dirname(rel)] + MPIDEFS


def compile_one(rel):
    src = os.path.join(REF, rel)
    obj = os.path.join(OBJ, rel.replace("/", "__")[:-2] + ".o")
    cmd = cflags(rel) + [src, "-o", obj]
    stamp = obj + ".cmd"
    key = hashlib.sha1((" ".join(cmd) + str(os.path.getmtime(src)) +
                        "".join(str(os.path.getmtime(os.path.join(CONF, f))) for f in sorted(os.listdir(CONF))) +
                        str(os.path.getmtime(os.path.join(MPIINC, "mpi.h")))
                        ).encode()).hexdigest()
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == key:
        return rel, obj, 0, ""
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode == 0:
        open(stamp, "w").write(key)
    return rel, obj, r.returncode, r.stderr[-2000:]


def main(verbose=True):
    if not os.path.isdir(REF):
        if os.path.exists(LIB):
            if verbose:
                print("reference tree absent; using prebuilt", LIB)
            return LIB
        raise SystemExit("reference tree %s absent and no prebuilt %s" % (REF, LIB))
    os.makedirs(OBJ, exist_ok=True)
    files = []
    for r in ROOTS:
        walk(r, files)
    files = sorted(set(files))
    if verbose:
        print("reference build: %d C files" % len(files))
    objs, failed = [], []
    with cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        for rel, obj, rc, err in ex.map(compile_one, files):
            if rc:
                failed.append((rel, err))
            else:
                objs.append(obj)
    if failed:
        for rel, err in failed:
            print("FAILED", rel, "\n", err, file=sys.stderr)
        raise SystemExit("%d reference files failed to compile" % len(failed))
    # our own pieces that complete the link: reference-BLAS loops + the test driver
    extra = []
    for name, opt in (("refblas.c", "-O3"), ("ref_driver.c", "-O2")):
        src = os.path.join(HERE, name)
        obj = os.path.join(OBJ, name[:-2] + ".o")
        if (not os.path.exists(obj)) or os.path.getmtime(obj) < os.path.getmtime(src):
            cmd = ["gcc", "-c", "-fPIC", opt, "-ffp-contract=off", "-w", "-I", CONF,
                   "-I", os.path.join(REF, "include"), "-I", MPIINC, "-I", REF,
                   "-D__SDIR__=\"oracle/\"", src, "-o", obj]
            subprocess.run(cmd, check=True)
        extra.append(obj)
    newest = max(os.path.getmtime(o) for o in objs + extra)
    if os.path.exists(LIB) and os.path.getmtime(LIB) >= newest:
        if verbose:
            print("up to date:", LIB)
        return LIB
    rsp = os.path.join(OUT, "link.rsp")
    open(rsp, "w").write("\n".join(objs + extra))
    subprocess.run(["gcc", "-shared", "-o", LIB, "@" + rsp, "-Wl,--no-undefined", "-lm", "-ldl"] + MPILIB, check=True)
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    if "--mpi" in sys.argv:
        use_mpi_shim()
    main()
