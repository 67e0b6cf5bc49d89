#!/usr/bin/env python3
"""Build the PETSc plugin (petsc-dev_b200/lib/libpetscb200plugin.so) and the reference's own example programs.

The plugin is host C compiled against the reference's headers (they only exist in this container, under
/root/reference) and linked against libpetscb200.so (the CUDA library, C ABI) and the PETSc it plugs into — on this box
the oracle build oracle/_ref/libpetsc_ref.so is the only PETSc there is.  The example programs are the reference's
UNMODIFIED sources (src/ksp/ksp/examples/tutorials/ex2.c, src/mat/examples/tests/ex5.c ...) compiled where they lie;
binaries go to oracle/_ref/bin/ (git-ignored, travels to the GPU box).  On the GPU box /root/reference is absent and
this script is a no-op that reports the prebuilt files.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PETSC_REFERENCE_DIR", "/root/reference")
CONF = os.path.join(ROOT, "oracle", "refconf")
REFLIBDIR = os.path.join(ROOT, "oracle", "_ref")
LIBDIR = os.path.join(ROOT, "petsc-dev_b200", "lib")
PLUGIN = os.path.join(LIBDIR, "libpetscb200plugin.so")
BINDIR = os.path.join(REFLIBDIR, "bin")

EXAMPLES = {  # unmodified reference programs used by tests/test_gpu_plugin.py
    "ksp_ex2": "src/ksp/ksp/examples/tutorials/ex2.c",
    "ksp_ex4": "src/ksp/ksp/examples/tutorials/ex4.c",
    "ksp_ex5": "src/ksp/ksp/examples/tutorials/ex5.c",
    "mat_ex5": "src/mat/examples/tests/ex5.c",
    "vec_tut_ex1": "src/vec/vec/examples/tutorials/ex1.c",
    "vec_ex43": "src/vec/vec/examples/tests/ex43.c",
    "vec_ex21": "src/vec/vec/examples/tests/ex21.c",
    # this repository's own driver (public PETSc API only): MatSetValuesBatch on the parallel type
    "b200_batch_mpiaij": os.path.join(ROOT, "tests", "c", "batch_mpiaij.c"),
}


def inc():
    return ["-I", CONF, "-I", os.path.join(REF, "include"), "-I", os.path.join(REF, "include/mpiuni"), "-I", REF,
            "-I", os.path.join(ROOT, "include")]


def newer(out, srcs):
    return (not os.path.exists(out)) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs if os.path.exists(s))


def main(verbose=True):
    if not os.path.isdir(REF):
        if verbose:
            print("reference headers absent; prebuilt plugin:", PLUGIN, os.path.exists(PLUGIN))
        return PLUGIN
    if not os.path.exists(os.path.join(REFLIBDIR, "libpetsc_ref.so")):
        raise SystemExit("build oracle/_ref/libpetsc_ref.so first (oracle/build_ref.py)")
    os.makedirs(LIBDIR, exist_ok=True)
    os.makedirs(BINDIR, exist_ok=True)
    srcs = [os.path.join(HERE, f) for f in ("b200vec.c", "b200mat.c", "b200ksp.c", "b200reg.c")]
    deps = srcs + [os.path.join(HERE, "petscb200plugin.h"), os.path.join(ROOT, "include", "petscb200.h"),
                   os.path.join(CONF, "petscconf.h")]
    if newer(PLUGIN, deps):
        cmd = ["gcc", "-shared", "-fPIC", "-O2", "-g", "-w",
               "-ffp-contract=off", "-D__INSDIR__=petsc-dev_b200/plugin/"] + inc() + srcs + \
              ["-o", PLUGIN, "-L", LIBDIR, "-lpetscb200", "-L", REFLIBDIR, "-lpetsc_ref",
               "-Wl,-rpath,$ORIGIN", "-Wl,-rpath,$ORIGIN/../../oracle/_ref", "-Wl,--no-undefined", "-lm"]
        subprocess.run(cmd, check=True)
        if verbose:
            print("built", PLUGIN)
    for name, rel in EXAMPLES.items():
        src = os.path.join(REF, rel)
        out = os.path.join(BINDIR, name)
        if not os.path.exists(src):
            continue
        if newer(out, [src, os.path.join(REFLIBDIR, "libpetsc_ref.so")]):
            cmd = ["gcc", "-O1", "-w", "-D__INSDIR__=%s/" % os.path.dirname(rel)] + inc() + [src, "-o", out,
                   "-L", REFLIBDIR, "-lpetsc_ref", "-Wl,-rpath,$ORIGIN/..", "-Wl,--export-dynamic", "-lm", "-ldl"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode:
                print("example %s failed to build:\n%s" % (name, r.stderr[-1500:]), file=sys.stderr)
            elif verbose:
                print("built", out)
    return PLUGIN


if __name__ == "__main__":
    main()
