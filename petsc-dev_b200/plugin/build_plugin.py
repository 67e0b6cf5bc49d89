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
    # columns rewritten in place under unchanged row offsets (MatZeroRows): the device copy must notice
    "b200_zerorows_refresh": os.path.join(ROOT, "tests", "c", "zerorows_refresh.c"),
    # VecScatter layouts between option-typed vectors (the device scatter behind PETSc's VecScatter object)
    "b200_scatter_cases": os.path.join(ROOT, "tests", "c", "scatter_cases.c"),
    # VecGetArrayRead between device operations: one download, no upload
    "b200_coherence_seam": os.path.join(ROOT, "tests", "c", "coherence_seam.c"),
    # MatSetValues into rows other ranks own: the stash path of MatAssemblyBegin/End under the new types
    "b200_stash_offproc": os.path.join(ROOT, "tests", "c", "stash_offproc.c"),
}


MPI_EXAMPLES = {  # additional unmodified reference programs whose goldens need more than one rank (built in the MPI flavour only)
    "vec_ex4": "src/vec/vec/examples/tests/ex4.c", "vec_ex9": "src/vec/vec/examples/tests/ex9.c",
    "vec_ex11": "src/vec/vec/examples/tests/ex11.c", "vec_ex12": "src/vec/vec/examples/tests/ex12.c",
    "vec_ex13": "src/vec/vec/examples/tests/ex13.c", "vec_ex14": "src/vec/vec/examples/tests/ex14.c",
    "vec_ex17": "src/vec/vec/examples/tests/ex17.c", "vec_ex22": "src/vec/vec/examples/tests/ex22.c",
    "vec_ex23": "src/vec/vec/examples/tests/ex23.c", "vec_ex24": "src/vec/vec/examples/tests/ex24.c",
    "vec_ex25": "src/vec/vec/examples/tests/ex25.c", "vec_ex33": "src/vec/vec/examples/tests/ex33.c",
    "vec_ex5": "src/vec/vec/examples/tests/ex5.c", "vec_ex10": "src/vec/vec/examples/tests/ex10.c",
    "vec_ex3": "src/vec/vec/examples/tests/ex3.c", "vec_ex1": "src/vec/vec/examples/tests/ex1.c",
    # this repository's own programs (public PETSc API only)
    "b200_dump_mpiaij": os.path.join(ROOT, "tests", "c", "dump_mpiaij.c"),
    "b200_ref_bench": os.path.join(ROOT, "oracle", "ref_bench.c"),
    "mpishim_selftest": os.path.join(ROOT, "tests", "c", "mpishim_selftest.c"),
}

SHIM = os.path.join(ROOT, "petsc-dev_b200", "mpishim")

# the two PETSc builds a plugin can be made for on this box: the reference against its own one-rank MPIUNI, and the same
# sources against the N-rank b200mpi subset (oracle/build_ref.py --mpi), run with petsc-dev_b200/lib/b200mpiexec -n P
FLAVOURS = {
    "uni": dict(refdir=os.path.join(ROOT, "oracle", "_ref"), reflib="petsc_ref", mpiinc=os.path.join(REF, "include/mpiuni"),
                plugdir=LIBDIR, extra_libs=[], rpaths=["$ORIGIN", "$ORIGIN/../../oracle/_ref"], defs=[]),
    "mpi": dict(refdir=os.path.join(ROOT, "oracle", "_ref_mpi"), reflib="petsc_ref_mpi", mpiinc=SHIM,
                plugdir=os.path.join(LIBDIR, "mpi"), extra_libs=["-lb200mpi"],
                rpaths=["$ORIGIN/..", "$ORIGIN/../../../oracle/_ref_mpi"], defs=["-DPETSC_HAVE_MPI_TYPE_DUP", "-DPETSC_HAVE_MPI_FINALIZED"]),
}


def inc(fl):
    return ["-I", CONF, "-I", os.path.join(REF, "include"), "-I", fl["mpiinc"], "-I", REF, "-I", os.path.join(ROOT, "include")] + fl["defs"]


def newer(out, srcs):
    return (not os.path.exists(out)) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs if os.path.exists(s))


def build_flavour(name, verbose=True):
    fl = FLAVOURS[name]
    reflib = os.path.join(fl["refdir"], "lib%s.so" % fl["reflib"])
    plugin = os.path.join(fl["plugdir"], "libpetscb200plugin.so")
    bindir = os.path.join(fl["refdir"], "bin")
    if not os.path.exists(reflib):
        if name == "uni":
            raise SystemExit("build oracle/_ref/libpetsc_ref.so first (oracle/build_ref.py)")
        if verbose:
            print("skipping the %s flavour: %s not built (oracle/build_ref.py --mpi)" % (name, reflib))
        return plugin
    os.makedirs(fl["plugdir"], exist_ok=True)
    os.makedirs(bindir, exist_ok=True)
    srcs = [os.path.join(HERE, f) for f in ("b200vec.c", "b200mat.c", "b200ksp.c", "b200reg.c", "b200scatter.c")]
    srcs = [f for f in srcs if os.path.exists(f)]
    deps = srcs + [os.path.join(HERE, "petscb200plugin.h"), os.path.join(ROOT, "include", "petscb200.h"),
                   os.path.join(CONF, "petscconf.h"), os.path.join(fl["mpiinc"], "mpi.h")]
    rp = ["-Wl,-rpath," + r for r in fl["rpaths"]]
    if newer(plugin, deps):
        cmd = ["gcc", "-shared", "-fPIC", "-O2", "-g", "-w",
               "-ffp-contract=off", "-D__INSDIR__=petsc-dev_b200/plugin/"] + inc(fl) + srcs + \
              ["-o", plugin, "-L", LIBDIR, "-lpetscb200", "-L", fl["refdir"], "-l" + fl["reflib"]] + fl["extra_libs"] + rp + \
              ["-Wl,--no-undefined", "-lm"]
        subprocess.run(cmd, check=True)
        if verbose:
            print("built", plugin)
    # the VecScatterCreate interposer (LD_PRELOAD): attaches the device scatter to every scatter an unmodified PETSc creates
    hook, hooksrc = os.path.join(fl["plugdir"], "libpetscb200scatterhook.so"), os.path.join(HERE, "b200scatterhook.c")
    if os.path.exists(hooksrc) and newer(hook, [hooksrc, plugin]):
        subprocess.run(["gcc", "-shared", "-fPIC", "-O2", "-w"] + inc(fl) + [hooksrc, "-o", hook, "-L", fl["plugdir"], "-lpetscb200plugin",
                        "-L", fl["refdir"], "-l" + fl["reflib"], "-L", LIBDIR] + fl["extra_libs"] + ["-Wl,-rpath,$ORIGIN"] + rp + ["-ldl"], check=True)
        if verbose:
            print("built", hook)
    examples = dict(EXAMPLES)
    if name == "mpi":
        examples.update(MPI_EXAMPLES)
    for prog, rel in examples.items():
        src = os.path.join(REF, rel)
        out = os.path.join(bindir, prog)
        if not os.path.exists(src):
            continue
        if newer(out, [src, reflib]):
            cmd = ["gcc", "-O1", "-w", "-D__INSDIR__=%s/" % os.path.dirname(rel)] + inc(fl) + [src, "-o", out,
                   "-L", fl["refdir"], "-l" + fl["reflib"], "-L", LIBDIR] + fl["extra_libs"] + \
                  ["-Wl,-rpath,$ORIGIN/..", "-Wl,-rpath,$ORIGIN/../../../petsc-dev_b200/lib", "-Wl,--export-dynamic", "-lm", "-ldl"]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode:
                print("example %s failed to build:\n%s" % (prog, r.stderr[-1500:]), file=sys.stderr)
            elif verbose:
                print("built", out)
    return plugin


def main(verbose=True):
    if not os.path.isdir(REF):
        if verbose:
            print("reference headers absent; prebuilt plugin:", PLUGIN, os.path.exists(PLUGIN))
        return PLUGIN
    build_flavour("uni", verbose)
    build_flavour("mpi", verbose)
    return PLUGIN


if __name__ == "__main__":
    main()
