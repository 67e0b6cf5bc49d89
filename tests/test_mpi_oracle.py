"""P > 1 parity pinned by the REFERENCE ITSELF: the unmodified reference (oracle/_ref_mpi, built against the b200mpi subset)
runs on P = 2, 3, 8 ranks under petsc-dev_b200/lib/b200mpiexec and dumps what its own MatSetValues_MPIAIJ / MatStash,
MatSetUpMultiply_MPIAIJ (mmaij.c:9-161), MatCreateColmap_MPIAIJ_Private (mpiaij.c:357-377) and VecScatterCreate_PtoS
(vpscat.c:1730-1924) built (tests/c/dump_mpiaij.c); libpetscb200's host restatements must equal them bit for bit, and
y = MatMult_MPIAIJ(x) must equal, per rank, MatMult_SeqAIJ on the diagonal block + MatMultAdd_SeqAIJ on the off-diagonal
block (the P > 1 numerical oracle the GPU tests use).  Also: the reference's own np > 1 golden outputs under the shim, which
is what pins the shim."""
import importlib
import os
import subprocess

import numpy as np
import pytest

from test_mpiaij_host import _gens, serial_plans

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MPIEXEC = os.path.join(ROOT, "petsc-dev_b200", "lib", "b200mpiexec")
BIN = os.path.join(ROOT, "oracle", "_ref_mpi", "bin")
REFOUT = os.path.join(ROOT, "tests", "golden", "ref_mpi_goldens.json")

needs_mpi_oracle = pytest.mark.skipif(not (os.path.exists(MPIEXEC) and os.path.exists(os.path.join(BIN, "b200_dump_mpiaij"))),
                                      reason="oracle/_ref_mpi not built (oracle/build_ref.py --mpi; petsc-dev_b200/plugin/build_plugin.py)")


def mpirun(n, prog, args=(), timeout=300, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([MPIEXEC, "-n", str(n), os.path.join(BIN, prog)] + list(args), capture_output=True, text=True, timeout=timeout, env=e)
    assert r.returncode == 0, "%s on %d ranks failed (rc %d):\n%s\n%s" % (prog, n, r.returncode, r.stdout[-3000:], r.stderr[-3000:])
    return r.stdout


def write_inputs(prefix, P, ranges, gen, N, bs, seed=2):
    for r in range(len(ranges) - 1):
        r0, r1 = int(ranges[r]), int(ranges[r + 1])
        ai, aj, aa = gen(r0, r1)
        ai = (ai - ai[0]).astype(np.int32)
        with open("%s.%d.in" % (prefix, r), "wb") as f:
            np.array([r1 - r0, N, bs, len(aj), r0], np.int32).tofile(f)
            ai.tofile(f); np.asarray(aj, np.int32).tofile(f); np.asarray(aa, np.float64).tofile(f)
            P.random_vector(r1 - r0, seed, r0=r0).tofile(f)


def read_dump(path):
    out = {}
    for line in open(path):
        w = line.split()
        name, n = w[0], int(w[1])
        if name in ("Aa", "Ba", "y"):
            out[name] = np.array([float.fromhex(v) for v in w[2:2 + n]], np.float64)
        else:
            out[name] = np.array([int(v) for v in w[2:2 + n]], np.int64)
    return out


@needs_mpi_oracle
@pytest.mark.parametrize("stash", [False, True])
@pytest.mark.parametrize("name", ["lap2d", "lap7", "s27", "bs3", "plaw"])
@pytest.mark.parametrize("size", [2, 3, 8])
def test_structures_equal_the_reference_run_on_P_ranks(pb, P, R, tmp_path, name, size, stash):
    par = importlib.import_module("petsc-dev_b200.parallel")
    N, bs, gen = _gens(P)[name]
    ranges, host = serial_plans(par, size, N, gen, bs)
    prefix = str(tmp_path / "case")
    write_inputs(prefix, P, ranges, gen, N, bs)
    mpirun(size, "b200_dump_mpiaij", ["-prefix", prefix] + (["-stash"] if stash else []))
    for r in range(size):
        d = read_dump("%s.%d.out" % (prefix, r))
        sp, f, t = host[r]["split"], host[r]["from_"], host[r]["to"]
        assert np.array_equal(d["range"], ranges)
        for k in ("Ai", "Aj", "Bi", "Bj", "garray"):
            assert np.array_equal(d[k], sp[k]), (name, size, r, k)
        assert np.array_equal(d["Aa"], sp["Aa"]) and np.array_equal(d["Ba"], sp["Ba"])
        assert np.array_equal(d["colmap"], par.colmap(N, sp["garray"]))
        assert int(d["scatter_bs"][0]) == host[r]["bs"]
        assert np.array_equal(d["from_procs"], f["from_procs"]) and np.array_equal(d["from_starts"], f["from_starts"])
        assert np.array_equal(d["from_indices"], f["from_indices"])
        assert np.array_equal(d["to_procs"], t["to_procs"]) and np.array_equal(d["to_starts"], t["to_starts"])
        assert np.array_equal(d["to_indices"], t["to_indices"])
        assert d["to_local"].size == f["nlocal"] == 0 and int(d["from_contiq"][0]) == 1
        # numerical oracle: MatMult_MPIAIJ on P ranks == per rank MatMult_SeqAIJ(A) then MatMultAdd_SeqAIJ(B, lvec)
        m = int(ranges[r + 1] - ranges[r])
        xl = P.random_vector(m, 2, r0=int(ranges[r]))
        lvec = P.random_vector_at(sp["garray"], 2)
        yd = R.RefMat(m, m, sp["Ai"], sp["Aj"], sp["Aa"], no_inode=True).mult(xl)
        y = R.RefMat(m, max(len(sp["garray"]), 1), sp["Bi"], sp["Bj"], sp["Ba"], no_inode=True).multadd(np.resize(lvec, max(len(lvec), 1)), yd) if len(sp["garray"]) else yd
        if bs == 1:
            assert np.array_equal(d["y"], y)          # rows with repeated patterns take the Inode kernel in the P-rank run
        assert np.all(np.abs(d["y"] - y) <= 1e-13 * np.maximum(np.abs(y), 1.0))


GOLDEN_RUNS = [  # (ranks, program, args, reference golden, order-free?)
    (2, "ksp_ex2", ["-ksp_monitor_short", "-m", "5", "-n", "5", "-ksp_gmres_cgs_refinement_type", "refine_always"], "src/ksp/ksp/examples/tutorials/output/ex2_2.out", False),
    (4, "ksp_ex2", ["-pc_type", "bjacobi", "-pc_bjacobi_blocks", "1", "-ksp_monitor_short", "-sub_pc_type", "jacobi", "-sub_ksp_type", "gmres"], "src/ksp/ksp/examples/tutorials/output/ex2_bjacobi.out", False),
    (3, "mat_ex5", ["-mat_type", "mpiaij"], "src/mat/examples/tests/output/ex5_23.out", False),
    (2, "vec_ex1", [], "src/vec/vec/examples/tests/output/ex1_2.out", False),
    (2, "vec_ex3", [], "src/vec/vec/examples/tests/output/ex3_1.out", False),
    (2, "vec_ex4", [], "src/vec/vec/examples/tests/output/ex4_1.out", False),
    (2, "vec_ex9", [], "src/vec/vec/examples/tests/output/ex9_1.out", False),
    (2, "vec_ex10", [], "src/vec/vec/examples/tests/output/ex10_1.out", False),
    (2, "vec_ex11", [], "src/vec/vec/examples/tests/output/ex11_1.out", False),
    (2, "vec_ex12", [], "src/vec/vec/examples/tests/output/ex12_1.out", False),
    (4, "vec_ex13", [], "src/vec/vec/examples/tests/output/ex13_1.out", False),
    (2, "vec_ex14", [], "src/vec/vec/examples/tests/output/ex14_1.out", False),
    (2, "vec_ex17", [], "src/vec/vec/examples/tests/output/ex17_1.out", False),
    (2, "vec_ex21", [], "src/vec/vec/examples/tests/output/ex21_2.out", False),
    (4, "vec_ex22", [], "src/vec/vec/examples/tests/output/ex22_1.out", False),
    (2, "vec_ex23", [], "src/vec/vec/examples/tests/output/ex23_1.out", False),
    (3, "vec_ex24", [], "src/vec/vec/examples/tests/output/ex24_1.out", True),     # every rank prints: line order is the OS's
    (3, "vec_ex25", [], "src/vec/vec/examples/tests/output/ex25_1.out", False),
    (4, "vec_ex33", [], "src/vec/vec/examples/tests/output/ex33_1.out", False),
]


def golden_text(rel):
    """The reference's golden output: read from /root/reference where it exists, else from the committed copy of its
    text (tests/golden/ref_mpi_goldens.json, written by tests/golden/make_mpi_golden.py) — the GPU box has no /root/reference."""
    import json
    p = os.path.join("/root/reference", rel)
    if os.path.exists(p):
        return open(p).read()
    return json.load(open(REFOUT))[rel]


@needs_mpi_oracle
@pytest.mark.parametrize("case", GOLDEN_RUNS, ids=lambda c: "%s_np%d" % (c[1], c[0]))
def test_reference_goldens_on_more_than_one_rank(case):
    n, prog, args, golden, orderfree = case
    got, want = mpirun(n, prog, args), golden_text(golden)
    if orderfree:
        assert sorted("".join(got.split())) == sorted("".join(want.split()))   # the ranks' writes interleave, even inside a line
    else:
        assert got == want


@needs_mpi_oracle
@pytest.mark.parametrize("n", [1, 2, 5, 8])
def test_shim_selftest(n):
    assert "b200mpi selftest ok %d" % n in mpirun(n, "mpishim_selftest")


@needs_mpi_oracle
@pytest.mark.parametrize("prog,ranks", [("b200_scatter_cases", (1, 2, 3)), ("b200_stash_offproc", (1, 2, 3, 4)), ("b200_coherence_seam", (1,))])
def test_plugin_test_programs_run_on_the_reference_types(prog, ranks):
    """The repository's own PETSc programs behind the GPU plugin tests (tests/c/*.c, public API only) must run under the
    reference's CPU types on every rank count the GPU tests use, and print the same text twice: they are the 'want' side of
    those tests (tests/test_gpu_plugin_mpi.py)."""
    if not os.path.exists(os.path.join(BIN, prog)):
        pytest.skip("%s not built (petsc-dev_b200/plugin/build_plugin.py)" % prog)
    for n in ranks:
        a = mpirun(n, prog, ["-mat_type", "aij"] if "stash" in prog else [])
        b = mpirun(n, prog, ["-mat_type", "aij"] if "stash" in prog else [])
        assert a == b and len(a.splitlines()) >= 2, (prog, n)
        if prog == "b200_scatter_cases":
            assert a.count("forward") + a.count("reverse") == 19          # the 19 scatters the device test counts
