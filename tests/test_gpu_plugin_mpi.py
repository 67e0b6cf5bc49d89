"""The drop-in boundary on a real multi-rank PETSc: the reference built against the b200mpi subset (oracle/_ref_mpi, every
source file unmodified), its example programs started by b200mpiexec on 1, 2, ... ranks with
`-dll_prepend petsc-dev_b200/lib/mpi/libpetscb200plugin.so -mat_type aijb200 -vec_type b200` — one rank per GPU — against the
same programs on the same number of ranks with the reference's CPU types, and against the reference's committed goldens.
This is where the plugin's `size > 1` code runs: MatB200CopyToDevice_MPIAIJ reading the VecScatter lists, the ghost exchange
GPU to GPU inside MatMult_MPIAIJB200, Vec{Dot,Norm,MDot}_MPIB200 all-reducing over MPI, KSPSolve_CGB200 across ranks.
Ranks beyond the number of GPUs on the box are skipped (the round-end driver box has one GPU: the -n 1 cases run there)."""
import json
import os
import subprocess

import pytest

from test_gpu_plugin import history, same_text

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "oracle", "_ref_mpi", "bin")
MPIEXEC = os.path.join(ROOT, "petsc-dev_b200", "lib", "b200mpiexec")
PLUGIN = os.path.join(ROOT, "petsc-dev_b200", "lib", "mpi", "libpetscb200plugin.so")
B200 = ["-dll_prepend", PLUGIN, "-mat_type", "aijb200", "-vec_type", "b200"]


def ngpus():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=60).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


def launch(np_, prog, args):
    exe = os.path.join(BIN, prog)
    if not (os.path.exists(exe) and os.path.exists(PLUGIN) and os.path.exists(MPIEXEC)):
        pytest.fail("MPI flavour not built (oracle/build_ref.py --mpi, petsc-dev_b200/plugin/build_plugin.py)")
    r = subprocess.run([MPIEXEC, "-n", str(np_), exe] + args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    return drop_nccl_banner(r.stdout)


def drop_nccl_banner(out):
    """NCCL prints its version on stdout when the box sets NCCL_DEBUG=VERSION: not program output"""
    return "\n".join(l for l in out.splitlines() if not l.startswith("NCCL version")) + "\n"


def need(np_):
    if ngpus() < np_:
        pytest.skip("box has fewer than %d GPUs" % np_)


RANKS = [1, 2, 4, 8]


@pytest.mark.parametrize("np_", RANKS)
def test_zerorows_rewrites_columns_in_place(np_):
    """MatZeroRows_MPIAIJ -> MatZeroRows_SeqAIJ on the diagonal block: i and nz unchanged, j changed (aij.c:1680-1683); the
    refresh decision of MatB200CopyToDevice_* hashes the column indices and must do a full upload."""
    need(np_)
    args = ["-n", str(6 * np_ + 3)]
    want = launch(np_, "b200_zerorows_refresh", args + ["-mat_type", "mpiaij"])
    got = launch(np_, "b200_zerorows_refresh", args + B200 + ["-info"])
    # one rank: MATAIJB200 resolves to the sequential type (MatRegisterBaseName, matreg.c:170), same decision there
    # PetscInfo on an object prints from the first rank of the object's communicator only
    mine = [l for l in got.splitlines() if "MatB200CopyToDevice_" in l]
    assert sum("full upload" in l for l in mine) == 2 and not any("values refreshed" in l for l in mine), mine
    plain = "\n".join(l for l in got.splitlines() if not l.startswith("["))          # -info lines start with "[rank]"
    same_text(plain, want)
    assert "after MatZeroRows" in want


@pytest.mark.parametrize("ksp", ["cg", "cgb200", "gmres", "gmresb200"])
@pytest.mark.parametrize("np_", RANKS)
def test_ksp_ex2_history_on_ranks(np_, ksp):
    need(np_)
    args = ["-ksp_monitor", "-m", "37", "-n", "41", "-pc_type", "jacobi", "-ksp_gmres_restart", "30"]
    want = launch(np_, "ksp_ex2", args + ["-ksp_type", ksp.replace("b200", "")])
    got = launch(np_, "ksp_ex2", args + B200 + ["-ksp_type", ksp])
    hg, hw = history(got), history(want)
    assert len(hg) == len(hw) and len(hg) > 10, (len(hg), len(hw))
    assert max(abs(a - b) / b for a, b in zip(hg, hw)) <= 1e-10
    assert got.strip().splitlines()[-1].split()[-1] == want.strip().splitlines()[-1].split()[-1]       # iterations N


@pytest.mark.parametrize("np_", RANKS)
def test_ref_bench_driver_on_ranks(np_):
    """oracle/ref_bench.c (PETSc API only) with the B200 types against itself with the reference's: y = A x by its norms, the
    20-iteration CG history, on a grid whose z planes split unevenly over the ranks."""
    need(np_)
    args = ["-nx", "24", "-ny", "20", "-nz", str(7 * np_ + 1), "-reps", "2", "-warm", "1", "-cg_its", "20"]

    def js(out):
        return json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    want = js(launch(np_, "b200_ref_bench", args))
    got = js(launch(np_, "b200_ref_bench", args + B200))
    assert got["mat_type"].endswith("b200"), got["mat_type"]
    for k in ("y_norm1", "y_norm2"):
        assert abs(got[k] - want[k]) <= 1e-13 * want[k], (k, got[k], want[k])
    assert len(got["hist"]) == len(want["hist"])
    assert max(abs(a - b) / b for a, b in zip(got["hist"], want["hist"])) <= 1e-10


GOLD = os.path.join(ROOT, "tests", "golden", "ref_mpi_goldens.json")      # the reference's golden texts, committed copy
VEC_GOLDENS = [  # (ranks, program, reference golden): those of tests/test_mpi_oracle.py's runs that call VecSetFromOptions
    (2, "vec_ex3", "src/vec/vec/examples/tests/output/ex3_1.out"),
    (2, "vec_ex4", "src/vec/vec/examples/tests/output/ex4_1.out"),
    (2, "vec_ex9", "src/vec/vec/examples/tests/output/ex9_1.out"),
    (2, "vec_ex10", "src/vec/vec/examples/tests/output/ex10_1.out"),
    (2, "vec_ex11", "src/vec/vec/examples/tests/output/ex11_1.out"),
    (2, "vec_ex12", "src/vec/vec/examples/tests/output/ex12_1.out"),
    (2, "vec_ex14", "src/vec/vec/examples/tests/output/ex14_1.out"),
    (2, "vec_ex17", "src/vec/vec/examples/tests/output/ex17_1.out"),
    (2, "vec_ex21", "src/vec/vec/examples/tests/output/ex21_2.out"),
    (2, "vec_ex23", "src/vec/vec/examples/tests/output/ex23_1.out"),
    (4, "vec_ex22", "src/vec/vec/examples/tests/output/ex22_1.out"),
]


@pytest.mark.parametrize("case", VEC_GOLDENS, ids=lambda c: "%s_np%d" % (c[1], c[0]))
def test_reference_vec_goldens_with_b200_vectors(case):
    """The reference's np > 1 vector tests (VecScatter with general / stride index sets, ghost updates, norms, VecSetValues
    to other ranks) with -vec_type b200 against their committed goldens."""
    np_, prog, rel = case
    need(np_)
    got = launch(np_, prog, ["-dll_prepend", PLUGIN, "-vec_type", "b200"])
    same_text(got, json.load(open(GOLD))[rel], ignore=("type:",))


HOOK = os.path.join(ROOT, "petsc-dev_b200", "lib", "mpi", "libpetscb200scatterhook.so")


def launch_hooked(np_, prog, args):
    """as launch(), with VecScatterCreate interposed (LD_PRELOAD) so that every scatter gets VecScatterB200Attach"""
    exe = os.path.join(BIN, prog)
    assert os.path.exists(HOOK), "scatter hook not built (petsc-dev_b200/plugin/build_plugin.py)"
    env = dict(os.environ, LD_PRELOAD=HOOK)
    r = subprocess.run([MPIEXEC, "-n", str(np_), exe] + args, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    return drop_nccl_banner(r.stdout)


def scatter_totals(out):
    """(on the device, through the host) summed over ranks from the plugin's -info totals line"""
    import re
    dev = host = n = 0
    for m in re.finditer(r"B200 totals: .*? (\d+) scatters on the device, (\d+) through the host", out):
        dev += int(m.group(1)); host += int(m.group(2)); n += 1
    return dev, host, n


@pytest.mark.parametrize("np_", [1, 2, 3, 4, 8])
def test_vecscatter_object_runs_on_the_device(np_):
    """PETSc's own VecScatter object (VecScatterCreate/Begin/End/Copy, vscat.c:920,1627,1698) between B200 vectors: all 19
    scatters of tests/c/scatter_cases.c (PtoS general/stride/block, PtoP, StoP, the four sequential layouts; forward and
    reverse, INSERT and ADD, repeated targets) go through b200_scatter_* — none through the host — and print what the
    reference's vector types print on the same number of ranks."""
    need(np_)
    want = launch(np_, "b200_scatter_cases", [])
    got = launch_hooked(np_, "b200_scatter_cases", ["-dll_prepend", PLUGIN, "-vec_type", "b200", "-info"])
    dev, host, n = scatter_totals(got)
    assert n == np_ and dev == 19 * np_ and host == 0, (dev, host, n)
    plain = "\n".join(l for l in got.splitlines() if not (l.startswith("[") and "] " in l and not l.split("] ", 1)[1][:1].isdigit()))
    same_text(plain, want)


def test_vecscatter_hook_leaves_host_vectors_alone():
    """the same program, hook preloaded, reference vector types: every scatter takes the reference's own code"""
    need(2)
    want = launch(2, "b200_scatter_cases", [])
    got = launch_hooked(2, "b200_scatter_cases", ["-dll_prepend", PLUGIN])
    same_text(got, want)


def test_vec_ex22_parallel_to_parallel_on_the_device():
    """the reference's own PtoP scatter test (both vectors from VecSetFromOptions) with the device scatter attached"""
    need(4)
    got = launch_hooked(4, "vec_ex22", ["-dll_prepend", PLUGIN, "-vec_type", "b200", "-info"])
    dev, host, n = scatter_totals(got)
    assert dev >= 4 and host == 0, (dev, host)
    plain = "\n".join(l for l in got.splitlines() if not (l.startswith("[") and "] " in l and not l.split("] ", 1)[1][:1].isdigit()))
    same_text(plain, json.load(open(GOLD))["src/vec/vec/examples/tests/output/ex22_1.out"], ignore=("type:",))


@pytest.mark.parametrize("np_", [1, 2, 3, 4])
def test_offprocess_matsetvalues_through_the_stash(np_):
    """tests/c/stash_offproc.c: ADD_VALUES into rows other ranks own (MatSetValues_MPIAIJ -> MatStash -> MatAssemblyEnd,
    mpiaij.c:496-570, inherited host code), two assemblies — the second one adds new nonzeros and ghost columns on three or
    more ranks, so the device copy must be rebuilt — MatMult and MatMultTranspose after each: same output as the reference types."""
    need(np_)
    want = launch(np_, "b200_stash_offproc", ["-mat_type", "aij"])
    got = launch(np_, "b200_stash_offproc", B200)
    same_text(got, want, ignore=("type:",))
    assert "after assembly 1" in want


def test_mat_ex5_on_three_ranks_golden():
    """the reference's own mat ex5 on 3 ranks (MatMult, MatMultAdd, MatMultTranspose[Add] — the reverse ADD exchange —,
    MatGetDiagonal, MatDiagonalScale) with the B200 types against its committed golden output ex5_23.out"""
    need(3)
    got = launch(3, "mat_ex5", ["-dll_prepend", PLUGIN, "-mat_type", "mpiaijb200", "-vec_type", "mpib200"])
    same_text(got, json.load(open(GOLD))["src/mat/examples/tests/output/ex5_23.out"], ignore=("type:",))
