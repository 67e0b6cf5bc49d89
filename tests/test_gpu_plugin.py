"""Drop-in test: the reference's own, UNMODIFIED example programs (oracle/_ref/bin/*, built from /root/reference where
they lie) are started with `-dll_prepend libpetscb200plugin.so -mat_type aijb200 -vec_type b200` and must print what
they print with the reference's CPU types — the reference's committed golden output where one exists
(tests/golden/plugin_goldens.json, made by tests/golden/make_plugin_golden.py), and the live reference arm otherwise.
Numbers printed with 6 significant digits are compared as text first and to 1e-5 relative where the last digit rounds
the other way; `-ksp_monitor` residual histories (12 digits) are held to the north-star tolerance of 1e-10 relative
and equal iteration counts."""
import json
import os
import re
import subprocess

import pytest

from plugin_cases import BIN, CASES, PLUGIN

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plugin_goldens.json")))
NUM = re.compile(r"^[-+]?(\d+\.?\d*|\.\d+)([eE][-+]?\d+)?$")


def run(prog, args):
    exe = os.path.join(BIN, prog)
    if not os.path.exists(exe) or not os.path.exists(PLUGIN):
        pytest.fail("plugin / reference example binaries not built (petsc-dev_b200/plugin/build_plugin.py)")
    r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-3000:])
    return r.stdout


def b200_run(case):
    c = CASES[case]
    return run(c["prog"], c["args"] + ["-dll_prepend", PLUGIN] + c["b200_args"])


def same_text(got, want, rtol=1e-5, ignore=("type:",)):
    gl, wl = got.strip().splitlines(), want.strip().splitlines()
    assert len(gl) == len(wl), "line counts differ:\n%s\n---\n%s" % (got, want)
    for a, b in zip(gl, wl):
        if a.strip() == b.strip() or any(t in b for t in ignore):
            continue
        ta, tb = a.split(), b.split()
        assert len(ta) == len(tb), (a, b)
        for x, y in zip(ta, tb):
            if x == y:
                continue
            assert NUM.match(x) and NUM.match(y), (a, b)
            fx, fy = float(x), float(y)
            assert abs(fx - fy) <= rtol * max(abs(fx), abs(fy), 1e-300) or max(abs(fx), abs(fy)) < 1e-12, (a, b)


def history(text):
    return [float(m.group(1)) for m in re.finditer(r"KSP Residual norm (\S+)", text)]


@pytest.mark.parametrize("case", [k for k, c in CASES.items() if not c.get("history")])
def test_reference_example_output(case):
    got = b200_run(case)
    c = CASES[case]
    if "golden" in GOLD[case]:
        same_text(got, GOLD[case]["golden"])
    same_text(got, GOLD[case]["reference"])
    live = run(c["prog"], c["args"] + c.get("ref_args", []))          # the reference arm, run here and now
    same_text(got, live)


@pytest.mark.parametrize("case", [k for k, c in CASES.items() if c.get("history")])
def test_reference_ksp_history(case):
    c = CASES[case]
    got = b200_run(case)
    live = run(c["prog"], c["args"] + c.get("ref_args", []))
    for want in (GOLD[case]["reference"], live):
        hg, hw = history(got), history(want)
        assert len(hg) == len(hw) and len(hg) > 3, (len(hg), len(hw))
        dev = max(abs(a - b) / b for a, b in zip(hg, hw))
        assert dev <= 1e-10, dev
        tail_g, tail_w = got.strip().splitlines()[-1], want.strip().splitlines()[-1]     # "Norm of error ... iterations N"
        assert tail_g.split()[-1] == tail_w.split()[-1]
        same_text(tail_g, tail_w, rtol=1e-4)


def test_types_really_are_b200():
    """-vec_view / -mat_view ::info print the type names: the programs above must have run on the new types."""
    c = CASES["ksp_ex2_cg_jacobi"]
    out = run(c["prog"], ["-m", "6", "-n", "6", "-ksp_type", "cg", "-pc_type", "jacobi", "-dll_prepend", PLUGIN] + c["b200_args"] + ["-ksp_view"])
    assert "type: seqaijb200" in out, out
    c = CASES["ksp_ex2_mpiaij"]
    out = run(c["prog"], ["-m", "6", "-n", "6", "-ksp_type", "cg", "-pc_type", "jacobi", "-dll_prepend", PLUGIN] + c["b200_args"] + ["-ksp_view"])
    assert "type: mpiaijb200" in out, out


def test_log_summary_counts_device_flops():
    """The ops log the reference's flop counts (2 nz - nonzerorow per MatMult, ...), so -log_summary stays comparable."""
    c = CASES["ksp_ex2_cg_jacobi"]
    a = run(c["prog"], c["args"] + ["-dll_prepend", PLUGIN] + c["b200_args"] + ["-log_summary"])
    b = run(c["prog"], c["args"] + ["-log_summary"])

    def flops(text, ev):
        for l in text.splitlines():
            if l.startswith(ev + " "):
                return float(l.split()[5])
        return None
    for ev in ("MatMult", "VecAXPY", "VecAYPX"):
        fa, fb = flops(a, ev), flops(b, ev)
        assert fa is not None and fb is not None and abs(fa - fb) <= 1e-3 * fb, (ev, fa, fb)


def matview_rows(text):
    return [l.strip() for l in text.splitlines() if re.match(r"row \d+:", l)]


def test_ex4_matsetvaluesbatch_on_device():
    """src/ksp/ksp/examples/tutorials/ex4.c ("Test MatSetValuesBatch") UNMODIFIED, CPU stage only (-gpu 0: the other
    stage hard-codes MATAIJCUSP): with -mat_type aijb200 MatSetValuesBatch dispatches to the plugin's device assembly,
    which — like the reference's MatSetValuesBatch_SeqAIJCUSP — creates the matrix from the element matrices alone.
    Its MatView must show exactly the rows the reference's device assembly printed in its golden output
    (ex4_0.out, "type: seqaijcusp": 41 entries), and the solve must follow the reference arm's residual history."""
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_goldens.json")))["ksp_ex4_0"]["rows"]
    want = [("row %d: %s" % (r, " ".join("(%d, %g) " % (c, v) for c, v in row))).strip() for r, row in enumerate(gold)]
    out = run("ksp_ex4", ["-gpu", "0", "-dll_prepend", PLUGIN, "-mat_type", "aijb200", "-vec_type", "b200"])
    assert "type: seqaijb200" in out
    assert matview_rows(out) == want, out
    # a larger grid, 40 iterations of CG + Jacobi on the Neumann problem with the constant null space removed
    # (KSPSetNullSpace; the random right-hand side is not consistent, so the solve is cut off instead of converging)
    args = ["-gpu", "0", "-view", "0", "-solve", "-da_grid_x", "24", "-da_grid_y", "19", "-ksp_type", "cg", "-pc_type", "jacobi", "-ksp_monitor", "-ksp_max_it", "40"]
    got = run("ksp_ex4", args + ["-dll_prepend", PLUGIN, "-mat_type", "aijb200", "-vec_type", "b200"])
    ref = run("ksp_ex4", args)
    hg, hw = history(got), history(ref)
    assert len(hg) == len(hw) and len(hg) > 5, (len(hg), len(hw))
    assert max(abs(a - b) / b for a, b in zip(hg, hw)) <= 1e-10


@pytest.mark.parametrize("case,uploads", [("ksp_ex5_1", "SeqAIJ"), ("ksp_ex5_mpiaij", "MPIAIJ")])
def test_same_pattern_reassembly_refreshes_values_only(case, uploads):
    """ex5 assembles a second matrix with the same nonzero structure: one full upload, then values only (-info lines of
    MatB200CopyToDevice_*); a MatSeqAIJSetPreallocation in between would have dropped the device copy instead."""
    c = CASES[case]
    out = run(c["prog"], c["args"] + ["-dll_prepend", PLUGIN] + c["b200_args"] + ["-info"])
    mine = [l for l in out.splitlines() if "MatB200CopyToDevice_%s" % uploads in l]
    assert sum("full upload" in l for l in mine) == 1, mine
    assert sum("values refreshed" in l for l in mine) >= 1, mine


def test_parallel_matsetvaluesbatch_takes_the_device_path():
    """tests/c/batch_mpiaij.c with -mat_type mpiaijb200: the -info line of MatSetValuesBatch_MPIAIJB200 shows the plugin's
    op ran (the output itself is compared with the reference arm by test_reference_example_output[batch_mpiaij]), and the
    3 x 3 grid gives the rows of the reference's golden ex4_0.out."""
    c = CASES["batch_mpiaij"]
    out = run(c["prog"], ["-X", "3", "-Y", "3", "-dll_prepend", PLUGIN] + c["b200_args"] + ["-info"])
    assert "MatSetValuesBatch_MPIAIJB200" in out and "on the device" in out
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_goldens.json")))["ksp_ex4_0"]["rows"]
    want = [("row %d: %s" % (r, " ".join("(%d, %g) " % (cc, v) for cc, v in row))).strip() for r, row in enumerate(gold)]
    assert matview_rows(out) == want, out


def test_read_access_does_not_invalidate_the_device_copy():
    """tests/c/coherence_seam.c: two VecGetArrayRead passes between device operations and one VecGetArray write — the
    plugin's -info totals must show 2 downloads and exactly 1 upload (the write), and the numbers must be the reference's."""
    want = run("b200_coherence_seam", [])
    got = run("b200_coherence_seam", ["-dll_prepend", PLUGIN, "-vec_type", "b200", "-info"])
    m = re.search(r"B200 totals: (\d+) vector uploads, (\d+) vector downloads", got)
    assert m, got[-1500:]
    assert (int(m.group(1)), int(m.group(2))) == (1, 2), m.group(0)
    plain = "\n".join(l for l in got.splitlines() if not l.startswith("["))
    same_text(plain, want)
    assert "after the write: norm1 23" in want
