"""Host-side checks of the PETSc plugin that need no GPU: the shared library exports the constructors / registration
entry points the reference's registry binds, the reference's unmodified ex2 still reproduces its golden
through this oracle build, and the new types FAIL LOUDLY (PETSC_ERR_LIB = 76) where there is no CUDA device — the
product path never computes on the host."""
import ctypes as C
import json
import os
import subprocess

import pytest

from plugin_cases import BIN, CASES, PLUGIN

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "plugin_goldens.json")))


def _need():
    if not (os.path.exists(PLUGIN) and os.path.exists(os.path.join(BIN, "ksp_ex2"))):
        pytest.skip("plugin / example binaries not built (needs /root/reference: petsc-dev_b200/plugin/build_plugin.py)")


def test_plugin_exports():
    _need()
    lib = C.CDLL(PLUGIN, mode=C.RTLD_GLOBAL)
    for s in ("PetscB200Register", "PetscDLLibraryRegister_petscb200plugin", "VecCreate_SeqB200", "VecCreate_MPIB200",
              "VecCreate_B200", "MatCreate_SeqAIJB200", "MatCreate_MPIAIJB200", "VecB200GetDeviceRead",
              "VecB200GetDeviceWrite", "VecB200GetDeviceReadWrite", "VecB200GetWhere", "PetscB200GetCounters"):
        assert hasattr(lib, s), s


def test_reference_arm_reproduces_goldens():
    _need()
    for name, c in CASES.items():
        if "golden" not in GOLD[name]:
            continue
        r = subprocess.run([os.path.join(BIN, c["prog"])] + c["args"] + c.get("ref_args", []), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1000:]
        assert r.stdout.strip() == GOLD[name]["golden"].strip(), name


def test_no_device_fails_loudly():
    _need()
    n = C.c_int()
    import importlib
    pb = importlib.import_module("petsc-dev_b200")
    if pb.lib.b200_device_count(C.byref(n)) == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    c = CASES["ksp_ex2_cg_jacobi"]
    r = subprocess.run([os.path.join(BIN, c["prog"])] + c["args"] + ["-dll_prepend", PLUGIN] + c["b200_args"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0
    assert "Error in external library" in r.stderr and "No usable CUDA device" in r.stderr
    assert "KSP Residual norm" not in r.stdout


def test_own_batch_driver_reference_arm_matches_the_ex4_golden():
    """tests/c/batch_mpiaij.c (this repository's driver for MatSetValuesBatch on the parallel type) with the REFERENCE's
    mpiaij type on the 3 x 3 grid must print the rows of the reference's golden ex4_0.out — the driver builds the same
    elements as ex4.c, so the GPU test that compares the plugin arm with this arm is anchored to the golden too.  The
    new KSP types and the batch ops are exported."""
    _need()
    exe = os.path.join(BIN, "b200_batch_mpiaij")
    if not os.path.exists(exe):
        pytest.skip("tests/c/batch_mpiaij.c not built")
    r = subprocess.run([exe, "-X", "3", "-Y", "3", "-mat_type", "mpiaij"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-1000:]
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_goldens.json")))["ksp_ex4_0"]["rows"]
    want = [("row %d: %s" % (k, " ".join("(%d, %g) " % (c, v) for c, v in row))).strip() for k, row in enumerate(gold)]
    got = [l.strip() for l in r.stdout.splitlines() if l.startswith("row ")]
    assert got == want
    lib = C.CDLL(PLUGIN, mode=C.RTLD_GLOBAL)
    for s in ("KSPCreate_CGB200", "KSPCreate_GMRESB200", "MatB200GetHandle", "MatSeqAIJSetPreallocation_SeqAIJB200"):
        assert hasattr(lib, s), s
