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
