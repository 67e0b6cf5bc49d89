"""The C-ABI library loads without a GPU and exports every symbol include/petscb200.h declares; host-only entry
points behave; compute entry points fail loudly (no CPU fallback) when there is no device."""
import ctypes as C
import re

import numpy as np
import pytest


def declared_symbols(pb):
    txt = open(pb.HEADER_PATH).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_header(pb):
    syms = declared_symbols(pb)
    assert len(syms) > 60
    missing = [s for s in syms if not hasattr(pb.lib, s)]
    assert not missing, missing


def test_version_and_error_string(pb):
    assert pb.lib.b200_version() >= 100
    assert isinstance(pb.lib.b200_last_error(), bytes)


def test_split_ownership_matches_reference_rule(pb, P):
    par = __import__("importlib").import_module("petsc-dev_b200.parallel")
    for N, size, bs in [(10, 3, 1), (1000000, 8, 1), (134217728, 8, 1), (6291456, 8, 3), (7, 8, 1), (0, 4, 1), (12, 5, 3)]:
        r = par.split_ownership(N, size, bs)
        assert np.array_equal(r, P.split_ownership(N, size, bs)), (N, size, bs)
    with pytest.raises(pb.PetscB200Error) as e:
        par.split_ownership(10, 2, 3)
    assert e.value.code == 60


def test_null_and_range_errors(pb):
    assert pb.lib.b200_split_ownership(10, 2, 1, None) == 85
    assert pb.lib.b200_ksp_default_options(None) == 85
    o = pb.ksp_options()
    assert (o.ksp_type, o.pc_type, o.max_it, o.gmres_restart) == (0, 1, 10000, 30)
    assert (o.rtol, o.abstol, o.dtol) == (1e-5, 1e-50, 1e5)


def test_no_cpu_fallback(pb):
    """Without a device every compute call must fail with PETSC_ERR_LIB, never compute on the host."""
    n = C.c_int()
    rc = pb.lib.b200_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pb.PetscB200Error) as e:
        pb.init(0)
    assert e.value.code == 76
    with pytest.raises(pb.PetscB200Error):
        pb.Vec(8)
    ip = np.array([0, 1], np.int32); j = np.array([0], np.int32); a = np.array([1.0])
    with pytest.raises(pb.PetscB200Error):
        pb.Mat.from_csr(1, 1, ip, j, a)


def test_noalloc_mdot_checks_its_limits_before_touching_a_device(pb):
    """b200_vec_mdot_noalloc validates sizes and pointers first: beyond the static scratch (16 vectors, 2^30 entries) it
    answers PETSC_ERR_ARG_SIZ with a message that names the limit; without a device the valid call fails loudly."""
    f = pb.lib.b200_vec_mdot_noalloc
    dummy = (pb.c_dp * 17)()
    x = C.cast(C.c_void_p(256), pb.c_dp)          # never dereferenced by the argument checks
    assert f(8, x, 17, dummy, x, C.c_void_p(None)) == 60 and b"exceeds the 16 vectors" in pb.lib.b200_last_error()
    assert f((1 << 30) + 1, x, 1, dummy, x, C.c_void_p(None)) == 60 and b"exceeds the maximum" in pb.lib.b200_last_error()
    assert f(-1, x, 1, dummy, x, C.c_void_p(None)) == 60
    assert f(8, x, -1, dummy, x, C.c_void_p(None)) == 63
    assert f(8, x, 0, None, None, C.c_void_p(None)) == 0
    assert f(8, x, 1, dummy, None, C.c_void_p(None)) == 85
    assert f(8, x, 1, dummy, x, C.c_void_p(None)) == 85          # y[0] is a null pointer
    n = C.c_int()
    if pb.lib.b200_device_count(C.byref(n)) == 0 and n.value > 0:
        return
    dummy[0] = x
    assert f(8, x, 1, dummy, x, C.c_void_p(None)) == 76          # no device: PETSC_ERR_LIB, nothing computed
