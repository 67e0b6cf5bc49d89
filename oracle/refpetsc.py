"""oracle/refpetsc.py — TEST INFRASTRUCTURE ONLY.

ctypes binding of oracle/_ref/libpetsc_ref.so: the UNMODIFIED reference (erdc/petsc-dev) compiled from
/root/reference by oracle/build_ref.py, driven through oracle/ref_driver.c.  Used by tests/ as the parity
oracle and by bench.py's cpu_baseline / `--impl reference` leg as the timed CPU baseline.  Nothing in the
product path imports this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libpetsc_ref.so")

_lib = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("reference oracle library %s not built (run oracle/build_ref.py)" % LIB_PATH)
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        rc = _lib.ref_initialize()
        if rc:
            raise RuntimeError("reference PetscInitialize failed: %d" % rc)
    return _lib


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def _chk(rc, what):
    if rc:
        raise RuntimeError("reference %s returned PetscErrorCode %d" % (what, rc))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def set_option(name, value=""):
    _chk(lib().ref_set_option(name.encode(), (value or "").encode()), "PetscOptionsSetValue")


def clear_option(name):
    _chk(lib().ref_set_option(name.encode(), b"\001"), "PetscOptionsClearValue")


class RefMat:
    """An assembled reference Mat (MATSEQAIJ, or MATMPIAIJ on the 1-rank MPIUNI communicator)."""

    def __init__(self, m, n, ai, aj, aa, kind="seqaij", bs=1, no_inode=False):
        self.kind = {"seqaij": 0, "mpiaij": 1}[kind]
        self.m, self.n = int(m), int(n)
        self._ai, self._aj, self._aa = _i32(ai), _i32(aj), _f64(aa)
        h = C.c_void_p()
        _chk(lib().ref_mat_create(self.kind, self.m, self.n, int(bs), _ip(self._ai), _ip(self._aj), _dp(self._aa),
                                  int(bool(no_inode)), C.byref(h)), "MatCreate/assembly")
        self.h = h

    def close(self):
        if self.h:
            lib().ref_mat_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def inode_info(self):
        use, cnt = C.c_int(), C.c_int()
        lib().ref_mat_inode_info(self.h, self.kind, C.byref(use), C.byref(cnt))
        return bool(use.value), cnt.value

    def csr(self):
        assert self.kind == 0
        m, nz = C.c_int(), C.c_int()
        lib().ref_mat_get_csr(self.h, C.byref(m), C.byref(nz), None, None, None)
        i = np.empty(m.value + 1, np.int32)
        j = np.empty(nz.value, np.int32)
        a = np.empty(nz.value, np.float64)
        _chk(lib().ref_mat_get_csr(self.h, C.byref(m), C.byref(nz), _ip(i), _ip(j), _dp(a)), "get_csr")
        return i, j, a

    def mult(self, x, reps=1, want_time=False):
        x = _f64(x)
        y = np.empty(self.m, np.float64)
        t = C.c_double()
        _chk(lib().ref_mat_mult(self.h, self.kind, _dp(x), _dp(y), int(reps), C.byref(t)), "MatMult")
        return (y, t.value) if want_time else y

    def multadd(self, x, y, inplace=False):
        x = _f64(x)
        y = _f64(y).copy()
        z = y if inplace else np.empty(self.m, np.float64)
        _chk(lib().ref_mat_multadd(self.h, self.kind, _dp(x), _dp(y), _dp(z)), "MatMultAdd")
        return z

    def multtranspose(self, x, y=None):
        """A^T x (MatMultTranspose) or y + A^T x (MatMultTransposeAdd)."""
        x = _f64(x)
        z = np.empty(self.n, np.float64)
        yy = _f64(y).copy() if y is not None else None
        _chk(lib().ref_mat_multtranspose(self.h, self.kind, _dp(x), _dp(yy) if yy is not None else None, _dp(z)), "MatMultTranspose")
        return z

    def diagonal(self):
        d = np.empty(self.m, np.float64)
        _chk(lib().ref_mat_get_diagonal(self.h, self.kind, _dp(d)), "MatGetDiagonal")
        return d

    def ksp_solve(self, b, ksptype="cg", pctype="jacobi", rtol=1e-5, abstol=1e-50, dtol=1e5, maxit=10000, x0=None,
                  options=None):
        """KSPSolve through the reference; returns dict(x, history, its, reason, seconds)."""
        options = options or {}
        for k, v in options.items():
            set_option(k, v)
        try:
            b = _f64(b)
            x = np.zeros(self.n, np.float64) if x0 is None else _f64(x0).copy()
            hist = np.zeros(maxit + 2, np.float64)
            nh, its, reason, secs = C.c_int(), C.c_int(), C.c_int(), C.c_double()
            _chk(lib().ref_ksp_solve(self.h, self.kind, ksptype.encode(), pctype.encode(), C.c_double(rtol),
                                     C.c_double(abstol), C.c_double(dtol), int(maxit), int(x0 is not None), _dp(b),
                                     _dp(x), _dp(hist), int(hist.size), C.byref(nh), C.byref(its), C.byref(reason),
                                     C.byref(secs)), "KSPSolve")
        finally:
            for k in options:
                clear_option(k)
        return dict(x=x, history=hist[:nh.value].copy(), its=its.value, reason=reason.value, seconds=secs.value)


# ---- Vec kernels -----------------------------------------------------------------------------------------------
NORM_1, NORM_2, NORM_INFINITY, NORM_1_AND_2 = 0, 1, 3, 4
(OP_AXPY, OP_AYPX, OP_WAXPY, OP_AXPBY, OP_AXPBYPCZ, OP_SCALE, OP_SET, OP_COPY, OP_SWAP, OP_PMULT, OP_PDIV,
 OP_RECIP) = range(12)


def vec_dot(x, y):
    x, y = _f64(x), _f64(y)
    r = C.c_double()
    _chk(lib().ref_vec_dot(x.size, _dp(x), _dp(y), C.byref(r)), "VecDot")
    return r.value


def vec_norm(x, ntype=NORM_2):
    x = _f64(x)
    r = (C.c_double * 2)()
    _chk(lib().ref_vec_norm(x.size, _dp(x), int(ntype), r), "VecNorm")
    return (r[0], r[1]) if ntype == NORM_1_AND_2 else r[0]


def _ptrs(vs):
    arr = (c_dp * max(len(vs), 1))()
    for k, v in enumerate(vs):
        arr[k] = _dp(v)
    return arr


def vec_mdot(x, ys):
    x = _f64(x)
    ys = [_f64(y) for y in ys]
    z = np.zeros(max(len(ys), 1), np.float64)
    _chk(lib().ref_vec_mdot(x.size, _dp(x), len(ys), _ptrs(ys), _dp(z)), "VecMDot")
    return z[:len(ys)]


def vec_maxpy(y, alphas, xs):
    y = _f64(y).copy()
    xs = [_f64(x) for x in xs]
    al = _f64(alphas) if len(xs) else np.zeros(1)
    _chk(lib().ref_vec_maxpy(y.size, _dp(y), len(xs), _dp(al), _ptrs(xs)), "VecMAXPY")
    return y


def vec_elementwise(op, n, a=0.0, b=0.0, c=0.0, x=None, y=None, w=None):
    """In-place on the given float64 arrays (caller passes copies); see ref_driver.c for operand roles."""
    _chk(lib().ref_vec_elementwise(int(op), int(n), C.c_double(a), C.c_double(b), C.c_double(c), _dp(x), _dp(y),
                                   _dp(w)), "Vec elementwise op %d" % op)


def mat_write_binary(mat, path):
    _chk(lib().ref_mat_write_binary(mat.h, str(path).encode()), "MatView (binary)")


def mat_load_binary(path):
    """MatLoad through the reference -> RefMat (sequential)."""
    h = C.c_void_p()
    _chk(lib().ref_mat_load_binary(str(path).encode(), C.byref(h)), "MatLoad")
    r = RefMat.__new__(RefMat)
    r.kind, r.h = 0, h
    m, nz = C.c_int(), C.c_int()
    lib().ref_mat_get_csr(h, C.byref(m), C.byref(nz), None, None, None)
    r.m = m.value
    r.n = None
    return r


def vec_write_binary(x, path):
    x = _f64(x)
    _chk(lib().ref_vec_write_binary(int(x.size), _dp(x), str(path).encode()), "VecView (binary)")


def vec_load_binary(path, capacity):
    x = np.empty(int(capacity), np.float64)
    n = C.c_int(int(capacity))
    _chk(lib().ref_vec_load_binary(str(path).encode(), C.byref(n), _dp(x)), "VecLoad")
    return x[:n.value].copy()


def mat_setvaluesbatch(m, n, elem_rows, elem_mats, nl):
    """MatSetValuesBatch (matrix.c:1698) on an empty SeqAIJ through the reference's host path (one MatSetValues with
    ADD_VALUES per element) -> RefMat.  elem_rows: ne*nl ids, elem_mats: ne*nl*nl values."""
    rows = _i32(elem_rows).copy()
    vals = _f64(elem_mats)
    ne = rows.size // nl
    assert rows.size == ne * nl and vals.size == ne * nl * nl
    ok = (rows >= 0) & (rows < m)
    hint = np.minimum(np.bincount(rows[ok], minlength=m).astype(np.int64) * nl, n).astype(np.int32)
    h = C.c_void_p()
    _chk(lib().ref_mat_setvaluesbatch(int(m), int(n), int(ne), int(nl), _ip(rows), _dp(vals), _ip(hint), C.byref(h)), "MatSetValuesBatch")
    r = RefMat.__new__(RefMat)
    r.kind, r.h, r.m, r.n = 0, h, int(m), int(n)
    return r
