"""petsc-dev_b200 — host-side mirror of the reference's Mat / Vec / KSP interface for the sparse-solve hot path,
bound to libpetscb200.so (include/petscb200.h) through ctypes.

The directory name is not a Python identifier; import it with
    import importlib; pb = importlib.import_module("petsc-dev_b200")

There is NO CPU fallback: if the CUDA library is missing this module raises at import, and every compute call
raises PetscB200Error when no B200 is present.  Nothing here imports oracle/.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libpetscb200.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "petscb200.h")

if not os.path.exists(LIB_PATH):
    raise ImportError("%s is not built: run `python -c 'import __graft_entry__ as g; g.build()'` or "
                      "`make -C petsc-dev_b200/csrc` (there is no CPU fallback)" % LIB_PATH)
lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
lib.b200_last_error.restype = C.c_char_p

NORM_1, NORM_2, NORM_FROBENIUS, NORM_INFINITY, NORM_1_AND_2 = 0, 1, 2, 3, 4
KSP_NORM_NONE, KSP_NORM_PRECONDITIONED, KSP_NORM_UNPRECONDITIONED, KSP_NORM_NATURAL = 0, 1, 2, 3
KSPCG, KSPGMRES = 0, 1
PCNONE, PCJACOBI = 0, 1

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


class PetscB200Error(RuntimeError):
    def __init__(self, code, what):
        self.code = code
        RuntimeError.__init__(self, "%s: PetscErrorCode %d: %s" % (what, code, lib.b200_last_error().decode()))


def _chk(rc, what):
    if rc:
        raise PetscB200Error(rc, what)


def _ip(a):
    return a.ctypes.data_as(c_ip) if a is not None else None


def _dp(a):
    return a.ctypes.data_as(c_dp) if a is not None else None


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def init(device=-1):
    _chk(lib.b200_init(int(device)), "b200_init")


def finalize():
    _chk(lib.b200_finalize(), "b200_finalize")


def synchronize():
    _chk(lib.b200_synchronize(), "b200_synchronize")


class Stream:
    """A caller-owned CUDA stream; `with stream:` makes the library run on it (b200_set_stream)."""

    def __init__(self):
        p = C.c_void_p()
        _chk(lib.b200_stream_create(C.byref(p)), "cudaStreamCreate")
        self.p = p

    def use(self):
        _chk(lib.b200_set_stream(self.p), "b200_set_stream")

    def synchronize(self):
        _chk(lib.b200_stream_synchronize(self.p), "cudaStreamSynchronize")

    def destroy(self):
        if self.p:
            lib.b200_stream_destroy(self.p)
            self.p = None


def use_default_stream():
    _chk(lib.b200_set_stream(None), "b200_set_stream")


def launch_count():
    n = C.c_longlong()
    lib.b200_launch_count(C.byref(n))
    return n.value


def timer_start():
    _chk(lib.b200_timer_start(), "timer_start")


def timer_stop():
    ms = C.c_double()
    _chk(lib.b200_timer_stop(C.byref(ms)), "timer_stop")
    return ms.value


def flush_l2():
    _chk(lib.b200_flush_l2(), "flush_l2")


class Vec:
    """A device-resident vector of n doubles (VECSEQB200 / the local part of VECMPIB200)."""

    def __init__(self, n, host=None, ptr=None):
        self.n = int(n)
        self._own = ptr is None
        if ptr is None:
            p = C.c_void_p()
            _chk(lib.b200_malloc(C.c_size_t(8 * max(self.n, 2)), C.byref(p)), "VecCreate")
            self.ptr = p.value
            if host is not None:
                self.set_array(host)
            else:
                self.set(0.0)
        else:
            self.ptr = ptr

    # --- host <-> device (VecCUSPCopyToGPU / FromGPU) ---
    def set_array(self, host):
        h = _f64(host)
        if h.size != self.n:
            raise PetscB200Error(60, "VecSetValues size")
        _chk(lib.b200_memcpy_h2d(C.c_void_p(self.ptr), h.ctypes.data_as(C.c_void_p), C.c_size_t(8 * self.n)), "H2D")
        return self

    def set_array_async(self, pinned):
        """Stream-ordered H2D from a pinned numpy array (pinned_array); returns at once."""
        _chk(lib.b200_memcpy_h2d_async(C.c_void_p(self.ptr), pinned.ctypes.data_as(C.c_void_p), C.c_size_t(8 * self.n)), "H2D async")

    def get_array_async(self, pinned):
        _chk(lib.b200_memcpy_d2h_async(pinned.ctypes.data_as(C.c_void_p), C.c_void_p(self.ptr), C.c_size_t(8 * self.n)), "D2H async")

    def get_array(self, out=None):
        out = np.empty(self.n, np.float64) if out is None else out
        _chk(lib.b200_memcpy_d2h(out.ctypes.data_as(C.c_void_p), C.c_void_p(self.ptr), C.c_size_t(8 * self.n)), "D2H")
        return out

    @property
    def p(self):
        return C.cast(C.c_void_p(self.ptr), c_dp)

    def offset(self, k):
        """A non-owning view starting at element k (tests use it for unaligned pointers)."""
        return Vec(self.n - k, ptr=self.ptr + 8 * k)

    def duplicate(self):
        return Vec(self.n)

    def destroy(self):
        if self._own and self.ptr:
            lib.b200_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # --- the reference's Vec API names ---
    def set(self, alpha):
        _chk(lib.b200_vec_set(self.n, self.p, C.c_double(alpha)), "VecSet")

    def copy(self, y):                      # VecCopy(x, y)
        _chk(lib.b200_vec_copy(self.n, self.p, y.p), "VecCopy")

    def swap(self, y):
        _chk(lib.b200_vec_swap(self.n, self.p, y.p), "VecSwap")

    def scale(self, alpha):
        _chk(lib.b200_vec_scale(self.n, self.p, C.c_double(alpha)), "VecScale")

    def axpy(self, alpha, x):               # VecAXPY(y, alpha, x)
        _chk(lib.b200_vec_axpy(self.n, self.p, C.c_double(alpha), x.p), "VecAXPY")

    def aypx(self, alpha, x):               # VecAYPX(y, alpha, x)
        _chk(lib.b200_vec_aypx(self.n, self.p, C.c_double(alpha), x.p), "VecAYPX")

    def axpby(self, alpha, beta, x):        # VecAXPBY(y, alpha, beta, x)
        _chk(lib.b200_vec_axpby(self.n, self.p, C.c_double(alpha), C.c_double(beta), x.p), "VecAXPBY")

    def waxpy(self, alpha, x, y):           # VecWAXPY(w, alpha, x, y)
        _chk(lib.b200_vec_waxpy(self.n, self.p, C.c_double(alpha), x.p, y.p), "VecWAXPY")

    def axpbypcz(self, alpha, beta, gamma, x, y):   # VecAXPBYPCZ(z, alpha, beta, gamma, x, y)
        _chk(lib.b200_vec_axpbypcz(self.n, self.p, C.c_double(alpha), C.c_double(beta), C.c_double(gamma), x.p, y.p),
             "VecAXPBYPCZ")

    def pointwise_mult(self, x, y):         # VecPointwiseMult(w, x, y)
        _chk(lib.b200_vec_pointwise_mult(self.n, self.p, x.p, y.p), "VecPointwiseMult")

    def pointwise_divide(self, x, y):
        _chk(lib.b200_vec_pointwise_divide(self.n, self.p, x.p, y.p), "VecPointwiseDivide")

    def reciprocal(self):
        _chk(lib.b200_vec_reciprocal(self.n, self.p), "VecReciprocal")

    def maxpy(self, alphas, xs):            # VecMAXPY(y, nv, alpha, x[])
        nv = len(xs)
        al = _f64(alphas) if nv else np.zeros(1)
        tab = (c_dp * max(nv, 1))(*[x.p for x in xs])
        _chk(lib.b200_vec_maxpy(self.n, self.p, nv, _dp(al), tab), "VecMAXPY")

    def dot(self, y, local=False):          # VecDot(x, y)
        r = C.c_double()
        f = lib.b200_vec_dot_local if local else lib.b200_vec_dot
        _chk(f(self.n, self.p, y.p, C.byref(r)), "VecDot")
        return r.value

    tdot = dot

    def norm(self, ntype=NORM_2, local=False):
        r = (C.c_double * 2)()
        f = lib.b200_vec_norm_local if local else lib.b200_vec_norm
        _chk(f(self.n, self.p, int(ntype), r), "VecNorm")
        return (r[0], r[1]) if ntype == NORM_1_AND_2 else r[0]

    def mdot(self, ys, local=False):        # VecMDot(x, nv, y[])
        nv = len(ys)
        z = np.zeros(max(nv, 1))
        tab = (c_dp * max(nv, 1))(*[y.p for y in ys])
        f = lib.b200_vec_mdot_local if local else lib.b200_vec_mdot
        _chk(f(self.n, self.p, nv, tab, _dp(z)), "VecMDot")
        return z[:nv]

    def mdot_noalloc(self, ys, z_dev, stream=None):
        """b200_vec_mdot_noalloc: VecMDot with static device scratch; z_dev is a device Vec of >= len(ys) entries.
        Queued on the library's stream (or the cudaStream_t given); nothing is read back here."""
        nv = len(ys)
        tab = (c_dp * max(nv, 1))(*[y.p for y in ys])
        _chk(lib.b200_vec_mdot_noalloc(self.n, self.p, nv, tab, z_dev.p, C.c_void_p(stream)), "VecMDot (no allocation)")

    def dotnorm2(self, t):                  # VecDotNorm2(s, t) -> (s.t, t.t)
        a, b = C.c_double(), C.c_double()
        _chk(lib.b200_vec_dotnorm2(self.n, self.p, t.p, C.byref(a), C.byref(b)), "VecDotNorm2")
        return a.value, b.value


class KSPResult(C.Structure):
    _fields_ = [("its", C.c_int), ("reason", C.c_int), ("nhist", C.c_int), ("rnorm", C.c_double),
                ("solve_ms", C.c_double)]


class KSPOptions(C.Structure):
    _fields_ = [("ksp_type", C.c_int), ("pc_type", C.c_int), ("rtol", C.c_double), ("abstol", C.c_double),
                ("dtol", C.c_double), ("max_it", C.c_int), ("norm_type", C.c_int), ("guess_nonzero", C.c_int),
                ("gmres_restart", C.c_int), ("gmres_cgs_refine", C.c_int), ("gmres_haptol", C.c_double),
                ("fused", C.c_int), ("check_every", C.c_int), ("gmres_modified_gs", C.c_int),
                ("pc_jacobi_inverse_diag", C.c_void_p)]


class Mat:
    """An assembled AIJ matrix on the device: MATSEQAIJB200 (from_csr) or MATMPIAIJB200 (parallel.create_mpiaij)."""

    def __init__(self, handle, m, n):
        self.h = handle
        self.m, self.n = int(m), int(n)

    @classmethod
    def from_csr(cls, m, n, ai, aj, aa):
        ai, aj, aa = _i32(ai), _i32(aj), _f64(aa)
        h = C.c_void_p()
        _chk(lib.b200_mat_create_seqaij(int(m), int(n), _ip(ai), _ip(aj), _dp(aa), C.byref(h)), "MatAssemblyEnd (upload)")
        return cls(h, m, n)

    def update_values(self, aa):
        aa = _f64(aa)
        _chk(lib.b200_mat_update_values(self.h, _dp(aa)), "MatAssemblyEnd (values)")

    def update_values_mpiaij(self, a_diag, a_offdiag):
        """Parallel type: new values of both blocks, same pattern and exchange plan (nothing collective)."""
        ad, ao = _f64(a_diag), _f64(a_offdiag)
        _chk(lib.b200_mat_update_values_mpiaij(self.h, _dp(ad), _dp(ao)), "MatAssemblyEnd_MPIAIJ (values)")

    @classmethod
    def set_values_batch(cls, m, n, nl, elem_rows, elem_mats, keep_plan=False):
        """MatSetValuesBatch (matrix.c:1698) on the device: ne element matrices of nl x nl values added into the rows and
        columns elem_rows[e*nl:(e+1)*nl] of a new m x n matrix.  Host arrays in, device matrix out."""
        rows, vals = _i32(elem_rows), _f64(elem_mats)
        ne = rows.size // int(nl)
        if rows.size != ne * nl or vals.size != ne * nl * nl:
            raise ValueError("elem_rows must hold ne*nl ids and elem_mats ne*nl*nl values")
        h = C.c_void_p()
        _chk(lib.b200_mat_setvaluesbatch(int(m), int(n), C.c_longlong(ne), int(nl), _ip(rows), _dp(vals), 0, int(bool(keep_plan)),
                                         C.byref(h)), "MatSetValuesBatch")
        return cls(h, m, n)

    @classmethod
    def set_values_batch_device(cls, m, n, ne, nl, elem_rows_dev, elem_mats_dev, keep_plan=False):
        """Same with the element arrays already on the device (raw device pointers: int32 ids, float64 values)."""
        h = C.c_void_p()
        _chk(lib.b200_mat_setvaluesbatch(int(m), int(n), C.c_longlong(int(ne)), int(nl), C.c_void_p(int(elem_rows_dev)),
                                         C.c_void_p(int(elem_mats_dev)), 1, int(bool(keep_plan)), C.byref(h)), "MatSetValuesBatch")
        return cls(h, m, n)

    def batch_values(self, elem_mats):
        """Same elem_rows, new element matrices: refill the values through the kept plan (host array)."""
        vals = _f64(elem_mats)
        _chk(lib.b200_mat_setvaluesbatch_values(self.h, _dp(vals), 0), "MatSetValuesBatch (values)")

    def batch_values_device(self, elem_mats_dev):
        _chk(lib.b200_mat_setvaluesbatch_values(self.h, C.c_void_p(int(elem_mats_dev)), 1), "MatSetValuesBatch (values)")

    def csr(self):
        """Device CSR copied back to the host: (i, j, a)."""
        nz = self.nnz()
        i, j, a = np.empty(self.m + 1, np.int32), np.empty(nz, np.int32), np.empty(nz, np.float64)
        _chk(lib.b200_mat_get_csr(self.h, _ip(i), _ip(j), _dp(a)), "MatCUSPCopyFromGPU")
        return i, j, a

    def destroy(self):
        if self.h:
            lib.b200_mat_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def get_vecs(self):                     # MatGetVecs -> (right, left)
        return Vec(self.n), Vec(self.m)

    def mult(self, x, y):                   # MatMult(A, x, y)
        _chk(lib.b200_mat_mult(self.h, x.p, y.p), "MatMult")

    def mult_add(self, x, y, z):            # MatMultAdd(A, x, y, z)
        _chk(lib.b200_mat_multadd(self.h, x.p, y.p, z.p), "MatMultAdd")

    def mult_transpose(self, x, y):         # MatMultTranspose(A, x, y)
        _chk(lib.b200_mat_multtranspose(self.h, x.p, y.p), "MatMultTranspose")

    def mult_transpose_add(self, x, y, z):  # MatMultTransposeAdd(A, x, y, z)
        _chk(lib.b200_mat_multtransposeadd(self.h, x.p, y.p, z.p), "MatMultTransposeAdd")

    def mult_dot(self, x, y):
        r = C.c_double()
        _chk(lib.b200_mat_mult_dot(self.h, x.p, y.p, C.byref(r)), "MatMult+VecDot")
        return r.value

    def get_diagonal(self, d):
        _chk(lib.b200_mat_get_diagonal(self.h, d.p), "MatGetDiagonal")

    def flops(self):
        f = C.c_double()
        _chk(lib.b200_mat_flops(self.h, C.byref(f)), "flops")
        return f.value

    def bytes(self):
        f = C.c_double()
        _chk(lib.b200_mat_bytes(self.h, C.byref(f)), "bytes")
        return f.value

    def nnz(self):
        nz = C.c_longlong()
        _chk(lib.b200_mat_get_size(self.h, None, None, C.byref(nz)), "MatGetInfo")
        return nz.value

    def strips(self):
        a, b = C.c_int(), C.c_int()
        _chk(lib.b200_mat_strips(self.h, C.byref(a), C.byref(b)), "strips")
        return a.value, b.value

    def bin_info(self):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        _chk(lib.b200_mat_bin_info(self.h, C.byref(a), C.byref(b), C.byref(c)), "bin_info")
        g, lt, gr, sh, pk = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _chk(lib.b200_mat_bin_info2(self.h, C.byref(g), C.byref(lt), C.byref(gr), C.byref(sh), C.byref(pk)), "bin_info2")
        return dict(ntiles=a.value, nlong=b.value, compressed=bool(c.value), lanes_per_row=g.value,
                    long_thresh=lt.value, grid=gr.value, shape=sh.value, park=bool(pk.value))


def spmv_tuning(shape=-1, lanes_per_row=-1, long_thresh=-1, park=-1):
    """Row-binning policy for matrices created after the call (-1 = automatic); lanes_per_row=1 makes every tile row
    a left-to-right sum, bit-identical to MatMult_SeqAIJ."""
    _chk(lib.b200_spmv_set_tuning(int(shape), int(lanes_per_row), int(long_thresh), int(park)), "spmv_tuning")


def spmv_strips(nstrips=-1):
    """Column strips for matrices created after the call: -1 automatic, 0 off, k forced (cache blocking of x)."""
    _chk(lib.b200_spmv_set_strips(int(nstrips)), "spmv_strips")


def ksp_options(**kw):
    o = KSPOptions()
    _chk(lib.b200_ksp_default_options(C.byref(o)), "KSPCreate")
    names = {"cg": KSPCG, "gmres": KSPGMRES, "none": PCNONE, "jacobi": PCJACOBI}
    for k, v in kw.items():
        if isinstance(v, str):
            v = names[v]
        setattr(o, k, v)
    return o


def ksp_solve(A, b, x, histmax=None, **kw):
    """KSPSolve(ksp, b, x): returns dict(its, reason, history, rnorm, solve_ms).  Keyword options are the fields of
    b200_ksp_options (ksp_type='cg'|'gmres', pc_type='jacobi'|'none', rtol, abstol, dtol, max_it, norm_type,
    guess_nonzero, gmres_restart, gmres_cgs_refine, gmres_modified_gs, fused, check_every)."""
    o = ksp_options(**kw)
    histmax = (o.max_it + 2) if histmax is None else int(histmax)
    hist = np.zeros(max(histmax, 1))
    res = KSPResult()
    _chk(lib.b200_ksp_solve(A.h, C.byref(o), b.p, x.p, _dp(hist), histmax, C.byref(res)), "KSPSolve")
    return dict(its=res.its, reason=res.reason, history=hist[:res.nhist].copy(), rnorm=res.rnorm, solve_ms=res.solve_ms)


def pinned_array(n):
    """n doubles of page-locked host memory as a numpy array (for the end-to-end path's H2D/D2H copies)."""
    p = C.c_void_p()
    _chk(lib.b200_host_alloc(C.c_size_t(8 * max(int(n), 1)), C.byref(p)), "cudaHostAlloc")
    return np.ctypeslib.as_array(C.cast(p, c_dp), shape=(int(n),))
