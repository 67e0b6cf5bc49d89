"""Vec kernels on the B200 through the C ABI against the reference's VECSEQ ops (oracle/_ref) on the same seeded
inputs.  Elementwise ops: BIT-EXACT.  Reductions: relative 1e-13 of sum|x_i y_i| (summation order differs)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = [0, 1, 2, 3, 31, 1000, 10007, 262144 + 5]


def _vecs(P, n, k):
    return [P.random_vector(n, 100 + s) for s in range(k)]


@pytest.mark.parametrize("n", SIZES)
def test_elementwise_bit_exact(gpu, R, P, n):
    pb = gpu
    xh, yh, wh = _vecs(P, n, 3)
    cases = [
        ("axpy", R.OP_AXPY, dict(a=0.37)), ("axpy0", R.OP_AXPY, dict(a=0.0)),
        ("aypx", R.OP_AYPX, dict(a=-0.81)), ("aypx0", R.OP_AYPX, dict(a=0.0)), ("aypx1", R.OP_AYPX, dict(a=1.0)),
        ("aypxm1", R.OP_AYPX, dict(a=-1.0)),
        ("waxpy", R.OP_WAXPY, dict(a=2.5)), ("waxpy1", R.OP_WAXPY, dict(a=1.0)), ("waxpym1", R.OP_WAXPY, dict(a=-1.0)),
        ("waxpy0", R.OP_WAXPY, dict(a=0.0)),
        ("axpby", R.OP_AXPBY, dict(a=0.3, b=-1.7)), ("axpby_a0", R.OP_AXPBY, dict(a=0.0, b=-1.7)),
        ("axpby_b1", R.OP_AXPBY, dict(a=0.3, b=1.0)), ("axpby_a1", R.OP_AXPBY, dict(a=1.0, b=0.4)),
        ("axpby_b0", R.OP_AXPBY, dict(a=0.3, b=0.0)),
        ("axpbypcz", R.OP_AXPBYPCZ, dict(a=0.3, b=-1.7, c=0.9)), ("axpbypcz_a1", R.OP_AXPBYPCZ, dict(a=1.0, b=-1.7, c=0.9)),
        ("axpbypcz_g1", R.OP_AXPBYPCZ, dict(a=0.3, b=-1.7, c=1.0)), ("axpbypcz_g0", R.OP_AXPBYPCZ, dict(a=0.3, b=-1.7, c=0.0)),
        ("scale", R.OP_SCALE, dict(a=1.0 / 3.0)), ("scale0", R.OP_SCALE, dict(a=0.0)), ("scale1", R.OP_SCALE, dict(a=1.0)),
        ("set", R.OP_SET, dict(a=-4.25)), ("copy", R.OP_COPY, {}), ("swap", R.OP_SWAP, {}),
        ("pmult", R.OP_PMULT, {}), ("pdiv", R.OP_PDIV, {}), ("recip", R.OP_RECIP, {}),
    ]
    for name, op, sc in cases:
        a, b, c = sc.get("a", 0.0), sc.get("b", 0.0), sc.get("c", 0.0)
        rx, ry, rw = xh.copy(), yh.copy(), wh.copy()
        if name == "recip" and n > 3:
            rx[3] = 0.0
        x, y, w = pb.Vec(n, rx), pb.Vec(n, ry), pb.Vec(n, rw)
        R.vec_elementwise(op, n, a, b, c, rx if n else None, ry if n else None, rw if n else None) if n else None
        if op == R.OP_AXPY: y.axpy(a, x)
        elif op == R.OP_AYPX: y.aypx(a, x)
        elif op == R.OP_WAXPY: w.waxpy(a, x, y)
        elif op == R.OP_AXPBY: y.axpby(a, b, x)
        elif op == R.OP_AXPBYPCZ: w.axpbypcz(a, b, c, x, y)
        elif op == R.OP_SCALE: x.scale(a)
        elif op == R.OP_SET: x.set(a)
        elif op == R.OP_COPY: x.copy(y)
        elif op == R.OP_SWAP: x.swap(y)
        elif op == R.OP_PMULT: w.pointwise_mult(x, y)
        elif op == R.OP_PDIV: w.pointwise_divide(x, y)
        elif op == R.OP_RECIP: x.reciprocal()
        for dv, hv in ((x, rx), (y, ry), (w, rw)):
            assert np.array_equal(dv.get_array(), hv), (name, n)


def test_elementwise_unaligned_pointers(gpu, R, P):
    """Pointers that are only 8-byte aligned take the scalar path and give the same bits."""
    n = 4099
    xh, yh = _vecs(P, n + 1, 2)
    X, Y = gpu.Vec(n + 1, xh), gpu.Vec(n + 1, yh)
    x, y = X.offset(1), Y.offset(1)
    y.axpy(0.77, x)
    ry = yh[1:].copy()
    R.vec_elementwise(R.OP_AXPY, n, 0.77, x=xh[1:].copy(), y=ry)
    assert np.array_equal(Y.get_array()[1:], ry)
    assert abs(x.dot(y) - R.vec_dot(xh[1:], ry)) <= 1e-13 * np.abs(xh[1:] * ry).sum()


@pytest.mark.parametrize("n", SIZES)
def test_reductions(gpu, R, P, n):
    xh, yh = _vecs(P, n, 2)
    x, y = gpu.Vec(n, xh), gpu.Vec(n, yh)
    tol = 1e-13
    assert abs(x.dot(y) - R.vec_dot(xh, yh)) <= tol * max(np.abs(xh * yh).sum(), 1e-300)
    assert abs(x.norm(gpu.NORM_2) - R.vec_norm(xh, R.NORM_2)) <= tol * max(np.linalg.norm(xh), 1e-300)
    assert abs(x.norm(gpu.NORM_1) - R.vec_norm(xh, R.NORM_1)) <= tol * max(np.abs(xh).sum(), 1e-300)
    assert x.norm(gpu.NORM_INFINITY) == R.vec_norm(xh, R.NORM_INFINITY)
    a12, r12 = x.norm(gpu.NORM_1_AND_2), R.vec_norm(xh, R.NORM_1_AND_2)
    assert abs(a12[0] - r12[0]) <= tol * max(r12[0], 1e-300) and abs(a12[1] - r12[1]) <= tol * max(r12[1], 1e-300)
    dn = x.dotnorm2(y)
    assert abs(dn[0] - np.dot(xh, yh)) <= 1e-12 * max(np.abs(xh * yh).sum(), 1e-300)
    assert abs(dn[1] - np.dot(yh, yh)) <= 1e-12 * max(np.dot(yh, yh), 1e-300)


def test_norm_nan_propagates(gpu, R, P):
    xh = P.random_vector(5000, 3)
    xh[1234] = np.nan
    x = gpu.Vec(5000, xh)
    # the kernel-level op (ops->norm, bvec2.c:619-631) propagates the NaN; the reference FRONT END then turns it into
    # PETSC_ERR_FP = 72 (rvector.c:214) -- which is what the oracle, driven through VecNorm(), reports
    assert np.isnan(x.norm(gpu.NORM_INFINITY))
    with pytest.raises(RuntimeError, match="PetscErrorCode 72"):
        R.vec_norm(xh, R.NORM_INFINITY)
    assert np.isnan(x.norm(gpu.NORM_2)) and np.isnan(x.dot(x))


@pytest.mark.parametrize("nv", [0, 1, 2, 3, 4, 5, 7, 8, 15, 16, 17, 30, 31, 33])
def test_mdot_maxpy(gpu, R, P, nv):
    n = 50021
    xh, yh = _vecs(P, n, 2)
    vs = [P.random_vector(n, 500 + k) for k in range(nv)]
    al = P.random_vector(max(nv, 1), 77)[:nv]
    x, y = gpu.Vec(n, xh), gpu.Vec(n, yh)
    dv = [gpu.Vec(n, v) for v in vs]
    z = x.mdot(dv)
    zr = R.vec_mdot(xh, vs)
    for k in range(nv):
        assert abs(z[k] - zr[k]) <= 1e-13 * np.abs(xh * vs[k]).sum()
    y.maxpy(al, dv)
    assert np.array_equal(y.get_array(), R.vec_maxpy(yh, al, vs))      # bit-exact: same grouping as dvec2.c:836


def test_fixtures(gpu, P, fixtures):
    """Against the committed outputs of the reference library (no oracle .so needed)."""
    n = int(fixtures["vec_n"])
    xh, yh, wh = (P.random_vector(n, s) for s in (11, 12, 13))
    x, y, w = gpu.Vec(n, xh), gpu.Vec(n, yh), gpu.Vec(n, wh)
    assert abs(x.dot(y) - float(fixtures["vec_dot"])) <= 1e-13 * np.abs(xh * yh).sum()
    assert abs(x.norm() - float(fixtures["vec_norm2"])) <= 1e-13 * float(fixtures["vec_norm2"])
    assert x.norm(gpu.NORM_INFINITY) == float(fixtures["vec_norminf"])
    ys = [P.random_vector(n, 20 + k) for k in range(7)]
    dv = [gpu.Vec(n, v) for v in ys]
    assert np.allclose(x.mdot(dv), fixtures["vec_mdot7"], rtol=0, atol=1e-13 * n)
    y2 = gpu.Vec(n, yh); y2.maxpy(P.random_vector(7, 99), dv)
    assert np.array_equal(y2.get_array(), fixtures["vec_maxpy7"])
    y3 = gpu.Vec(n, yh); y3.axpy(0.37, x); assert np.array_equal(y3.get_array(), fixtures["vec_axpy"])
    y4 = gpu.Vec(n, yh); y4.aypx(-0.81, x); assert np.array_equal(y4.get_array(), fixtures["vec_aypx"])
    w.pointwise_mult(x, y); assert np.array_equal(w.get_array(), fixtures["vec_pmult"])


def test_large_vector_properties(gpu, P):
    """BASELINE size (256^3 rows): size-independent properties — linearity of dot, norm^2 == dot(x,x), axpy inverse."""
    n = 256 ** 3
    x, y = gpu.Vec(n), gpu.Vec(n)
    x.set(0.5); y.set(-2.0)
    assert x.dot(y) == -float(n)                        # exactly representable
    assert x.norm(gpu.NORM_1) == 0.5 * n and x.norm(gpu.NORM_INFINITY) == 0.5
    assert abs(x.norm() - 0.5 * np.sqrt(n)) <= 1e-12 * np.sqrt(n)
    y.axpy(4.0, x)                                      # -2 + 4*0.5 = 0
    assert y.norm(gpu.NORM_INFINITY) == 0.0


def test_reductions_return_when_the_input_carries_the_poll_stamp(gpu):
    """b200_reduce_finish polls the mapped result slots for a change from a signalling-NaN stamp (b200_vec.cu): input that
    holds the very same bit pattern must still come back (as NaN), and the slots must be usable afterwards."""
    n = 4099
    h = np.full(n, 0.25)
    h.view(np.uint64)[17] = 0x7FF4B200DEAD5EED
    x, y = gpu.Vec(n, h), gpu.Vec(n)
    y.set(1.0)
    for t in (gpu.NORM_1, gpu.NORM_2, gpu.NORM_INFINITY):
        assert np.isnan(x.norm(t))                          # bvec2.c:626: a NaN entry makes every norm NaN
    assert all(np.isnan(v) for v in x.norm(gpu.NORM_1_AND_2))
    assert np.isnan(x.dot(y)) and np.isnan(x.mdot([y, x, y])).all()
    assert y.norm(gpu.NORM_1) == float(n) and y.dot(y) == float(n)
    assert np.array_equal(y.mdot([y] * 20), np.full(20, float(n)))      # two launches, 20 slots
