"""AIJ MatMult / MatMultAdd / MatGetDiagonal on the B200 against the reference's MatMult_SeqAIJ (oracle/_ref).
Tile rows handled by ONE lane (lanes_per_row == 1: the automatic choice for short rows such as the 5/7-point stencils,
or forced with spmv_tuning) are summed left to right like aij.h:384 -> BIT-EXACT; rows shared by several lanes and long
rows use a fixed tree -> |dy_i| <= 1e-12 * sum_j |a_ij x_j| (the tolerance BASELINE.json states)."""
import numpy as np
import pytest

from conftest import ex5_matrix

pytestmark = pytest.mark.gpu


def _check(gpu, R, P, m, n, ip, j, a, seed=2, exact=True, lanes=-1, park=-1, long_thresh=-1):
    gpu.spmv_tuning(lanes_per_row=lanes, park=park, long_thresh=long_thresh)
    try:
        A = gpu.Mat.from_csr(m, n, ip, j, a)
    finally:
        gpu.spmv_tuning()
    info = A.bin_info()
    if lanes > 0:
        assert info["lanes_per_row"] == lanes
    exact = exact and info["lanes_per_row"] == 1 and info["nlong"] == 0
    ref = R.RefMat(m, n, ip, j, a, no_inode=True)
    xh, yh = P.random_vector(n, seed), P.random_vector(m, seed + 1)
    x, y, z = gpu.Vec(n, xh), gpu.Vec(m, yh), gpu.Vec(m)
    A.mult(x, z)
    zr = ref.mult(xh)
    rows = np.repeat(np.arange(m), np.diff(ip))
    scale = np.bincount(rows, weights=np.abs(a * xh[j]), minlength=m) if len(a) else np.zeros(m)
    got = z.get_array()
    if exact:
        assert np.array_equal(got, zr)
    assert np.all(np.abs(got - zr) <= 1e-12 * scale)
    A.mult_add(x, y, z)
    zr2 = ref.multadd(xh, yh)
    got = z.get_array()
    if exact:
        assert np.array_equal(got, zr2)
    assert np.all(np.abs(got - zr2) <= 1e-12 * (scale + np.abs(yh)))
    A.mult_add(x, y, y)                                   # z aliases y (aij.c:1306)
    assert np.array_equal(y.get_array(), got)
    if m == n:
        d = gpu.Vec(m)
        A.get_diagonal(d)
        assert np.array_equal(d.get_array(), ref.diagonal())
        d.set_array(np.full(m, -7.0))
        A.get_diagonal(d)                                 # second call: gather through the recorded positions
        assert np.array_equal(d.get_array(), ref.diagonal())
    assert A.flops() == 2.0 * len(a) - np.count_nonzero(np.diff(ip))
    return A


def test_ex5_golden(gpu, goldens):
    m, n, ip, j, a = ex5_matrix()
    A = gpu.Mat.from_csr(m, n, ip, j, a)
    x, y = gpu.Vec(n, np.arange(n, dtype=np.float64)), gpu.Vec(m)
    A.mult(x, y)
    assert [float("%G" % v) for v in y.get_array()] == goldens["mat_ex5_11_A"]["matmult"]
    sq = [k for k in range(ip[-1]) if j[k] < m]
    S = gpu.Mat.from_csr(m, m, np.arange(0, (m + 1) * m, m, dtype=np.int32), j[sq], a[sq])
    d = gpu.Vec(m); S.get_diagonal(d)
    assert [float("%G" % v) for v in d.get_array()] == goldens["mat_ex5_11_A"]["diagonal"]


@pytest.mark.parametrize("case", ["lap2d_ex2", "lap7", "s27", "bs3", "rect", "empty_rows", "one_row", "no_rows", "all_empty"])
def test_structured_bit_exact(gpu, R, P, case):
    if case == "lap2d_ex2":
        ip, j, a = P.laplace2d_5pt(57, 43); m = n = 57 * 43
    elif case == "lap7":
        ip, j, a = P.laplace3d_7pt(21); m = n = 21 ** 3
    elif case == "s27":
        ip, j, a = P.stencil3d_27pt(17); m = n = 17 ** 3
    elif case == "bs3":
        ip, j, a = P.elasticity_bs3(9); m = n = 3 * 9 ** 3
    elif case == "rect":
        m, n, ip, j, a = ex5_matrix()
    elif case == "empty_rows":                            # ragged with many empty rows but below the compression ratio
        ip0, j, a = P.laplace2d_5pt(30, 30)
        keep = (np.arange(900) % 3) != 1
        cnt = np.diff(ip0) * keep
        sel = np.repeat(keep, np.diff(ip0))
        ip, j, a, m, n = np.concatenate([[0], np.cumsum(cnt)]), j[sel], a[sel], 900, 900
    elif case == "one_row":
        ip, j, a, m, n = np.array([0, 3]), np.array([0, 2, 5], np.int32), np.array([1.5, -2.0, 0.25]), 1, 6
    elif case == "no_rows":
        ip, j, a, m, n = np.array([0]), np.zeros(0, np.int32), np.zeros(0), 0, 5
    elif case == "all_empty":
        ip, j, a, m, n = np.zeros(11, np.int64), np.zeros(0, np.int32), np.zeros(0), 10, 10
    ip, j = np.asarray(ip, np.int64), np.asarray(j, np.int32)
    A = _check(gpu, R, P, m, n, ip, j, a)                       # automatic lanes per row (tolerance when > 1)
    if case in ("lap2d_ex2", "lap7", "rect", "empty_rows"):
        assert A.bin_info()["lanes_per_row"] == 1               # short rows: reference order by default
    _check(gpu, R, P, m, n, ip, j, a, lanes=1)                  # one lane per row: bit-exact for every pattern
    _check(gpu, R, P, m, n, ip, j, a, lanes=1, park=1)          # ... also when the products are parked first
    for lanes in (2, 8, 32):
        _check(gpu, R, P, m, n, ip, j, a, lanes=lanes, park=lanes == 8)
    if len(a) > 64:                                             # every row becomes a chain of chunk tiles
        B = _check(gpu, R, P, m, n, ip, j, a, long_thresh=2)
        assert B.bin_info()["nlong"] > 0


def test_compressed_rows_offdiag_like(gpu, R, P):
    """A block that is mostly empty rows (the MPIAIJ off-diagonal block) is visited through the non-empty-row list."""
    m, n = 5000, 300
    rng = np.random.default_rng(5)
    rows = np.sort(rng.choice(m, 400, replace=False))
    cnt = np.zeros(m, np.int64); cnt[rows] = rng.integers(1, 9, rows.size)
    ip = np.concatenate([[0], np.cumsum(cnt)])
    j = np.concatenate([np.sort(rng.choice(n, c, replace=False)) for c in cnt[rows]]).astype(np.int32)
    a = P.random_vector(j.size, 8)
    A = _check(gpu, R, P, m, n, ip, j, a)
    assert A.bin_info()["compressed"]


def test_irregular_powerlaw(gpu, R, P):
    """C4-like: power-law row lengths incl. rows far above the long-row threshold."""
    N = 20000
    ip, j, a = P.powerlaw(N, seed=4, xm=17.0, lmax=5000)
    assert np.diff(ip).max() > 1000
    A = _check(gpu, R, P, N, N, ip, j, a, exact=False)
    info = A.bin_info()
    assert info["nlong"] > 0 and info["ntiles"] > 0 and info["park"]
    _check(gpu, R, P, N, N, ip, j, a, exact=False, park=0)
    # with one lane per row, rows at or below the long-row threshold are still bit-exact
    A = _check(gpu, R, P, N, N, ip, j, a, exact=False, lanes=1)
    ref = R.RefMat(N, N, ip, j, a)
    xh = P.random_vector(N, 2)
    x, y = gpu.Vec(N, xh), gpu.Vec(N)
    A.mult(x, y)
    short = np.diff(ip) <= A.bin_info()["long_thresh"]
    assert np.array_equal(y.get_array()[short], ref.mult(xh)[short])


def test_deterministic_run_to_run(gpu, P):
    ip, j, a = P.powerlaw(20000, seed=4, xm=17.0, lmax=5000)
    A = gpu.Mat.from_csr(20000, 20000, ip, j, a)
    x, y1, y2 = gpu.Vec(20000, P.random_vector(20000, 2)), gpu.Vec(20000), gpu.Vec(20000)
    A.mult(x, y1); A.mult(x, y2)
    assert np.array_equal(y1.get_array(), y2.get_array())
    assert A.mult_dot(x, y1) == A.mult_dot(x, y2)


def test_mult_dot_fused(gpu, R, P):
    ip, j, a = P.laplace3d_7pt(30)
    n = 27000
    A = gpu.Mat.from_csr(n, n, ip, j, a)
    xh = P.random_vector(n, 2)
    x, y = gpu.Vec(n, xh), gpu.Vec(n)
    d = A.mult_dot(x, y)
    yr = R.RefMat(n, n, ip, j, a).mult(xh)
    assert np.array_equal(y.get_array(), yr)
    assert abs(d - R.vec_dot(xh, yr)) <= 1e-13 * np.abs(xh * yr).sum()


def test_update_values_and_errors(gpu, R, P):
    ip, j, a = P.laplace2d_5pt(20, 20)
    A = gpu.Mat.from_csr(400, 400, ip, j, a)
    A.update_values(2.0 * a)
    xh = P.random_vector(400, 2)
    x, y = gpu.Vec(400, xh), gpu.Vec(400)
    A.mult(x, y)
    assert np.array_equal(y.get_array(), R.RefMat(400, 400, ip, j, 2.0 * a).mult(xh))
    with pytest.raises(gpu.PetscB200Error) as e:
        A.mult(x, x)                                      # matrix.c:2143 "x and y must be different vectors"
    assert e.value.code == 61


def test_fixtures(gpu, P, fixtures):
    ip, j, a = P.laplace3d_7pt(12)
    N = 12 ** 3
    A = gpu.Mat.from_csr(N, N, ip, j, a)
    x, y, z = gpu.Vec(N, P.random_vector(N, 2)), gpu.Vec(N, P.random_vector(N, 3)), gpu.Vec(N)
    A.mult(x, z); assert np.array_equal(z.get_array(), fixtures["lap7_12_mult"])
    A.mult_add(x, y, z); assert np.array_equal(z.get_array(), fixtures["lap7_12_multadd"])
    ip, j, a = P.powerlaw(3000, seed=4, xm=5.0, lmax=800)
    A = gpu.Mat.from_csr(3000, 3000, ip, j, a)
    x, z = gpu.Vec(3000, P.random_vector(3000, 5)), gpu.Vec(3000)
    A.mult(x, z)
    ref = fixtures["plaw3000_mult"]
    assert np.all(np.abs(z.get_array() - ref) <= 1e-12 * np.maximum(np.abs(ref), 1.0))


def test_full_size_c2_properties(gpu, P):
    """BASELINE config 2 (7-point Laplacian on 256^3): size-independent properties.  A*1 is the exact row sum
    (0 in the interior, 1..3 on faces/edges/corners), A is symmetric (x.Ay == y.Ax up to rounding), linearity."""
    nx = 256
    N = nx ** 3
    ip, j, a = P.laplace3d_7pt(nx)
    A = gpu.Mat.from_csr(N, N, ip, j, a)
    one, y = gpu.Vec(N), gpu.Vec(N)
    one.set(1.0)
    A.mult(one, y)
    rs = y.get_array()
    g = np.arange(nx)
    b1 = ((g == 0) | (g == nx - 1)).astype(np.float64)
    expect = (b1[None, None, :] + b1[None, :, None] + b1[:, None, None]).reshape(-1)
    assert np.array_equal(rs, expect)
    xh, wh = P.random_vector(N, 2), P.random_vector(N, 3)
    x, w, ax, aw = gpu.Vec(N, xh), gpu.Vec(N, wh), gpu.Vec(N), gpu.Vec(N)
    A.mult(x, ax); A.mult(w, aw)
    s1, s2 = w.dot(ax), x.dot(aw)
    assert abs(s1 - s2) <= 1e-11 * N
    x.axpy(2.0, w)                                         # A(x + 2w) == Ax + 2Aw up to rounding
    A.mult(x, y)
    ax.axpy(2.0, aw); ax.axpy(-1.0, y)
    assert ax.norm(gpu.NORM_INFINITY) <= 1e-12 * 12 * 3


@pytest.mark.parametrize("strips", [2, 3, 7])
@pytest.mark.parametrize("name", ["lap7", "plaw", "lap2d_rect"])
def test_column_strips(gpu, R, P, name, strips, monkeypatch):
    """Column-strip storage (cache blocking of x for matrices with scattered columns): y = S_0 x; y += S_1 x; ... With one
    lane per row the row sum is still ONE left-to-right chain -> bit-identical to MatMult_SeqAIJ; the diagonal is
    collected from the strip that owns it; KSP on a stripped operator follows the reference history.  The strips are cut
    on row ranges by several host threads (forced here on a small matrix)."""
    monkeypatch.setenv("B200_HOST_THREADS", "3" if strips % 2 else "5")
    if name == "lap7":
        m = n = 12 ** 3
        ip, j, a = P.laplace3d_7pt(12)
        lanes = 1
    elif name == "plaw":
        m = n = 3000
        ip, j, a = P.powerlaw(3000, seed=4, xm=4.0, lmax=500)
        lanes = -1
    else:                                   # rectangular: 600 rows of the 30 x 40 grid operator
        ip, j, a = P.laplace2d_5pt(30, 40, 0, 600)
        m, n = 600, 1200
        lanes = 1
    gpu.spmv_strips(strips)
    try:
        A = _check(gpu, R, P, m, n, ip, j, a, lanes=lanes, exact=(lanes == 1))
    finally:
        gpu.spmv_strips(-1)
    assert A.strips()[0] == strips
    if name == "lap7":
        ref = R.RefMat(m, n, ip, j, a, no_inode=True)
        bh = ref.mult(np.ones(n))
        b, sol = gpu.Vec(m, bh), gpu.Vec(m)
        for fused in (0, 1):
            r = gpu.ksp_solve(A, b, sol, ksp_type="cg", pc_type="jacobi", rtol=1e-8, fused=fused)
            rr = ref.ksp_solve(bh, "cg", "jacobi", rtol=1e-8)
            assert abs(r["its"] - rr["its"]) <= 1
            k = min(len(r["history"]), len(rr["history"]))
            assert np.max(np.abs(r["history"][:k] - rr["history"][:k]) / rr["history"][:k]) < 1e-10


def test_strips_automatic_choice(gpu, P):
    """Automatic rule: no strips for banded matrices or small x; the tuning knob forces them."""
    ip, j, a = P.laplace3d_7pt(16)
    A = gpu.Mat.from_csr(16 ** 3, 16 ** 3, ip, j, a)
    assert A.strips() == (1, 0)


@pytest.mark.parametrize("name", ["ex5", "lap7", "s27", "rect", "plaw", "bs3"])
def test_mult_transpose(gpu, R, P, name):
    """MatMultTranspose / MatMultTransposeAdd against MatMultTranspose_SeqAIJ (aij.c:1078-1135): the explicit transpose lists
    source rows ascending, so with one lane per row every y[j] receives its terms in the reference's order -> bit-exact;
    several lanes per row (bs3, plaw) -> 1e-12 of the term magnitudes."""
    lanes = 1
    if name == "ex5":
        m, n, ip, j, a = ex5_matrix()
    elif name == "lap7":
        m = n = 14 ** 3; ip, j, a = P.laplace3d_7pt(14)
    elif name == "s27":
        m = n = 10 ** 3; ip, j, a = P.stencil3d_27pt(10)
    elif name == "rect":
        ip, j, a = P.laplace2d_5pt(30, 40, 100, 700); ip = ip - ip[0]; m, n = 600, 1200
    elif name == "plaw":
        m = n = 3000; ip, j, a = P.powerlaw(3000, seed=4, xm=4.0, lmax=500); lanes = -1
    else:
        m = n = 3 * 6 ** 3; ip, j, a = P.elasticity_bs3(6); lanes = -1
    gpu.spmv_tuning(lanes_per_row=lanes)
    try:
        A = gpu.Mat.from_csr(m, n, ip, j, a)
    finally:
        gpu.spmv_tuning()
    ref = R.RefMat(m, n, ip, j, a, no_inode=True)
    xh, yh = P.random_vector(m, 5), P.random_vector(n, 6)
    x, y, z = gpu.Vec(m, xh), gpu.Vec(n, yh), gpu.Vec(n)
    A.mult_transpose(x, z)
    want = ref.multtranspose(xh)
    cols = np.asarray(j)
    rows = np.repeat(np.arange(m), np.diff(ip))
    scale = np.bincount(cols, weights=np.abs(a * xh[rows]), minlength=n)
    got = z.get_array()
    if lanes == 1:
        assert np.array_equal(got, want)
    assert np.all(np.abs(got - want) <= 1e-12 * scale)
    A.mult_transpose_add(x, y, z)
    want2 = ref.multtranspose(xh, yh)
    got = z.get_array()
    if lanes == 1:
        assert np.array_equal(got, want2)
    assert np.all(np.abs(got - want2) <= 1e-12 * (scale + np.abs(yh)))
    A.mult_transpose_add(x, y, y)                          # z aliases y
    assert np.array_equal(y.get_array(), got)
    # the forward product is untouched by the transpose's existence
    u, v = gpu.Vec(n, yh), gpu.Vec(m)
    A.mult(u, v)
    if lanes == 1:
        assert np.array_equal(v.get_array(), ref.mult(yh))


def test_mult_transpose_strips_and_refresh(gpu, R, P):
    """Transpose of a column-stripped matrix, and value refresh drops the cached transpose."""
    m = n = 12 ** 3
    ip, j, a = P.laplace3d_7pt(12)
    gpu.spmv_strips(3); gpu.spmv_tuning(lanes_per_row=1)
    try:
        A = gpu.Mat.from_csr(m, n, ip, j, a)
    finally:
        gpu.spmv_strips(-1); gpu.spmv_tuning()
    xh = P.random_vector(m, 7)
    x, z = gpu.Vec(m, xh), gpu.Vec(n)
    A.mult_transpose(x, z)
    assert np.array_equal(z.get_array(), R.RefMat(m, n, ip, j, a, no_inode=True).multtranspose(xh))
    B = gpu.Mat.from_csr(m, n, ip, j, a)
    B.mult_transpose(x, z)
    a2 = a * 1.5 + 0.25
    B.update_values(a2)
    B.mult_transpose(x, z)
    assert np.array_equal(z.get_array(), R.RefMat(m, n, ip, j, a2, no_inode=True).multtranspose(xh))


def test_against_the_inode_kernel(gpu, R, P):
    """C5 (bs = 3): by default the reference multiplies this matrix with MatMult_SeqAIJ_Inode (inode.c:392-580), which
    adds two products per step and so rounds differently from MatMult_SeqAIJ.  The device types switch inodes off like
    the reference's own GPU type (aijcusp.cu:665) and are bit-identical to the plain kernel (tests above); against
    the Inode kernel the results must still agree within the north-star tolerance 1e-12 * sum |a_ij x_j|."""
    nx = 10
    n = 3 * nx ** 3
    ip, j, a = P.elasticity_bs3(nx)
    ref = R.RefMat(n, n, ip, j, a, bs=3)                      # inode path (default options)
    assert ref.inode_info()[0]
    A = gpu.Mat.from_csr(n, n, ip, j, a)
    xh, yh = P.random_vector(n, 2), P.random_vector(n, 3)
    x, y, z = gpu.Vec(n, xh), gpu.Vec(n, yh), gpu.Vec(n)
    A.mult(x, z)
    scale = np.bincount(np.repeat(np.arange(n), np.diff(ip)), weights=np.abs(a * xh[j]), minlength=n)
    assert np.all(np.abs(z.get_array() - ref.mult(xh)) <= 1e-12 * scale)
    A.mult_add(x, y, z)
    assert np.all(np.abs(z.get_array() - ref.multadd(xh, yh)) <= 1e-12 * (scale + np.abs(yh)))
    # and the solver on top of it: same iteration count, history within 1e-10
    b = ref.mult(np.ones(n))
    rr = ref.ksp_solve(b, "cg", "jacobi", rtol=1e-8)
    bv, xv = gpu.Vec(n, b), gpu.Vec(n)
    r = gpu.ksp_solve(A, bv, xv, ksp_type="cg", pc_type="jacobi", rtol=1e-8)
    assert abs(r["its"] - rr["its"]) <= 1
    k = min(len(r["history"]), len(rr["history"]))
    h, hr = np.asarray(r["history"][:k]), np.asarray(rr["history"][:k])
    gate = hr > 1e-6 * hr[0]
    assert np.max(np.abs(h[gate] - hr[gate]) / hr[gate]) <= 1e-10
