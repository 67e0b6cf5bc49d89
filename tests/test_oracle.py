"""The oracle (the UNMODIFIED reference CPU library, oracle/_ref) against the reference's own golden outputs for this
path, and against the committed fixtures.  CPU only."""
import numpy as np

from conftest import ex5_matrix, monitor_short, tridiag


def test_mat_ex5_golden(R, goldens):
    m, n, ip, j, a = ex5_matrix()
    A = R.RefMat(m, n, ip, j, a)
    y = A.mult(np.arange(n, dtype=np.float64))
    assert [float("%G" % v) for v in y] == goldens["mat_ex5_11_A"]["matmult"]
    # MatMultAdd consistency check of ex5.c: ||x + z - w|| <= 1e-8
    z = 100.0 * (np.arange(m) + 1)
    w = A.multadd(np.arange(n, dtype=np.float64), z)
    assert np.linalg.norm(y + z - w) <= 1e-8


def test_mat_ex5_mpiaij_np1(R, goldens):
    """ex5_21: the same numbers through MATMPIAIJ on one rank (makefile:774-778)."""
    m, n, ip, j, a = ex5_matrix()
    sq = [k for k in range(ip[-1]) if j[k] < m]          # square part so the 1-rank MPIAIJ layout is valid
    ip2 = np.arange(0, (m + 1) * m, m, dtype=np.int32)
    A = R.RefMat(m, m, ip2, j[sq], a[sq], kind="mpiaij")
    S = R.RefMat(m, m, ip2, j[sq], a[sq])
    x = np.arange(m, dtype=np.float64)
    assert np.array_equal(A.mult(x), S.mult(x))
    d = S.diagonal()
    assert [float("%G" % v) for v in d] == goldens["mat_ex5_11_A"]["diagonal"]


def test_ksp_ex2_golden(R, P, goldens):
    """ksp/tutorials runex2: -m 5 -n 5 -ksp_gmres_cgs_refinement_type refine_always, default GMRES(30)+ILU(0)."""
    m = n = 5
    ip, j, a = P.laplace2d_5pt(m, n)
    A = R.RefMat(m * n, m * n, ip, j, a)
    u = np.ones(m * n)
    b = A.mult(u)
    r = A.ksp_solve(b, "gmres", "ilu", rtol=1e-2 / ((m + 1) * (n + 1)),
                    options={"-ksp_gmres_cgs_refinement_type": "refine_always"})
    g = goldens["ksp_ex2_1"]
    assert [monitor_short(v) for v in r["history"]] == g["history"]
    assert r["its"] == g["iterations"]
    assert "%G" % np.linalg.norm(r["x"] - u) == g["error_norm"]


def test_ksp_ex5_golden(R, goldens):
    """ksp/tutorials runex5 (first solve): m = 3, n = 2, GMRES + Jacobi, refine_always."""
    m, n = 3, 2
    N = m * n
    rows = []
    for Ii in range(N):
        i, jj = divmod(Ii, n)
        e = {Ii: 4.0}
        if i > 0: e[Ii - n] = -1.0
        if i < m - 1: e[Ii + n] = -1.0
        if jj > 0: e[Ii - 1] = -1.0
        if jj < n - 1: e[Ii + 1] = -1.0
        rows.append(sorted(e.items()))
    ip = np.cumsum([0] + [len(r) for r in rows]).astype(np.int32)
    j = np.array([c for r in rows for c, _ in r], np.int32)
    a = np.array([v for r in rows for _, v in r])
    A = R.RefMat(N, N, ip, j, a)
    b = A.mult(np.arange(N, dtype=np.float64))
    r = A.ksp_solve(b, "gmres", "jacobi", options={"-ksp_gmres_cgs_refinement_type": "refine_always"})
    assert [monitor_short(v) for v in r["history"]] == goldens["ksp_ex5_1"]["history"]


def test_pc_ex2_cg_golden(R, goldens):
    """ksp/pc/tests runex2: CG + PCNONE on tridiag(-1,2,-1), n = 10."""
    ip, j, a = tridiag(10)
    A = R.RefMat(10, 10, ip, j, a)
    b = A.mult(np.ones(10))
    r = A.ksp_solve(b, "cg", "none")
    assert [monitor_short(v) for v in r["history"]] == goldens["pc_ex2_1"]["history"]


def test_fixtures_reproduce(R, P, fixtures):
    """The committed oracle fixtures are what the reference library produces today (guards against a stale file)."""
    n = int(fixtures["vec_n"])
    x, y = P.random_vector(n, 11), P.random_vector(n, 12)
    assert R.vec_dot(x, y) == float(fixtures["vec_dot"])
    assert R.vec_norm(x, R.NORM_2) == float(fixtures["vec_norm2"])
    ip, j, a = P.laplace3d_7pt(12)
    A = R.RefMat(12 ** 3, 12 ** 3, ip, j, a)
    assert np.array_equal(A.mult(P.random_vector(12 ** 3, 2)), fixtures["lap7_12_mult"])
    r = A.ksp_solve(A.mult(np.ones(12 ** 3)), "cg", "jacobi", rtol=1e-8)
    assert np.array_equal(r["history"], fixtures["lap7_12_cg_hist"])


def test_oracle_matches_plain_loops(R, P):
    """The reference MatMult is the textbook left-to-right row sum (aij.h:384): check against a numpy restatement."""
    ip, j, a = P.powerlaw(500, seed=4, xm=5.0, lmax=200)
    A = R.RefMat(500, 500, ip, j, a)
    x = P.random_vector(500, 7)
    y = A.mult(x)
    for r in range(500):
        s = 0.0
        for k in range(ip[r], ip[r + 1]):
            s += a[k] * x[j[k]]
        assert s == y[r]
    assert A.inode_info()[0] is False


def test_inode_path_for_bs3(R, P):
    """C5: the blocked 27-point operator takes the reference's Inode MatMult (inode.c:392) unless -mat_no_inode."""
    ip, j, a = P.elasticity_bs3(4)
    n = 3 * 64
    A = R.RefMat(n, n, ip, j, a, bs=3)
    use, cnt = A.inode_info()
    assert use and cnt == 64
    B = R.RefMat(n, n, ip, j, a, bs=3, no_inode=True)
    assert not B.inode_info()[0]
    x = P.random_vector(n, 3)
    ya, yb = A.mult(x), B.mult(x)
    scale = np.abs(a).max() * np.abs(x).max() * 81
    assert np.max(np.abs(ya - yb)) <= 1e-13 * scale
