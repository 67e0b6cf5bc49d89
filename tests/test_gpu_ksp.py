"""KSPSolve (CG, GMRES(30); PCJACOBI / PCNONE) on the B200 against the reference's KSPSolve_CG / KSPSolve_GMRES
(oracle/_ref) on identical matrices: iteration counts within +-1, residual histories within 1e-10 relative
(BASELINE.json tolerances).  Both CG forms (op-by-op and fused/device-scalar) are checked."""
import numpy as np
import pytest

from conftest import monitor_short, tridiag

pytestmark = pytest.mark.gpu


def _hist_dev(h, hr, window=1e-5):
    """Deviation of a residual history from the reference's: the larger of
         max_k |h_k - hr_k| / hr_k   over the entries with hr_k >= window * hr_0   (relative, BASELINE's 1e-10), and
         max_k |h_k - hr_k| / hr_0   over ALL entries (rounding in a Krylov recurrence is absolute in ||r_0||:
                                     once the residual has dropped by 1e5 a 1e-16 perturbation of a dot product is
                                     already 1e-11 relative to r_k, for the reference against itself with another
                                     BLAS as much as for us; see DESIGN.md "KSP tolerances")."""
    k = min(len(h), len(hr))
    if not k:
        return 0.0
    h, hr = np.asarray(h[:k]), np.asarray(hr[:k])
    d = np.abs(h - hr)
    win = hr >= window * hr[0]
    return max(np.max(d[win] / hr[win]) if win.any() else 0.0, np.max(d) / hr[0])


def _solve_both(gpu, R, ip, j, a, n, bh, ksp, pc, rtol=1e-8, tol=1e-10, **kw):
    A = gpu.Mat.from_csr(n, n, ip, j, a)
    ref = R.RefMat(n, n, ip, j, a, no_inode=True)
    opts = {}
    if kw.get("norm_type") == gpu.KSP_NORM_UNPRECONDITIONED: opts["-ksp_norm_type"] = "unpreconditioned"
    if kw.get("norm_type") == gpu.KSP_NORM_NATURAL: opts["-ksp_norm_type"] = "natural"
    if kw.get("gmres_cgs_refine") == 2: opts["-ksp_gmres_cgs_refinement_type"] = "refine_always"
    if kw.get("gmres_cgs_refine") == 1: opts["-ksp_gmres_cgs_refinement_type"] = "refine_ifneeded"
    if "gmres_restart" in kw: opts["-ksp_gmres_restart"] = str(kw["gmres_restart"])
    if kw.get("gmres_modified_gs"): opts["-ksp_gmres_modifiedgramschmidt"] = ""
    x0 = kw.pop("x0", None)
    rr = ref.ksp_solve(bh, ksp, pc, rtol=rtol, maxit=kw.get("max_it", 10000), x0=x0, options=opts)
    out = []
    for fused in ((0, 1) if ksp == "cg" else (0,)):
        b, x = gpu.Vec(n, bh), gpu.Vec(n, x0 if x0 is not None else None)
        r = gpu.ksp_solve(A, b, x, ksp_type=ksp, pc_type=pc, rtol=rtol, fused=fused, check_every=7,
                          guess_nonzero=int(x0 is not None), **kw)
        assert abs(r["its"] - rr["its"]) <= 1, (fused, r["its"], rr["its"])
        assert r["reason"] == rr["reason"], (fused, r["reason"], rr["reason"])
        assert abs(len(r["history"]) - len(rr["history"])) <= 1
        assert _hist_dev(r["history"], rr["history"]) <= tol, (fused, _hist_dev(r["history"], rr["history"]))
        xs = x.get_array()
        assert np.linalg.norm(xs - rr["x"]) <= 1e-8 * max(np.linalg.norm(rr["x"]), 1.0)
        out.append(r)
    if len(out) == 2:                                    # the two CG forms do the same arithmetic
        assert out[0]["its"] == out[1]["its"]
        assert _hist_dev(out[0]["history"], out[1]["history"]) <= 1e-12
    return out[-1], rr


def test_pc_ex2_golden_cg_none(gpu, goldens):
    """ksp/pc/tests ex2 -ksp_type cg: the reference's own golden history, 6 digits."""
    ip, j, a = tridiag(10)
    A = gpu.Mat.from_csr(10, 10, ip, j, a)
    one, b, x = gpu.Vec(10), gpu.Vec(10), gpu.Vec(10)
    one.set(1.0); A.mult(one, b)
    for fused in (0, 1):
        r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="none", fused=fused)
        assert [monitor_short(v) for v in r["history"]] == goldens["pc_ex2_1"]["history"]


def test_ksp_ex5_golden_gmres_jacobi(gpu, goldens):
    m, n = 3, 2
    N = 6
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
    A = gpu.Mat.from_csr(N, N, ip, j, a)
    u, b, x = gpu.Vec(N, np.arange(N, dtype=np.float64)), gpu.Vec(N), gpu.Vec(N)
    A.mult(u, b)
    r = gpu.ksp_solve(A, b, x, ksp_type="gmres", pc_type="jacobi", gmres_cgs_refine=2)
    assert [monitor_short(v) for v in r["history"]] == goldens["ksp_ex5_1"]["history"]


def test_c1_ex2_cg_jacobi(gpu, R, P):
    """BASELINE config 1 shape (ex2 5-point Laplacian, CG + Jacobi, rtol = 1e-2/((m+1)(n+1))) at m = n = 100."""
    m = n = 100
    ip, j, a = P.laplace2d_5pt(m, n)
    bh = R.RefMat(m * n, m * n, ip, j, a).mult(np.ones(m * n))
    _solve_both(gpu, R, ip, j, a, m * n, bh, "cg", "jacobi", rtol=1e-2 / ((m + 1) * (n + 1)))


@pytest.mark.parametrize("pc", ["jacobi", "none"])
@pytest.mark.parametrize("norm", ["preconditioned", "unpreconditioned", "natural"])
def test_cg_variants(gpu, R, P, pc, norm):
    ip, j, a = P.laplace3d_7pt(16)
    n = 16 ** 3
    a = a * (1.0 + 0.5 * (np.repeat(np.arange(n), np.diff(ip)) == j) * P.random_vector(len(a), 9, 0.0, 1.0))  # non-constant diagonal
    bh = P.random_vector(n, 4)
    nt = dict(preconditioned=gpu.KSP_NORM_PRECONDITIONED, unpreconditioned=gpu.KSP_NORM_UNPRECONDITIONED, natural=gpu.KSP_NORM_NATURAL)[norm]
    _solve_both(gpu, R, ip, j, a, n, bh, "cg", pc, norm_type=nt)


def test_cg_nonzero_guess_and_maxit(gpu, R, P):
    ip, j, a = P.laplace3d_7pt(14)
    n = 14 ** 3
    bh = P.random_vector(n, 4)
    _solve_both(gpu, R, ip, j, a, n, bh, "cg", "jacobi", x0=P.random_vector(n, 6))
    r, rr = _solve_both(gpu, R, ip, j, a, n, bh, "cg", "jacobi", rtol=1e-14, max_it=9)
    assert r["reason"] == -3 and r["its"] == 9


def test_cg_indefinite_matrix_detected(gpu, R, P):
    ip, j, a = P.laplace2d_5pt(12, 12)
    a = -a                                               # negative definite: dpi < 0 at i = 0 is allowed, product test at i > 0
    a[ip[5] + 2] = 7.0
    bh = P.random_vector(144, 4)
    A = gpu.Mat.from_csr(144, 144, ip, j, a)
    rr = R.RefMat(144, 144, ip, j, a).ksp_solve(bh, "cg", "none", rtol=1e-10)
    for fused in (0, 1):
        b, x = gpu.Vec(144, bh), gpu.Vec(144)
        r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="none", rtol=1e-10, fused=fused)
        assert r["reason"] == rr["reason"] and r["its"] == rr["its"]


@pytest.mark.parametrize("refine", [0, 1, 2])
def test_gmres_jacobi(gpu, R, P, refine):
    ip, j, a = P.powerlaw(4000, seed=4, xm=6.0, lmax=600)
    n = 4000
    bh = R.RefMat(n, n, ip, j, a).mult(np.ones(n))
    _solve_both(gpu, R, ip, j, a, n, bh, "gmres", "jacobi", gmres_cgs_refine=refine)


def test_gmres_modified_gram_schmidt(gpu, R, P):
    """KSPGMRESModifiedGramSchmidtOrthogonalization (borthog.c:30-52): one VecDot + VecAXPY per basis vector."""
    ip, j, a = P.powerlaw(4000, seed=4, xm=6.0, lmax=600)
    n = 4000
    bh = R.RefMat(n, n, ip, j, a).mult(np.ones(n))
    _solve_both(gpu, R, ip, j, a, n, bh, "gmres", "jacobi", gmres_modified_gs=1)
    ip, j, a = P.laplace2d_5pt(30, 30)
    _solve_both(gpu, R, ip, j, a, 900, P.random_vector(900, 4), "gmres", "jacobi", rtol=1e-6, gmres_restart=12, gmres_modified_gs=1, tol=1e-8)


def test_gmres_restarts_and_guess(gpu, R, P):
    ip, j, a = P.laplace2d_5pt(40, 40)
    n = 1600
    bh = P.random_vector(n, 4)
    r, rr = _solve_both(gpu, R, ip, j, a, n, bh, "gmres", "jacobi", rtol=1e-6, gmres_restart=10, tol=1e-8)
    assert r["its"] > 30                                  # several restart cycles
    _solve_both(gpu, R, ip, j, a, n, bh, "gmres", "none", rtol=1e-6, x0=P.random_vector(n, 6), tol=1e-8)
    r, rr = _solve_both(gpu, R, ip, j, a, n, bh, "gmres", "jacobi", rtol=1e-12, max_it=45, gmres_restart=30, tol=1e-8)
    assert r["reason"] == -3


def test_fixture_histories(gpu, P, fixtures):
    ip, j, a = P.laplace3d_7pt(12)
    N = 12 ** 3
    A = gpu.Mat.from_csr(N, N, ip, j, a)
    one, b, x = gpu.Vec(N), gpu.Vec(N), gpu.Vec(N)
    one.set(1.0); A.mult(one, b)
    r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-8)
    assert abs(r["its"] - int(fixtures["lap7_12_cg_its"])) <= 1
    assert _hist_dev(r["history"], fixtures["lap7_12_cg_hist"]) <= 1e-10
    r = gpu.ksp_solve(A, b, x, ksp_type="gmres", pc_type="jacobi", rtol=1e-8)
    assert abs(r["its"] - int(fixtures["lap7_12_gmres_its"])) <= 1
    assert _hist_dev(r["history"], fixtures["lap7_12_gmres_hist"]) <= 1e-10


def test_c2_cg_256_cubed_properties(gpu, P):
    """BASELINE config 2 at full size: CG + Jacobi on the 256^3 7-point Laplacian, b = A*1.  Properties: monotone
    decrease of the recorded natural quantities is not guaranteed, so check convergence, true residual and error."""
    nx = 256
    N = nx ** 3
    ip, j, a = P.laplace3d_7pt(nx)
    A = gpu.Mat.from_csr(N, N, ip, j, a)
    one, b, x, r = gpu.Vec(N), gpu.Vec(N), gpu.Vec(N), gpu.Vec(N)
    one.set(1.0); A.mult(one, b)
    res = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-8, max_it=3000)
    assert res["reason"] == 2 and res["its"] < 1500
    A.mult(x, r); r.aypx(-1.0, b)                        # r = b - A x
    bn = b.norm()
    assert r.norm() <= 1e-6 * bn
    res0 = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-8, max_it=3000, fused=0)
    assert res0["its"] == res["its"]
    assert _hist_dev(res["history"], res0["history"]) <= 1e-9


def test_ex10_style_binary_file_solve(gpu, R, P, tmp_path):
    """A matrix file written by the REFERENCE's binary viewer, loaded with petsc-dev_b200/binaryio.py, solved on the B200
    by tools/ex10_b200.py: iteration count and residual history follow the reference's KSPSolve on the same matrix."""
    import subprocess, sys, os, re
    n = 10 ** 3
    ip, j, a = P.laplace3d_7pt(10)
    ref = R.RefMat(n, n, ip, j, a, no_inode=True)
    f = str(tmp_path / "lap7.bin")
    R.mat_write_binary(ref, f)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "ex10_b200.py"), "-f0", f, "-ksp_type", "cg", "-pc_type", "jacobi",
                          "-ksp_rtol", "1e-8", "-ksp_monitor"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    hist = np.array([float(m.group(1)) for m in re.finditer(r"KSP Residual norm (\S+)", out.stdout)])
    its = int(re.search(r"Number of iterations = +(\d+)", out.stdout).group(1))
    rr = ref.ksp_solve(ref.mult(np.ones(n)), "cg", "jacobi", rtol=1e-8)
    assert abs(its - rr["its"]) <= 1
    k = min(len(hist), len(rr["history"]))
    assert np.max(np.abs(hist[:k] - rr["history"][:k]) / rr["history"][:k]) < 1e-10


@pytest.mark.parametrize("mode", ["graph", "launches", "persistent"])
def test_fused_cg_block_replay_modes(gpu, R, P, mode, monkeypatch):
    """Several hundred iterations on a small matrix, so that the fused CG driver gets past its first block of iterations:
    from the second block on the iterations are a CUDA-graph replay (default for small single-GPU problems), plain
    launches (B200_CG_GRAPH=0) or the persistent cooperative kernel (B200_CG_PERSISTENT=1) — same arithmetic, so the same
    history as the reference's KSPSolve_CG within 1e-10 and the same iteration count in all three."""
    if mode == "launches":
        monkeypatch.setenv("B200_CG_GRAPH", "0")
    if mode == "persistent":
        monkeypatch.setenv("B200_CG_PERSISTENT", "1")
    ip, j, a = P.laplace2d_5pt(70, 64)
    n = 70 * 64
    ref = R.RefMat(n, n, ip, j, a, no_inode=True)
    bh = ref.mult(P.random_vector(n, 3))
    rr = ref.ksp_solve(bh, "cg", "jacobi", rtol=1e-12, maxit=10000)
    assert rr["its"] > 150, rr["its"]                   # at least three blocks of 64
    A = gpu.Mat.from_csr(n, n, ip, j, a)
    b, x = gpu.Vec(n, bh), gpu.Vec(n)
    r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-12, fused=1)
    assert r["its"] == rr["its"] and r["reason"] == rr["reason"], (r["its"], rr["its"], r["reason"], rr["reason"])
    assert len(r["history"]) == len(rr["history"])
    assert _hist_dev(r["history"], rr["history"], window=1e-7) <= 1e-10
    assert np.linalg.norm(x.get_array() - rr["x"]) <= 1e-8 * np.linalg.norm(rr["x"])
    # a second solve on the same operator (a fresh graph is captured per solve) with a different right-hand side
    bh2 = ref.mult(P.random_vector(n, 4))
    rr2 = ref.ksp_solve(bh2, "cg", "jacobi", rtol=1e-12, maxit=10000)
    b.set_array(bh2)
    r2 = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-12, fused=1)
    assert r2["its"] == rr2["its"]
    assert _hist_dev(r2["history"], rr2["history"], window=1e-7) <= 1e-10
