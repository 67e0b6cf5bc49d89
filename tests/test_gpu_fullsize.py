"""Parity at the BASELINE.json sizes against the ORACLE (the unmodified reference CPU library, oracle/_ref) — not
size-independent properties: the device MatMult / MatMultAdd of every configuration is compared entry by entry with
the reference's MatMult_SeqAIJ (src/mat/impls/aij/seq/aij.c:1225-1358) on the same arrays, and CG's residual history
with the reference's KSPSolve_CG (src/ksp/ksp/impls/cg/cg.c:92-286).

  C2  7-point Laplacian 256^3            one lane per row -> BIT-EXACT
  C3  27-point stencil at 256^3 (one GPU's share of the 512^3 grid on 8 GPUs): one lane per row -> BIT-EXACT
  C5  bs = 3 27-point, 3 x 128^3 unknowns, 81 entries per row -> |dy_i| <= 1e-12 * sum_j |a_ij x_j|
  C4  power-law rows, 2.5 M rows (a quarter of the 10 M-row matrix; ~125 M entries) -> same tolerance

The error scale sum_j |a_ij| |x_j| comes from the reference itself (|A| |x|), so no nnz-sized temporaries are made."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _compare(gpu, R, P, m, ip, j, a, exact, seed=2):
    A = gpu.Mat.from_csr(m, m, ip, j, a)
    info = A.bin_info()
    ref = R.RefMat(m, m, ip, j, a, no_inode=True)
    xh, yh = P.random_vector(m, seed), P.random_vector(m, seed + 1)
    x, y, z = gpu.Vec(m, xh), gpu.Vec(m, yh), gpu.Vec(m)
    A.mult(x, z)
    got = z.get_array()
    zr = ref.mult(xh)
    if exact:
        assert info["lanes_per_row"] == 1 and info["nlong"] == 0
        assert np.array_equal(got, zr), "MatMult differs from MatMult_SeqAIJ (max abs %g)" % np.max(np.abs(got - zr))
    absref = R.RefMat(m, m, ip, j, np.abs(a), no_inode=True)
    scale = absref.mult(np.abs(xh))
    absref.close()
    assert np.all(np.abs(got - zr) <= 1e-12 * scale)
    A.mult_add(x, y, z)
    got = z.get_array()
    zr = ref.multadd(xh, yh)
    if exact:
        assert np.array_equal(got, zr), "MatMultAdd differs from MatMultAdd_SeqAIJ"
    assert np.all(np.abs(got - zr) <= 1e-12 * (scale + np.abs(yh)))
    d = gpu.Vec(m)
    A.get_diagonal(d)
    assert np.array_equal(d.get_array(), ref.diagonal())
    return A, ref


def _hist_dev(h, hr):
    k = min(len(h), len(hr))
    return float(np.max(np.abs(h[:k] - hr[:k]) / hr[:k]))


def test_c2_lap7_256_against_the_reference(gpu, R, P):
    nx = 256
    N = nx ** 3
    ip, j, a = P.laplace3d_7pt(nx)
    A, ref = _compare(gpu, R, P, N, ip, j, a, exact=True)
    # KSPCG + PCJACOBI, 25 iterations of the full-size problem, history against KSPSolve_CG
    bh = ref.mult(np.ones(N))
    b, x = gpu.Vec(N, bh), gpu.Vec(N)
    rr = ref.ksp_solve(bh, "cg", "jacobi", rtol=1e-50, maxit=25)
    for fused in (1, 0):
        r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-50, max_it=25, fused=fused)
        assert r["its"] == rr["its"] and r["reason"] == rr["reason"]
        assert len(r["history"]) == len(rr["history"])
        assert _hist_dev(r["history"], rr["history"]) <= 1e-10
        assert np.max(np.abs(x.get_array() - rr["x"])) <= 1e-10 * np.max(np.abs(rr["x"]))


def test_c3_stencil27_256_against_the_reference(gpu, R, P):
    nx = 256
    ip, j, a = P.stencil3d_27pt(nx)
    _compare(gpu, R, P, nx ** 3, ip, j, a, exact=True)


def test_c5_elasticity_bs3_128_against_the_reference(gpu, R, P):
    nx = 128
    ip, j, a = P.elasticity_bs3(nx)
    _compare(gpu, R, P, 3 * nx ** 3, ip, j, a, exact=False)


def test_c4_powerlaw_2p5M_rows_against_the_reference(gpu, R, P):
    """A quarter of C4's rows with the SAME column space (N = 10^7 hashed columns would make the block rectangular, so the
    square 2.5 M-row matrix is generated with its own N; x = 20 MB, 125 M entries, longest rows ~ 10^5)."""
    N = 2_500_000
    ip, j, a = P.powerlaw(N)
    A, ref = _compare(gpu, R, P, N, ip, j, a, exact=False)
    # GMRES(30) + Jacobi to rtol 1e-8 against KSPSolve_GMRES
    bh = ref.mult(np.ones(N))
    b, x = gpu.Vec(N, bh), gpu.Vec(N)
    rr = ref.ksp_solve(bh, "gmres", "jacobi", rtol=1e-8, maxit=200)
    r = gpu.ksp_solve(A, b, x, ksp_type="gmres", pc_type="jacobi", rtol=1e-8, max_it=200)
    assert rr["reason"] == 2 and r["reason"] == 2
    assert abs(r["its"] - rr["its"]) <= 1
    assert _hist_dev(r["history"], rr["history"]) <= 1e-10
