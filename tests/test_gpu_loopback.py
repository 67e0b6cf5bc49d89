"""The multi-GPU MatMult path on ONE GPU: a rank whose only neighbour is itself (b200_comm_enable_p2p(2), "loop-back").
pack/put over the peer arena, arrival flags, credits, the one-launch MatMult_MPIAIJ kernel (b200_spmv_mpi_kernels.cuh) and
the four-step path run exactly as between GPUs — only the stores stay on the device — so the driver's single-GPU test run
covers the kernels that the 2/4/8-GPU runs use.

Operator: the 7-point Laplacian, periodic in z, held as diagonal block (the non-periodic stencil) + off-diagonal block (the
wrap-around couplings, ghost ids) + "ghost exchange" (the top and bottom planes copied into the ghost vector).
Oracle: the unmodified reference kernels, MatMult_SeqAIJ on the diagonal block then MatMultAdd_SeqAIJ on the off-diagonal
block (mpiaij.c:1111-1114) -> bit-exact; and the reference's KSPSolve_CG on the same operator assembled as one SeqAIJ."""
import importlib
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def loop(gpu):
    par = importlib.import_module("petsc-dev_b200.parallel")
    par.enable_loopback()
    yield par
    gpu.synchronize()
    gpu.lib.b200_comm_destroy()


def _operator(par, P, nx, ny, nz, fused):
    d = P.periodic_z_loopback(nx, ny, nz)
    old = os.environ.get("B200_MPI_FUSED")
    os.environ["B200_MPI_FUSED"] = "1" if fused else "0"
    try:
        ng = d["nghost"]
        A = par.create_mpiaij_raw(d["m"], d["m"], *d["Ad"], *d["B"], np.arange(ng), [0], [0, ng], d["to_idx"], [0], [0, ng])
    finally:
        if old is None:
            del os.environ["B200_MPI_FUSED"]
        else:
            os.environ["B200_MPI_FUSED"] = old
    return d, A


def _oracle(R, d, xh, yh=None):
    m = d["m"]
    ad = R.RefMat(m, m, *d["Ad"], no_inode=True)
    yd = ad.mult(xh) if yh is None else ad.multadd(xh, yh)
    return R.RefMat(m, d["nghost"], *d["B"], no_inode=True).multadd(xh[d["to_idx"]], yd)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("grid", [(16, 16, 16), (64, 64, 40), (33, 17, 5)])
def test_matmult_bit_exact(gpu, loop, R, P, grid, fused):
    d, A = _operator(loop, P, *grid, fused=fused)
    isf, nb = loop.is_fused(A)
    assert isf == fused and (nb > 0) == fused
    m = d["m"]
    x, y, z = gpu.Vec(m), gpu.Vec(m), gpu.Vec(m)
    for rep in range(5):                                  # several exchanges: both parities of the ghost buffers, the credits
        xh, yh = P.random_vector(m, 10 + rep), P.random_vector(m, 20 + rep)
        x.set_array(xh); y.set_array(yh)
        A.mult(x, z)
        assert np.array_equal(z.get_array(), _oracle(R, d, xh)), (grid, fused, rep)
        A.mult_add(x, y, z)
        assert np.array_equal(z.get_array(), _oracle(R, d, xh, yh))
        A.mult_add(x, y, y)                               # z aliases y
        assert np.array_equal(y.get_array(), _oracle(R, d, xh, yh))
    A.destroy()


@pytest.mark.parametrize("lanes", [2, 8])
def test_matmult_several_lanes_per_row(gpu, loop, R, P, lanes):
    gpu.spmv_tuning(lanes_per_row=lanes)
    try:
        d, A = _operator(loop, P, 20, 20, 12, fused=True)
    finally:
        gpu.spmv_tuning()
    assert loop.is_fused(A)[0]
    m = d["m"]
    xh = P.random_vector(m, 3)
    x, z = gpu.Vec(m, xh), gpu.Vec(m)
    A.mult(x, z)
    want = _oracle(R, d, xh)
    assert np.all(np.abs(z.get_array() - want) <= 1e-12 * 12.0)          # |a_ij x_j| <= 6 + 6
    A.destroy()


@pytest.mark.parametrize("fused_spmv", [True, False])
def test_cg_on_the_loopback_operator(gpu, loop, R, P, fused_spmv):
    d, A = _operator(loop, P, 24, 24, 24, fused=fused_spmv)
    m = d["m"]
    ref = R.RefMat(m, m, *d["full"], no_inode=True)
    bh = ref.mult(P.random_vector(m, 7))      # (an all-ones solution on this symmetric grid makes the CG history move by
                                              #  0.5% under a 1-ulp change of b: nothing can be compared on it)
    b, x = gpu.Vec(m, bh), gpu.Vec(m)
    rr = ref.ksp_solve(bh, "cg", "jacobi", rtol=1e-8)
    for fused in (1, 0):
        r = gpu.ksp_solve(A, b, x, ksp_type="cg", pc_type="jacobi", rtol=1e-8, fused=fused)
        assert abs(r["its"] - rr["its"]) <= 1 and r["reason"] == rr["reason"]
        k = min(len(r["history"]), len(rr["history"]))
        assert np.max(np.abs(r["history"][:k] - rr["history"][:k]) / rr["history"][:k]) <= 1e-10
    A.destroy()


def test_transpose_and_general_scatter_on_the_loopback(gpu, loop, R, P):
    """The reverse (ADD) exchange through the same arena: z = A^T x for the loop-back operator (symmetric, so A^T x = A x up
    to the order of the additions)."""
    d, A = _operator(loop, P, 12, 10, 6, fused=True)
    m = d["m"]
    xh = P.random_vector(m, 5)
    x, z = gpu.Vec(m, xh), gpu.Vec(m)
    A.mult_transpose(x, z)
    want = R.RefMat(m, m, *d["full"], no_inode=True).multtranspose(xh)
    assert np.all(np.abs(z.get_array() - want) <= 1e-12 * 12.0)
    A.destroy()


def test_arena_ranges_return_in_any_order(gpu, loop, R, P):
    """Operators (and attached VecScatters) are destroyed in any order: every range goes back to the peer arena's free list
    and is reused, so a long create/destroy sequence never falls back to the NCCL transport (is_fused stays true) and the
    products stay bit-exact.  The arena is 512 MiB; 600 operators of ~1 MiB each would not fit without reuse."""
    keep = []
    for rep in range(600):
        d, A = _operator(loop, P, 128, 128, 3 + rep % 3, fused=True)     # ghost vectors of 2 x 128 x 128 entries, 2 parities, 2 directions
        assert loop.is_fused(A)[0], "operator %d fell off the peer transport: the arena leaks" % rep
        keep.append((d, A))
        if len(keep) == 3:
            keep.pop(rep % 2)[1].destroy()                               # not the most recent one
    d, A = keep[-1]
    m = d["m"]
    xh = P.random_vector(m, 5)
    x, z = gpu.Vec(m, xh), gpu.Vec(m)
    A.mult(x, z)
    assert np.array_equal(z.get_array(), _oracle(R, d, xh))
    for _, B in keep:
        B.destroy()
