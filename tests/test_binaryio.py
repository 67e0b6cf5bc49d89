"""petsc-dev_b200/binaryio.py against the reference's own binary viewer (oracle/_ref): files the reference writes
(MatView_SeqAIJ_Binary aij.c:414, VecView) must parse to the same CSR / values, and files this module writes must load
in the reference (MatLoad_SeqAIJ aij.c:4093, VecLoad) to the same matrix — byte-identical files both ways."""
import importlib
import os

import numpy as np
import pytest

from conftest import ex5_matrix


@pytest.fixture(scope="module")
def bio():
    return importlib.import_module("petsc-dev_b200.binaryio")


def _cases(P):
    m, n, ip, j, a = ex5_matrix()
    yield "ex5_rect", m, n, ip, j, a
    ip, j, a = P.laplace3d_7pt(7)
    yield "lap7", 343, 343, ip, j, a
    ip, j, a = P.powerlaw(500, seed=4, xm=4.0, lmax=100)
    yield "plaw", 500, 500, ip, j, a
    yield "empty_rows", 4, 5, np.array([0, 0, 2, 2, 3], np.int32), np.array([1, 4, 0], np.int32), np.array([1.5, -2.0, 1e-300])


def test_mat_files_both_ways(bio, R, P, tmp_path):
    for name, m, n, ip, j, a in _cases(P):
        ref = R.RefMat(m, n, ip, j, a, no_inode=True)
        f_ref, f_ours = str(tmp_path / (name + "_ref.bin")), str(tmp_path / (name + "_ours.bin"))
        R.mat_write_binary(ref, f_ref)
        M, N, ai, aj, aa = bio.read_mat(f_ref)
        assert (M, N) == (m, n) and np.array_equal(ai, ip) and np.array_equal(aj, j) and np.array_equal(aa, a), name
        bio.write_mat(f_ours, m, n, ip, j, a)
        assert open(f_ours, "rb").read() == open(f_ref, "rb").read(), name            # byte-identical files
        back = R.mat_load_binary(f_ours)
        bi, bj, ba = back.csr()
        assert np.array_equal(bi, ip) and np.array_equal(bj, j) and np.array_equal(ba, a), name


def test_vec_files_both_ways(bio, R, P, tmp_path):
    x = P.random_vector(1234, 3) * 1e7 - 5e6
    x[:3] = [0.0, -0.0, 1e-310]
    f_ref, f_ours = str(tmp_path / "v_ref.bin"), str(tmp_path / "v_ours.bin")
    R.vec_write_binary(x, f_ref)
    assert np.array_equal(bio.read_vec(f_ref), x)
    bio.write_vec(f_ours, x)
    assert open(f_ours, "rb").read() == open(f_ref, "rb").read()
    assert np.array_equal(R.vec_load_binary(f_ours, 2000), x)


def test_system_file_and_errors(bio, P, tmp_path):
    ip, j, a = P.laplace2d_5pt(6, 5)
    b = np.arange(30, dtype=np.float64)
    f = str(tmp_path / "sys.bin")
    bio.write_mat(f, 30, 30, ip, j, a)
    with open(f, "ab") as fh:
        np.array([bio.VEC_FILE_CLASSID, 30], dtype=">i4").tofile(fh)
        b.astype(">f8").tofile(fh)
    M, N, ai, aj, aa, bb = bio.read_system(f)
    assert (M, N) == (30, 30) and np.array_equal(ai, ip) and np.array_equal(bb, b)
    bio.write_mat(f, 30, 30, ip, j, a)
    assert bio.read_system(f)[5] is None                      # no right-hand side in the file
    with pytest.raises(bio.PetscBinaryError):
        bio.read_vec(f)                                       # a matrix file is not a vector file
    with open(f, "r+b") as fh:
        fh.truncate(200)
    with pytest.raises(bio.PetscBinaryError):
        bio.read_mat(f)


def test_striped_reads_match_the_whole_file(bio, tmp_path, R, P):
    """Every rank reads its own row slab with seeks: slabs of a file the REFERENCE wrote (matrix followed by a vector)
    stack up to what the whole-file reader and the reference's MatLoad return."""
    n = 1000
    ip, j, a = P.powerlaw(n, seed=4, xm=3.0, lmax=200)
    ref = R.RefMat(n, n, ip, j, a, no_inode=True)
    path = str(tmp_path / "sys.bin")
    R.mat_write_binary(ref, path)
    bh = P.random_vector(n, 9)
    vpath = str(tmp_path / "v.bin")
    R.vec_write_binary(bh, vpath)
    with open(path, "ab") as f:                              # ex10-style file: the vector follows the matrix
        f.write(open(vpath, "rb").read())
    M, N, ai, aj, aa, b = bio.read_system(path)
    assert bio.mat_file_size(path) == 16 + 4 * M + 12 * len(aa)
    par = importlib.import_module("petsc-dev_b200.parallel")
    for size in (1, 3, 7):
        ranges = par.split_ownership(M, size)
        rows, cols, vals, rhs = [0], [], [], []
        for r in range(size):
            r0, r1 = int(ranges[r]), int(ranges[r + 1])
            m2, n2, si, sj, sa = bio.read_mat_rows(path, r0, r1)
            assert (m2, n2) == (M, N) and si[0] == 0 and si.size == r1 - r0 + 1
            rows += list(si[1:] + rows[-1]); cols.append(sj); vals.append(sa)
            rhs.append(bio.read_vec_rows(path, r0, r1, offset=bio.mat_file_size(path)))
        assert np.array_equal(np.array(rows, np.int32), ai)
        assert np.array_equal(np.concatenate(cols), aj) and np.array_equal(np.concatenate(vals), aa)
        assert np.array_equal(np.concatenate(rhs), b) and np.array_equal(b, bh)
    assert bio.read_vec_rows(vpath, 0, 10, offset=4) is None          # no vector header there
    with pytest.raises(bio.PetscBinaryError):
        bio.read_mat_rows(path, 5, M + 1)
