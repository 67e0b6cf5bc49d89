"""PETSc binary Mat / Vec files (the on-disk format of the reference: MatView_SeqAIJ_Binary src/mat/impls/aij/seq/aij.c:414-452,
MatLoad_SeqAIJ aij.c:4093-4159, VecView_Seq_Binary / VecLoad; bin/pythonscripts/PetscBinaryIO.py documents the same layout).

Everything is big-endian.  Mat:  int32 classid 1211216, M, N, nz | int32 row lengths [M] | int32 column ids [nz] | float64
values [nz].  Vec:  int32 classid 1211214, n | float64 values [n].  Host-side I/O only (numpy), feeding Mat.from_csr /
Vec; pinned against the reference's own viewer in tests/test_binaryio.py."""
import numpy as np

MAT_FILE_CLASSID = 1211216      # include/petscmat.h:136
VEC_FILE_CLASSID = 1211214      # include/petscvec.h:113


class PetscBinaryError(ValueError):
    pass


def read_mat(path_or_file):
    """-> (M, N, ai[M+1] int32, aj[nz] int32, aa[nz] float64)."""
    f = open(path_or_file, "rb") if isinstance(path_or_file, (str, bytes)) else path_or_file
    try:
        hdr = np.fromfile(f, dtype=">i4", count=4)
        if hdr.size != 4 or hdr[0] != MAT_FILE_CLASSID:
            raise PetscBinaryError("not a PETSc matrix file (classid %s)" % (hdr[0] if hdr.size else "missing"))
        M, N, nz = int(hdr[1]), int(hdr[2]), int(hdr[3])
        if nz < 0:
            raise PetscBinaryError("dense-format matrix files (nz = -1) are not supported")      # MATRIX_BINARY_FORMAT_DENSE
        lens = np.fromfile(f, dtype=">i4", count=M)
        if lens.size != M or int(lens.sum()) != nz:
            raise PetscBinaryError("row lengths do not add up to nz (Inconsistant matrix data in file, aij.c:4139)")
        aj = np.fromfile(f, dtype=">i4", count=nz)
        aa = np.fromfile(f, dtype=">f8", count=nz)
        if aj.size != nz or aa.size != nz:
            raise PetscBinaryError("file ends before %d entries were read" % nz)
        ai = np.zeros(M + 1, np.int32)
        np.cumsum(lens, out=ai[1:])
        return M, N, ai, aj.astype(np.int32), aa.astype(np.float64)
    finally:
        if f is not path_or_file:
            f.close()


def write_mat(path, M, N, ai, aj, aa):
    ai = np.asarray(ai, np.int64)
    nz = int(ai[M])
    with open(path, "wb") as f:
        np.array([MAT_FILE_CLASSID, M, N, nz], dtype=">i4").tofile(f)
        np.diff(ai[:M + 1]).astype(">i4").tofile(f)
        np.asarray(aj[:nz]).astype(">i4").tofile(f)
        np.asarray(aa[:nz], np.float64).astype(">f8").tofile(f)


def read_vec(path_or_file):
    f = open(path_or_file, "rb") if isinstance(path_or_file, (str, bytes)) else path_or_file
    try:
        hdr = np.fromfile(f, dtype=">i4", count=2)
        if hdr.size != 2 or hdr[0] != VEC_FILE_CLASSID:
            raise PetscBinaryError("not a PETSc vector file (classid %s)" % (hdr[0] if hdr.size else "missing"))
        n = int(hdr[1])
        x = np.fromfile(f, dtype=">f8", count=n)
        if x.size != n:
            raise PetscBinaryError("file ends before %d entries were read" % n)
        return x.astype(np.float64)
    finally:
        if f is not path_or_file:
            f.close()


def write_vec(path, x):
    x = np.asarray(x, np.float64)
    with open(path, "wb") as f:
        np.array([VEC_FILE_CLASSID, x.size], dtype=">i4").tofile(f)
        x.astype(">f8").tofile(f)


def read_system(path):
    """An ex10-style file: a matrix followed (optionally) by a right-hand side vector in the same file
    (src/ksp/ksp/examples/tutorials/ex10.c:100-160).  -> (M, N, ai, aj, aa, b or None)."""
    with open(path, "rb") as f:
        M, N, ai, aj, aa = read_mat(f)
        b = None
        try:
            b = read_vec(f)
        except PetscBinaryError:
            b = None
    return M, N, ai, aj, aa, b
