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


# ---- striped reads: every rank fetches its own row slab (MatLoad_MPIAIJ, src/mat/impls/aij/mpi/mpiaij.c:3416-3600, has
# rank 0 read the file and send the slabs out; with one process per GPU and a shared file every rank seeks instead) ----
def read_mat_header(path):
    """-> (M, N, nz, offset of the row-length block)."""
    with open(path, "rb") as f:
        hdr = np.fromfile(f, dtype=">i4", count=4)
    if hdr.size != 4 or hdr[0] != MAT_FILE_CLASSID:
        raise PetscBinaryError("not a PETSc matrix file (classid %s)" % (hdr[0] if hdr.size else "missing"))
    if hdr[3] < 0:
        raise PetscBinaryError("dense-format matrix files (nz = -1) are not supported")
    return int(hdr[1]), int(hdr[2]), int(hdr[3]), 16


def read_mat_rows(path, r0, r1):
    """Rows [r0, r1) only -> (M, N, ai[r1-r0+1] starting at 0, aj (global column ids), aa).  Reads the row lengths up to r1
    (to locate the slab) and then exactly the slab's column ids and values."""
    M, N, nz, off = read_mat_header(path)
    if not (0 <= r0 <= r1 <= M):
        raise PetscBinaryError("rows %d..%d are not inside 0..%d" % (r0, r1, M))
    with open(path, "rb") as f:
        f.seek(off)
        lens = np.fromfile(f, dtype=">i4", count=r1).astype(np.int64)
        if lens.size != r1:
            raise PetscBinaryError("file ends inside the row lengths")
        k0, k1 = int(lens[:r0].sum()), int(lens.sum())
        if k1 > nz:
            raise PetscBinaryError("row lengths exceed nz (Inconsistant matrix data in file, aij.c:4139)")
        f.seek(off + 4 * M + 4 * k0)
        aj = np.fromfile(f, dtype=">i4", count=k1 - k0)
        f.seek(off + 4 * M + 4 * nz + 8 * k0)
        aa = np.fromfile(f, dtype=">f8", count=k1 - k0)
    if aj.size != k1 - k0 or aa.size != k1 - k0:
        raise PetscBinaryError("file ends before the slab's %d entries were read" % (k1 - k0))
    ai = np.zeros(r1 - r0 + 1, np.int32)
    np.cumsum(lens[r0:r1], out=ai[1:])
    return M, N, ai, aj.astype(np.int32), aa.astype(np.float64)


def mat_file_size(path):
    """Bytes the matrix occupies at the start of the file (a right-hand side, if any, follows)."""
    M, N, nz, off = read_mat_header(path)
    return off + 4 * M + 12 * nz


def read_vec_rows(path, r0, r1, offset=0):
    """Entries [r0, r1) of the vector that starts `offset` bytes into the file; None when no vector starts there."""
    with open(path, "rb") as f:
        f.seek(offset)
        hdr = np.fromfile(f, dtype=">i4", count=2)
        if hdr.size != 2 or hdr[0] != VEC_FILE_CLASSID:
            return None
        n = int(hdr[1])
        if not (0 <= r0 <= r1 <= n):
            raise PetscBinaryError("entries %d..%d are not inside 0..%d" % (r0, r1, n))
        f.seek(offset + 8 + 8 * r0)
        x = np.fromfile(f, dtype=">f8", count=r1 - r0)
    if x.size != r1 - r0:
        raise PetscBinaryError("file ends before %d entries were read" % (r1 - r0))
    return x.astype(np.float64)
