"""oracle/assembly_restate.py — TEST INFRASTRUCTURE ONLY (never imported by petsc-dev_b200/).

numpy restatement of MatSetValuesBatch on an empty SeqAIJ matrix through the reference's host path
(src/mat/interface/matrix.c:1712-1716: one MatSetValues(A, nl, rows_e, nl, rows_e, v_e, ADD_VALUES) per element;
MatSetValues_SeqAIJ src/mat/impls/aij/seq/aij.c:293-368):
  * the contributions to nonzero (r, c) arrive ordered by (element, row-in-element, column-in-element); the first one
    creates the entry with that value (aij.c:350-366), every later one is added to it (aij.c:343) — one left-to-right sum
  * negative row ids (aij.c:311) and negative column ids (aij.c:325) are skipped; explicit zeros stay (ignorezeroentries
    is off by default, aij.c:329)
  * rows keep their columns ascending (aij.c:333-360), MatAssemblyEnd_SeqAIJ (aij.c:811-866) squeezes the slack out.
Pinned against the reference library itself (oracle/_ref, tests/test_assembly.py) and against the matrix printed in the
reference's golden output src/ksp/ksp/examples/tutorials/output/ex4_0.out.
"""
import numpy as np


def setvaluesbatch(m, n, elem_rows, elem_mats, nl, rstart=None):
    """rstart=None: the sequential m x n matrix.  rstart given: the local rows rstart .. rstart+m of a parallel matrix
    with n global columns — entries of rows owned elsewhere are left out (MatSetValues_MPIAIJ stashes them for their
    owner, mpiaij.c:496-570; the owner gets them by assembling the same element), column ids stay global."""
    rows = np.asarray(elem_rows, dtype=np.int64).reshape(-1, nl)
    vals = np.asarray(elem_mats, dtype=np.float64).reshape(-1, nl, nl)
    ne = rows.shape[0]
    r = np.repeat(rows, nl, axis=1).reshape(-1)                  # row of fat entry (e, a, b) = rows[e, a]
    c = np.tile(rows, (1, nl)).reshape(-1)                       # column            = rows[e, b]
    v = vals.reshape(-1)
    keep = (r >= 0) & (c >= 0)
    if rstart is not None:
        if np.any(r[keep] >= n):
            raise IndexError("row too large")
        keep &= (r >= rstart) & (r < rstart + m)
        r = r - rstart
    if np.any(r[keep] >= m) or np.any(c[keep] >= n):
        raise IndexError("row/column too large")
    r, c, v = r[keep], c[keep], v[keep]
    order = np.lexsort((c, r))                                   # stable: equal (r, c) keep contribution order
    r, c, v = r[order], c[order], v[order]
    head = np.ones(r.size, dtype=bool)
    head[1:] = (r[1:] != r[:-1]) | (c[1:] != c[:-1])
    seg = np.cumsum(head) - 1
    nz = int(head.sum())
    a = np.zeros(nz)
    a[...] = v[head]                                             # the entry starts as its first contribution
    np.add.at(a, seg[~head], v[~head])                           # unbuffered, in index order: left-to-right sums
    j = c[head].astype(np.int32)
    i = np.zeros(m + 1, dtype=np.int32)
    np.add.at(i, r[head] + 1, 1)
    return np.cumsum(i, dtype=np.int64).astype(np.int32), j, a
