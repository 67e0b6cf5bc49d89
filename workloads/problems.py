"""Synthetic AIJ (CSR) test/bench matrices for the BASELINE.json configurations.

Every generator is deterministic and partition independent: it can produce any contiguous global row range
[r0, r1) on its own (global column ids), so each rank of a row-partitioned MPIAIJ matrix builds its own rows with no
communication, and the CPU oracle and the GPU path are fed the same arrays.  Column ids are ascending within each
row, the invariant the reference's MatSetValues_SeqAIJ insertion maintains (src/mat/impls/aij/seq/aij.c:333-360).

  laplace2d_5pt   C1  src/ksp/ksp/examples/tutorials/ex2.c:88-96 (diag 4, -1 to the four neighbours, Ii = i*n + j)
  laplace3d_7pt   C2 / C3'  7-point Laplacian, natural ordering (x fastest), diag 6, -1 neighbours
  stencil3d_27pt  C3  27-point box stencil, diag 26, -1 neighbours
  elasticity_bs3  C5  L27 (x) M3 Kronecker product, bs = 3 interleaved unknowns
  powerlaw        C4  irregular rows, power-law row lengths, strictly diagonally dominant
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """Vectorised SplitMix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    with np.errstate(over="ignore"):
        z = (np.asarray(x, dtype=np.uint64) + np.uint64(0x9E3779B97F4A7C15))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def hash3(seed, a, b):
    """hash(seed, a, b) -> uint64, a/b broadcastable integer arrays."""
    with np.errstate(over="ignore"):
        h = splitmix64(np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + np.asarray(a, dtype=np.uint64))
        return splitmix64(h ^ (np.asarray(b, dtype=np.uint64) * np.uint64(0xA0761D6478BD642F)))


def u01(h):
    """uint64 hash -> float64 uniform in (0, 1)."""
    return ((h >> np.uint64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def random_vector(n, seed, lo=-1.0, hi=1.0, r0=0):
    """Entry g of the global vector is a pure function of (seed, g): rows [r0, r0+n)."""
    g = np.arange(r0, r0 + n, dtype=np.uint64)
    return lo + (hi - lo) * u01(hash3(seed, g, 0))


def _stencil(dims, offsets, diag, offv, r0, r1, bs=1, blk=None, chunk=1 << 22):
    """Generic structured-grid stencil rows [r0, r1) of an nx*ny*nz grid, natural ordering (x fastest).

    offsets: list of (dx, dy, dz) in ASCENDING linear-offset order (so columns come out sorted).
    With bs > 1 the scalar stencil is Kronecker-multiplied by the dense bs x bs block `blk` (row-major),
    unknown id = bs*node + c; r0/r1 then count scalar rows and must be multiples of bs.
    """
    nx, ny, nz = dims
    N = nx * ny * nz
    if r1 is None:
        r1 = N * bs
    assert r0 % bs == 0 and r1 % bs == 0
    n0, n1 = r0 // bs, r1 // bs
    offs = np.asarray(offsets, dtype=np.int64)
    lin = offs[:, 0] + nx * (offs[:, 1] + ny * offs[:, 2])
    assert np.all(np.diff(lin) > 0)
    indptr_parts, ind_parts, val_parts = [], [], []
    base = 0
    for c0 in range(n0, n1, chunk):
        c1 = min(n1, c0 + chunk)
        node = np.arange(c0, c1, dtype=np.int64)
        ix = node % nx
        iy = (node // nx) % ny
        iz = node // (nx * ny)
        X = ix[:, None] + offs[None, :, 0]
        Y = iy[:, None] + offs[None, :, 1]
        Z = iz[:, None] + offs[None, :, 2]
        ok = (X >= 0) & (X < nx) & (Y >= 0) & (Y < ny) & (Z >= 0) & (Z < nz)
        cols = node[:, None] + lin[None, :]
        isdiag = (lin == 0)[None, :]
        vals = np.where(isdiag, float(diag), float(offv)) * np.ones_like(cols, dtype=np.float64)
        cnt = ok.sum(axis=1)
        if bs == 1:
            ind_parts.append(cols[ok].astype(np.int32))
            val_parts.append(vals[ok])
            ip = np.empty(c1 - c0, dtype=np.int64)
            np.cumsum(cnt, out=ip)
            indptr_parts.append(ip + base)
            base += int(cnt.sum())
        else:
            B = np.asarray(blk, dtype=np.float64).reshape(bs, bs)
            # expand (node, off) -> (node, r, off, c): row-major flattening under the mask is exactly
            # "scalar rows in order, columns ascending"
            v4 = vals[:, None, :, None] * B[None, :, None, :]
            c4 = np.broadcast_to(cols[:, None, :, None] * bs + np.arange(bs)[None, None, None, :], v4.shape)
            m4 = np.broadcast_to(ok[:, None, :, None], v4.shape)
            ind_parts.append(c4[m4].astype(np.int32))
            val_parts.append(v4[m4])
            scnt = np.repeat(cnt * bs, bs)
            ip = np.empty(scnt.size, dtype=np.int64)
            np.cumsum(scnt, out=ip)
            indptr_parts.append(ip + base)
            base += int(scnt.sum())
    indptr = np.concatenate([np.zeros(1, np.int64)] + indptr_parts)
    indices = np.concatenate(ind_parts) if ind_parts else np.zeros(0, np.int32)
    data = np.concatenate(val_parts) if val_parts else np.zeros(0, np.float64)
    return indptr, indices, data


def laplace2d_5pt(m, n, r0=0, r1=None):
    """ex2.c: N = m*n, row Ii = i*n + j; neighbours Ii-n (i>0), Ii-1 (j>0), Ii+1 (j<n-1), Ii+n (i<m-1)."""
    # grid with x = j (fastest, size n), y = i (size m)
    return _stencil((n, m, 1), [(0, -1, 0), (-1, 0, 0), (0, 0, 0), (1, 0, 0), (0, 1, 0)], 4.0, -1.0, r0, r1)


_OFF7 = [(0, 0, -1), (0, -1, 0), (-1, 0, 0), (0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1)]
_OFF27 = [(dx, dy, dz) for dz in (-1, 0, 1) for dy in (-1, 0, 1) for dx in (-1, 0, 1)]


def laplace3d_7pt(nx, ny=None, nz=None, r0=0, r1=None):
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _stencil((nx, ny, nz), _OFF7, 6.0, -1.0, r0, r1)


def stencil3d_27pt(nx, ny=None, nz=None, r0=0, r1=None):
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _stencil((nx, ny, nz), _OFF27, 26.0, -1.0, r0, r1)


M3 = np.array([[2.0, 0.5, 0.25], [0.5, 2.0, 0.5], [0.25, 0.5, 2.0]])


def elasticity_bs3(nx, ny=None, nz=None, r0=0, r1=None):
    """A = L27 (x) M3 (SPD Kronecker product), unknown id = 3*node + c."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _stencil((nx, ny, nz), _OFF27, 26.0, -1.0, r0, r1, bs=3, blk=M3, chunk=1 << 16)


def powerlaw_row_lengths(N, r0, r1, seed=4, xm=17.0, a=1.5, lmax=100000):
    r = np.arange(r0, r1, dtype=np.uint64)
    U = u01(hash3(seed, r, 0))
    L = np.floor(xm * U ** (-1.0 / a))
    return np.clip(L, 1, min(lmax, N)).astype(np.int64)


def powerlaw(N, r0=0, r1=None, seed=4, xm=17.0, a=1.5, lmax=100000, chunk=1 << 18):
    """Irregular matrix: row r has the diagonal plus (L_r - 1) hashed columns (duplicates merged), sorted ascending;
    off-diagonal values uniform(-1, 0)/L_r, diagonal 1 + sum |offdiag| (strictly diagonally dominant)."""
    if r1 is None:
        r1 = N
    indptr_parts, ind_parts, val_parts = [], [], []
    base = 0
    for c0 in range(r0, r1, chunk):
        c1 = min(r1, c0 + chunk)
        L = powerlaw_row_lengths(N, c0, c1, seed, xm, a, lmax)
        rows = np.repeat(np.arange(c0, c1, dtype=np.int64), L)
        start = np.concatenate([[0], np.cumsum(L)[:-1]])
        k = np.arange(rows.size, dtype=np.int64) - np.repeat(start, L)
        cols = (hash3(seed + 1, rows, k) % np.uint64(N)).astype(np.int64)
        cols[k == 0] = rows[k == 0]                       # slot 0 is the diagonal
        key = rows * np.int64(N) + cols if N < (1 << 31) else None
        order = np.lexsort((cols, rows)) if key is None else np.argsort(key, kind="stable")
        rows, cols, k = rows[order], cols[order], k[order]
        keep = np.ones(rows.size, dtype=bool)
        keep[1:] = (rows[1:] != rows[:-1]) | (cols[1:] != cols[:-1])
        # a hashed column that collides with the diagonal must not displace it: mark the diagonal entry
        rows, cols = rows[keep], cols[keep]
        isdiag = rows == cols
        Lrow = L[rows - c0].astype(np.float64)
        v = -u01(hash3(seed + 2, rows, cols)) / Lrow
        v[isdiag] = 0.0
        cnt = np.bincount(rows - c0, minlength=c1 - c0)
        offsum = np.bincount(rows - c0, weights=np.abs(v), minlength=c1 - c0)
        v[isdiag] = 1.0 + offsum[rows[isdiag] - c0]
        ind_parts.append(cols.astype(np.int32))
        val_parts.append(v)
        indptr_parts.append(np.cumsum(cnt) + base)
        base += int(cnt.sum())
    indptr = np.concatenate([np.zeros(1, np.int64)] + indptr_parts)
    return indptr, np.concatenate(ind_parts), np.concatenate(val_parts)


def split_ownership(N, size, bs=1):
    """PetscSplitOwnership / PetscLayoutSetUp (src/sys/utils/psplit.c:88, src/vec/vec/impls/mpi/pmap.c:142-159):
    n_r = N/size + ((N % size) > r), on block counts when bs > 1.  Returns range[size+1]."""
    nb = N // bs
    n = np.array([nb // size + (1 if (nb % size) > r else 0) for r in range(size)], dtype=np.int64) * bs
    return np.concatenate([[0], np.cumsum(n)])


# ---- element batches for MatSetValuesBatch (src/mat/interface/matrix.c:1698) ---------------------------------------
def elements_tri2d(X, Y):
    """The mesh of src/ksp/ksp/examples/tutorials/ex4.c:46-111 on an X x Y vertex grid: every cell is cut into two
    triangles with the element matrix [[1, -.5, -.5], [-.5, .5, 0], [-.5, 0, .5]] (explicit zeros included).
    Returns (N, nl, elem_rows int32[ne*3], elem_mats float64[ne*9])."""
    j, i = np.meshgrid(np.arange(Y - 1), np.arange(X - 1), indexing="ij")
    a = (j * X + i).reshape(-1)
    b, c, d = a + 1, a + X, a + X + 1
    rows = np.stack([a, b, c, d, c, b], axis=1).astype(np.int32).reshape(-1)      # lower (A,B,C) then upper (D,C,B)
    em = np.array([1.0, -0.5, -0.5, -0.5, 0.5, 0.0, -0.5, 0.0, 0.5])
    mats = np.tile(em, 2 * a.size)
    return X * Y, 3, rows, mats


def elements_hex3d(nx, seed=11, scale_values=True):
    """Q1 hexahedra on an nx^3 vertex grid (nl = 8): a fixed symmetric 8 x 8 pattern matrix times a per-element
    coefficient hashed from the element id, so the contributions that meet in a nonzero all differ (the summation
    order matters for the last bit).  Returns (N, nl, elem_rows, elem_mats)."""
    ne1 = nx - 1
    k, j, i = np.meshgrid(np.arange(ne1), np.arange(ne1), np.arange(ne1), indexing="ij")
    v0 = ((k * nx + j) * nx + i).reshape(-1).astype(np.int64)
    corner = np.array([0, 1, nx, nx + 1, nx * nx, nx * nx + 1, nx * nx + nx, nx * nx + nx + 1], dtype=np.int64)
    rows = (v0[:, None] + corner[None, :]).astype(np.int32).reshape(-1)
    d = np.abs(np.arange(8)[:, None] ^ np.arange(8)[None, :])
    hamming = (d & 1) + ((d >> 1) & 1) + ((d >> 2) & 1)
    base = np.array([4.0, 0.0, -1.0, -1.0])[hamming] / 3.0                          # Laplacian-like: row sums are 0
    ne = v0.size
    coef = 0.5 + u01(hash3(seed, np.arange(ne, dtype=np.uint64), 7)) if scale_values else np.ones(ne)
    mats = (coef[:, None] * base.reshape(1, 64)).reshape(-1)
    return nx ** 3, 8, rows, mats


def random_vector_at(idx, seed, lo=-1.0, hi=1.0):
    """Entries idx (global ids) of the vector random_vector(N, seed) — what a rank's ghost vector holds."""
    return lo + (hi - lo) * u01(hash3(seed, np.asarray(idx, dtype=np.uint64), 0))


def periodic_z_loopback(nx, ny, nz):
    """The 7-point Laplacian, periodic in z, written the way ONE rank of a parallel matrix would hold it if its only
    neighbour were itself: A_d = the non-periodic stencil (local columns), B = the wrap-around couplings (-1) with ghost
    ids, to_idx = the local entry every ghost is a copy of (ghosts [0, nx*ny) mirror the top plane, the rest the bottom
    plane).  Returns dict(m, Ad=(i,j,a), B=(i,j,a), nghost, to_idx, full=(i,j,a) — the same operator as one SeqAIJ matrix)."""
    import scipy.sparse as sp
    m, npl = nx * ny * nz, nx * ny
    ai, aj, aa = laplace3d_7pt(nx, ny, nz)
    rows_bot, rows_top = np.arange(npl), np.arange(m - npl, m)
    bcnt = np.zeros(m, np.int64)
    bcnt[rows_bot] = 1
    bcnt[rows_top] += 1                                   # nz == 1 would double up: not used
    bi = np.concatenate([[0], np.cumsum(bcnt)])
    bj = np.concatenate([np.arange(npl), np.arange(npl, 2 * npl)]).astype(np.int32)     # bottom rows read ghosts 0.., top rows the rest
    ba = -np.ones(2 * npl)
    to_idx = np.concatenate([rows_top, rows_bot]).astype(np.int32)
    A = sp.csr_matrix((aa, aj, ai), shape=(m, m))
    W = sp.csr_matrix((ba, to_idx[bj], bi), shape=(m, m))
    F = (A + W).tocsr()
    F.sort_indices()
    return dict(m=m, Ad=(ai, aj, aa), B=(bi, bj, ba), nghost=2 * npl, to_idx=to_idx,
                full=(F.indptr.astype(np.int64), F.indices.astype(np.int32), F.data.astype(np.float64)))
