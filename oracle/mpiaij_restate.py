"""oracle/mpiaij_restate.py — TEST INFRASTRUCTURE ONLY (never imported by the product).

Multi-rank oracle for the row-partitioned (MPIAIJ) operator.  The reference library built in oracle/_ref is a
one-rank MPIUNI build (no MPI exists on this box), so for P > 1 the INTEGER structures are restated here in numpy,
independently of the C implementation under test, by looping over virtual ranks and following
  PetscSplitOwnership / PetscLayoutSetUp     src/sys/utils/psplit.c:88, src/vec/vec/impls/mpi/pmap.c:142-159
  MatSetValues_MPIAIJ column routing         src/mat/impls/aij/mpi/mpiaij.c:507-565
  MatSetUpMultiply_MPIAIJ                    src/mat/impls/aij/mpi/mmaij.c:28-64,106-143 (garray, renumbered B->j, block IS)
  VecScatterCreate_PtoS                      src/vec/vec/utils/vpscat.c:1759-1903 (owner scan, to/from lists)
Parity status of these integer lists for P > 1: restated, pinned by uniqueness (garray = sorted distinct ghost ids and
the renumbering are uniquely determined by the definition; the to/from lists are determined up to the orderings the
reference fixes at vpscat.c:1780 and :1798-1806, which are followed here) and by the P = 1 reference run; NOT pinned
by a multi-rank run of the reference itself.

The NUMERICAL oracle for P > 1 is the real reference code: rank r's product is MatMult_SeqAIJ on its diagonal block
followed by MatMultAdd_SeqAIJ on its off-diagonal block with the gathered ghost values (mpiaij.c:1111-1114), both
executed by oracle/_ref through oracle/refpetsc.py.
"""
import numpy as np


def split_ownership(N, size, bs=1):
    nb = N // bs
    n = [(nb // size + (1 if (nb % size) > r else 0)) * bs for r in range(size)]
    return np.concatenate([[0], np.cumsum(n)]).astype(np.int64)


def rank_structures(rank, ranges, ai, aj, aa, bs=1):
    """ai/aj/aa: rows [ranges[rank], ranges[rank+1]) with global column ids."""
    rs, re_ = int(ranges[rank]), int(ranges[rank + 1])
    m = re_ - rs
    ai = np.asarray(ai, np.int64); aj = np.asarray(aj, np.int64); aa = np.asarray(aa, np.float64)
    rows = np.repeat(np.arange(m), np.diff(ai))
    isA = (aj >= rs) & (aj < re_)
    Ai = np.concatenate([[0], np.cumsum(np.bincount(rows[isA], minlength=m))])
    Bi = np.concatenate([[0], np.cumsum(np.bincount(rows[~isA], minlength=m))])
    garray = np.unique(aj[~isA])
    Bj = np.searchsorted(garray, aj[~isA])
    out = dict(m=m, Ai=Ai.astype(np.int32), Aj=(aj[isA] - rs).astype(np.int32), Aa=aa[isA],
               Bi=Bi.astype(np.int32), Bj=Bj.astype(np.int32), Ba=aa[~isA], garray=garray.astype(np.int32))
    # block index set?
    blk = False
    if bs > 1 and garray.size % bs == 0:
        g2 = garray.reshape(-1, bs)
        blk = bool(np.all(g2[:, 0] % bs == 0) and np.all(g2 == g2[:, :1] + np.arange(bs)))
    out["block_is"] = blk
    pbs = bs if blk else 1
    inidx = garray[::pbs] // pbs
    inidy = np.arange(inidx.size)
    owner = np.searchsorted(np.asarray(ranges), pbs * inidx, side="right") - 1
    procs = np.unique(owner)
    out["bs"] = pbs
    out["from_procs"] = procs.astype(np.int32)
    out["from_starts"] = np.concatenate([[0], np.cumsum([np.sum(owner == p) for p in procs])]).astype(np.int32)
    out["from_indices"] = np.concatenate([pbs * inidy[owner == p] for p in procs]).astype(np.int32) if procs.size else np.zeros(0, np.int32)
    out["requests"] = {int(p): (pbs * inidx[owner == p]).astype(np.int32) for p in procs}
    return out


def all_structures(size, N, row_generator, bs=1):
    """row_generator(r0, r1) -> (ai, aj, aa).  Returns (ranges, [per-rank dict incl. the `to` lists])."""
    ranges = split_ownership(N, size, bs)
    S = []
    for r in range(size):
        ai, aj, aa = row_generator(int(ranges[r]), int(ranges[r + 1]))
        S.append(rank_structures(r, ranges, ai, aj, aa, bs))
    for p in range(size):
        procs, starts, idx = [], [0], []
        for r in range(size):                                      # ascending requester (vpscat.c:1780)
            if r != p and p in S[r]["requests"]:
                q = S[r]["requests"][p]
                procs.append(r); starts.append(starts[-1] + q.size); idx.append(q - int(ranges[p]))
        S[p]["to_procs"] = np.array(procs, np.int32)
        S[p]["to_starts"] = np.array(starts, np.int32)
        S[p]["to_indices"] = np.concatenate(idx).astype(np.int32) if idx else np.zeros(0, np.int32)
    return ranges, S


def mult(R, ranges, S, x, add_y=None):
    """MatMult_MPIAIJ (or MatMultAdd with add_y) over all virtual ranks with the reference's sequential kernels."""
    size = len(S)
    y = np.empty(int(ranges[-1]))
    for r in range(size):
        s = S[r]
        rs, re_ = int(ranges[r]), int(ranges[r + 1])
        m = s["m"]
        A = R.RefMat(m, m, s["Ai"], s["Aj"], s["Aa"], no_inode=True)
        lvec = x[s["garray"]]                                       # the scatter is a pure copy (INSERT_VALUES)
        if add_y is None:
            yr = A.mult(x[rs:re_])
        else:
            yr = A.multadd(x[rs:re_], add_y[rs:re_])
        if s["garray"].size:
            B = R.RefMat(m, s["garray"].size, s["Bi"], s["Bj"], s["Ba"], no_inode=True)
            yr = B.multadd(lvec, yr, inplace=True)
        y[rs:re_] = yr
    return y


def mult_transpose(R, ranges, S, x, add_y=None):
    """MatMultTranspose_MPIAIJ / MatMultTransposeAdd_MPIAIJ (mpiaij.c:1147-1195) over all virtual ranks with the reference's
    sequential kernels: lvec_r = B_r^T x_r, y_r = A_r^T x_r (+ y_r), then the reverse scatter with ADD_VALUES; the messages
    an owner receives are added in ascending source-rank order (the reference adds them in arrival order)."""
    size = len(S)
    y = np.empty(int(ranges[-1]))
    lv = []
    for r in range(size):
        s = S[r]
        rs, re_ = int(ranges[r]), int(ranges[r + 1])
        m = s["m"]
        A = R.RefMat(m, m, s["Ai"], s["Aj"], s["Aa"], no_inode=True)
        y[rs:re_] = A.multtranspose(x[rs:re_], None if add_y is None else add_y[rs:re_])
        if s["garray"].size:
            B = R.RefMat(m, s["garray"].size, s["Bi"], s["Bj"], s["Ba"], no_inode=True)
            lv.append(B.multtranspose(x[rs:re_]))
        else:
            lv.append(np.zeros(0))
    for p in range(size):
        ps, pe = int(ranges[p]), int(ranges[p + 1])
        for r in range(size):
            if r == p or not S[r]["garray"].size:
                continue
            g = S[r]["garray"]
            sel = (g >= ps) & (g < pe)
            y[g[sel]] = y[g[sel]] + lv[r][sel]
    return y
