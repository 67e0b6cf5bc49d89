"""Row-partitioned (MPIAIJ) assembly over one-process-per-GPU ranks.

Host-side plumbing only: the integer structures come from the C ABI (b200_mpiaij_split, b200_scatter_plan_from/_to —
restatements of MatSetUpMultiply_MPIAIJ, src/mat/impls/aij/mpi/mmaij.c:9-161, and VecScatterCreate_PtoS,
src/vec/vec/utils/vpscat.c:1730-1924); the request lists travel between ranks with torch.distributed (gloo on CPU,
nccl on GPUs), which stands in for the Isend/Irecv of vpscat.c:1778-1822.
"""
import ctypes as C

import numpy as np

from . import Mat, PetscB200Error, _chk, _dp, _f64, _i32, _ip, lib


def split_ownership(N, size, bs=1):
    """PetscSplitOwnership / PetscLayoutSetUp -> range[size+1]."""
    r = np.zeros(size + 1, np.int32)
    _chk(lib.b200_split_ownership(int(N), int(size), int(bs), _ip(r)), "PetscLayoutSetUp")
    return r


def mpiaij_split(m, cstart, cend, ai, aj, aa):
    """Local rows (global column ids) -> dict(Ai, Aj, Aa, Bi, Bj, Ba, garray)."""
    ai, aj, aa = _i32(ai), _i32(aj), _f64(aa)
    nzA, nzB, ng = C.c_longlong(), C.c_longlong(), C.c_int()
    _chk(lib.b200_mpiaij_split(int(m), int(cstart), int(cend), _ip(ai), _ip(aj), _dp(aa), None, None, None, None, None,
                               None, None, C.byref(nzA), C.byref(nzB), C.byref(ng)), "MatSetUpMultiply (sizes)")
    out = dict(Ai=np.zeros(m + 1, np.int32), Aj=np.zeros(max(nzA.value, 1), np.int32), Aa=np.zeros(max(nzA.value, 1)),
               Bi=np.zeros(m + 1, np.int32), Bj=np.zeros(max(nzB.value, 1), np.int32), Ba=np.zeros(max(nzB.value, 1)),
               garray=np.zeros(max(ng.value, 1), np.int32))
    _chk(lib.b200_mpiaij_split(int(m), int(cstart), int(cend), _ip(ai), _ip(aj), _dp(aa), _ip(out["Ai"]), _ip(out["Aj"]),
                               _dp(out["Aa"]), _ip(out["Bi"]), _ip(out["Bj"]), _dp(out["Ba"]), _ip(out["garray"]),
                               C.byref(nzA), C.byref(nzB), C.byref(ng)), "MatSetUpMultiply")
    out["Aj"], out["Aa"] = out["Aj"][:nzA.value], out["Aa"][:nzA.value]
    out["Bj"], out["Ba"] = out["Bj"][:nzB.value], out["Ba"][:nzB.value]
    out["garray"] = out["garray"][:ng.value]
    return out


def colmap(N, garray):
    cm = np.zeros(N, np.int32)
    g = _i32(garray)
    _chk(lib.b200_mpiaij_colmap(int(N), int(g.size), _ip(g), _ip(cm)), "MatCreateColmap")
    return cm


def use_block_is(garray, bs):
    g = _i32(garray)
    u = C.c_int()
    _chk(lib.b200_mpiaij_use_block_is(int(g.size), _ip(g), int(bs), C.byref(u)), "useblockis")
    return bool(u.value)


def scatter_plan_from(inidx, inidy, bs, ranges, rank, lengthy):
    """Receive side + request lists of VecScatterCreate_PtoS for this rank."""
    inidx, inidy, ranges = _i32(inidx), _i32(inidy), _i32(ranges)
    nx, size = inidx.size, ranges.size - 1
    nfrom, nlocal = C.c_int(), C.c_int()
    fp, fs, fi = np.zeros(size, np.int32), np.zeros(size + 1, np.int32), np.zeros(max(nx, 1), np.int32)
    rc, rv = np.zeros(size, np.int32), np.zeros(max(nx, 1), np.int32)
    tl, fl = np.zeros(max(nx, 1), np.int32), np.zeros(max(nx, 1), np.int32)
    _chk(lib.b200_scatter_plan_from(nx, _ip(inidx), _ip(inidy), int(bs), size, _ip(ranges), int(rank), int(lengthy),
                                    C.byref(nfrom), _ip(fp), _ip(fs), _ip(fi), _ip(rc), _ip(rv), C.byref(nlocal),
                                    _ip(tl), _ip(fl)), "VecScatterCreate_PtoS (from)")
    nf, nl = nfrom.value, nlocal.value
    nreq = int(rc.sum())
    return dict(nfrom=nf, from_procs=fp[:nf].copy(), from_starts=fs[:nf + 1].copy(), from_indices=fi[:nreq].copy(),
                req_counts=rc, req_values=rv[:nreq].copy(), nlocal=nl, to_local=tl[:nl].copy(), from_local=fl[:nl].copy())


def scatter_plan_to(procs, lengths, requested, rstart, rend):
    lengths, requested = _i32(lengths), _i32(requested)
    nto = lengths.size
    ts, ti = np.zeros(nto + 1, np.int32), np.zeros(max(requested.size, 1), np.int32)
    _chk(lib.b200_scatter_plan_to(nto, _ip(lengths), _ip(requested), int(rstart), int(rend), _ip(ts), _ip(ti)),
         "VecScatterCreate_PtoS (to)")
    return dict(nto=nto, to_procs=_i32(procs), to_starts=ts, to_indices=ti[:requested.size].copy())


def exchange_requests(plan, rank, size, dist=None, device="cpu"):
    """PetscGatherMessageLengths + the Isend/Irecv of the wanted-index lists (vpscat.c:1778-1822) over
    torch.distributed: returns (procs ascending, lengths, concatenated requested ids) as seen by the OWNER."""
    import torch
    import torch.distributed as td
    dist = dist or td
    counts = torch.tensor(plan["req_counts"], dtype=torch.int32, device=device)
    allc = [torch.zeros(size, dtype=torch.int32, device=device) for _ in range(size)]
    dist.all_gather(allc, counts)
    allc = torch.stack(allc).cpu().numpy()                 # allc[r][p] = what rank r wants from p
    # index lists: all_gather of the padded request arrays (setup-time, tiny next to the matrix)
    maxlen = int(allc.sum(axis=1).max()) if size else 0
    mine = torch.zeros(max(maxlen, 1), dtype=torch.int32, device=device)
    rv = plan["req_values"]
    if rv.size:
        mine[:rv.size] = torch.tensor(rv, dtype=torch.int32, device=device)
    allv = [torch.zeros_like(mine) for _ in range(size)]
    dist.all_gather(allv, mine)
    procs, lengths, req = [], [], []
    for r in range(size):                                  # ascending requester rank = PetscSortMPIIntWithArray order
        c = int(allc[r][rank])
        if r == rank or c == 0:
            continue
        off = int(allc[r][:rank].sum())
        procs.append(r); lengths.append(c)
        req.append(allv[r][off:off + c].cpu().numpy())
    req = np.concatenate(req).astype(np.int32) if req else np.zeros(0, np.int32)
    return np.array(procs, np.int32), np.array(lengths, np.int32), req


def setup_mpiaij_host(rank, size, ranges, ai, aj, aa, N, bs=1, dist=None, device="cpu"):
    """Everything MatAssemblyEnd_MPIAIJ + MatSetUpMultiply_MPIAIJ decide on the host for this rank."""
    rstart, rend = int(ranges[rank]), int(ranges[rank + 1])
    m = rend - rstart
    sp = mpiaij_split(m, rstart, rend, ai, aj, aa)
    return dict(plan_from_garray(rank, size, ranges, sp["garray"], N, bs=bs, dist=dist, device=device), split=sp)


def plan_from_garray(rank, size, ranges, garray, N, bs=1, dist=None, device="cpu"):
    """The VecScatter of MatSetUpMultiply_MPIAIJ (mmaij.c:106-161) from the ghost column list alone."""
    rstart, rend = int(ranges[rank]), int(ranges[rank + 1])
    m = rend - rstart
    g = _i32(garray)
    blk = use_block_is(g, bs)
    if blk:                                                # ISCreateBlock(bs, ec/bs, garray[i*bs]/bs), mmaij.c:130-136
        inidx = (g[::bs] // bs).astype(np.int32)
        inidy = np.arange(g.size // bs, dtype=np.int32)
        pbs = bs
    else:
        inidx, inidy, pbs = g, np.arange(g.size, dtype=np.int32), 1
    frm = scatter_plan_from(inidx, inidy, pbs, ranges, rank, g.size)
    if size > 1:
        procs, lengths, req = exchange_requests(frm, rank, size, dist, device)
    else:
        procs, lengths, req = np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32)
    to = scatter_plan_to(procs, lengths, req, rstart, rend)
    return dict(from_=frm, to=to, bs=pbs, m=m, rstart=rstart, rend=rend, N=int(N), block_is=blk)


def expand_bs(starts, indices, bs):
    """Block lists -> scalar lists (what VecScatterInitializeForGPU does, vscatcusp.c:60-100)."""
    if bs == 1:
        return _i32(starts), _i32(indices)
    idx = (np.asarray(indices, np.int64)[:, None] + np.arange(bs)[None, :]).reshape(-1)
    return _i32(np.asarray(starts, np.int64) * bs), _i32(idx)


def create_mpiaij(host):
    """Upload: the device-resident MATMPIAIJB200 of this rank."""
    sp, frm, to, bs = host["split"], host["from_"], host["to"], host["bs"]
    ts, ti = expand_bs(to["to_starts"], to["to_indices"], bs)
    fs, _ = expand_bs(frm["from_starts"], frm["from_indices"], bs)
    h = C.c_void_p()
    m = host["m"]
    g = _i32(sp["garray"])
    tp, fp = _i32(to["to_procs"]), _i32(frm["from_procs"])
    _chk(lib.b200_mat_create_mpiaij(int(m), int(m), int(host["N"]), _ip(sp["Ai"]), _ip(_i32(sp["Aj"])), _dp(_f64(sp["Aa"])),
                                    int(g.size), _ip(sp["Bi"]), _ip(_i32(sp["Bj"])), _dp(_f64(sp["Ba"])), _ip(g),
                                    int(to["nto"]), _ip(tp), _ip(ts), _ip(ti),
                                    int(frm["nfrom"]), _ip(fp), _ip(fs), C.byref(h)), "MatAssemblyEnd_MPIAIJ (upload)")
    return Mat(h, m, m)


def comm_init(rank, size, dist=None, device=None):
    """ncclCommInitRank for the library's own communicator; the id is broadcast with torch.distributed."""
    import torch
    import torch.distributed as td
    dist = dist or td
    buf = (C.c_ubyte * 128)()
    if size > 1:
        if rank == 0:
            _chk(lib.b200_comm_get_unique_id(buf), "ncclGetUniqueId")
        t = torch.tensor(list(bytes(buf)), dtype=torch.uint8, device=device or "cpu")
        dist.broadcast(t, 0)
        raw = bytes(t.cpu().numpy().tolist())
        buf = (C.c_ubyte * 128).from_buffer_copy(raw)
    _chk(lib.b200_comm_init(int(rank), int(size), buf), "b200_comm_init")


def enable_p2p(rank=None, size=None, dist=None, strict=True):
    """Switch the ghost exchange and the scalar allreduces to NVLink peer memory (collective; after comm_init).  The
    CUDA IPC handles travel over the library's own NCCL communicator, so nothing is needed from the host program.  The
    library agrees on the outcome across ranks: either every rank is switched, or none is and every rank gets the same
    error (strict=False: return False instead of raising, the NCCL transport stays in use everywhere)."""
    rc = lib.b200_comm_enable_p2p(1)
    if rc and not strict:
        return False
    _chk(rc, "b200_comm_enable_p2p")
    return True


def enable_loopback():
    """Peer-memory transport on a ONE-rank communicator: the rank is its own neighbour (b200_comm_enable_p2p(2)).  The
    exchange kernels, arrival flags and credits run as between GPUs; used by the single-GPU tests of the multi-GPU path."""
    _chk(lib.b200_comm_enable_p2p(2), "b200_comm_enable_p2p (loop-back)")


def create_mpiaij_raw(m, N, Ai, Aj, Aa, Bi, Bj, Ba, garray, to_procs, to_starts, to_indices, from_procs, from_starts):
    """b200_mat_create_mpiaij from explicit pieces (tests)."""
    Ai, Aj, Aa, Bi, Bj, Ba = _i32(Ai), _i32(Aj), _f64(Aa), _i32(Bi), _i32(Bj), _f64(Ba)
    g, tp, ts, ti, fp, fs = _i32(garray), _i32(to_procs), _i32(to_starts), _i32(to_indices), _i32(from_procs), _i32(from_starts)
    h = C.c_void_p()
    _chk(lib.b200_mat_create_mpiaij(int(m), int(m), int(N), _ip(Ai), _ip(Aj), _dp(Aa), int(g.size), _ip(Bi), _ip(Bj), _dp(Ba), _ip(g),
                                    int(tp.size), _ip(tp), _ip(ts), _ip(ti), int(fp.size), _ip(fp), _ip(fs), C.byref(h)),
         "MatAssemblyEnd_MPIAIJ (upload)")
    return Mat(h, m, m)


def is_fused(A):
    """(one-launch MatMult?, number of boundary tiles)"""
    f, nb = C.c_int(), C.c_int()
    _chk(lib.b200_mat_mpiaij_is_fused(A.h, C.byref(f), C.byref(nb)), "b200_mat_mpiaij_is_fused")
    return bool(f.value), nb.value


def p2p_enabled():
    e = C.c_int()
    _chk(lib.b200_comm_p2p_enabled(C.byref(e)), "b200_comm_p2p_enabled")
    return bool(e.value)


class Scatter:
    """Device-resident VecScatter between this rank's part of a parallel vector and a sequential vector
    (VecScatterCreate_PtoS + VecScatterBegin/End, both modes and InsertModes)."""

    def __init__(self, handle, nx, ny):
        self.h, self.nx, self.ny = handle, nx, ny

    def begin(self, x, y, add=False, reverse=False):
        _chk(lib.b200_scatter_begin(self.h, x.p, y.p, int(bool(add)), int(bool(reverse))), "VecScatterBegin")

    def end(self, x, y, add=False, reverse=False):
        _chk(lib.b200_scatter_end(self.h, x.p, y.p, int(bool(add)), int(bool(reverse))), "VecScatterEnd")

    def scatter(self, x, y, add=False, reverse=False):
        self.begin(x, y, add, reverse)
        self.end(x, y, add, reverse)

    def destroy(self):
        if self.h:
            lib.b200_scatter_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


def create_scatter(rank, size, ranges, inidx, inidy, ny, bs=1, dist=None, device="cpu"):
    """VecScatterCreate(x parallel, ix = inidx (global, block ids when bs > 1), y sequential of length ny, iy = inidy)."""
    rstart, rend = int(ranges[rank]), int(ranges[rank + 1])
    frm = scatter_plan_from(inidx, inidy, bs, ranges, rank, ny)
    if size > 1:
        procs, lengths, req = exchange_requests(frm, rank, size, dist, device)
    else:
        procs, lengths, req = np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32)
    to = scatter_plan_to(procs, lengths, req, rstart, rend)
    ts, ti = expand_bs(to["to_starts"], to["to_indices"], bs)          # block lists -> scalar entries
    fs, fi = expand_bs(frm["from_starts"], frm["from_indices"], bs)
    _, tl = expand_bs(np.zeros(1), frm["to_local"], bs)
    _, fl = expand_bs(np.zeros(1), frm["from_local"], bs)
    h = C.c_void_p()
    tp, fp = _i32(to["to_procs"]), _i32(frm["from_procs"])
    _chk(lib.b200_scatter_create(rend - rstart, int(ny), int(to["nto"]), _ip(tp), _ip(ts), _ip(ti),
                                 int(frm["nfrom"]), _ip(fp), _ip(fs), _ip(fi), int(tl.size), _ip(tl), _ip(fl), C.byref(h)),
         "VecScatterCreate")
    return Scatter(h, rend - rstart, int(ny))


# ---- MatSetValuesBatch on the parallel type (src/mat/impls/aij/mpi/mpicusp/mpiaijAssemble.cu:200-460) ----------------
def batch_exchange(rank, size, ranges, elem_rows, elem_mats, nl, dist=None, device="cpu"):
    """Hand every rank the elements that touch its rows.  The reference ships the (i, j, v) entries of rows a rank does
    not own through the stash (MatAssemblyBegin_MPIAIJ, mpiaij.c:630-650; MPI_Alltoallv of rows/cols/values in
    mpiaijAssemble.cu:326-352); whole elements are sent here instead, so the owner assembles them with the same
    kernel and simply passes over the rows that are not its own.
    Returns (rows, mats): this rank's own elements first, then what ranks 0, 1, ... sent, in their order."""
    import torch
    import torch.distributed as td
    dist = dist or td
    rows = np.ascontiguousarray(np.asarray(elem_rows, dtype=np.int32).reshape(-1, nl))
    mats = np.ascontiguousarray(np.asarray(elem_mats, dtype=np.float64).reshape(-1, nl * nl))
    if size == 1:
        return rows.reshape(-1), mats.reshape(-1)
    rg = _i32(ranges)
    ne = rows.shape[0]
    cnt = np.zeros(size, np.int64)
    llp = C.POINTER(C.c_longlong)
    _chk(lib.b200_batch_send_plan(int(size), int(rank), _ip(rg), C.c_longlong(ne), int(nl), _ip(rows), cnt.ctypes.data_as(llp), None),
         "MatSetValuesBatch (send plan)")
    order = np.zeros(max(int(cnt.sum()), 1), np.int64)
    _chk(lib.b200_batch_send_plan(int(size), int(rank), _ip(rg), C.c_longlong(ne), int(nl), _ip(rows), cnt.ctypes.data_as(llp),
                                  order.ctypes.data_as(llp)), "MatSetValuesBatch (send plan)")
    order = order[:int(cnt.sum())]
    counts = torch.tensor(cnt, dtype=torch.int64, device=device)
    allc = [torch.zeros(size, dtype=torch.int64, device=device) for _ in range(size)]
    dist.all_gather(allc, counts)
    allc = np.ascontiguousarray(torch.stack(allc).cpu().numpy())                  # allc[s][d] = elements s sends to d
    maxlen = max(int(allc.sum(axis=1).max()), 1)
    myr = torch.zeros((maxlen, nl), dtype=torch.int32, device=device)           # one padded buffer per rank, groups in
    mym = torch.zeros((maxlen, nl * nl), dtype=torch.float64, device=device)    # destination order
    if order.size:
        myr[:order.size] = torch.from_numpy(rows[order]).to(device)
        mym[:order.size] = torch.from_numpy(mats[order]).to(device)
    allr = [torch.zeros_like(myr) for _ in range(size)]
    allm = [torch.zeros_like(mym) for _ in range(size)]
    dist.all_gather(allr, myr)
    dist.all_gather(allm, mym)
    off, num = np.zeros(size, np.int64), np.zeros(size, np.int64)
    _chk(lib.b200_batch_recv_plan(int(size), int(rank), allc.ctypes.data_as(llp), off.ctypes.data_as(llp), num.ctypes.data_as(llp)),
         "MatSetValuesBatch (receive plan)")
    out_r, out_m = [rows], [mats]
    for s in range(size):
        if num[s]:
            out_r.append(allr[s][int(off[s]):int(off[s] + num[s])].cpu().numpy())
            out_m.append(allm[s][int(off[s]):int(off[s] + num[s])].cpu().numpy())
    return np.concatenate(out_r).reshape(-1), np.concatenate(out_m).reshape(-1)


def split_device(L, cstart, cend):
    """b200_mat_split_device: (handle, nzA, nzB, garray) of the local-rows matrix L, blocks stay on the device."""
    S, ng = C.c_void_p(), C.c_int()
    _chk(lib.b200_mat_split_device(L.h, int(cstart), int(cend), C.byref(ng), C.byref(S)), "MatSetUpMultiply (device split)")
    nzA, nzB = C.c_longlong(), C.c_longlong()
    g = np.zeros(max(ng.value, 1), np.int32)
    _chk(lib.b200_split_get(S, C.byref(nzA), C.byref(nzB), C.byref(ng), _ip(g)), "b200_split_get")
    return S, nzA.value, nzB.value, g[:ng.value]


def create_mpiaij_from_split(S, plan):
    """b200_mat_create_mpiaij_split: the operator from the device blocks of `S` (consumed) and the scatter plan."""
    frm, to, bs = plan["from_"], plan["to"], plan["bs"]
    ts, ti = expand_bs(to["to_starts"], to["to_indices"], bs)
    fs, _ = expand_bs(frm["from_starts"], frm["from_indices"], bs)
    tp, fp = _i32(to["to_procs"]), _i32(frm["from_procs"])
    h = C.c_void_p()
    _chk(lib.b200_mat_create_mpiaij_split(S, int(plan["N"]), int(to["nto"]), _ip(tp), _ip(ts), _ip(ti),
                                          int(frm["nfrom"]), _ip(fp), _ip(fs), C.byref(h)), "MatAssemblyEnd_MPIAIJ (device blocks)")
    return Mat(h, plan["m"], plan["m"])


def set_values_batch_mpiaij(rank, size, ranges, N, nl, elem_rows, elem_mats, bs=1, dist=None, device="cpu", return_host=False,
                            device_split=True):
    """MatSetValuesBatch + MatAssemblyBegin/End on a parallel matrix: exchange the interface elements, assemble this
    rank's rows on the device (b200_mat_setvaluesbatch_rows: local rows x global columns), split into the diagonal
    and off-diagonal blocks and build the ghost exchange exactly as for any other MPIAIJ matrix.  device_split: the
    split happens on the device too (b200_mat_split_device; only garray and row offsets reach the host); otherwise the
    local rows are read back and split by b200_mpiaij_split."""
    rows, mats = batch_exchange(rank, size, ranges, elem_rows, elem_mats, nl, dist, device)
    rstart, rend = int(ranges[rank]), int(ranges[rank + 1])
    h = C.c_void_p()
    _chk(lib.b200_mat_setvaluesbatch_rows(rend - rstart, rstart, int(N), C.c_longlong(rows.size // nl), int(nl), _ip(rows), _dp(mats),
                                          0, 0, C.byref(h)), "MatSetValuesBatch (local rows)")
    L = Mat(h, rend - rstart, int(N))
    if device_split:
        S, _, _, g = split_device(L, rstart, rend)
        host_csr = L.csr() if return_host else None
        L.destroy()
        plan = plan_from_garray(rank, size, ranges, g, N, bs=bs, dist=dist, device=device)
        A = create_mpiaij_from_split(S, plan)
        return (A, host_csr) if return_host else A
    ai, aj, aa = L.csr()
    L.destroy()
    host = setup_mpiaij_host(rank, size, ranges, ai, aj, aa, N, bs=bs, dist=dist, device=device)
    A = create_mpiaij(host)
    return (A, (ai, aj, aa)) if return_host else A


def create_scatter_local(nx, ny, src_idx, dst_idx):
    """VecScatterCreate between two sequential vectors (vscat.c:1030-1260), index sets already expanded to lists:
    forward y[dst_idx[i]] (+)= x[src_idx[i]].  Not collective."""
    s, d = _i32(src_idx), _i32(dst_idx)
    h = C.c_void_p()
    _chk(lib.b200_scatter_create_local(int(nx), int(ny), int(s.size), _ip(s), _ip(d), C.byref(h)), "VecScatterCreate")
    return Scatter(h, int(nx), int(ny))
