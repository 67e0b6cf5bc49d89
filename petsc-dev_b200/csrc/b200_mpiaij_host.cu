/* b200_mpiaij_host.cu — host-side INTEGER work of the row-partitioned (MPIAIJ) operator.  No device code and no
 * CUDA calls: these functions run on a box without a GPU (the world_size-2 gloo tests use them) and must reproduce
 * the reference's integer structures bit for bit:
 *   ownership ranges   PetscSplitOwnership (src/sys/utils/psplit.c:75-97), PetscLayoutSetUp (src/vec/vec/impls/mpi/pmap.c:130-165)
 *   A/B split, garray  MatSetValues_MPIAIJ (src/mat/impls/aij/mpi/mpiaij.c:507-565), MatSetUpMultiply_MPIAIJ (mmaij.c:9-161)
 *   colmap             MatCreateColmap_MPIAIJ_Private (mpiaij.c:357-377)
 *   scatter lists      VecScatterCreate_PtoS (src/vec/vec/utils/vpscat.c:1730-1924)
 * The rank-to-rank exchange of the request lists (PetscGatherMessageLengths + Isend/Irecv, vpscat.c:1778-1822) is
 * done by the host program between b200_scatter_plan_from and b200_scatter_plan_to.
 */
#include "b200_internal.h"
#include <vector>
#include <thread>
#include <functional>
#include <algorithm>

extern "C" int b200_split_ownership(int N, int size, int bs, int *range)
{
  if (!range) return b200_seterr(B200_ERR_ARG_NULL, "null range");
  if (size < 1 || bs < 1 || N < 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "bad N/size/bs %d/%d/%d", N, size, bs);
  if (N % bs) return b200_seterr(B200_ERR_ARG_SIZ, "global size %d not divisible by block size %d", N, bs);
  /* pmap.c:142-146: split block counts, n = N/size + ((N % size) > rank)  (psplit.c:88) */
  const int Nb = N / bs;
  range[0] = 0;
  for (int r = 0; r < size; r++) {
    const int nb = Nb / size + ((Nb % size) > r);
    range[r + 1] = range[r] + nb * bs;
  }
  return 0;
}

extern "C" int b200_mpiaij_split(int m, int cstart, int cend, const int *i, const int *j, const double *a,
                                 int *Ai, int *Aj, double *Aa, int *Bi, int *Bj, double *Ba, int *garray,
                                 long long *nzA_out, long long *nzB_out, int *nghost_out)
{
  if (m < 0 || !i) return b200_seterr(B200_ERR_ARG_NULL, "bad arguments");
  const long long nz = i[m];
  /* Row ranges of about equal entry counts, one host thread each: the split of a 512^3 27-point slab (450 M entries per
   * rank) is 4 passes over the entries, which one thread takes seconds for.  Every pass below is independent per row, so
   * the result does not depend on the number of threads. */
  int T = (int)std::thread::hardware_concurrency();
  if (T < 1 || nz < (1 << 22)) T = 1;
  { const char *e = getenv("B200_HOST_THREADS"); if (e && atoi(e) > 0) T = atoi(e); }   /* also forces threads on small inputs (tests) */
  if (T > 16) T = 16;
  if (T > m) T = m > 0 ? m : 1;
  std::vector<int> cut((size_t)T + 1, m);
  cut[0] = 0;
  for (int t = 1; t < T; t++) cut[t] = (int)(std::lower_bound(i, i + m + 1, (int)(nz * t / T)) - i);
  for (int t = 1; t <= T; t++) if (cut[t] < cut[t - 1]) cut[t] = cut[t - 1];
  auto par = [&](const std::function<void(int)> &f) {
    if (T == 1) { f(0); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(f, t);
    f(0);
    for (auto &x : th) x.join();
  };
  /* pass 1: entries per row in the diagonal block (Bcount = row length - Acount), and the distinct off-process columns
   * of each row range (sorted; rows are ascending, chunks are sorted, made unique and merged) */
  std::vector<int> acnt((size_t)m + 1, 0);
  std::vector<std::vector<int>> gt((size_t)T);
  par([&](int t) {
    std::vector<int> &g = gt[t], buf;
    buf.reserve(1 << 20);
    auto flush = [&]() {
      std::sort(buf.begin(), buf.end());
      buf.erase(std::unique(buf.begin(), buf.end()), buf.end());
      std::vector<int> merged;
      merged.reserve(g.size() + buf.size());
      std::set_union(g.begin(), g.end(), buf.begin(), buf.end(), std::back_inserter(merged));
      g.swap(merged);
      buf.clear();
    };
    for (int r = cut[t]; r < cut[t + 1]; r++) {
      int na = 0;
      for (int k = i[r]; k < i[r + 1]; k++) {
        const int c = j[k];
        if (c >= cstart && c < cend) na++;
        else { buf.push_back(c); if (buf.size() >= (size_t)(1 << 20)) flush(); }
      }
      acnt[r] = na;
    }
    flush();
  });
  /* distinct off-process columns, ascending (mmaij.c:42-50 sorts; :80-84 scans ascending — same array) */
  std::vector<int> g;
  for (int t = 0; t < T; t++) {
    if (gt[t].empty()) continue;
    std::vector<int> merged;
    merged.reserve(g.size() + gt[t].size());
    std::set_union(g.begin(), g.end(), gt[t].begin(), gt[t].end(), std::back_inserter(merged));
    g.swap(merged);
    std::vector<int>().swap(gt[t]);
  }
  long long nzA = 0;
  for (int r = 0; r < m; r++) nzA += acnt[r];
  const long long nzB = nz - nzA;
  const int nghost = (int)g.size();
  if (nzA_out) *nzA_out = nzA;
  if (nzB_out) *nzB_out = nzB;
  if (nghost_out) *nghost_out = nghost;
  if (!Ai) return 0;                                   /* size query */
  if (!Aj || !Aa || !Bi || !Bj || !Ba || !garray) return b200_seterr(B200_ERR_ARG_NULL, "null output arrays");
  for (int k = 0; k < nghost; k++) garray[k] = g[k];
  Ai[0] = 0; Bi[0] = 0;
  for (int r = 0; r < m; r++) { Ai[r + 1] = Ai[r] + acnt[r]; Bi[r + 1] = Bi[r] + (i[r + 1] - i[r] - acnt[r]); }
  /* pass 2: every row writes at its own offsets */
  par([&](int t) {
    for (int r = cut[t]; r < cut[t + 1]; r++) {
      long long pa = Ai[r], pb = Bi[r];
      for (int k = i[r]; k < i[r + 1]; k++) {
        const int c = j[k];
        if (c >= cstart && c < cend) { Aj[pa] = c - cstart; Aa[pa] = a[k]; pa++; }   /* mpiaij.c:515-520 */
        else {
          /* compacted id = position in garray (mmaij.c:56-63) */
          Bj[pb] = (int)(std::lower_bound(g.begin(), g.end(), c) - g.begin()); Ba[pb] = a[k]; pb++;
        }
      }
    }
  });
  return 0;
}

extern "C" int b200_mpiaij_colmap(int N, int nghost, const int *garray, int *colmap)
{
  if (!colmap || (nghost && !garray)) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  memset(colmap, 0, sizeof(int) * (size_t)N);
  for (int k = 0; k < nghost; k++) {
    if (garray[k] < 0 || garray[k] >= N) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "garray[%d] = %d outside [0,%d)", k, garray[k], N);
    colmap[garray[k]] = k + 1;                        /* mpiaij.c:373 */
  }
  return 0;
}

/* Receive side + request lists of VecScatterCreate_PtoS.  inidx[nx] are (block) indices into the parallel vector,
 * inidy[nx] (block) positions in the sequential one; every array the caller passes must hold nx (resp. size,
 * size+1) entries.  Outputs:
 *   nfrom, from_procs[nfrom], from_starts[nfrom+1], from_indices[...]        the `from` context (vpscat.c:1850-1879)
 *   req_counts[size] (own rank zeroed, :1775), req_values[...] grouped by owner ascending (:1798-1806)
 *   nlocal, to_local[nlocal], from_local[nlocal]                              the self part (:1890-1903)       */
extern "C" int b200_scatter_plan_from(int nx, const int *inidx, const int *inidy, int bs, int size, const int *range, int rank,
                                      int lengthy, int *nfrom_out, int *from_procs, int *from_starts, int *from_indices,
                                      int *req_counts, int *req_values, int *nlocal_out, int *to_local, int *from_local)
{
  if (nx < 0 || size < 1 || !range) return b200_seterr(B200_ERR_ARG_NULL, "bad arguments");
  if (nx && (!inidx || !inidy)) return b200_seterr(B200_ERR_ARG_NULL, "null index arrays");
  std::vector<int> nprocs(size, 0), owner(nx);
  int j = 0, nsends = 0;
  for (int i = 0; i < nx; i++) {                       /* vpscat.c:1759-1772 */
    const int idx = bs * inidx[i];
    if (idx < range[j]) j = 0;
    for (; j < size; j++) {
      if (idx < range[j + 1]) { if (!nprocs[j]++) nsends++; owner[i] = j; break; }
    }
    if (j == size) return b200_seterr(B200_ERR_PLIB, "ith %d block entry %d not owned by any process, upper bound %d", i, idx, range[size]);
  }
  const int nprocslocal = nprocs[rank];
  nprocs[rank] = 0;
  if (nprocslocal) nsends--;
  if (nfrom_out) *nfrom_out = nsends;
  if (nlocal_out) *nlocal_out = nprocslocal;
  if (!from_procs) return 0;                           /* size query */
  std::vector<int> starts(size + 1);
  starts[0] = 0;
  for (int p = 1; p <= size; p++) starts[p] = starts[p - 1] + nprocs[p - 1];
  if (req_counts) for (int p = 0; p < size; p++) req_counts[p] = nprocs[p];
  if (req_values) {
    std::vector<int> pos(starts.begin(), starts.end() - 1);
    for (int i = 0; i < nx; i++) if (owner[i] != rank) req_values[pos[owner[i]]++] = bs * inidx[i];
  }
  std::vector<int> lowner(size, 0), start(nsends + 1, 0);
  int count = 0;
  from_starts[0] = 0;
  for (int p = 0; p < size; p++) {
    if (nprocs[p]) {
      lowner[p] = count;
      from_procs[count++] = p;
      from_starts[count] = start[count] = start[count - 1] + nprocs[p];
    }
  }
  for (int i = 0; i < nx; i++) {
    if (owner[i] != rank) {
      if (bs * inidy[i] >= lengthy) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "Scattering past end of TO vector");
      from_indices[start[lowner[owner[i]]]++] = bs * inidy[i];
    }
  }
  if (nprocslocal && to_local && from_local) {
    int nt = 0;
    for (int i = 0; i < nx; i++) {
      const int idx = bs * inidx[i];
      if (idx >= range[rank] && idx < range[rank + 1]) { to_local[nt] = idx - range[rank]; from_local[nt++] = bs * inidy[i]; }
    }
  }
  return 0;
}

/* Send side: nto requesting ranks (ascending, PetscSortMPIIntWithArray vpscat.c:1780), lengths[nto], and their
 * concatenated requested GLOBAL ids; to_starts[nto+1], to_indices = id - range[rank] (vpscat.c:1836-1848). */
extern "C" int b200_scatter_plan_to(int nto, const int *lengths, const int *requested, int rstart, int rend,
                                    int *to_starts, int *to_indices)
{
  if (nto < 0 || !to_starts) return b200_seterr(B200_ERR_ARG_NULL, "bad arguments");
  to_starts[0] = 0;
  for (int p = 0; p < nto; p++) {
    to_starts[p + 1] = to_starts[p] + lengths[p];
    for (int k = 0; k < lengths[p]; k++) {
      const int v = requested[to_starts[p] + k];
      if (v < rstart || v >= rend) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "requested entry %d is not owned here [%d,%d)", v, rstart, rend);
      to_indices[to_starts[p] + k] = v - rstart;
    }
  }
  return 0;
}

/* mmaij.c:106-128: may the ghost list be described by a block index set? */
extern "C" int b200_mpiaij_use_block_is(int nghost, const int *garray, int bs, int *use)
{
  if (!use) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *use = 0;
  if (bs <= 1) return 0;
  if (nghost % bs) return 0;
  for (int i = 0; i < nghost / bs; i++) {
    const int ga = garray[i * bs];
    if (ga % bs) return 0;
    for (int j = 1; j < bs; j++) if (garray[i * bs + j] != ga + j) return 0;
  }
  *use = 1;
  return 0;
}

/* ---- MatSetValuesBatch on the parallel type: who gets which element -------------------------------------------------
 * The reference routes the (i, j, v) entries of rows a rank does not own to their owners at assembly time (stash:
 * MatAssemblyBegin_MPIAIJ mpiaij.c:630-650; MatSetValuesBatch_MPIAIJCUSP sends rows/cols/values with MPI_Alltoallv,
 * mpiaijAssemble.cu:326-352).  Here whole elements travel: an element goes to every OTHER rank that owns one of its
 * rows; the receiver assembles own + received elements and passes over the rows that are not its own.
 * Pure integer work, shared by the PETSc plugin (MPI) and petsc-dev_b200/parallel.py (torch.distributed). */

/* counts[d] = how many of my elements go to rank d (0 for d == rank).  send_elems (may be NULL on a sizing call)
 * receives the element numbers grouped by destination rank ascending, element order kept inside a group. */
extern "C" int b200_batch_send_plan(int size, int rank, const int *range, long long ne, int nl, const int *elem_rows,
                                    long long *counts, long long *send_elems)
{
  if (!range || !counts || (ne > 0 && !elem_rows)) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (size < 1 || rank < 0 || rank >= size || ne < 0 || nl < 1) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "bad size/rank/ne/nl %d/%d/%lld/%d", size, rank, ne, nl);
  const int N = range[size];
  std::vector<char> hit((size_t)size);
  std::vector<long long> pos((size_t)size + 1, 0);
  for (int pass = 0; pass < (send_elems ? 2 : 1); pass++) {
    if (pass == 1) {                                           /* group starts from the counts of pass 0 */
      long long acc = 0;
      for (int d = 0; d < size; d++) { pos[d] = acc; acc += counts[d]; }
    } else {
      for (int d = 0; d < size; d++) counts[d] = 0;
    }
    for (long long e = 0; e < ne; e++) {
      std::fill(hit.begin(), hit.end(), 0);
      for (int a = 0; a < nl; a++) {
        const int row = elem_rows[e * nl + a];
        if (row < 0) continue;                                 /* ignored ids, aij.c:311 */
        if (row >= N) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "Row too large: row %d max %d", row, N - 1);
        const int owner = (int)(std::upper_bound(range, range + size + 1, row) - range) - 1;   /* PetscLayoutFindOwner */
        if (owner != rank) hit[owner] = 1;
      }
      for (int d = 0; d < size; d++) {
        if (!hit[d]) continue;
        if (pass == 0) counts[d]++;
        else send_elems[pos[d]++] = e;
      }
    }
  }
  return 0;
}

/* all_counts[s*size + d] = elements rank s sends to rank d (the gathered counts of every rank).  Every rank publishes
 * ONE buffer holding its groups in destination order; for the receiving rank: src_off[s] = first element of its group
 * inside rank s's buffer, src_cnt[s] = its length (0 for s == rank).  The receiver appends the groups of s = 0, 1, ...
 * after its own elements. */
extern "C" int b200_batch_recv_plan(int size, int rank, const long long *all_counts, long long *src_off, long long *src_cnt)
{
  if (!all_counts || !src_off || !src_cnt) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (size < 1 || rank < 0 || rank >= size) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "bad size/rank %d/%d", size, rank);
  for (int s = 0; s < size; s++) {
    long long before = 0;
    for (int d = 0; d < rank; d++) before += all_counts[(size_t)s * size + d];
    src_off[s] = before;
    src_cnt[s] = s == rank ? 0 : all_counts[(size_t)s * size + rank];
  }
  return 0;
}
