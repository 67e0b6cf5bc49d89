/* b200_comm.cu — rank-to-rank traffic of the hot path: the tiny allreduces of VecDot/VecNorm/VecMDot
 * (MPI_Allreduce in src/vec/vec/impls/mpi/pbvec.c:16, pvec2.c:19,61) and the MPIAIJ ghost exchange
 * (VecScatterBegin_1/End_1, src/vec/vec/utils/vpscat.h:14-233), device-resident, one rank per GPU.
 *
 * Transport 1 (default): NCCL, resolved at run time from the process (torch's bundled libnccl when the host program
 * is Python, the system one otherwise) — grouped ncclSend/ncclRecv per neighbour straight into the ghost vector
 * (receives are contiguous because garray is sorted: `from->contiq`, vpscat.c:1951-1960), on a side stream that
 * overlaps the diagonal-block product.
 * Transport 2 (b200_comm_enable_p2p): NVLink peer memory.  Every rank owns an "arena" (one cudaMalloc, exported with
 * cudaIpcGetMemHandle, opened by every peer) holding its ghost vectors, arrival flags, credits and an allreduce inbox.
 *   - ghost exchange = ONE kernel: gather x[indices] and store straight into the neighbour's ghost vector through the
 *     NVSwitch (pack + send + "unpack" fused; the payload never stops in a send buffer), system-scope fence, then the
 *     last CTA raises the arrival flag on each neighbour.  The stores drain while the diagonal-block product runs; the
 *     off-diagonal product is preceded by a one-warp wait on the local flags.  NCCL cannot overlap here because the
 *     diagonal product is a persistent kernel that fills every SM (its send/recv kernel would queue behind it).
 *   - ghost vectors are double-buffered by exchange parity and guarded by credits (the receiver tells its senders which
 *     exchange it has consumed), so non-symmetric neighbour graphs cannot overrun a buffer.
 *   - allreduce of the Krylov scalars = one tiny kernel: every rank stores its values into every peer's inbox slot,
 *     fences, stores the sequence number; each rank then sums the slots in rank order (bitwise identical everywhere).
 *   - the same engine runs backwards (reverse_begin / reverse_wait): ghost-side contributions travel to their owners'
 *     return buffers and are added neighbour by neighbour in ascending rank order — MatMultTranspose_MPIAIJ
 *     (mpiaij.c:1147) and SCATTER_REVERSE/ADD_VALUES of the general b200_scatter (end of this file), which is an
 *     operator-shaped exchange object without matrix blocks plus unpack kernels for arbitrary index lists.
 *   - every wait on a peer is bounded (b200_spin_guard): a dead rank ends in a trap, not in a GPU that spins for ever.
 */
#include "b200_internal.h"
#include <dlfcn.h>
#include <nccl.h>
#include <vector>
#include <algorithm>
#include "b200_p2p_dev.cuh"
#include "b200_spmv_mpi_kernels.cuh"

struct NcclApi {
  void *h;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *);
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)(void);
  ncclResult_t (*GroupEnd)(void);
  const char *(*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl = {};

struct ArenaRange { size_t off, len; };
struct B200P2P {
  size_t bytes, bump;                    /* arena size, end of the part handed out so far                           */
  ArenaRange *freelist; int nfree, capfree;   /* returned ranges below `bump`, sorted by offset, neighbours merged  */
  char  *arena;                          /* this rank's arena                                                       */
  char  *peer[B200_MAX_RANKS];           /* every rank's arena as mapped here ([rank] == arena)                     */
  unsigned long long ar_seq;             /* allreduce sequence number (same on all ranks: collective calls)        */
  unsigned long long **d_inbox;          /* device table: inbox base of every rank                                  */
};
#define P2P_CTRL_BYTES ((size_t)2 * B200_MAX_RANKS * P2P_AR_SLOT * 8)   /* allreduce inbox: [parity][sender][slot]  */

static int nccl_load(void)
{
  if (g_nccl.h) return 0;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return b200_seterr(B200_ERR_LIB, "cannot load libnccl.so.2: %s", dlerror());
#define SYM(f) do { *(void **)(&g_nccl.f) = dlsym(h, "nccl" #f); if (!g_nccl.f) return b200_seterr(B200_ERR_LIB, "libnccl lacks nccl" #f); } while (0)
  SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(AllReduce); SYM(AllGather); SYM(Send); SYM(Recv); SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
#undef SYM
  g_nccl.h = h;
  return 0;
}

#define B200_NCCL(call) do { ncclResult_t r__ = (call); if (r__ != ncclSuccess) return b200_seterr(B200_ERR_LIB, "NCCL error %d (%s) in %s", (int)r__, g_nccl.GetErrorString(r__), #call); } while (0)

extern "C" int b200_comm_get_unique_id(void *id128)
{
  if (!id128) return b200_seterr(B200_ERR_ARG_NULL, "null id");
  B200_CHK(nccl_load());
  ncclUniqueId id;
  B200_NCCL(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) <= B200_COMM_ID_BYTES, "id size");
  memset(id128, 0, B200_COMM_ID_BYTES);
  memcpy(id128, &id, sizeof(id));
  return 0;
}

extern "C" int b200_comm_init(int rank, int size, const void *id128)
{
  B200_NEED_INIT();
  if (size < 1 || rank < 0 || rank >= size || size > B200_MAX_RANKS) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "rank %d of %d", rank, size);
  if (g_b200.nccl) return b200_seterr(B200_ERR_ARG_WRONGSTATE, "communicator already initialised");
  g_b200.rank = rank; g_b200.size = size;
  if (size == 1) return 0;
  if (!id128) return b200_seterr(B200_ERR_ARG_NULL, "null id");
  B200_CHK(nccl_load());
  ncclUniqueId id;
  memcpy(&id, id128, sizeof(id));
  ncclComm_t comm;
  B200_NCCL(g_nccl.CommInitRank(&comm, size, id, rank));
  g_b200.nccl = (void *)comm;
  return 0;
}

extern "C" int b200_comm_rank(int *rank, int *size)
{
  if (rank) *rank = g_b200.rank;
  if (size) *size = g_b200.inited ? g_b200.size : 1;
  return 0;
}

extern "C" int b200_comm_destroy(void)
{
  if (g_b200.peer) {
    B200P2P *P = g_b200.peer;
    g_b200.p2p = false;
    cudaDeviceSynchronize();
    if (g_b200.nccl && g_b200.size > 1) b200_comm_barrier();        /* nobody may still be storing into an arena */
    for (int r = 0; r < g_b200.size; r++) if (r != g_b200.rank && P->peer[r]) cudaIpcCloseMemHandle(P->peer[r]);
    cudaFree(P->d_inbox); cudaFree(P->arena);
    free(P->freelist);
    free(P);
    g_b200.peer = NULL;
  }
  g_b200.p2p = false;
  if (g_b200.nccl) { g_nccl.CommDestroy((ncclComm_t)g_b200.nccl); g_b200.nccl = NULL; }
  g_b200.rank = 0; g_b200.size = 1;
  return 0;
}

static int p2p_allreduce(double *dvals, int n, int op_max);

int b200_comm_allreduce_dev(double *dvals, int n, int op_max)
{
  if (g_b200.size <= 1) return 0;
  if (!g_b200.nccl) return b200_seterr(B200_ERR_ORDER, "communicator not initialised");
  if (g_b200.p2p && n <= P2P_AR_MAXV) return p2p_allreduce(dvals, n, op_max);
  B200_NCCL(g_nccl.AllReduce(dvals, dvals, (size_t)n, ncclDouble, op_max ? ncclMax : ncclSum, (ncclComm_t)g_b200.nccl, g_b200.stream));
  return 0;
}

extern "C" int b200_comm_allreduce_sum(double *dvals, int n)
{
  B200_NEED_INIT();
  return b200_comm_allreduce_dev(dvals, n, 0);
}

extern "C" int b200_comm_barrier(void)
{
  B200_NEED_INIT();
  if (g_b200.size > 1) {
    B200_CHK(b200_comm_allreduce_dev(g_b200.d_scal + B200_RES_SLOTS - 1, 1, 0));
  }
  B200_CUDA(cudaStreamSynchronize(g_b200.stream));
  return 0;
}

/* ------------------------------------------------------------------------------------------------ peer memory */

__global__ void __launch_bounds__(64) p2p_allreduce_kernel(double *vals, int n, int op_max, unsigned long long seq, int rank, int size,
                                                           unsigned long long *const *__restrict__ inbox)
{
  b200_p2p_allreduce_cta(vals, n, op_max, seq, rank, size, inbox);
}

static int p2p_allreduce(double *dvals, int n, int op_max)
{
  B200Ctx &c = g_b200;
  B200P2P *P = c.peer;
  if (n <= 0) return 0;
  P->ar_seq++;
  p2p_allreduce_kernel<<<1, 64, 0, c.stream>>>(dvals, n, op_max, P->ar_seq, c.rank, c.size, P->d_inbox);
  B200_LAUNCHED(1);
  return 0;
}

/* the fused CG folds the allreduce into its own scalar kernels: hand out the next sequence number and the inbox table */
bool b200_p2p_next(unsigned long long *seq, unsigned long long *const **inbox)
{
  if (!g_b200.p2p || !g_b200.peer) return false;
  *seq = ++g_b200.peer->ar_seq;
  *inbox = g_b200.peer->d_inbox;
  return true;
}

/* Arena ranges (256-byte granules): first fit in the list of returned ranges, else from the top.  Matrices and scatters
 * come and go in any order (every attached VecScatter owns a range), so a range goes back to the list on destroy;
 * neighbours merge, and a range ending at `bump` lowers it.  Purely local bookkeeping: ranks exchange the offsets they
 * chose (peerplan_build), they need not choose the same ones. */
static int p2p_alloc(size_t bytes, size_t *off)
{
  B200P2P *P = g_b200.peer;
  bytes = (bytes + 255) & ~(size_t)255;
  for (int k = 0; k < P->nfree; k++)
    if (P->freelist[k].len >= bytes) {
      *off = P->freelist[k].off;
      P->freelist[k].off += bytes; P->freelist[k].len -= bytes;
      if (!P->freelist[k].len) { memmove(P->freelist + k, P->freelist + k + 1, sizeof(ArenaRange) * (size_t)(P->nfree - k - 1)); P->nfree--; }
      return 0;
    }
  const size_t a = (P->bump + 255) & ~(size_t)255;
  if (a + bytes > P->bytes) return 1;
  *off = a; P->bump = a + bytes;
  return 0;
}

static void p2p_free(size_t off, size_t bytes)
{
  B200P2P *P = g_b200.peer;
  if (!P || !bytes) return;
  bytes = (bytes + 255) & ~(size_t)255;
  if (P->nfree == P->capfree) {
    const int cap = P->capfree ? 2 * P->capfree : 16;
    ArenaRange *q = (ArenaRange *)realloc(P->freelist, sizeof(ArenaRange) * (size_t)cap);
    if (!q) return;                                        /* the range stays lost; nothing else breaks */
    P->freelist = q; P->capfree = cap;
  }
  int k = 0;
  while (k < P->nfree && P->freelist[k].off < off) k++;
  memmove(P->freelist + k + 1, P->freelist + k, sizeof(ArenaRange) * (size_t)(P->nfree - k));
  P->freelist[k].off = off; P->freelist[k].len = bytes; P->nfree++;
  if (k + 1 < P->nfree && P->freelist[k].off + P->freelist[k].len == P->freelist[k + 1].off) {
    P->freelist[k].len += P->freelist[k + 1].len;
    memmove(P->freelist + k + 1, P->freelist + k + 2, sizeof(ArenaRange) * (size_t)(P->nfree - k - 2)); P->nfree--;
  }
  if (k > 0 && P->freelist[k - 1].off + P->freelist[k - 1].len == P->freelist[k].off) {
    P->freelist[k - 1].len += P->freelist[k].len;
    memmove(P->freelist + k, P->freelist + k + 1, sizeof(ArenaRange) * (size_t)(P->nfree - k - 1)); P->nfree--;
  }
  if (P->nfree && P->freelist[P->nfree - 1].off + P->freelist[P->nfree - 1].len == P->bump) { P->bump = P->freelist[P->nfree - 1].off; P->nfree--; }
}

extern "C" int b200_comm_enable_p2p(int enable)
{
  B200_NEED_INIT();
  B200Ctx &c = g_b200;
  if (!enable) {
    if (c.peer) return b200_seterr(B200_ERR_ARG_WRONGSTATE, "the peer-memory transport stays on until b200_comm_destroy");
    return 0;
  }
  if (c.p2p) return 0;
  if (c.size <= 1 && enable != 2) return 0;
  if (c.size > 1 && !c.nccl) return b200_seterr(B200_ERR_ORDER, "b200_comm_enable_p2p needs b200_comm_init first (the handles travel over it)");
  B200P2P *P = (B200P2P *)calloc(1, sizeof(B200P2P));
  if (!P) return b200_seterr(B200_ERR_MEM, "out of host memory");
  size_t mb = 512;
  const char *e = getenv("B200_P2P_ARENA_MB");
  if (e && atol(e) > 0) mb = (size_t)atol(e);
  P->bytes = mb << 20;
  B200_CUDA(cudaMalloc((void **)&P->arena, P->bytes));
  B200_CUDA(cudaMemset(P->arena, 0, P->bytes));
  P->bump = P2P_CTRL_BYTES;
  if (c.size <= 1) {
    /* loop-back (enable == 2): one rank whose only peer is itself.  The exchange kernels, flags and credits run exactly
     * as between GPUs, only the stores stay on this device — the single-GPU test of the multi-GPU path */
    unsigned long long *tab[1] = {reinterpret_cast<unsigned long long *>(P->arena)};
    P->peer[0] = P->arena;
    B200_CUDA(cudaMalloc((void **)&P->d_inbox, sizeof(tab)));
    B200_CUDA(cudaMemcpy(P->d_inbox, tab, sizeof(tab), cudaMemcpyHostToDevice));
    c.peer = P; c.p2p = true;
    return 0;
  }
  /* exchange the IPC handles through the NCCL communicator */
  cudaIpcMemHandle_t mine, all[B200_MAX_RANKS];
  B200_CUDA(cudaIpcGetMemHandle(&mine, P->arena));
  char *dh = NULL;
  B200_CUDA(cudaMalloc((void **)&dh, sizeof(mine) * (size_t)(c.size + 1)));
  B200_CUDA(cudaMemcpy(dh, &mine, sizeof(mine), cudaMemcpyHostToDevice));
  B200_NCCL(g_nccl.AllGather(dh, dh + sizeof(mine), sizeof(mine), ncclChar, (ncclComm_t)c.nccl, c.stream));
  B200_CUDA(cudaStreamSynchronize(c.stream));
  B200_CUDA(cudaMemcpy(all, dh + sizeof(mine), sizeof(mine) * (size_t)c.size, cudaMemcpyDeviceToHost));
  cudaFree(dh);
  /* the kernels of this transport spin on flags other ranks raise: every rank must own its GPU (two ranks sharing a
   * device could wait on each other without ever being co-scheduled) */
  {
    char busid[B200_MAX_RANKS + 1][32], *db = NULL;
    memset(busid, 0, sizeof(busid));
    B200_CUDA(cudaDeviceGetPCIBusId(busid[0], 32, c.device));
    B200_CUDA(cudaMalloc((void **)&db, 32 * (size_t)(c.size + 1)));
    B200_CUDA(cudaMemcpy(db, busid[0], 32, cudaMemcpyHostToDevice));
    B200_NCCL(g_nccl.AllGather(db, db + 32, 32, ncclChar, (ncclComm_t)c.nccl, c.stream));
    B200_CUDA(cudaStreamSynchronize(c.stream));
    B200_CUDA(cudaMemcpy(busid[1], db + 32, 32 * (size_t)c.size, cudaMemcpyDeviceToHost));
    cudaFree(db);
    for (int a = 0; a < c.size; a++)
      for (int b = a + 1; b < c.size; b++)
        if (!strncmp(busid[1 + a], busid[1 + b], 32)) {
          cudaFree(P->arena); free(P);
          return b200_seterr(B200_ERR_SUP, "ranks %d and %d share GPU %s: the peer-memory transport needs one GPU per rank (use the NCCL transport)", a, b, busid[1 + a]);
        }
  }
  /* map every peer's arena; the outcome is agreed on collectively (a rank that cannot map a peer must not leave the
   * others waiting in a barrier it never reaches): all ranks succeed, or all ranks undo and report the same error */
  int okflag = 1, badrank = -1;
  cudaError_t bad = cudaSuccess;
  for (int r = 0; r < c.size; r++) {
    if (r == c.rank) { P->peer[r] = P->arena; continue; }
    cudaError_t ce = cudaIpcOpenMemHandle((void **)&P->peer[r], all[r], cudaIpcMemLazyEnablePeerAccess);
    if (ce != cudaSuccess) { okflag = 0; bad = ce; badrank = r; P->peer[r] = NULL; cudaGetLastError(); break; }
  }
  if (okflag) {
    unsigned long long *tab[B200_MAX_RANKS];
    for (int r = 0; r < c.size; r++) tab[r] = reinterpret_cast<unsigned long long *>(P->peer[r]);
    if (cudaMalloc((void **)&P->d_inbox, sizeof(tab)) != cudaSuccess ||
        cudaMemcpy(P->d_inbox, tab, sizeof(void *) * (size_t)c.size, cudaMemcpyHostToDevice) != cudaSuccess) { okflag = 0; cudaGetLastError(); }
  }
  {
    int *dflag = NULL, agreed = 0;
    B200_CUDA(cudaMalloc((void **)&dflag, sizeof(int)));
    B200_CUDA(cudaMemcpy(dflag, &okflag, sizeof(int), cudaMemcpyHostToDevice));
    B200_NCCL(g_nccl.AllReduce(dflag, dflag, 1, ncclInt, ncclMin, (ncclComm_t)c.nccl, c.stream));
    B200_CUDA(cudaStreamSynchronize(c.stream));                   /* doubles as the barrier: every arena is zeroed and mapped */
    B200_CUDA(cudaMemcpy(&agreed, dflag, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(dflag);
    if (!agreed) {
      for (int r = 0; r < c.size; r++) if (r != c.rank && P->peer[r]) cudaIpcCloseMemHandle(P->peer[r]);
      cudaFree(P->d_inbox); cudaFree(P->arena); free(P);
      if (!okflag && badrank >= 0) return b200_seterr(B200_ERR_LIB, "cudaIpcOpenMemHandle for rank %d: %s (no NVLink peer access?); the NCCL transport stays in use", badrank, cudaGetErrorString(bad));
      return b200_seterr(B200_ERR_LIB, "another rank could not map the peer arenas; the NCCL transport stays in use");
    }
  }
  c.peer = P;
  c.p2p = true;
  return 0;
}

extern "C" int b200_comm_p2p_enabled(int *enabled)
{
  if (enabled) *enabled = g_b200.p2p ? 1 : 0;
  return 0;
}

/* ------------------------------------------------------------------------------------------------ ghost exchange */
/* Pack_1 (vpscat.c:493): buf[i] = x[indices[i]] */
__global__ void __launch_bounds__(256) pack_kernel(int n, const int *__restrict__ idx, const double *__restrict__ x, double *__restrict__ buf)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) buf[i] = __ldg(&x[__ldg(&idx[i])]);
}

struct PeerPlan {
  PeerPlanDev dev;                               /* forward: x -> neighbours' ghost vectors (MatMult)                */
  PeerPlanDev rev;                               /* reverse: my ghost contributions -> their owners (MatMultTranspose) */
  void       *d_blob;                            /* one device allocation behind the tables above                    */
  void       *d_blob_rev;
  double     *rbuf[2];                           /* arena: what the neighbours return to me (nsend entries, by parity) */
  unsigned   *d_ticket_rev;
  size_t      arena_lo, arena_hi;                /* arena range owned by this matrix                                 */
  unsigned   *d_ticket;
};

/* VecScatterBegin as one kernel: Pack_1 (vpscat.c:493) + the send + UnPack_1 fused — x[indices[i]] is stored straight
 * into the neighbour's ghost vector over NVLink.  Before the first store each CTA checks the credits (the neighbour has
 * consumed the exchange that last used this parity's buffer); CTA 0 also returns this rank's own credits.  After a
 * system-scope fence the last CTA raises the arrival flags. */
__global__ void __launch_bounds__(256) pack_put_kernel(int nsend, const int *__restrict__ idx, const double *__restrict__ x,
                                                       const PeerPlanDev plan, int parity, unsigned long long epoch, unsigned *ticket)
{
  __shared__ int s_starts[B200_MAX_RANKS + 1];
  __shared__ double *s_dst[B200_MAX_RANKS];
  __shared__ bool s_last;
  const int tid = threadIdx.x;
  /* credits first (stream order: this rank's off-diagonal product of exchange epoch-1 is complete), then wait */
  if (blockIdx.x == 0 && tid < plan.nfrom) st_release_sys(plan.rcredit[tid], epoch - 1);
  if (tid <= plan.nto) s_starts[tid] = plan.to_starts[tid];
  if (tid < plan.nto) {
    s_dst[tid] = plan.dst[parity * plan.nto + tid];
    { const long long t0 = clock64(); while (ld_acquire_sys(plan.lcredit[tid]) + 2 < epoch) b200_spin_guard(t0); }
  }
  __syncthreads();
  const int stride = gridDim.x * blockDim.x;
  int k = 0;
  for (int i = blockIdx.x * blockDim.x + tid; i < nsend; i += stride) {
    while (i >= s_starts[k + 1]) k++;
    s_dst[k][i - s_starts[k]] = __ldg(&x[idx ? __ldg(&idx[i]) : i]);
  }
  /* completion: CTA barrier, then ONE system-scope fence per CTA (cumulative over the CTA's stores through the barrier,
   * the cooperative-groups grid-sync pattern) — a fence in every thread serialised 65 k MEMBAR.SYS and cost 15 us */
  __syncthreads();
  if (tid == 0) {
    __threadfence_system();
    s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  if (tid < plan.nto) {
    __threadfence_system();
    st_release_sys(plan.rflag[tid], epoch);
  }
  if (tid == 0) *ticket = 0;
}

/* VecScatterEnd: one warp waits until every sender's flag shows this exchange */
__global__ void __launch_bounds__(32) wait_arrivals_kernel(const PeerPlanDev plan, unsigned long long epoch)
{
  if ((int)threadIdx.x < plan.nfrom) { const long long t0 = clock64(); while (ld_acquire_sys(plan.lflag[threadIdx.x]) < epoch) b200_spin_guard(t0); }
}

static int fusedplan_build(b200_mat M, const int *Bi);

static void peerplan_free(b200_mat M)
{
  PeerPlan *pp = (PeerPlan *)M->peerplan;
  if (!pp) return;
  cudaFree(pp->d_blob); cudaFree(pp->d_ticket); cudaFree(pp->d_blob_rev); cudaFree(pp->d_ticket_rev);
  p2p_free(pp->arena_lo, pp->arena_hi - pp->arena_lo);            /* back to the arena's free list, whatever the order */
  free(pp);
  M->peerplan = NULL;
  M->d_lvec[0] = M->d_lvec[1] = NULL;                             /* they lived in the arena                          */
}
void b200_mpiaij_peer_free(b200_mat M) { peerplan_free(M); }

/* device tables of one direction: starts[ns+1] | dst[2*ns] | rflag[ns] | lcredit[ns] | rcredit[nr] | lflag[nr] */
static int plan_blob(PeerPlanDev *dev, void **d_blob, int ns, const int *starts, void *const *dst0, void *const *dst1,
                     void *const *rflag, void *const *lcredit, int nr, void *const *rcredit, void *const *lflag)
{
  const size_t n_int = (size_t)((ns + 1 + 1) & ~1);
  const size_t nptr = (size_t)(4 * ns + 2 * nr);
  const size_t bytes = n_int * sizeof(int) + nptr * sizeof(void *) + 16;
  char *hb = (char *)calloc(1, bytes);
  if (!hb) return b200_seterr(B200_ERR_MEM, "out of host memory");
  int *h_starts = (int *)hb;
  void **hp = (void **)(hb + n_int * sizeof(int));
  for (int k = 0; k <= ns; k++) h_starts[k] = starts[k];
  for (int k = 0; k < ns; k++) { hp[k] = dst0[k]; hp[ns + k] = dst1[k]; hp[2 * ns + k] = rflag[k]; hp[3 * ns + k] = lcredit[k]; }
  for (int k = 0; k < nr; k++) { hp[4 * ns + k] = rcredit[k]; hp[4 * ns + nr + k] = lflag[k]; }
  B200_CUDA(cudaMalloc(d_blob, bytes));
  B200_CUDA(cudaMemcpy(*d_blob, hb, bytes, cudaMemcpyHostToDevice));
  free(hb);
  char *db = (char *)*d_blob;
  void **dp = (void **)(db + n_int * sizeof(int));
  dev->nto = ns; dev->nfrom = nr;
  dev->to_starts = (const int *)db;
  dev->dst     = (double *const *)(dp);
  dev->rflag   = (unsigned long long *const *)(dp + 2 * ns);
  dev->lcredit = (const unsigned long long *const *)(dp + 3 * ns);
  dev->rcredit = (unsigned long long *const *)(dp + 4 * ns);
  dev->lflag   = (const unsigned long long *const *)(dp + 4 * ns + nr);
  return 0;
}

/* collective: carve the ghost vectors / return buffers / flags / credits out of the arena and learn where every
 * neighbour put theirs */
static int peerplan_build(b200_mat M)
{
  B200Ctx &c = g_b200;
  B200P2P *P = c.peer;
  const int size = c.size, rank = c.rank, nghost = M->nghost, R = B200_MAX_RANKS;
  const int nto = M->nto, nfrom = M->nfrom, nsend = M->to_starts[nto];
  /* meta = { ok, lvec0, lvec1, flag, credit, rbuf0, rbuf1, rflag, rcredit, recv_start[R], send_start[R] } */
  enum { OK = 0, LV0, LV1, FLG, CRD, RB0, RB1, RFLG, RCRD, RECV0 };
  const int W = RECV0 + 2 * B200_MAX_RANKS;
  long long mine[RECV0 + 2 * B200_MAX_RANKS], all[B200_MAX_RANKS][RECV0 + 2 * B200_MAX_RANKS];
  size_t off[RECV0] = {0};
  const size_t want[RECV0] = {0, sizeof(double) * (size_t)(nghost + 2), sizeof(double) * (size_t)(nghost + 2), 8 * (size_t)R, 8 * (size_t)R,
                              sizeof(double) * (size_t)(nsend + 2), sizeof(double) * (size_t)(nsend + 2), 8 * (size_t)R, 8 * (size_t)R};
  /* one range for everything this operator keeps in the arena */
  size_t total = 0, lo = 0;
  for (int k = 1; k < RECV0; k++) { off[k] = total; total += (want[k] + 255) & ~(size_t)255; }
  const int fail = p2p_alloc(total, &lo);
  const size_t hi = lo + total;
  for (int k = 1; k < RECV0; k++) off[k] += lo;
  mine[OK] = fail ? 0 : 1;
  for (int k = 1; k < RECV0; k++) mine[k] = (long long)off[k];
  for (int r = 0; r < 2 * R; r++) mine[RECV0 + r] = -1;
  for (int k = 0; k < nfrom; k++) mine[RECV0 + M->from_procs[k]] = M->from_starts[k];
  for (int k = 0; k < nto; k++)   mine[RECV0 + R + M->to_procs[k]] = M->to_starts[k];
  if (size == 1) memcpy(all[0], mine, sizeof(mine));
  else {
    long long *dm = NULL;
    B200_CUDA(cudaMalloc((void **)&dm, sizeof(long long) * (size_t)W * (size_t)(size + 1)));
    B200_CUDA(cudaMemcpy(dm, mine, sizeof(mine), cudaMemcpyHostToDevice));
    B200_NCCL(g_nccl.AllGather(dm, dm + W, (size_t)W, ncclInt64, (ncclComm_t)c.nccl, c.stream));
    B200_CUDA(cudaStreamSynchronize(c.stream));
    B200_CUDA(cudaMemcpy(all, dm + W, sizeof(long long) * (size_t)W * (size_t)size, cudaMemcpyDeviceToHost));
    cudaFree(dm);
  }
  bool ok = true;
  for (int r = 0; r < size; r++) ok = ok && all[r][OK] == 1;
  if (!ok) { if (!fail) p2p_free(lo, total); return 1; }          /* every rank sees the same table -> same decision  */
  /* the lists must mirror each other; every rank checks the WHOLE table, so all of them reach the same verdict and none
     is left alone in the closing barrier */
  for (int r = 0; r < size; r++)
    for (int q = 0; q < size; q++) {
      const bool sends = all[r][RECV0 + R + q] >= 0, expects = all[q][RECV0 + r] >= 0;
      if (sends != expects) {
        p2p_free(lo, total);
        return b200_seterr(B200_ERR_PLIB, sends ? "rank %d does not expect the entries rank %d sends" : "rank %d expects entries rank %d does not send", q, r);
      }
    }
  for (int k = 1; k < RECV0; k++) B200_CUDA(cudaMemsetAsync(P->arena + off[k], 0, want[k], c.stream));
  PeerPlan *pp = (PeerPlan *)calloc(1, sizeof(PeerPlan));
  if (!pp) return b200_seterr(B200_ERR_MEM, "out of host memory");
  pp->arena_lo = lo; pp->arena_hi = hi;
  void *d0[B200_MAX_RANKS] = {0}, *d1[B200_MAX_RANKS] = {0}, *rf[B200_MAX_RANKS] = {0}, *lc[B200_MAX_RANKS] = {0}, *rc[B200_MAX_RANKS] = {0}, *lf[B200_MAX_RANKS] = {0};
  /* forward: I send to to_procs (their ghost vectors), I receive from from_procs */
  for (int k = 0; k < nto; k++) {
    const int q = M->to_procs[k];
    const long long rs = all[q][RECV0 + rank];
    d0[k] = P->peer[q] + all[q][LV0] + 8 * rs;
    d1[k] = P->peer[q] + all[q][LV1] + 8 * rs;
    rf[k] = P->peer[q] + all[q][FLG] + 8 * (size_t)rank;
    lc[k] = P->arena + off[CRD] + 8 * (size_t)q;
  }
  for (int k = 0; k < nfrom; k++) {
    const int q = M->from_procs[k];
    rc[k] = P->peer[q] + all[q][CRD] + 8 * (size_t)rank;
    lf[k] = P->arena + off[FLG] + 8 * (size_t)q;
  }
  B200_CHK(plan_blob(&pp->dev, &pp->d_blob, nto, M->to_starts, d0, d1, rf, lc, nfrom, rc, lf));
  /* reverse: I send my ghost contributions to from_procs (their return buffers), I receive from to_procs */
  for (int k = 0; k < nfrom; k++) {
    const int q = M->from_procs[k];
    const long long ss = all[q][RECV0 + R + rank];
    d0[k] = P->peer[q] + all[q][RB0] + 8 * ss;
    d1[k] = P->peer[q] + all[q][RB1] + 8 * ss;
    rf[k] = P->peer[q] + all[q][RFLG] + 8 * (size_t)rank;
    lc[k] = P->arena + off[RCRD] + 8 * (size_t)q;
  }
  for (int k = 0; k < nto; k++) {
    const int q = M->to_procs[k];
    rc[k] = P->peer[q] + all[q][RCRD] + 8 * (size_t)rank;
    lf[k] = P->arena + off[RFLG] + 8 * (size_t)q;
  }
  B200_CHK(plan_blob(&pp->rev, &pp->d_blob_rev, nfrom, M->from_starts, d0, d1, rf, lc, nto, rc, lf));
  B200_CUDA(cudaMalloc((void **)&pp->d_ticket, sizeof(unsigned)));
  B200_CUDA(cudaMemset(pp->d_ticket, 0, sizeof(unsigned)));
  B200_CUDA(cudaMalloc((void **)&pp->d_ticket_rev, sizeof(unsigned)));
  B200_CUDA(cudaMemset(pp->d_ticket_rev, 0, sizeof(unsigned)));
  M->d_lvec[0] = (double *)(P->arena + off[LV0]);
  M->d_lvec[1] = (double *)(P->arena + off[LV1]);
  pp->rbuf[0] = (double *)(P->arena + off[RB0]);
  pp->rbuf[1] = (double *)(P->arena + off[RB1]);
  M->peerplan = pp;
  /* the zeroing above must be complete on every rank before anybody's first store can land */
  B200_CUDA(cudaStreamSynchronize(c.stream));
  return b200_comm_barrier();
}

/* the part of MatCreate_MPIAIJ + MatSetUpMultiply that follows the two blocks: scatter lists, ghost buffers (peer arena or
 * plain), pack list, the one-launch plan.  M->A and M->B are in place (uploaded or adopted); M is destroyed on failure. */
int b200_mpiaij_finish(b200_mat M, int N, int nghost, const int *Bi, const int *garray,
                       int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                       int nfrom, const int *from_procs, const int *from_starts, b200_mat *Aout)
{
  int rc = 0;
  M->kind = 1; M->N = N; M->nghost = nghost; M->nto = nto; M->nfrom = nfrom;
  if ((nto > 0 && (!to_procs || !to_starts || !to_indices)) || (nfrom > 0 && (!from_procs || !from_starts)) || (nghost > 0 && !garray)) {
    b200_mat_destroy(M);
    return b200_seterr(B200_ERR_ARG_NULL, "null scatter list with nto %d, nfrom %d, nghost %d", nto, nfrom, nghost);
  }
  /* a single rank (or a rank without neighbours) may pass NULL lists: the copies are then all-zero */
  auto dup = [](const int *p, int n) { int *q = (int *)calloc((size_t)(n > 0 ? n : 1), sizeof(int)); if (q && p && n > 0) memcpy(q, p, sizeof(int) * (size_t)n); return q; };
  M->to_procs = dup(to_procs, nto); M->to_starts = dup(to_starts, nto + 1);
  M->from_procs = dup(from_procs, nfrom); M->from_starts = dup(from_starts, nfrom + 1);
  if (nto == 0) M->to_starts[0] = 0;
  if (nfrom == 0) M->from_starts[0] = 0;
  const int nsend = M->to_starts[nto];
  if (M->from_starts[nfrom] != nghost) { b200_mat_destroy(M); return b200_seterr(B200_ERR_ARG_SIZ, "receive lists cover %d of %d ghosts", M->from_starts[nfrom], nghost); }
  B200_CUDA(cudaMalloc((void **)&M->d_garray, sizeof(int) * (size_t)(nghost + 1)));
  if (g_b200.p2p && (g_b200.size > 1 || nto + nfrom > 0)) {
    /* collective; falls back to the NCCL exchange for this matrix (on every rank alike) when an arena is full */
    const int prc = peerplan_build(M);
    if (prc > 1) { b200_mat_destroy(M); return prc; }
  }
  if (!M->peerplan) {
    B200_CUDA(cudaMalloc((void **)&M->d_lvec[0], sizeof(double) * (size_t)(nghost + 2)));
    B200_CUDA(cudaMalloc((void **)&M->d_lvec[1], sizeof(double) * (size_t)(nghost + 2)));
    B200_CUDA(cudaMemset(M->d_lvec[0], 0, sizeof(double) * (size_t)(nghost + 2)));
    B200_CUDA(cudaMemset(M->d_lvec[1], 0, sizeof(double) * (size_t)(nghost + 2)));
  }
  B200_CUDA(cudaMalloc((void **)&M->d_to_indices, sizeof(int) * (size_t)(nsend + 1)));
  B200_CUDA(cudaMalloc((void **)&M->d_sendbuf, sizeof(double) * (size_t)(nsend + 2)));
  if (nghost) B200_CUDA(cudaMemcpy(M->d_garray, garray, sizeof(int) * (size_t)nghost, cudaMemcpyHostToDevice));
  if (nsend) B200_CUDA(cudaMemcpy(M->d_to_indices, to_indices, sizeof(int) * (size_t)nsend, cudaMemcpyHostToDevice));
  M->epoch = 0;
  if (M->peerplan) { rc = fusedplan_build(M, Bi); if (rc) { b200_mat_destroy(M); return rc; } }
  *Aout = M;
  return 0;
}

extern "C" int b200_mat_create_mpiaij(int m, int n_local, int N,
                                      const int *Ai, const int *Aj, const double *Aa,
                                      int nghost, const int *Bi, const int *Bj, const double *Ba, const int *garray,
                                      int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                                      int nfrom, const int *from_procs, const int *from_starts,
                                      b200_mat *Aout)
{
  B200_NEED_INIT();
  if (!Aout) return b200_seterr(B200_ERR_ARG_NULL, "null output handle");
  *Aout = NULL;
  b200_mat M = (b200_mat)calloc(1, sizeof(*M));
  if (!M) return b200_seterr(B200_ERR_MEM, "out of host memory");
  M->kind = 1; M->N = N; M->nghost = nghost; M->nto = nto; M->nfrom = nfrom;
  int rc = b200_seq_upload(&M->A, m, n_local, Ai, Aj, Aa);
  if (!rc) rc = b200_seq_upload(&M->B, m, nghost, Bi, Bj, Ba);
  if (rc) { b200_mat_destroy(M); return rc; }
  return b200_mpiaij_finish(M, N, nghost, Bi, garray, nto, to_procs, to_starts, to_indices, nfrom, from_procs, from_starts, Aout);
}

/* ------------------------------------------------------------------------------------------------ one-launch MatMult */
struct FusedPlan {
  int    ntiles, grid, nboundary, threads, tile_nnz;
  int   *d_bt_start, *d_bt_list;     /* boundary tiles of every CTA's walk (phase 2 of the kernel)                    */
  double *d_dotpart;
  unsigned *d_ticket_dot;
};

void b200_mpiaij_fused_free(b200_mat M)
{
  FusedPlan *fp = (FusedPlan *)M->fusedplan;
  if (!fp) return;
  cudaFree(fp->d_bt_start); cudaFree(fp->d_bt_list); cudaFree(fp->d_dotpart); cudaFree(fp->d_ticket_dot);
  free(fp);
  M->fusedplan = NULL;
}

bool b200_mpiaij_fused(b200_mat A) { return A && A->fusedplan != NULL; }

void b200_mpiaij_fused_dotparts(b200_mat A, const double **p, int *n)
{
  FusedPlan *fp = (FusedPlan *)A->fusedplan;
  *p = fp->d_dotpart; *n = fp->grid;
}

typedef void (*fused_kern_t)(const MpiSpmvArgs);
template <int G> static fused_kern_t fused_pick(bool add, bool dot)
{
  if (add) return dot ? spmv_mpiaij_kernel<G, true, true, 256, 2048> : spmv_mpiaij_kernel<G, true, false, 256, 2048>;
  return dot ? spmv_mpiaij_kernel<G, false, true, 256, 2048> : spmv_mpiaij_kernel<G, false, false, 256, 2048>;
}
static fused_kern_t fused_kernel(int G, bool add, bool dot)
{
  switch (G) {
    case 1:  return fused_pick<1>(add, dot);
    case 2:  return fused_pick<2>(add, dot);
    case 4:  return fused_pick<4>(add, dot);
    case 8:  return fused_pick<8>(add, dot);
    case 16: return fused_pick<16>(add, dot);
    default: return fused_pick<32>(add, dot);
  }
}

/* Decide whether this operator takes the one-launch path and, if so, lay out its tile table: the diagonal block's row
 * tiles, those without off-diagonal entries first.  Bi = host row offsets of the off-diagonal block. */
static int fusedplan_build(b200_mat M, const int *Bi)
{
  const B200Seq *A = &M->A, *B = &M->B;
  const char *e = getenv("B200_MPI_FUSED");
  if (e && !atoi(e)) return 0;
  if (!M->peerplan || A->m == 0 || A->ntiles == 0) return 0;
  if (A->nlong || A->compressed || A->nstrips > 1 || B->nstrips > 1 || A->cfg != 0 || A->park) return 0;
  if (A->m != A->n) return 0;
  if (B->nz * 4 > A->nz) return 0;                         /* a fat off-diagonal block streams better through its own TMA pipeline */
  std::vector<int4> tiles((size_t)A->ntiles);
  B200_CUDA(cudaMemcpy(tiles.data(), A->d_tiles, sizeof(int4) * (size_t)A->ntiles, cudaMemcpyDeviceToHost));
  int nb = 0;
  for (size_t t = 0; t < tiles.size(); t++) {
    if (tiles[t].y < 0) return 0;                          /* chunk of a long row: not on this path */
    nb += Bi[tiles[t].y] > Bi[tiles[t].x];
  }
  FusedPlan *fp = (FusedPlan *)calloc(1, sizeof(FusedPlan));
  if (!fp) return b200_seterr(B200_ERR_MEM, "out of host memory");
  fp->ntiles = A->ntiles; fp->nboundary = nb; fp->threads = 256; fp->tile_nnz = 2048;
  const size_t smem = SpmvSmem<256, 2048, 1>::BYTES;
  int occ = 0;
  for (int add = 0; add < 2; add++) for (int dot = 0; dot < 2; dot++) {
    fused_kern_t k = fused_kernel(A->G, add, dot);
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int o = 0;
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, k, 256, smem));
    if (!occ || o < occ) occ = o;
  }
  if (occ < 1) { free(fp); return 0; }
  /* every CTA of the grid is resident at once (waits on arrival flags never hold up a CTA that still has to send) */
  fp->grid = g_b200.nsm * occ;
  if (fp->grid > fp->ntiles) fp->grid = fp->ntiles;
  /* which tiles of each CTA's walk (t = c, c + grid, ...) hold rows with off-diagonal entries: phase 2 of the kernel */
  {
    std::vector<int> start((size_t)fp->grid + 1, 0), list;
    list.reserve((size_t)nb);
    for (int c = 0; c < fp->grid; c++) {
      for (int t = c; t < A->ntiles; t += fp->grid)
        if (Bi[tiles[t].y] > Bi[tiles[t].x]) list.push_back(t);
      start[(size_t)c + 1] = (int)list.size();
    }
    B200_CUDA(cudaMalloc((void **)&fp->d_bt_start, sizeof(int) * start.size()));
    B200_CUDA(cudaMemcpy(fp->d_bt_start, start.data(), sizeof(int) * start.size(), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMalloc((void **)&fp->d_bt_list, sizeof(int) * (list.size() + 1)));
    if (!list.empty()) B200_CUDA(cudaMemcpy(fp->d_bt_list, list.data(), sizeof(int) * list.size(), cudaMemcpyHostToDevice));
  }
  B200_CUDA(cudaMalloc((void **)&fp->d_dotpart, sizeof(double) * (size_t)(fp->grid + 1)));
  B200_CUDA(cudaMalloc((void **)&fp->d_ticket_dot, sizeof(unsigned)));
  B200_CUDA(cudaMemset(fp->d_ticket_dot, 0, sizeof(unsigned)));
  M->fusedplan = fp;
  return 0;
}

extern "C" int b200_mat_mpiaij_is_fused(b200_mat A, int *fused, int *boundary_tiles)
{
  if (!A || A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  if (fused) *fused = A->fusedplan ? 1 : 0;
  if (boundary_tiles) *boundary_tiles = A->fusedplan ? ((FusedPlan *)A->fusedplan)->nboundary : 0;
  return 0;
}

/* z = (add ? yin : 0) + A_d x + B ghost(x), exchange included, in ONE launch.  want_dot: per-CTA partials of x . (z - yin)
 * are left for the caller (b200_mpiaij_fused_dotparts). */
int b200_mpiaij_mult_fused(b200_mat M, const double *x, const double *yin, double *z, bool add, bool want_dot) { return b200_mpiaij_mult_fused_cg(M, x, yin, z, add, want_dot, NULL, NULL); }

/* cg != NULL (with want_dot): the kernel's last CTA also finishes x . z -> cg->red[3], all-reduces it (ar) and runs the CG
 * scalar step that follows p.w */
int b200_mpiaij_mult_fused_cg(b200_mat M, const double *x, const double *yin, double *z, bool add, bool want_dot, CgState *cg, const CgAr *ar)
{
  B200Ctx &c = g_b200;
  FusedPlan *fp = (FusedPlan *)M->fusedplan;
  PeerPlan *pp = (PeerPlan *)M->peerplan;
  M->epoch++;
  MpiSpmvArgs a;
  memset(&a, 0, sizeof(a));
  a.ntiles = fp->ntiles; a.tiles = M->A.d_tiles; a.bt_start = fp->d_bt_start; a.bt_list = fp->d_bt_list;
  a.ai = M->A.d_i; a.aj = M->A.d_j; a.aa = M->A.d_a;
  a.bi = M->B.d_i; a.bj = M->B.d_j; a.ba = M->B.d_a;
  a.x = x; a.yin = yin; a.z = z;
  a.lvec = M->d_lvec[M->epoch & 1];
  a.dotpart = fp->d_dotpart; a.dotx = x;
  a.nsend = M->to_starts[M->nto];
  a.npack = (M->nto + M->nfrom) ? (a.nsend + fp->threads - 1) / fp->threads : 0;
  if ((M->nto + M->nfrom) && a.npack < 1) a.npack = 1;
  if (a.npack > fp->grid) a.npack = fp->grid;
  a.to_idx = M->d_to_indices;
  a.plan = pp->dev;
  a.parity = (int)(M->epoch & 1); a.epoch = (unsigned long long)M->epoch; a.ticket = pp->d_ticket;
  a.cg = want_dot ? cg : NULL; a.ticket_dot = fp->d_ticket_dot;
  if (a.cg && ar) a.ar = *ar; else a.ar.size = 1;
  fused_kernel(M->A.G, add, want_dot)<<<fp->grid, fp->threads, SpmvSmem<256, 2048, 1>::BYTES, c.stream>>>(a);
  B200_LAUNCHED(1);
  return 0;
}

/* MatMult_MPIAIJ / MatMultAdd_MPIAIJ (mpiaij.c:1102-1143) */
int b200_mpiaij_mult(b200_mat A, const double *x, const double *yin, double *z, bool add)
{
  B200Ctx &c = g_b200;
  if (c.multi_stream) {                                   /* a caller that alternates streams: products of one matrix stay ordered */
    if (!A->ev_done) B200_CUDA(cudaEventCreateWithFlags(&A->ev_done, cudaEventDisableTiming));
    if (A->last_stream && A->last_stream != c.stream) B200_CUDA(cudaStreamWaitEvent(c.stream, A->ev_done, 0));
  }
  if (A->fusedplan) { B200_CHK(b200_mpiaij_mult_fused(A, x, yin, z, add, false)); }
  else {
    B200_CHK(b200_mpiaij_exchange_begin(A, x));
    B200_CHK(b200_seq_mult(&A->A, x, yin, z, add, NULL));
    B200_CHK(b200_mpiaij_exchange_end(A));
    B200_CHK(b200_seq_mult(&A->B, A->d_lvec[A->epoch & 1], z, z, true, NULL));
  }
  if (c.multi_stream) { B200_CUDA(cudaEventRecord(A->ev_done, c.stream)); A->last_stream = c.stream; }
  return 0;
}

extern "C" int b200_mat_mpiaij_get_lvec(b200_mat A, double **lvec, int *nghost)
{
  if (!A || A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  if (lvec) *lvec = A->d_lvec[A->epoch & 1];
  if (nghost) *nghost = A->nghost;
  return 0;
}

/* fill the ghost vector directly (single-process emulation of P ranks: the caller plays the scatter) */
extern "C" int b200_mat_mpiaij_set_lvec(b200_mat A, const double *ghost_values_dev)
{
  B200_NEED_INIT();
  if (!A || A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  for (int k = 0; k < 2; k++)
    if (A->nghost) B200_CUDA(cudaMemcpyAsync(A->d_lvec[k], ghost_values_dev, sizeof(double) * (size_t)A->nghost, cudaMemcpyDeviceToDevice, g_b200.stream));
  return 0;
}

/* VecScatterBegin: pack on the compute stream (x was produced there), then the sends/receives on the side stream so
 * the diagonal-block product that follows on the compute stream overlaps them. */
int b200_mpiaij_exchange_begin(b200_mat A, const double *x)
{
  B200Ctx &c = g_b200;
  A->epoch++;
  if ((c.size <= 1 && !A->peerplan) || (A->nto == 0 && A->nfrom == 0)) return 0;
  if (!c.nccl && !A->peerplan) return b200_seterr(B200_ERR_ORDER, "MPIAIJ MatMult on %d ranks needs b200_comm_init", c.size);
  const int nsend = A->to_starts[A->nto];
  double *lvec = A->d_lvec[A->epoch & 1];
  if (A->peerplan) {
    PeerPlan *pp = (PeerPlan *)A->peerplan;
    int g = (nsend + 1023) / 1024; if (g > B200_NSM * 2) g = B200_NSM * 2; if (g < 1) g = 1;
    /* on the side stream, so that the diagonal-block product (launched next on the compute stream) starts at once and
     * the two run side by side: the pack kernel is small and short, the persistent product fills the rest of the GPU */
    B200_CUDA(cudaEventRecord(c.ev_fork, c.stream));
    B200_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_fork, 0));
    pack_put_kernel<<<g, 256, 0, c.comm_stream>>>(nsend, A->d_to_indices, x, pp->dev, (int)(A->epoch & 1), (unsigned long long)A->epoch, pp->d_ticket);
    B200_LAUNCHED(1);
    B200_CUDA(cudaEventRecord(c.ev_join, c.comm_stream));
    return 0;
  }
  if (nsend) {
    int g = (nsend + 255) / 256; if (g > B200_NSM * 4) g = B200_NSM * 4;
    pack_kernel<<<g, 256, 0, c.stream>>>(nsend, A->d_to_indices, x, A->d_sendbuf);
    B200_LAUNCHED(1);
  }
  B200_CUDA(cudaEventRecord(c.ev_fork, c.stream));
  B200_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_fork, 0));
  B200_NCCL(g_nccl.GroupStart());
  for (int p = 0; p < A->nfrom; p++)
    B200_NCCL(g_nccl.Recv(lvec + A->from_starts[p], (size_t)(A->from_starts[p + 1] - A->from_starts[p]), ncclDouble, A->from_procs[p], (ncclComm_t)c.nccl, c.comm_stream));
  for (int p = 0; p < A->nto; p++)
    B200_NCCL(g_nccl.Send(A->d_sendbuf + A->to_starts[p], (size_t)(A->to_starts[p + 1] - A->to_starts[p]), ncclDouble, A->to_procs[p], (ncclComm_t)c.nccl, c.comm_stream));
  B200_NCCL(g_nccl.GroupEnd());
  B200_CUDA(cudaEventRecord(c.ev_join, c.comm_stream));
  return 0;
}

/* VecScatterEnd: the compute stream waits for the arrivals; no unpack (receives landed in place) */
int b200_mpiaij_exchange_end(b200_mat A)
{
  B200Ctx &c = g_b200;
  if ((c.size <= 1 && !A->peerplan) || (A->nto == 0 && A->nfrom == 0)) return 0;
  if (A->peerplan) {
    B200_CUDA(cudaStreamWaitEvent(c.stream, c.ev_join, 0));        /* x may be overwritten only after the pack kernel read it */
    if (A->nfrom) {
      wait_arrivals_kernel<<<1, 32, 0, c.stream>>>(((PeerPlan *)A->peerplan)->dev, (unsigned long long)A->epoch);
      B200_LAUNCHED(1);
    }
    return 0;
  }
  B200_CUDA(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
  return 0;
}

extern "C" int b200_mat_mpiaij_scatter(b200_mat A, const double *x)
{
  B200_NEED_INIT();
  if (!A || A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  B200_CHK(b200_mpiaij_exchange_begin(A, x));
  return b200_mpiaij_exchange_end(A);
}

/* ------------------------------------------------------------------------------------------------ reverse exchange */
/* UnPack_1 with ADD_VALUES (vpscat.c:520-570) for ONE neighbour's message: y[indices[i]] += buf[i].  The messages are
 * applied neighbour by neighbour in ascending rank order (the reference applies them in MPI_Waitany arrival order), so
 * entries that several neighbours contribute to are summed in a fixed order. */
__global__ void __launch_bounds__(256) unpack_add_kernel(int n, const int *__restrict__ idx, const double *__restrict__ buf, double *y)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) { const int r = __ldg(&idx[i]); y[r] = y[r] + buf[i]; }
}

/* reverse exchange, first half: every rank's ghost-side values (nghost entries, grouped by owner like lvec) travel to
 * their owners; *buf_out is where this rank will find what its neighbours returned (nsend entries, to_starts order) */
static int reverse_begin(b200_mat A, const double *ghost_values, const double **buf_out)
{
  B200Ctx &c = g_b200;
  const int nghost = A->nghost, nto = A->nto, nfrom = A->nfrom;
  *buf_out = A->d_sendbuf;
  A->repoch++;
  if ((c.size <= 1 && !A->peerplan) || (nto == 0 && nfrom == 0)) return 0;
  if (!c.nccl && !A->peerplan) return b200_seterr(B200_ERR_ORDER, "a reverse ghost exchange on %d ranks needs b200_comm_init", c.size);
  if (A->peerplan) {
    PeerPlan *pp = (PeerPlan *)A->peerplan;
    int g = (nghost + 1023) / 1024; if (g > B200_NSM * 2) g = B200_NSM * 2; if (g < 1) g = 1;
    pack_put_kernel<<<g, 256, 0, c.stream>>>(nghost, NULL, ghost_values, pp->rev, (int)(A->repoch & 1), (unsigned long long)A->repoch, pp->d_ticket_rev);
    B200_LAUNCHED(1);
    *buf_out = pp->rbuf[A->repoch & 1];
    return 0;
  }
  B200_CUDA(cudaEventRecord(c.ev_fork, c.stream));
  B200_CUDA(cudaStreamWaitEvent(c.comm_stream, c.ev_fork, 0));
  B200_NCCL(g_nccl.GroupStart());
  for (int k = 0; k < nto; k++)
    B200_NCCL(g_nccl.Recv(A->d_sendbuf + A->to_starts[k], (size_t)(A->to_starts[k + 1] - A->to_starts[k]), ncclDouble, A->to_procs[k], (ncclComm_t)c.nccl, c.comm_stream));
  for (int k = 0; k < nfrom; k++)
    B200_NCCL(g_nccl.Send(ghost_values + A->from_starts[k], (size_t)(A->from_starts[k + 1] - A->from_starts[k]), ncclDouble, A->from_procs[k], (ncclComm_t)c.nccl, c.comm_stream));
  B200_NCCL(g_nccl.GroupEnd());
  B200_CUDA(cudaEventRecord(c.ev_join, c.comm_stream));
  return 0;
}

/* second half: the compute stream waits until every neighbour's message is in the buffer */
static int reverse_wait(b200_mat A)
{
  B200Ctx &c = g_b200;
  if ((c.size <= 1 && !A->peerplan) || (A->nto == 0 && A->nfrom == 0)) return 0;
  if (A->peerplan) {
    if (A->nto) {
      wait_arrivals_kernel<<<1, 32, 0, c.stream>>>(((PeerPlan *)A->peerplan)->rev, (unsigned long long)A->repoch);
      B200_LAUNCHED(1);
    }
    return 0;
  }
  B200_CUDA(cudaStreamWaitEvent(c.stream, c.ev_join, 0));
  return 0;
}

/* MatMultTranspose_MPIAIJ / MatMultTransposeAdd_MPIAIJ (mpiaij.c:1147-1195): lvec = B^T x ; reverse scatter (ADD) begins ;
 * z = A^T x (+ y) ; scatter ends: every rank adds what its neighbours computed for the rows it owns. */
int b200_mpiaij_multtranspose(b200_mat A, const double *x, const double *yin, double *z)
{
  B200Ctx &c = g_b200;
  if (!A || A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  B200_CHK(b200_seq_transpose(&A->A));
  B200_CHK(b200_seq_transpose(&A->B));
  const int nghost = A->nghost, nto = A->nto;
  if (!A->d_ghost_t) {
    B200_CUDA(cudaMalloc((void **)&A->d_ghost_t, sizeof(double) * (size_t)(nghost + 2)));
    B200_CUDA(cudaMemset(A->d_ghost_t, 0, sizeof(double) * (size_t)(nghost + 2)));
  }
  if (nghost) { B200_CHK(b200_seq_mult(A->B.T, x, NULL, A->d_ghost_t, false, NULL)); }
  const double *buf = NULL;
  B200_CHK(reverse_begin(A, A->d_ghost_t, &buf));
  /* local part */
  if (A->A.m == 0) { if (yin && z != yin && A->A.n > 0) B200_CHK(b200_vec_copy(A->A.n, yin, z)); }
  else if (yin) { B200_CHK(b200_seq_mult(A->A.T, x, yin, z, true, NULL)); }
  else { B200_CHK(b200_seq_mult(A->A.T, x, NULL, z, false, NULL)); }
  if (c.size <= 1 && !A->peerplan) return 0;
  B200_CHK(reverse_wait(A));
  for (int k = 0; k < nto; k++) {
    const int len = A->to_starts[k + 1] - A->to_starts[k];
    if (len <= 0) continue;
    int g = (len + 255) / 256; if (g > B200_NSM * 4) g = B200_NSM * 4;
    unpack_add_kernel<<<g, 256, 0, c.stream>>>(len, A->d_to_indices + A->to_starts[k], buf + A->to_starts[k], z);
    B200_LAUNCHED(1);
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------ general VecScatter */
/* A device-resident VecScatter between a parallel vector (x, this rank's part) and a sequential vector (y): the general
 * MPI -> Seq case of VecScatterCreate_PtoS (vpscat.c:1730-1924) with VecScatterBegin_1 / End_1 (vpscat.h:14-233) in both
 * modes (SCATTER_FORWARD x -> y, SCATTER_REVERSE y -> x) and both InsertModes.  The exchange engine is the one the
 * MPIAIJ operator uses (an operator-shaped object without matrix blocks): forward messages land contiguously in its
 * ghost vector and are then unpacked to from_indices; reverse messages land in its return buffer and are unpacked to
 * to_indices.  The self part (entries this rank owns) is a device copy done in Begin like vpscat.h:100-128. */
/* A target list that names an entry more than once (ADD_VALUES into shared ghost points, or INSERT_VALUES where the last
 * writer wins) is applied through a segment plan: entries grouped by target (stable, so positions ascend inside a
 * group), one thread per distinct target, contributions taken in list order — the order of the reference's sequential
 * loop `y[idx[i]] += v[i]` (vpscat.c:520-570, vscat.c:316-330).  No floating-point atomics: results are reproducible. */
struct SegPlan { int nseg; int *d_target, *d_start, *d_pos; };

struct b200_scatter_s {
  b200_mat ex;
  int      nx, ny, nrecv, nsend, nlocal;
  int     *d_from_indices, *d_to_local, *d_from_local;
  double  *d_stage;            /* [nrecv] reverse mode: y[from_indices] gathered before it travels                 */
  SegPlan  seg_from, seg_to, seg_local_from, seg_local_to;   /* nseg > 0: the list repeats targets               */
};

template <bool ADD>
__global__ void __launch_bounds__(256) scatter_store_kernel(int n, const int *__restrict__ dst_idx, const int *__restrict__ src_idx,
                                                            const double *__restrict__ src, double *dst)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double v = src[src_idx ? __ldg(&src_idx[i]) : i];
    const int r = dst_idx ? __ldg(&dst_idx[i]) : i;
    if (!ADD) dst[r] = v;
    else dst[r] = dst[r] + v;
  }
}

template <bool ADD>
__global__ void __launch_bounds__(256) scatter_seg_kernel(int nseg, const int *__restrict__ target, const int *__restrict__ start,
                                                          const int *__restrict__ pos, const int *__restrict__ src_idx,
                                                          const double *__restrict__ src, double *dst)
{
  const int stride = gridDim.x * blockDim.x;
  for (int s = blockIdx.x * blockDim.x + threadIdx.x; s < nseg; s += stride) {
    const int t = __ldg(&target[s]), e = __ldg(&start[s + 1]);
    int k = ADD ? __ldg(&start[s]) : e - 1;                /* INSERT_VALUES: the last entry of the list wins       */
    double acc = ADD ? dst[t] : 0.0;
    for (; k < e; k++) {
      const int i = __ldg(&pos[k]);
      const double v = src[src_idx ? __ldg(&src_idx[i]) : i];
      acc = ADD ? acc + v : v;
    }
    dst[t] = acc;
  }
}

static int scatter_store(int n, const int *dst_idx, const int *src_idx, const double *src, double *dst, bool add, const SegPlan *seg)
{
  if (n <= 0) return 0;
  cudaStream_t st = g_b200.stream;
  if (seg && seg->nseg > 0) {
    int g = (seg->nseg + 255) / 256; if (g > B200_NSM * 4) g = B200_NSM * 4;
    if (add) scatter_seg_kernel<true><<<g, 256, 0, st>>>(seg->nseg, seg->d_target, seg->d_start, seg->d_pos, src_idx, src, dst);
    else scatter_seg_kernel<false><<<g, 256, 0, st>>>(seg->nseg, seg->d_target, seg->d_start, seg->d_pos, src_idx, src, dst);
    B200_LAUNCHED(1);
    return 0;
  }
  int g = (n + 255) / 256; if (g > B200_NSM * 4) g = B200_NSM * 4;
  if (!add) scatter_store_kernel<false><<<g, 256, 0, st>>>(n, dst_idx, src_idx, src, dst);
  else scatter_store_kernel<true><<<g, 256, 0, st>>>(n, dst_idx, src_idx, src, dst);
  B200_LAUNCHED(1);
  return 0;
}

/* builds the plan only when `tgt` repeats an entry (nseg stays 0 otherwise) */
static int segplan_build(SegPlan *P, const int *tgt, int n)
{
  P->nseg = 0; P->d_target = P->d_start = P->d_pos = NULL;
  if (n < 2) return 0;
  std::vector<int> pos((size_t)n);
  for (int i = 0; i < n; i++) pos[i] = i;
  std::stable_sort(pos.begin(), pos.end(), [tgt](int a, int b) { return tgt[a] < tgt[b]; });
  bool dup = false;
  for (int i = 1; i < n && !dup; i++) dup = tgt[pos[i]] == tgt[pos[i - 1]];
  if (!dup) return 0;
  std::vector<int> target, start;
  for (int i = 0; i < n; i++) if (i == 0 || tgt[pos[i]] != tgt[pos[i - 1]]) { target.push_back(tgt[pos[i]]); start.push_back(i); }
  start.push_back(n);
  const int nseg = (int)target.size();
  B200_CUDA(cudaMalloc((void **)&P->d_target, sizeof(int) * (size_t)nseg));
  B200_CUDA(cudaMalloc((void **)&P->d_start, sizeof(int) * (size_t)(nseg + 1)));
  B200_CUDA(cudaMalloc((void **)&P->d_pos, sizeof(int) * (size_t)n));
  B200_CUDA(cudaMemcpy(P->d_target, target.data(), sizeof(int) * (size_t)nseg, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemcpy(P->d_start, start.data(), sizeof(int) * (size_t)(nseg + 1), cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemcpy(P->d_pos, pos.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
  P->nseg = nseg;
  return 0;
}

static void segplan_free(SegPlan *P) { cudaFree(P->d_target); cudaFree(P->d_start); cudaFree(P->d_pos); P->nseg = 0; P->d_target = P->d_start = P->d_pos = NULL; }

extern "C" int b200_scatter_create(int nx_local, int ny,
                                   int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                                   int nfrom, const int *from_procs, const int *from_starts, const int *from_indices,
                                   int nlocal, const int *to_local, const int *from_local, b200_scatter *out)
{
  B200_NEED_INIT();
  if (!out) return b200_seterr(B200_ERR_ARG_NULL, "null output handle");
  *out = NULL;
  if (nto < 0 || nfrom < 0 || nlocal < 0 || nx_local < 0 || ny < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative size");
  const int nrecv = nfrom ? from_starts[nfrom] : 0, nsend = nto ? to_starts[nto] : 0;
  for (int i = 0; i < nsend; i++) if (to_indices[i] < 0 || to_indices[i] >= nx_local) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "send index %d outside the local part [0,%d)", to_indices[i], nx_local);
  for (int i = 0; i < nrecv; i++) if (from_indices[i] < 0 || from_indices[i] >= ny) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "Scattering past end of TO vector");
  for (int i = 0; i < nlocal; i++)
    if (to_local[i] < 0 || to_local[i] >= nx_local || from_local[i] < 0 || from_local[i] >= ny) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "local scatter index out of range");
  b200_scatter S = (b200_scatter)calloc(1, sizeof(*S));
  if (!S) return b200_seterr(B200_ERR_MEM, "out of host memory");
  S->nx = nx_local; S->ny = ny; S->nrecv = nrecv; S->nsend = nsend; S->nlocal = nlocal;
  static const int zero_i[2] = {0, 0};
  std::vector<int> ghost_ids((size_t)(nrecv > 0 ? nrecv : 1));
  for (int i = 0; i < nrecv; i++) ghost_ids[i] = i;
  int rc = b200_mat_create_mpiaij(0, 0, 0, zero_i, NULL, NULL, nrecv, zero_i, NULL, NULL, ghost_ids.data(),
                                  nto, to_procs, to_starts, to_indices, nfrom, from_procs, from_starts, &S->ex);
  if (rc) { free(S); return rc; }
  auto up = [](int **d, const int *h, int n) -> cudaError_t {
    cudaError_t e = cudaMalloc((void **)d, sizeof(int) * (size_t)(n + 1));
    if (e == cudaSuccess && n) e = cudaMemcpy(*d, h, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice);
    return e;
  };
  B200_CUDA(up(&S->d_from_indices, from_indices, nrecv));
  B200_CUDA(up(&S->d_to_local, to_local, nlocal));
  B200_CUDA(up(&S->d_from_local, from_local, nlocal));
  B200_CUDA(cudaMalloc((void **)&S->d_stage, sizeof(double) * (size_t)(nrecv + 2)));
  B200_CHK(segplan_build(&S->seg_from, from_indices, nrecv));
  B200_CHK(segplan_build(&S->seg_local_from, from_local, nlocal));
  B200_CHK(segplan_build(&S->seg_local_to, to_local, nlocal));
  B200_CHK(segplan_build(&S->seg_to, to_indices, nsend));     /* over all neighbours: ascending rank, then list order */
  *out = S;
  return 0;
}

/* Sequential vector -> sequential vector (VecScatterCreate's SGtoSG / SGtoSS / SStoSG / SStoSS cases, vscat.c:316-700):
 * y[dst_idx[i]] (+)= x[src_idx[i]], nothing travels, not collective. */
extern "C" int b200_scatter_create_local(int nx, int ny, int n, const int *src_idx, const int *dst_idx, b200_scatter *out)
{
  B200_NEED_INIT();
  if (!out) return b200_seterr(B200_ERR_ARG_NULL, "null output handle");
  *out = NULL;
  if (n < 0 || nx < 0 || ny < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative size");
  for (int i = 0; i < n; i++)
    if (src_idx[i] < 0 || src_idx[i] >= nx || dst_idx[i] < 0 || dst_idx[i] >= ny) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "local scatter index out of range");
  b200_scatter S = (b200_scatter)calloc(1, sizeof(*S));
  if (!S) return b200_seterr(B200_ERR_MEM, "out of host memory");
  S->nx = nx; S->ny = ny; S->nlocal = n;
  B200_CUDA(cudaMalloc((void **)&S->d_to_local, sizeof(int) * (size_t)(n + 1)));
  B200_CUDA(cudaMalloc((void **)&S->d_from_local, sizeof(int) * (size_t)(n + 1)));
  if (n) {
    B200_CUDA(cudaMemcpy(S->d_to_local, src_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(S->d_from_local, dst_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
  }
  B200_CHK(segplan_build(&S->seg_local_from, dst_idx, n));
  B200_CHK(segplan_build(&S->seg_local_to, src_idx, n));
  *out = S;
  return 0;
}

extern "C" int b200_scatter_destroy(b200_scatter S)
{
  if (!S) return 0;
  if (S->ex) b200_mat_destroy(S->ex);
  cudaFree(S->d_from_indices); cudaFree(S->d_to_local); cudaFree(S->d_from_local); cudaFree(S->d_stage);
  segplan_free(&S->seg_from); segplan_free(&S->seg_to); segplan_free(&S->seg_local_from); segplan_free(&S->seg_local_to);
  free(S);
  return 0;
}

/* x: this rank's part of the parallel vector (nx_local), y: the sequential vector (ny).  add: 0 INSERT_VALUES, 1 ADD_VALUES;
 * reverse: 0 SCATTER_FORWARD (x -> y), 1 SCATTER_REVERSE (y -> x). */
extern "C" int b200_scatter_begin(b200_scatter S, double *x, double *y, int add, int reverse)
{
  B200_NEED_INIT();
  if (!S || !x || !y) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (!S->ex) {                                            /* sequential -> sequential: the whole scatter is the self part */
    if (!reverse) return scatter_store(S->nlocal, S->d_from_local, S->d_to_local, x, y, add != 0, &S->seg_local_from);
    return scatter_store(S->nlocal, S->d_to_local, S->d_from_local, y, x, add != 0, &S->seg_local_to);
  }
  if (!reverse) {
    B200_CHK(b200_mpiaij_exchange_begin(S->ex, x));
    return scatter_store(S->nlocal, S->d_from_local, S->d_to_local, x, y, add != 0, &S->seg_local_from);
  }
  B200_CHK(scatter_store(S->nrecv, NULL, S->d_from_indices, y, S->d_stage, false, NULL));     /* Pack: stage[i] = y[from_indices[i]] */
  const double *buf = NULL;
  B200_CHK(reverse_begin(S->ex, S->d_stage, &buf));
  return scatter_store(S->nlocal, S->d_to_local, S->d_from_local, y, x, add != 0, &S->seg_local_to);
}

extern "C" int b200_scatter_end(b200_scatter S, double *x, double *y, int add, int reverse)
{
  B200_NEED_INIT();
  if (!S || !x || !y) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  b200_mat E = S->ex;
  if (!E) return 0;
  if (!reverse) {
    B200_CHK(b200_mpiaij_exchange_end(E));
    if (g_b200.size <= 1 && !E->peerplan) return 0;
    return scatter_store(S->nrecv, S->d_from_indices, NULL, E->d_lvec[E->epoch & 1], y, add != 0, &S->seg_from);
  }
  if (g_b200.size <= 1 && !E->peerplan) return 0;
  B200_CHK(reverse_wait(E));
  const double *buf = E->peerplan ? ((PeerPlan *)E->peerplan)->rbuf[E->repoch & 1] : E->d_sendbuf;
  /* one pass over every neighbour's message: entries several neighbours contribute to are summed by the segment plan in
     ascending rank order (the reference takes them in MPI_Waitany arrival order, vpscat.h:200-216) */
  return scatter_store(S->nsend, E->d_to_indices, NULL, buf, x, add != 0, &S->seg_to);
}
