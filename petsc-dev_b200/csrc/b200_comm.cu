/* b200_comm.cu — rank-to-rank traffic of the hot path: the tiny allreduces of VecDot/VecNorm/VecMDot
 * (MPI_Allreduce in src/vec/vec/impls/mpi/pbvec.c:16, pvec2.c:19,61) and the MPIAIJ ghost exchange
 * (VecScatterBegin_1/End_1, src/vec/vec/utils/vpscat.h:14-233), device-resident, one rank per GPU.
 *
 * Transport 1 (default): NCCL, resolved at run time from the process (torch's bundled libnccl when the host program
 * is Python, the system one otherwise) — grouped ncclSend/ncclRecv per neighbour straight into the ghost vector
 * (receives are contiguous because garray is sorted: `from->contiq`, vpscat.c:1951-1960), on a side stream that
 * overlaps the diagonal-block product.
 * Transport 2 (b200_comm_enable_p2p): NVLink peer stores — see the second half of this file.
 */
#include "b200_internal.h"
#include <dlfcn.h>
#include <nccl.h>

struct NcclApi {
  void *h;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *);
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
  ncclResult_t (*CommDestroy)(ncclComm_t);
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
  ncclResult_t (*GroupStart)(void);
  ncclResult_t (*GroupEnd)(void);
  const char *(*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl = {};

static int nccl_load(void)
{
  if (g_nccl.h) return 0;
  void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!h) return b200_seterr(B200_ERR_LIB, "cannot load libnccl.so.2: %s", dlerror());
#define SYM(f) do { *(void **)(&g_nccl.f) = dlsym(h, "nccl" #f); if (!g_nccl.f) return b200_seterr(B200_ERR_LIB, "libnccl lacks nccl" #f); } while (0)
  SYM(GetUniqueId); SYM(CommInitRank); SYM(CommDestroy); SYM(AllReduce); SYM(Send); SYM(Recv); SYM(GroupStart); SYM(GroupEnd); SYM(GetErrorString);
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
  if (g_b200.nccl) { g_nccl.CommDestroy((ncclComm_t)g_b200.nccl); g_b200.nccl = NULL; }
  g_b200.rank = 0; g_b200.size = 1; g_b200.p2p = false;
  return 0;
}

int b200_comm_allreduce_dev(double *dvals, int n, int op_max)
{
  if (g_b200.size <= 1) return 0;
  if (!g_b200.nccl) return b200_seterr(B200_ERR_ORDER, "communicator not initialised");
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

extern "C" int b200_comm_enable_p2p(int enable)
{
  if (enable) return b200_seterr(B200_ERR_SUP, "peer-store transport is not built in this version");
  g_b200.p2p = false;
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
  auto dup = [](const int *p, int n) { int *q = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1)); if (n > 0) memcpy(q, p, sizeof(int) * (size_t)n); return q; };
  M->to_procs = dup(to_procs, nto); M->to_starts = dup(to_starts, nto + 1);
  M->from_procs = dup(from_procs, nfrom); M->from_starts = dup(from_starts, nfrom + 1);
  if (nto == 0) M->to_starts[0] = 0;
  if (nfrom == 0) M->from_starts[0] = 0;
  const int nsend = M->to_starts[nto];
  if (M->from_starts[nfrom] != nghost) { b200_mat_destroy(M); return b200_seterr(B200_ERR_ARG_SIZ, "receive lists cover %d of %d ghosts", M->from_starts[nfrom], nghost); }
  B200_CUDA(cudaMalloc((void **)&M->d_garray, sizeof(int) * (size_t)(nghost + 1)));
  B200_CUDA(cudaMalloc((void **)&M->d_lvec[0], sizeof(double) * (size_t)(nghost + 2)));
  B200_CUDA(cudaMalloc((void **)&M->d_lvec[1], sizeof(double) * (size_t)(nghost + 2)));
  B200_CUDA(cudaMalloc((void **)&M->d_to_indices, sizeof(int) * (size_t)(nsend + 1)));
  B200_CUDA(cudaMalloc((void **)&M->d_sendbuf, sizeof(double) * (size_t)(nsend + 2)));
  if (nghost) B200_CUDA(cudaMemcpy(M->d_garray, garray, sizeof(int) * (size_t)nghost, cudaMemcpyHostToDevice));
  if (nsend) B200_CUDA(cudaMemcpy(M->d_to_indices, to_indices, sizeof(int) * (size_t)nsend, cudaMemcpyHostToDevice));
  B200_CUDA(cudaMemset(M->d_lvec[0], 0, sizeof(double) * (size_t)(nghost + 2)));
  B200_CUDA(cudaMemset(M->d_lvec[1], 0, sizeof(double) * (size_t)(nghost + 2)));
  M->epoch = 0;
  *Aout = M;
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
  if (c.size <= 1 || (A->nto == 0 && A->nfrom == 0)) return 0;
  if (!c.nccl) return b200_seterr(B200_ERR_ORDER, "MPIAIJ MatMult on %d ranks needs b200_comm_init", c.size);
  const int nsend = A->to_starts[A->nto];
  double *lvec = A->d_lvec[A->epoch & 1];
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
  if (c.size <= 1 || (A->nto == 0 && A->nfrom == 0)) return 0;
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
