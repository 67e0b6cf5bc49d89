/* b200_internal.h — shared state of libpetscb200 (one process = one rank = one B200). */
#ifndef B200_INTERNAL_H
#define B200_INTERNAL_H

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>
#include "petscb200.h"

#define B200_MAX_RANKS   16
#define B200_NSM         148           /* B200: 148 SMs; grids are sized in multiples of it                        */
#define B200_RED_MAXV    32            /* values one reduction launch can produce (VecMDot chunk)                  */
#define B200_RED_MAXBLK  (B200_NSM * 8)
#define P2P_AR_SLOT      64            /* u64 words per allreduce inbox slot: sequence number + up to 63 values   */
#define P2P_AR_MAXV      (P2P_AR_SLOT - 1)
#define B200_RES_SLOTS   256           /* doubles in the pinned/mapped result buffer                              */

struct B200P2P;                         /* peer-memory state, b200_comm.cu */

struct B200Ctx {
  bool         inited;
  int          device;
  int          nsm;
  cudaStream_t own_stream, stream, comm_stream;
  cudaEvent_t  ev_t0, ev_t1, ev_fork, ev_join;
  double      *h_res;        /* pinned + mapped host memory: reductions land here                                  */
  double      *d_res;        /* device alias of h_res                                                              */
  double      *d_partials;   /* [B200_RED_MAXBLK * B200_RED_MAXV]                                                  */
  unsigned    *d_ticket;     /* last-block tickets (one per concurrent reduction slot)                             */
  double      *d_scal;       /* device scalars (allreduce staging, fused-CG state)                                 */
  void        *d_ptrtab;     /* device copy of pointer tables for MDot/MAXPY                                        */
  void        *h_ptrtab;     /* pinned staging for the same                                                        */
  void        *l2flush;
  size_t       l2bytes;
  long long    launches;
  /* communicator */
  int          rank, size;
  void        *nccl;         /* ncclComm_t                                                                         */
  bool         p2p;
  B200P2P     *peer;
  bool         multi_stream;  /* the caller has handed in a stream of its own at least once (b200_set_stream)          */
};

extern B200Ctx g_b200;

int  b200_seterr(int code, const char *fmt, ...);
int  b200_cuda_err(cudaError_t e, const char *what, const char *file, int line);
bool b200_ready(void);
/* host -> device copy of a large pageable array through parallel pinned staging (b200_runtime.cu); synchronous */
int  b200_upload_pageable(void *dst, const void *src, size_t bytes);
void b200_upload_release(void);

#define B200_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return b200_cuda_err(e__, #call, __FILE__, __LINE__); } while (0)
#define B200_CHK(call)  do { int r__ = (call); if (r__) return r__; } while (0)
#define B200_NEED_INIT() do { if (!g_b200.inited) { int r__ = b200_init(-1); if (r__) return r__; } } while (0)
#define B200_LAUNCHED(k) do { g_b200.launches += (k); cudaError_t e__ = cudaGetLastError(); if (e__ != cudaSuccess) return b200_cuda_err(e__, "kernel launch", __FILE__, __LINE__); } while (0)

/* ---- matrices ------------------------------------------------------------------------------------------------ */
struct B200Seq {              /* one CSR block on the device + its row binning                                        */
  int        m, n;
  long long  nz;
  int        nonzerorow;      /* rows with at least one entry (flop count 2 nz - nonzerorow, aij.c:1281)            */
  int       *d_i;             /* [m+1] row offsets                                                                   */
  int       *d_j;             /* [nz + pad] column ids                                                               */
  double    *d_a;             /* [nz + pad] values                                                                   */
  bool       compressed;      /* rows are visited through the non-empty-row list (compressedrow, aij.h:99-115)       */
  int        ncrows;          /* rows in the visiting order (m, or the number of non-empty rows)                     */
  int       *d_ci;            /* [ncrows+1] offsets in visiting order (== d_i when not compressed)                   */
  int       *d_rindex;        /* [ncrows] visiting order -> row id (NULL when not compressed)                        */
  int        cfg, G, grid;    /* tile-kernel launch shape, lanes per row (1..32), persistent grid size               */
  int        long_thresh;     /* rows longer than this go to the long-row kernel                                     */
  int        ntiles;          /* row tiles: contiguous runs of rows, <= tile_nnz entries and <= threads/G rows each  */
  int4      *d_tiles;         /* {row_begin, row_end, k_begin, k_end} in visiting order                              */
  int        park;            /* 1: products are parked in shared memory before the row sums (irregular matrices)    */
  int        nlong;           /* rows longer than long_thresh: cut into chunk tiles                                   */
  int        nchunks;
  int       *d_long;          /* their positions in visiting order                                                   */
  int       *d_longstart;     /* [nlong+1] first chunk slot of each long row                                          */
  double    *d_longpart;      /* [nchunks] per-chunk partial sums                                                     */
  double    *d_dotpart;       /* [grid + nlong] per-CTA partials of the fused x.y                                    */
  /* warp-tile form of an irregular block (spmv_warp_kernel): its own tile table and long-row chunks (256 entries each)     */
  int        nwtiles, nwlong, nwchunks;
  int4      *d_wtiles;
  int       *d_wlong, *d_wlongstart;
  double    *d_wlongpart;
  int        park_cap;        /* parked variant: entries a row's own lane sums before its warp takes over (0: all)  */
  unsigned  *d_ticket;        /* last-CTA ticket of the DOT variants' CG tail                                         */
  int       *d_diagpos;       /* [m] position of a_ii in d_a or -1 (Mat_SeqAIJ::diag, aij.c:1486 MatMarkDiagonal_SeqAIJ);
                                 recorded by the first MatGetDiagonal, kept across value-only refreshes               */
  /* column strips (cache blocking for matrices whose x gathers would not stay in L2): when nstrips > 1 the block is
   * stored as nstrips CSR blocks over disjoint column ranges (global column ids kept) and the product is
   * y = S_0 x; y += S_1 x; ...  — for one lane per row the row sum is still one left-to-right chain */
  int        nstrips;
  B200Seq   *strips;
  int        req_G;           /* lanes-per-row request in force when the block was created (-1 automatic)            */
  B200Seq   *T;               /* explicit transpose (n x m CSR, source rows ascending), built at the first MatMultTranspose */
  int        perm_pad;
  long long *strip_nzoff;     /* unused on device; host: first entry of each strip in strip-major order (nstrips+1)  */
};

struct b200_mat_s {
  int      kind;              /* 0 = SeqAIJ, 1 = MPIAIJ                                                               */
  B200Seq  A;                 /* the matrix (seq) or the diagonal block (mpi)                                         */
  /* MPIAIJ part (Mat_MPIAIJ, src/mat/impls/aij/mpi/mpiaij.h:34-77) */
  B200Seq  B;                 /* off-diagonal block, columns index garray / lvec                                      */
  int      N;                 /* global columns                                                                       */
  int      nghost;
  int     *d_garray;          /* [nghost] sorted global ids of the ghost columns                                      */
  double  *d_lvec[2];         /* ghost values (double-buffered for the peer-store exchange)                           */
  int      nto, nfrom;
  int     *to_procs, *to_starts, *from_procs, *from_starts;   /* host copies of the VecScatter lists                 */
  int     *d_to_indices;      /* [to_starts[nto]] local offsets to pack                                               */
  double  *d_sendbuf;         /* packed send values                                                                   */
  long long epoch;            /* exchange counter                                                                     */
  long long repoch;           /* reverse-exchange counter (MatMultTranspose)                                          */
  double  *d_ghost_t;         /* [nghost] B^T x, what this rank contributes to rows its neighbours own                */
  void    *peerplan;          /* per-matrix peer-store plan (b200_comm.cu)                                            */
  void    *batchplan;         /* element -> nonzero plan kept by the device assembly (b200_assembly.cu), or NULL       */
  void    *fusedplan;         /* one-launch MatMult (exchange + both blocks in one kernel, b200_comm.cu), or NULL      */
  cudaEvent_t  ev_done;       /* end of the last product, for callers that alternate between streams                   */
  cudaStream_t last_stream;
};

int  b200_seq_upload(B200Seq *S, int m, int n, const int *i, const int *j, const double *a);
int  b200_seq_adopt(B200Seq *S, int m, int n, const int *i_host, int *d_i, int *d_j, double *d_a);
int  b200_mpiaij_finish(b200_mat M, int N, int nghost, const int *Bi, const int *garray, int nto, const int *to_procs, const int *to_starts,
                        const int *to_indices, int nfrom, const int *from_procs, const int *from_starts, b200_mat *Aout);
void *b200_pinned_scratch(size_t bytes);   /* one growing pinned host buffer for short-lived read-backs (single host thread per rank) */
void b200_seq_free(B200Seq *S);
int  b200_seq_mult(const B200Seq *S, const double *x, const double *yin, double *z, bool add, double *dot_out);
struct CgState;
int  b200_seq_mult_ex(const B200Seq *S, const double *x, const double *yin, double *z, bool add, bool want_dot, const double *dotx, double *dot_out,
                      CgState *cg = NULL);   /* cg: the kernel's last CTA also takes the CG scalar step after p.w (fused CG on one GPU) */
int  b200_seq_transpose(B200Seq *S);
void b200_seq_dotparts(const B200Seq *S, const double **p, int *n);
int  b200_mpiaij_mult_dot(b200_mat A, const double *x, double *y, double *dot_out);
int  b200_mpiaij_mult(b200_mat A, const double *x, const double *yin, double *z, bool add);   /* MatMult / MatMultAdd of the parallel type */
bool b200_mpiaij_fused(b200_mat A);
int  b200_mpiaij_mult_fused(b200_mat A, const double *x, const double *yin, double *z, bool add, bool want_dot);
struct CgState; struct CgAr;
int  b200_mpiaij_mult_fused_cg(b200_mat A, const double *x, const double *yin, double *z, bool add, bool want_dot, CgState *cg, const CgAr *ar);
void b200_mpiaij_fused_dotparts(b200_mat A, const double **p, int *n);
void b200_mpiaij_fused_free(b200_mat A);

/* internal cross-file entry points */
void b200_reduce_arm(int nvals);   /* before the launch: stamp nvals result slots so that b200_reduce_finish can poll them (0 = do not poll) */
int b200_reduce_finish(int nvals, double *result_host, bool allreduce, int op_max /*0 sum, 1 max*/);
int b200_comm_allreduce_dev(double *dvals, int n, int op_max);
int b200_mpiaij_exchange_begin(b200_mat A, const double *x);
int b200_mpiaij_exchange_end(b200_mat A);
int b200_mpiaij_multtranspose(b200_mat A, const double *x, const double *yin, double *z);   /* yin NULL: plain transpose product */
void b200_mpiaij_peer_free(b200_mat A);
void b200_batchplan_free(void *plan);
bool b200_p2p_next(unsigned long long *seq, unsigned long long *const **inbox);

struct CgState;
bool b200_cg_persistent_ok(b200_mat A);
int  b200_cg_persistent_run(b200_mat A, int jacobi, const double *d, double *X, double *R, double *ZW, double *P, CgState *ds, double *dhist, int iters);

/* reduction grid helper */
static inline int b200_red_grid(long long n, int threads, int per_thread)
{
  long long g = (n + (long long)threads * per_thread - 1) / ((long long)threads * per_thread);
  if (g < 1) g = 1;
  if (g > B200_RED_MAXBLK) g = B200_RED_MAXBLK;
  return (int)g;
}

#endif
