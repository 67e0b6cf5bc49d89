/* petscb200.h — the C ABI of libpetscb200.so: a B200 (sm_100a) implementation of PETSc's sparse-solve hot path
 * (AIJ MatMult/MatMultAdd, the MPIAIJ ghost exchange, the Krylov vector kernels, KSPSolve_CG / KSPSolve_GMRES).
 *
 * This header is the drop-in boundary.  Everything is `extern "C"`, plain pointers and sizes; `double *` vector
 * arguments are DEVICE pointers unless the name says `host`.  PetscInt is 32-bit `int`, PetscScalar is `double`
 * (the reference's default build, include/petscsys.h).  Every function returns a PetscErrorCode-compatible int
 * (0 = success, otherwise one of the PETSC_ERR_* values of reference include/petscerror.h:41-92, repeated below);
 * b200_last_error() gives the message.  Reference citations (file:line, relative to the erdc/petsc-dev tree) name
 * the interface each entry point replaces; INTEGRATION.md shows the Mat/Vec plugin that binds them.
 *
 * There is no CPU fallback: a compute entry point called without a usable CUDA device returns B200_ERR_LIB.
 */
#ifndef PETSCB200_H
#define PETSCB200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes (== reference include/petscerror.h) ------------------------------------------------------- */
#define B200_ERR_MEM             55
#define B200_ERR_SUP             56
#define B200_ERR_ORDER           58
#define B200_ERR_ARG_SIZ         60
#define B200_ERR_ARG_IDN         61
#define B200_ERR_ARG_WRONG       62
#define B200_ERR_ARG_OUTOFRANGE  63
#define B200_ERR_FP              72
#define B200_ERR_ARG_WRONGSTATE  73
#define B200_ERR_LIB             76
#define B200_ERR_PLIB            77
#define B200_ERR_ARG_NULL        85

/* NormType, reference include/petscvec.h:155 */
#define B200_NORM_1         0
#define B200_NORM_2         1
#define B200_NORM_FROBENIUS 2
#define B200_NORM_INFINITY  3
#define B200_NORM_1_AND_2   4

/* KSPConvergedReason, reference include/petscksp.h:404-424 */
#define B200_KSP_CONVERGED_RTOL            2
#define B200_KSP_CONVERGED_ATOL            3
#define B200_KSP_CONVERGED_ITERATING       0
#define B200_KSP_DIVERGED_NULL            -2
#define B200_KSP_DIVERGED_ITS             -3
#define B200_KSP_DIVERGED_DTOL            -4
#define B200_KSP_DIVERGED_BREAKDOWN       -5
#define B200_KSP_DIVERGED_INDEFINITE_PC   -8
#define B200_KSP_DIVERGED_NAN             -9
#define B200_KSP_DIVERGED_INDEFINITE_MAT -10

/* KSPNormType, reference include/petscksp.h (KSP_NORM_NONE..NATURAL) */
#define B200_KSP_NORM_NONE             0
#define B200_KSP_NORM_PRECONDITIONED   1
#define B200_KSP_NORM_UNPRECONDITIONED 2
#define B200_KSP_NORM_NATURAL          3

typedef struct b200_mat_s     *b200_mat;      /* MATSEQAIJ-on-device, or the MPIAIJ pair (diag, off-diag, ghost plan) */
typedef struct b200_scatter_s *b200_scatter;  /* device-resident VecScatter (MPI -> Seq ghost gather)               */

/* ---- runtime ------------------------------------------------------------------------------------------------
 * The reference's GPU types initialise their library in PetscInitialize (src/sys/objects/pinit.c:838-846) and never
 * pick a device; here the caller does (one process per GPU). */
int         b200_init(int device);                 /* cudaSetDevice + streams + scratch; idempotent                 */
int         b200_finalize(void);
const char *b200_last_error(void);
int         b200_version(void);
int         b200_device_count(int *n);
int         b200_set_stream(void *cuda_stream);    /* run on the caller's stream (NULL = the library's own)         */
int         b200_get_stream(void **cuda_stream);
int         b200_stream_create(void **cuda_stream);   /* caller-owned non-blocking stream for b200_set_stream (double buffering) */
int         b200_stream_destroy(void *cuda_stream);
int         b200_stream_synchronize(void *cuda_stream);
int         b200_synchronize(void);                /* WaitForGPU, src/vec/vec/impls/seq/seqcusp/cuspvecimpl.h:88    */
int         b200_launch_count(long long *n);       /* number of kernels this library has launched so far           */
int         b200_timer_start(void);                /* cudaEventRecord on the compute stream                         */
int         b200_timer_stop(double *milliseconds); /* record + synchronize + elapsed                                */
int         b200_flush_l2(void);                   /* overwrite a buffer larger than L2 (benchmark hygiene)        */

/* device / pinned memory (VecCUSPAllocateCheck, VecCUSPCopyToGPU, VecCUSPCopyFromGPU:
 * src/vec/vec/impls/seq/seqcusp/veccusp.cu:68,109,173) */
int b200_malloc(size_t bytes, void **dptr);
int b200_free(void *dptr);
int b200_host_alloc(size_t bytes, void **hptr);    /* pinned */
int b200_host_free(void *hptr);
int b200_memcpy_h2d(void *dst, const void *src_host, size_t bytes);   /* stream-ordered, returns when complete    */
int b200_memcpy_d2h(void *dst_host, const void *src, size_t bytes);
int b200_memcpy_h2d_async(void *dst, const void *src_host, size_t bytes);
int b200_memcpy_d2h_async(void *dst_host, const void *src, size_t bytes);
int b200_memset(void *dst, int byte, size_t bytes);

/* ---- communicator (one rank per GPU) -------------------------------------------------------------------------
 * Replaces the MPI communicator of VecDot_MPI / VecScatterBegin (src/vec/vec/impls/mpi/pbvec.c:16,
 * src/vec/vec/utils/vpscat.h:97).  The unique id is produced on rank 0 and handed to the other ranks by the host
 * program (torch.distributed, MPI_Bcast, a file ...). */
#define B200_COMM_ID_BYTES 128
int b200_comm_get_unique_id(void *id128);
int b200_comm_init(int rank, int size, const void *id128);
int b200_comm_rank(int *rank, int *size);
int b200_comm_destroy(void);
int b200_comm_allreduce_sum(double *dvals, int n);              /* in place, device buffer, stream-ordered        */
int b200_comm_barrier(void);
/* peer-memory mode: reductions and the ghost exchange go through NVLink peer stores instead of NCCL */
int b200_comm_enable_p2p(int enable);      /* 1: on (size > 1).  2: also on a ONE-rank communicator ("loop-back": the rank is its own
                                              neighbour; the same kernels, flags and credits, stores stay on the device — single-GPU tests) */
int b200_comm_p2p_enabled(int *enabled);

/* ---- Vec kernels (device arrays of n doubles) ----------------------------------------------------------------
 * Each replaces the VECSEQ op named (src/vec/vec/impls/seq/{bvec1,bvec2,dvec2}.c) and the VECSEQCUSP op that
 * shadows it (src/vec/vec/impls/seq/seqcusp/veccusp.cu); reductions return through HOST pointers and are valid on
 * return.  Functions without `_local` reduce over the communicator when one is initialised with size > 1
 * (VecDot_MPI pbvec.c:9, VecNorm_MPI pvec2.c:49, VecMDot_MPI pvec2.c:10). */
int b200_vec_set(int n, double *x, double alpha);                                   /* VecSet_Seq dvec2.c:722      */
int b200_vec_copy(int n, const double *x, double *y);                               /* VecCopy_Seq bvec2.c:464     */
int b200_vec_swap(int n, double *x, double *y);                                     /* VecSwap_Seq bvec2.c:519     */
int b200_vec_scale(int n, double *x, double alpha);                                 /* VecScale_Seq bvec1.c:183    */
int b200_vec_axpy(int n, double *y, double alpha, const double *x);                 /* VecAXPY_Seq bvec1.c:244     */
int b200_vec_aypx(int n, double *y, double alpha, const double *x);                 /* VecAYPX_Seq dvec2.c:971     */
int b200_vec_axpby(int n, double *y, double alpha, double beta, const double *x);   /* VecAXPBY_Seq bvec1.c:320    */
int b200_vec_waxpy(int n, double *w, double alpha, const double *x, const double *y); /* VecWAXPY_Seq dvec2.c:1082 */
int b200_vec_axpbypcz(int n, double *z, double alpha, double beta, double gamma, const double *x, const double *y); /* bvec1.c:418 */
int b200_vec_pointwise_mult(int n, double *w, const double *x, const double *y);    /* bvec2.c:234                 */
int b200_vec_pointwise_divide(int n, double *w, const double *x, const double *y);  /* bvec2.c:298                 */
int b200_vec_reciprocal(int n, double *x);                                          /* VecReciprocal_Default       */
int b200_vec_maxpy(int n, double *y, int nv, const double *alpha_host, const double *const *x_host_ptrs); /* VecMAXPY_Seq dvec2.c:836 */

int b200_vec_dot(int n, const double *x, const double *y, double *result_host);     /* VecDot_Seq bvec1.c:57 / _MPI */
int b200_vec_dot_local(int n, const double *x, const double *y, double *result_host);
int b200_vec_norm(int n, const double *x, int normtype, double *result_host);       /* VecNorm_Seq bvec2.c:605 / _MPI; 2 values for NORM_1_AND_2 */
int b200_vec_norm_local(int n, const double *x, int normtype, double *result_host);
int b200_vec_mdot(int n, const double *x, int nv, const double *const *y_host_ptrs, double *result_host); /* VecMDot_Seq dvec2.c:146 / _MPI */
int b200_vec_mdot_local(int n, const double *x, int nv, const double *const *y_host_ptrs, double *result_host);
int b200_vec_dotnorm2(int n, const double *s, const double *t, double *dp_host, double *nm_host);  /* VecDotNorm2, veccusp.cu:1816 */
/* b200_vec_mdot_noalloc: NO counterpart in the reference — added for callers
 * that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The b200_vec_mdot_local
 * kernel with its partial-sum scratch in a static device array (148.25 KB, held by every GPU that loads the library
 * until exit): it calls no cudaMalloc/cudaFree or allocating routine and does not need b200_init.  z_dev[0..nv) is a
 * DEVICE buffer of the caller; the work is queued on cuda_stream (a cudaStream_t; NULL = the library's stream after
 * b200_init, else the default stream) and nothing synchronises.  Limits: nv <= 16, n <= 2^30 — beyond them
 * B200_ERR_ARG_SIZ.  All calls on a device share the scratch: calls that overlap on the GPU (streams with no
 * ordering between them) race; calls on one stream do not.  Same grid and summation order as b200_vec_mdot_local. */
int b200_vec_mdot_noalloc(int n, const double *x, int nv, const double *const *y_host_ptrs, double *z_dev, void *cuda_stream);

/* ---- Mat (AIJ / CSR) ------------------------------------------------------------------------------------------
 * b200_mat_create_seqaij uploads an ASSEMBLED host CSR (Mat_SeqAIJ i/j/a, src/mat/impls/aij/seq/aij.h:10-39; column
 * ids ascending within a row) and bins its rows — what MatAssemblyEnd_SeqAIJCUSP -> MatCUSPCopyToGPU does
 * (src/mat/impls/aij/seq/seqcusp/aijcusp.cu:126,458).  The host arrays are not retained. */
int b200_mat_create_seqaij(int m, int n, const int *i_host, const int *j_host, const double *a_host, b200_mat *A);
int b200_mat_update_values(b200_mat A, const double *a_host);    /* same pattern, new values (value-only re-upload) */
/* The same for the parallel type: new values of the diagonal and of the off-diagonal block (Mat_MPIAIJ::A, ::B in their
 * assembled order), pattern, ghost columns and exchange plan unchanged — nothing collective happens. */
int b200_mat_update_values_mpiaij(b200_mat A, const double *a_diag_host, const double *a_offdiag_host);
/* Device assembly: MatSetValuesBatch (src/mat/interface/matrix.c:1698; the reference's device version is
 * MatSetValuesBatch_SeqAIJCUSP, src/mat/impls/aij/seq/seqcusp/aijAssemble.cu:157-290).  ne element matrices of nl x nl
 * values (row-major, concatenated in elem_mats) are added into the rows/columns elem_rows[e*nl .. e*nl+nl) of a new
 * m x n SeqAIJ matrix.  The result is what the reference's host path builds with one MatSetValues(..., ADD_VALUES)
 * per element on an empty matrix (matrix.c:1712-1716): same i/j (explicit zeros kept, aij.c:293-368), and every value
 * is the sum of its contributions in element order, left to right — bit-identical.  Negative row / column ids are
 * skipped (aij.c:311,325); ids beyond the matrix are an error (B200_ERR_ARG_OUTOFRANGE).
 * on_device != 0: elem_rows / elem_mats are device pointers (no host<->device traffic except the m+1 row offsets).
 * keep_plan != 0: the element -> nonzero plan stays with the matrix and b200_mat_setvaluesbatch_values refills the
 * values from a new elem_mats (same elem_rows) without sorting again — the "same pattern, new values" fast path of a
 * time-dependent or nonlinear loop (replaces the full re-upload of aijcusp.cu:138-152).
 * Limit: ne*nl*nl < 2^31 - 1 per matrix (per GPU). */
int b200_mat_setvaluesbatch(int m, int n, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                            int on_device, int keep_plan, b200_mat *A);
int b200_mat_setvaluesbatch_values(b200_mat A, const double *elem_mats, int on_device);
/* The local rows of a parallel matrix (MatSetValuesBatch on MPIAIJ; the reference's device version is
 * MatSetValuesBatch_MPIAIJCUSP, src/mat/impls/aij/mpi/mpicusp/mpiaijAssemble.cu:200-460): rows rstart .. rstart+m_local
 * of the global N x N matrix are assembled from the element list, ids of rows owned elsewhere are passed over as rows
 * and kept as (global) column ids — what MatSetValues_MPIAIJ does with owned rows (mpiaij.c:496-570).  The result is a
 * sequential m_local x N matrix with global column ids: the input of the diagonal / off-diagonal split
 * (b200_mpiaij_split).  The caller first hands every rank the elements that touch its rows (the reference ships such
 * entries through the stash at MatAssemblyBegin, mpiaij.c:630-650; petsc-dev_b200/parallel.py does it with one
 * exchange of whole elements). */
int b200_mat_setvaluesbatch_rows(int m_local, int rstart, int N, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                                 int on_device, int keep_plan, b200_mat *A);
/* The diagonal/off-diagonal split of such local rows ON THE DEVICE (MatSetValues_MPIAIJ routing mpiaij.c:507-565 +
 * MatSetUpMultiply_MPIAIJ compaction mmaij.c:28-64; the reference's device assembly sorts two COO lists for it,
 * mpiaijAssemble.cu:200-460): L = the matrix b200_mat_setvaluesbatch_rows returned (kept), [cstart, cend) this rank's
 * columns.  Only garray (nghost ints) and the two row-offset arrays reach the host.  b200_split_get reads the sizes and
 * garray (the caller builds the scatter lists from it with b200_scatter_plan_from/_to); b200_mat_create_mpiaij_split is
 * b200_mat_create_mpiaij with the device blocks of the split, which it consumes. */
typedef struct b200_split_s *b200_split;
int b200_mat_split_device(b200_mat L, int cstart, int cend, int *nghost, b200_split *S);
int b200_split_get(b200_split S, long long *nzA, long long *nzB, int *nghost, int *garray_host);
int b200_split_destroy(b200_split S);
int b200_mat_create_mpiaij_split(b200_split S, int N,
                                 int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                                 int nfrom, const int *from_procs, const int *from_starts, b200_mat *A);
/* The exchange plan of that step (host integer work, no GPU needed).  b200_batch_send_plan: counts[d] = how many of my
 * elements touch rows of rank d != rank; send_elems (NULL on a sizing call, else sum(counts) entries) = their element
 * numbers grouped by destination ascending.  Every rank publishes one buffer with its groups in that order and the
 * counts; b200_batch_recv_plan turns the gathered counts all_counts[s*size+d] into where this rank's group starts in
 * rank s's buffer (src_off[s], in elements) and how long it is (src_cnt[s]).  The receiver assembles its own elements
 * followed by the groups of s = 0, 1, ... */
int b200_batch_send_plan(int size, int rank, const int *range /* size+1 */, long long ne, int nl, const int *elem_rows,
                         long long *counts /* size */, long long *send_elems);
int b200_batch_recv_plan(int size, int rank, const long long *all_counts /* size*size */, long long *src_off /* size */, long long *src_cnt /* size */);
/* Copy the device CSR of a sequential matrix back to host arrays (i: m+1, j/a: nz; any may be NULL) —
 * MatCUSPCopyFromGPU, aijcusp.cu:187. */
int b200_mat_get_csr(b200_mat A, int *i_host, int *j_host, double *a_host);
int b200_mat_destroy(b200_mat A);
int b200_mat_get_size(b200_mat A, int *m, int *n, long long *nz);
int b200_mat_mult(b200_mat A, const double *x, double *y);                    /* MatMult_SeqAIJ aij.c:1225 / MatMult_MPIAIJ mpiaij.c:1102       */
int b200_mat_multadd(b200_mat A, const double *x, const double *y, double *z); /* MatMultAdd_SeqAIJ aij.c:1291 / _MPIAIJ mpiaij.c:1132 ; z may alias y */
/* Transposed products.  The explicit transpose is built once at the first call (rows = source rows ascending), so every
 * y[j] receives its terms in the order MatMultTranspose_SeqAIJ adds them (aij.c:1108-1118) and the product streams. */
int b200_mat_multtranspose(b200_mat A, const double *x, double *y);                     /* MatMultTranspose_SeqAIJ aij.c:1127 / _MPIAIJ mpiaij.c:1147 */
int b200_mat_multtransposeadd(b200_mat A, const double *x, const double *y, double *z); /* MatMultTransposeAdd_SeqAIJ aij.c:1078 / _MPIAIJ mpiaij.c:1178 ; z may alias y */
int b200_mat_get_diagonal(b200_mat A, double *d);                             /* MatGetDiagonal_SeqAIJ aij.c:1040 / _MPIAIJ mpiaij.c:1246       */
int b200_mat_flops(b200_mat A, double *mult_flops);                           /* 2*nz - nonzerorow, aij.c:1281                                 */
int b200_mat_bytes(b200_mat A, double *mult_bytes);                           /* algorithmic bytes of one MatMult                  */
int b200_mat_bin_info(b200_mat A, int *ntiles, int *nlong, int *compressed);  /* row-binning summary                                           */
int b200_mat_bin_info2(b200_mat A, int *lanes_per_row, int *long_thresh, int *grid, int *shape, int *park); /* tile-kernel launch description */
/* Row-binning policy for matrices created AFTER the call (-1 = automatic): launch shape index, lanes per row
 * (1, 2, 4, ... 32; 1 = every tile row is summed left to right exactly like aij.h:384-386), long-row threshold,
 * and whether products are parked in shared memory before the row sums (the choice for irregular row lengths). */
int b200_spmv_set_tuning(int shape, int lanes_per_row, int long_thresh, int park);
/* Parked (irregular-matrix) variant: how many entries of a row its own lane sums left to right before the row's warp sums
 * the rest together (default 32; <= 0: the lane sums the whole row, the round-1 behaviour).  For matrices created AFTER
 * the call, and only where the binning chose one lane per row by itself: lanes_per_row = 1 REQUESTED through
 * b200_spmv_set_tuning always means the reference's left-to-right order for every row. */
int b200_spmv_set_park_cap(int cap);
/* Column strips (cache blocking of the x gathers) for matrices created AFTER the call: -1 = automatic (strips of ~24 MB
 * of x when x exceeds 48 MB and the mean row spans more than a strip), 0/1 = off, k = force k strips.  With one lane
 * per row the row sums remain one left-to-right chain (bit-identical to MatMult_SeqAIJ). */
int b200_spmv_set_strips(int nstrips);
int b200_mat_strips(b200_mat A, int *diag_strips, int *offdiag_strips);
/* the four fused entry points the CG driver uses (kept public so a KSP plugin can call them) */
int b200_mat_mult_dot(b200_mat A, const double *x, double *y, double *xdoty_host);   /* y = A x and x.y in one pass (seq, m == n)      */

/* ---- MPIAIJ host-side setup (pure integer work, no GPU needed; bit-exact with the reference) ------------------ */
/* PetscSplitOwnership / PetscLayoutSetUp: src/sys/utils/psplit.c:88, src/vec/vec/impls/mpi/pmap.c:142-159 */
int b200_split_ownership(int N, int size, int bs, int *range /* size+1 */);
/* MatSetValues_MPIAIJ column routing + MatSetUpMultiply_MPIAIJ (src/mat/impls/aij/mpi/mpiaij.c:507-565,
 * src/mat/impls/aij/mpi/mmaij.c:28-64): split local rows (global column ids, ascending) into the diagonal block
 * (local ids = col - cstart) and the off-diagonal block (ids into the sorted distinct `garray`).
 * Call once with NULL outputs to get sizes (nzA, nzB, nghost), then with arrays. */
int b200_mpiaij_split(int m, int cstart, int cend, const int *i, const int *j, const double *a,
                      int *Ai, int *Aj, double *Aa, int *Bi, int *Bj, double *Ba, int *garray,
                      long long *nzA, long long *nzB, int *nghost);
/* MatCreateColmap_MPIAIJ_Private (mpiaij.c:357-377, dense form): colmap[garray[i]] = i+1, zero elsewhere */
int b200_mpiaij_colmap(int N, int nghost, const int *garray, int *colmap);
/* VecScatterCreate_PtoS (src/vec/vec/utils/vpscat.c:1730-1924), receive side and request lists.  inidx[nx] are
 * (block) indices into the parallel vector (garray for MatMult, mmaij.c:132-143), inidy[nx] the (block) positions in
 * the sequential vector, bs the block size, range[size+1] the ownership ranges, lengthy the sequential length.
 * Call with from_procs == NULL to query nfrom / nlocal; arrays must hold size (+1) resp. nx entries.
 *   from_procs/from_starts/from_indices : the `from` context (vpscat.c:1850-1879)
 *   req_counts[size] (own rank zeroed), req_values : wanted global ids grouped by owner, ascending (vpscat.c:1798-1806)
 *   to_local/from_local[nlocal] : the self part (vpscat.c:1890-1903)                                              */
int b200_scatter_plan_from(int nx, const int *inidx, const int *inidy, int bs, int size, const int *range, int rank,
                           int lengthy, int *nfrom, int *from_procs, int *from_starts, int *from_indices,
                           int *req_counts, int *req_values, int *nlocal, int *to_local, int *from_local);
/* send side (vpscat.c:1827-1848): nto requesting ranks in ascending order (PetscSortMPIIntWithArray, :1780) with
 * lengths[nto] and their concatenated requested GLOBAL ids; outputs to_starts[nto+1], to_indices = id - rstart. */
int b200_scatter_plan_to(int nto, const int *lengths, const int *requested_global, int rstart, int rend,
                         int *to_starts, int *to_indices);
/* mmaij.c:106-128: may the ghost list be described by a block index set of size bs? */
int b200_mpiaij_use_block_is(int nghost, const int *garray, int bs, int *use);
/* assemble the device-resident MPIAIJ operator from the pieces above.  The ghost exchange goes rank-to-rank over
 * NCCL (or peer stores when p2p is enabled) on a side stream, overlapped with the diagonal-block product
 * (MatMult_MPIAIJ mpiaij.c:1102-1116; MatMult_MPIAIJCUSP mpiaijcusp.cu:85-112). */
int b200_mat_create_mpiaij(int m, int n_local, int N,
                           const int *Ai, const int *Aj, const double *Aa,
                           int nghost, const int *Bi, const int *Bj, const double *Ba, const int *garray,
                           int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                           int nfrom, const int *from_procs, const int *from_starts,
                           b200_mat *A);
int b200_mat_mult_phases(b200_mat A, const double *x, double *y, double *ms4_host);  /* diagnostic: device ms of exchange-begin | diagonal | exchange-end | off-diagonal */
/* 1 when MatMult/MatMultAdd of this operator run as ONE kernel (ghost exchange, diagonal and off-diagonal block fused:
 * peer-memory transport, thin off-diagonal block, no long rows); B200_MPI_FUSED=0 in the environment keeps the four-step path */
int b200_mat_mpiaij_is_fused(b200_mat A, int *fused, int *boundary_tiles);
int b200_mat_mpiaij_get_lvec(b200_mat A, double **lvec, int *nghost);         /* Mat_MPIAIJ::lvec, mpiaij.h:60   */
int b200_mat_mpiaij_scatter(b200_mat A, const double *x);                     /* VecScatterBegin+End only        */
int b200_mat_mpiaij_set_lvec(b200_mat A, const double *ghost_values_dev);     /* caller-performed scatter (P virtual ranks in one process) */

/* ---- device-resident VecScatter (parallel vector <-> sequential vector) ---------------------------------------------
 * The general MPI -> Seq scatter of VecScatterCreate_PtoS (vpscat.c:1730-1924) with VecScatterBegin_1 / End_1
 * (vpscat.h:14-233): built from the lists b200_scatter_plan_from/_to produce (block lists already expanded to scalar
 * entries).  x is this rank's part of the parallel vector, y the sequential vector; add 0 = INSERT_VALUES, 1 = ADD_VALUES;
 * reverse 0 = SCATTER_FORWARD (y[from_indices] <- x of the owners), 1 = SCATTER_REVERSE (x of the owners <- y).  Messages
 * travel GPU to GPU (NVLink peer stores or NCCL); received messages are applied neighbour by neighbour in ascending rank
 * order; the self part is applied in Begin (vpscat.h:100-128).  Collective. */
int b200_scatter_create(int nx_local, int ny,
                        int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                        int nfrom, const int *from_procs, const int *from_starts, const int *from_indices,
                        int nlocal, const int *to_local, const int *from_local, b200_scatter *ctx);
/* sequential -> sequential scatters (vscat.c:316-700: general/stride index sets on both sides, expanded to lists):
 * forward y[dst_idx[i]] (+)= x[src_idx[i]], reverse x[src_idx[i]] (+)= y[dst_idx[i]]; repeated targets are applied in
 * list order like the reference's loop.  Not collective. */
int b200_scatter_create_local(int nx, int ny, int n, const int *src_idx, const int *dst_idx, b200_scatter *ctx);
int b200_scatter_begin(b200_scatter ctx, double *x, double *y, int add, int reverse);
int b200_scatter_end(b200_scatter ctx, double *x, double *y, int add, int reverse);
int b200_scatter_destroy(b200_scatter ctx);

/* ---- KSP drivers ----------------------------------------------------------------------------------------------
 * Restatements of KSPSolve_CG (src/ksp/ksp/impls/cg/cg.c:92-286) and KSPSolve_GMRES / KSPGMRESCycle
 * (src/ksp/ksp/impls/gmres/gmres.c:118-238, borthog2.c:35-118) with PCJACOBI / PCNONE
 * (src/ksp/pc/impls/jacobi/jacobi.c:125-206,266-278) and KSPDefaultConverged (src/ksp/ksp/interface/iterativ.c:702).
 * The scalar recurrences stay on the host exactly as in the reference; the vector work is the kernels above plus
 * fused forms of them. */
typedef struct {
  int    ksp_type;           /* 0 = cg, 1 = gmres                                                                */
  int    pc_type;            /* 0 = none, 1 = jacobi                                                             */
  double rtol, abstol, dtol; /* KSPSetTolerances; defaults 1e-5, 1e-50, 1e5                                      */
  int    max_it;             /* default 10000                                                                    */
  int    norm_type;          /* B200_KSP_NORM_*; CG default PRECONDITIONED (cg.c:439), GMRES PRECONDITIONED      */
  int    guess_nonzero;      /* KSPSetInitialGuessNonzero                                                        */
  int    gmres_restart;      /* default 30 (gmres.c:947)                                                         */
  int    gmres_cgs_refine;   /* 0 never (default), 1 ifneeded, 2 always (borthog2.c:43,85-108)                   */
  double gmres_haptol;       /* default 1e-30                                                                    */
  int    fused;              /* 0 = op-by-op exactly like the reference's call sequence; 1 = fused kernels       */
  int    check_every;        /* fused CG only: host looks at the device-side convergence flag every k iterations */
  int    gmres_modified_gs;  /* 1 = modified Gram-Schmidt (borthog.c:30-52: one VecDot + VecAXPY per basis vector)
                                instead of the default classical one (borthog2.c:35-118); -ksp_gmres_modifiedgramschmidt */
  const double *pc_jacobi_inverse_diag;   /* device vector or NULL.  NULL: PCJACOBI is set up from the matrix diagonal
                                (jacobi.c:160-190, default variant).  Otherwise the inverse diagonal the caller's PC already
                                holds — the plugin passes PCApply(pc, ones), so -pc_jacobi_rowmax/_rowsum/_abs mean what
                                they mean in the reference                                                            */
} b200_ksp_options;
int b200_ksp_default_options(b200_ksp_options *o);
typedef struct {
  int    its;                /* KSPGetIterationNumber                                                            */
  int    reason;             /* KSPGetConvergedReason                                                            */
  int    nhist;              /* entries written to the residual history                                          */
  double rnorm;              /* last residual norm                                                               */
  double solve_ms;           /* device time of the solve (CUDA events), setup excluded                           */
} b200_ksp_result;
/* x, b device vectors of the matrix's local row count; hist_host may be NULL.  PCSetUp (diagonal extraction) is
 * done before the timed region like KSPSetUp (itfunc.c:376). */
int b200_ksp_solve(b200_mat A, const b200_ksp_options *o, const double *b, double *x,
                   double *hist_host, int histmax, b200_ksp_result *result);

#ifdef __cplusplus
}
#endif
#endif /* PETSCB200_H */
