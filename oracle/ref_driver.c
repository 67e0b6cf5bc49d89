/* oracle/ref_driver.c — TEST INFRASTRUCTURE ONLY.
 *
 * A flat, ctypes-callable driver over the UNMODIFIED reference library (oracle/_ref/libpetsc_ref.so, built by
 * oracle/build_ref.py from /root/reference).  Every function here goes through the reference's own public API
 * (MatCreate/MatSetType/Mat*SetPreallocationCSR/MatMult, VecCreateSeqWithArray/VecDot/..., KSPSolve), so what the
 * tests compare against is the reference's real MatMult_SeqAIJ (src/mat/impls/aij/seq/aij.c:1225),
 * MatMult_MPIAIJ (src/mat/impls/aij/mpi/mpiaij.c:1102), Vec kernels (src/vec/vec/impls/seq/{bvec1,bvec2,dvec2}.c),
 * KSPSolve_CG (src/ksp/ksp/impls/cg/cg.c:92), KSPSolve_GMRES (src/ksp/ksp/impls/gmres/gmres.c:213) and
 * PCApply_Jacobi (src/ksp/pc/impls/jacobi/jacobi.c:266) — not a restatement.
 * Never linked into the product (petsc-dev_b200/).
 */
#include <petscksp.h>
#include <petsctime.h>
#include <../src/mat/impls/aij/seq/aij.h>
#include <../src/mat/impls/aij/mpi/mpiaij.h>

#define REFCHK(ierr) do { if (ierr) { return (int)(ierr); } } while (0)

static int ref_inited = 0;

int ref_initialize(void)
{
  PetscErrorCode ierr;
  if (ref_inited) return 0;
  ierr = PetscInitializeNoArguments(); REFCHK(ierr);
  /* errors come back as return codes to the Python caller instead of aborting the interpreter */
  ierr = PetscPushErrorHandler(PetscReturnErrorHandler, PETSC_NULL); REFCHK(ierr);
  ref_inited = 1;
  return 0;
}

int ref_set_option(const char *name, const char *value)
{
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  if (value && value[0] == '\001') { ierr = PetscOptionsClearValue(name); REFCHK(ierr); }
  else { ierr = PetscOptionsSetValue(name, (value && value[0]) ? value : PETSC_NULL); REFCHK(ierr); }
  return 0;
}

/* kind 0: MATSEQAIJ; kind 1: MATMPIAIJ on the (size-1, MPIUNI) world communicator.  The matrix is assembled by the
 * reference's own MatSetValues path (Mat*SetPreallocationCSR inserts row by row), bs = block size (>=1). */
int ref_mat_create(int kind, int m, int n, int bs, const int *i, const int *j, const double *a, int no_inode, void **out)
{
  Mat            A;
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  if (no_inode) { ierr = PetscOptionsSetValue("-mat_no_inode", PETSC_NULL); REFCHK(ierr); }
  else { ierr = PetscOptionsClearValue("-mat_no_inode"); REFCHK(ierr); }
  ierr = MatCreate(kind ? PETSC_COMM_WORLD : PETSC_COMM_SELF, &A); REFCHK(ierr);
  ierr = MatSetSizes(A, m, n, m, n); REFCHK(ierr);
  ierr = MatSetType(A, kind ? MATMPIAIJ : MATSEQAIJ); REFCHK(ierr);
  if (bs > 1) { ierr = MatSetBlockSize(A, bs); REFCHK(ierr); }
  ierr = MatSetFromOptions(A); REFCHK(ierr);
  if (kind) { ierr = MatMPIAIJSetPreallocationCSR(A, i, j, a); REFCHK(ierr); }
  else { ierr = MatSeqAIJSetPreallocationCSR(A, i, j, a); REFCHK(ierr); }
  *out = (void *)A;
  return 0;
}

int ref_mat_destroy(void *h)
{
  Mat A = (Mat)h;
  return (int)MatDestroy(&A);
}

/* 1 if the assembled SeqAIJ matrix (or the diagonal block of the MPIAIJ one) runs the Inode MatMult
 * (src/mat/impls/aij/seq/inode.c:392), with its node count */
int ref_mat_inode_info(void *h, int kind, int *use, int *node_count)
{
  Mat         A = (Mat)h;
  Mat_SeqAIJ *a = kind ? (Mat_SeqAIJ *)((Mat_MPIAIJ *)A->data)->A->data : (Mat_SeqAIJ *)A->data;
  *use        = (int)a->inode.use;
  *node_count = (int)a->inode.node_count;
  return 0;
}

/* copy out the assembled CSR of the SeqAIJ matrix: pass NULL arrays to query sizes */
int ref_mat_get_csr(void *h, int *m, int *nz, int *i, int *j, double *a)
{
  Mat            A  = (Mat)h;
  Mat_SeqAIJ    *aa = (Mat_SeqAIJ *)A->data;
  PetscErrorCode ierr;
  *m  = A->rmap->n;
  *nz = aa->nz;
  if (i) { ierr = PetscMemcpy(i, aa->i, (A->rmap->n + 1) * sizeof(PetscInt)); REFCHK(ierr); }
  if (j) { ierr = PetscMemcpy(j, aa->j, aa->nz * sizeof(PetscInt)); REFCHK(ierr); }
  if (a) { ierr = PetscMemcpy(a, aa->a, aa->nz * sizeof(PetscScalar)); REFCHK(ierr); }
  return 0;
}

static PetscErrorCode wrapvec(Mat A, int kind, PetscInt n, const double *arr, Vec *v)
{
  if (kind) return VecCreateMPIWithArray(PETSC_COMM_WORLD, 1, n, n, arr, v);
  return VecCreateSeqWithArray(PETSC_COMM_SELF, 1, n, arr, v);
}

/* y = A x  (reps >= 1 repetitions; returns the minimum wall time of one MatMult in *tmin when non-NULL) */
int ref_mat_mult(void *h, int kind, const double *x, double *y, int reps, double *tmin)
{
  Mat            A = (Mat)h;
  Vec            xv, yv;
  PetscErrorCode ierr;
  PetscInt       m, n, r;
  double         best = 1e300;
  ierr = MatGetLocalSize(A, &m, &n); REFCHK(ierr);
  ierr = wrapvec(A, kind, n, x, &xv); REFCHK(ierr);
  ierr = wrapvec(A, kind, m, y, &yv); REFCHK(ierr);
  for (r = 0; r < (reps > 0 ? reps : 1); r++) {
    PetscLogDouble t0, t1;
    ierr = PetscGetTime(&t0); REFCHK(ierr);
    ierr = MatMult(A, xv, yv); REFCHK(ierr);
    ierr = PetscGetTime(&t1); REFCHK(ierr);
    if (t1 - t0 < best) best = t1 - t0;
  }
  if (tmin) *tmin = best;
  ierr = VecDestroy(&xv); REFCHK(ierr);
  ierr = VecDestroy(&yv); REFCHK(ierr);
  return 0;
}

/* z = y + A x ; z may alias y (pass the same pointer) */
int ref_mat_multadd(void *h, int kind, const double *x, double *y, double *z)
{
  Mat            A = (Mat)h;
  Vec            xv, yv, zv;
  PetscErrorCode ierr;
  PetscInt       m, n;
  ierr = MatGetLocalSize(A, &m, &n); REFCHK(ierr);
  ierr = wrapvec(A, kind, n, x, &xv); REFCHK(ierr);
  ierr = wrapvec(A, kind, m, y, &yv); REFCHK(ierr);
  if (z == y) zv = yv;
  else { ierr = wrapvec(A, kind, m, z, &zv); REFCHK(ierr); }
  ierr = MatMultAdd(A, xv, yv, zv); REFCHK(ierr);
  if (zv != yv) { ierr = VecDestroy(&zv); REFCHK(ierr); }
  ierr = VecDestroy(&xv); REFCHK(ierr);
  ierr = VecDestroy(&yv); REFCHK(ierr);
  return 0;
}

/* y = A^T x (MatMultTranspose) ; with yin != NULL z = yin + A^T x (MatMultTransposeAdd), z may alias yin */
int ref_mat_multtranspose(void *h, int kind, const double *x, double *yin, double *z)
{
  Mat            A = (Mat)h;
  Vec            xv, yv = 0, zv;
  PetscErrorCode ierr;
  PetscInt       m, n;
  ierr = MatGetLocalSize(A, &m, &n); REFCHK(ierr);
  ierr = wrapvec(A, kind, m, x, &xv); REFCHK(ierr);
  ierr = wrapvec(A, kind, n, z, &zv); REFCHK(ierr);
  if (!yin) { ierr = MatMultTranspose(A, xv, zv); REFCHK(ierr); }
  else {
    if (yin == z) yv = zv;
    else { ierr = wrapvec(A, kind, n, yin, &yv); REFCHK(ierr); }
    ierr = MatMultTransposeAdd(A, xv, yv, zv); REFCHK(ierr);
    if (yv != zv) { ierr = VecDestroy(&yv); REFCHK(ierr); }
  }
  ierr = VecDestroy(&xv); REFCHK(ierr);
  ierr = VecDestroy(&zv); REFCHK(ierr);
  return 0;
}

int ref_mat_get_diagonal(void *h, int kind, double *d)
{
  Mat            A = (Mat)h;
  Vec            dv;
  PetscErrorCode ierr;
  PetscInt       m, n;
  ierr = MatGetLocalSize(A, &m, &n); REFCHK(ierr);
  ierr = wrapvec(A, kind, m, d, &dv); REFCHK(ierr);
  ierr = MatGetDiagonal(A, dv); REFCHK(ierr);
  ierr = VecDestroy(&dv); REFCHK(ierr);
  return 0;
}

/* ---------------------------------------------------------------- Vec kernels (reference VecXxx front-ends on
 * VECSEQ vectors that wrap the caller's arrays; src/vec/vec/interface/rvector.c -> impls/seq) */
#define WRAP(v, n, arr) ierr = VecCreateSeqWithArray(PETSC_COMM_SELF, 1, (n), (arr), &(v)); REFCHK(ierr)
#define DROP(v) ierr = VecDestroy(&(v)); REFCHK(ierr)

int ref_vec_dot(int n, const double *x, const double *y, double *r)
{
  Vec xv, yv; PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  WRAP(xv, n, x); WRAP(yv, n, y);
  ierr = VecDot(xv, yv, r); REFCHK(ierr);
  DROP(xv); DROP(yv);
  return 0;
}

/* type: 0 NORM_1, 1 NORM_2, 3 NORM_INFINITY, 4 NORM_1_AND_2 (writes r[0], r[1]) — the reference's NormType values */
int ref_vec_norm(int n, const double *x, int type, double *r)
{
  Vec xv; PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  WRAP(xv, n, x);
  ierr = VecNorm(xv, (NormType)type, r); REFCHK(ierr);
  DROP(xv);
  return 0;
}

int ref_vec_mdot(int n, const double *x, int nv, const double *const *y, double *z)
{
  Vec xv, *yv; PetscErrorCode ierr; int k;
  ierr = ref_initialize(); REFCHK(ierr);
  WRAP(xv, n, x);
  ierr = PetscMalloc((nv + 1) * sizeof(Vec), &yv); REFCHK(ierr);
  for (k = 0; k < nv; k++) { WRAP(yv[k], n, y[k]); }
  ierr = VecMDot(xv, nv, yv, z); REFCHK(ierr);
  for (k = 0; k < nv; k++) { DROP(yv[k]); }
  ierr = PetscFree(yv); REFCHK(ierr);
  DROP(xv);
  return 0;
}

int ref_vec_maxpy(int n, double *y, int nv, const double *alpha, const double *const *x)
{
  Vec yv, *xv; PetscErrorCode ierr; int k;
  ierr = ref_initialize(); REFCHK(ierr);
  WRAP(yv, n, y);
  ierr = PetscMalloc((nv + 1) * sizeof(Vec), &xv); REFCHK(ierr);
  for (k = 0; k < nv; k++) { WRAP(xv[k], n, x[k]); }
  ierr = VecMAXPY(yv, nv, alpha, xv); REFCHK(ierr);
  for (k = 0; k < nv; k++) { DROP(xv[k]); }
  ierr = PetscFree(xv); REFCHK(ierr);
  DROP(yv);
  return 0;
}

/* elementwise family: op selects the reference front-end; unused operands may be NULL.
 *  0 AXPY   y += a x            1 AYPX  y = x + a y         2 WAXPY w = a x + y     3 AXPBY y = a x + b y
 *  4 AXPBYPCZ z = a x + b y + c z   5 SCALE x *= a          6 SET   x = a           7 COPY  y = x
 *  8 SWAP x<->y                 9 POINTWISEMULT w = x.*y    10 POINTWISEDIVIDE w = x./y   11 RECIPROCAL x = 1/x
 * operand convention: x (in), y (in/out), w (out; z for AXPBYPCZ) */
int ref_vec_elementwise(int op, int n, double a, double b, double c, double *x, double *y, double *w)
{
  Vec xv = 0, yv = 0, wv = 0; PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  if (x) { WRAP(xv, n, x); }
  if (y) { WRAP(yv, n, y); }
  if (w) { WRAP(wv, n, w); }
  switch (op) {
  case 0: ierr = VecAXPY(yv, a, xv); break;
  case 1: ierr = VecAYPX(yv, a, xv); break;
  case 2: ierr = VecWAXPY(wv, a, xv, yv); break;
  case 3: ierr = VecAXPBY(yv, a, b, xv); break;
  case 4: ierr = VecAXPBYPCZ(wv, a, b, c, xv, yv); break;
  case 5: ierr = VecScale(xv, a); break;
  case 6: ierr = VecSet(xv, a); break;
  case 7: ierr = VecCopy(xv, yv); break;
  case 8: ierr = VecSwap(xv, yv); break;
  case 9: ierr = VecPointwiseMult(wv, xv, yv); break;
  case 10: ierr = VecPointwiseDivide(wv, xv, yv); break;
  case 11: ierr = VecReciprocal(xv); break;
  default: ierr = PETSC_ERR_ARG_OUTOFRANGE;
  }
  REFCHK(ierr);
  if (xv) { DROP(xv); }
  if (yv) { DROP(yv); }
  if (wv) { DROP(wv); }
  return 0;
}

/* ---------------------------------------------------------------- KSPSolve through the reference's own driver
 * (src/ksp/ksp/interface/itfunc.c:335).  ksptype "cg"/"gmres", pctype "jacobi"/"none"; further behaviour
 * (-ksp_norm_type, -ksp_gmres_restart, -ksp_cg_single_reduction, ...) through ref_set_option() beforehand.
 * nonzero_guess: x holds the initial guess.  hist receives the residual history (KSPSetResidualHistory). */
int ref_ksp_solve(void *h, int kind, const char *ksptype, const char *pctype, double rtol, double abstol, double dtol,
                  int maxit, int nonzero_guess, const double *b, double *x, double *hist, int histmax, int *nhist,
                  int *its, int *reason, double *solve_seconds)
{
  Mat            A = (Mat)h;
  KSP            ksp;
  PC             pc;
  Vec            bv, xv;
  PetscErrorCode ierr;
  PetscInt       m, n, na, it;
  KSPConvergedReason rs;
  PetscLogDouble t0, t1;

  ierr = MatGetLocalSize(A, &m, &n); REFCHK(ierr);
  ierr = wrapvec(A, kind, m, b, &bv); REFCHK(ierr);
  ierr = wrapvec(A, kind, n, x, &xv); REFCHK(ierr);
  ierr = KSPCreate(kind ? PETSC_COMM_WORLD : PETSC_COMM_SELF, &ksp); REFCHK(ierr);
  ierr = KSPSetOperators(ksp, A, A, SAME_NONZERO_PATTERN); REFCHK(ierr);
  ierr = KSPSetType(ksp, ksptype); REFCHK(ierr);
  ierr = KSPGetPC(ksp, &pc); REFCHK(ierr);
  ierr = PCSetType(pc, pctype); REFCHK(ierr);
  ierr = KSPSetTolerances(ksp, rtol, abstol, dtol, maxit); REFCHK(ierr);
  ierr = KSPSetInitialGuessNonzero(ksp, nonzero_guess ? PETSC_TRUE : PETSC_FALSE); REFCHK(ierr);
  ierr = KSPSetResidualHistory(ksp, hist, histmax, PETSC_TRUE); REFCHK(ierr);
  ierr = KSPSetFromOptions(ksp); REFCHK(ierr);
  ierr = KSPSetUp(ksp); REFCHK(ierr);
  ierr = PetscGetTime(&t0); REFCHK(ierr);
  ierr = KSPSolve(ksp, bv, xv); REFCHK(ierr);
  ierr = PetscGetTime(&t1); REFCHK(ierr);
  ierr = KSPGetResidualHistory(ksp, PETSC_NULL, &na); REFCHK(ierr);
  ierr = KSPGetIterationNumber(ksp, &it); REFCHK(ierr);
  ierr = KSPGetConvergedReason(ksp, &rs); REFCHK(ierr);
  *nhist = (int)na; *its = (int)it; *reason = (int)rs;
  if (solve_seconds) *solve_seconds = t1 - t0;
  ierr = KSPDestroy(&ksp); REFCHK(ierr);
  ierr = VecDestroy(&bv); REFCHK(ierr);
  ierr = VecDestroy(&xv); REFCHK(ierr);
  return 0;
}

/* ---- PETSc binary files through the reference's own viewer (MatView_SeqAIJ_Binary aij.c:414, MatLoad_SeqAIJ aij.c:4093,
 * VecView / VecLoad): used to pin petsc-dev_b200/binaryio.py against the real format ---- */
int ref_mat_write_binary(void *h, const char *path)
{
  Mat            A = (Mat)h;
  PetscViewer    v;
  PetscErrorCode ierr;
  ierr = PetscViewerBinaryOpen(PETSC_COMM_SELF, path, FILE_MODE_WRITE, &v); REFCHK(ierr);
  ierr = MatView(A, v); REFCHK(ierr);
  ierr = PetscViewerDestroy(&v); REFCHK(ierr);
  return 0;
}

int ref_mat_load_binary(const char *path, void **out)
{
  Mat            A;
  PetscViewer    v;
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  ierr = PetscViewerBinaryOpen(PETSC_COMM_SELF, path, FILE_MODE_READ, &v); REFCHK(ierr);
  ierr = MatCreate(PETSC_COMM_SELF, &A); REFCHK(ierr);
  ierr = MatSetType(A, MATSEQAIJ); REFCHK(ierr);
  ierr = MatLoad(A, v); REFCHK(ierr);
  ierr = PetscViewerDestroy(&v); REFCHK(ierr);
  *out = (void *)A;
  return 0;
}

int ref_vec_write_binary(int n, const double *x, const char *path)
{
  Vec            xv;
  PetscViewer    v;
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  ierr = VecCreateSeqWithArray(PETSC_COMM_SELF, 1, n, x, &xv); REFCHK(ierr);
  ierr = PetscViewerBinaryOpen(PETSC_COMM_SELF, path, FILE_MODE_WRITE, &v); REFCHK(ierr);
  ierr = VecView(xv, v); REFCHK(ierr);
  ierr = PetscViewerDestroy(&v); REFCHK(ierr);
  ierr = VecDestroy(&xv); REFCHK(ierr);
  return 0;
}

/* n_inout: capacity on entry, the vector's length on return */
int ref_vec_load_binary(const char *path, int *n_inout, double *x)
{
  Vec            xv;
  PetscViewer    v;
  PetscScalar    *arr;
  PetscInt       n;
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  ierr = PetscViewerBinaryOpen(PETSC_COMM_SELF, path, FILE_MODE_READ, &v); REFCHK(ierr);
  ierr = VecCreate(PETSC_COMM_SELF, &xv); REFCHK(ierr);
  ierr = VecSetType(xv, VECSEQ); REFCHK(ierr);
  ierr = VecLoad(xv, v); REFCHK(ierr);
  ierr = PetscViewerDestroy(&v); REFCHK(ierr);
  ierr = VecGetLocalSize(xv, &n); REFCHK(ierr);
  if (n > *n_inout) { VecDestroy(&xv); *n_inout = (int)n; return 60; }
  ierr = VecGetArray(xv, &arr); REFCHK(ierr);
  memcpy(x, arr, sizeof(double) * (size_t)n);
  ierr = VecRestoreArray(xv, &arr); REFCHK(ierr);
  *n_inout = (int)n;
  ierr = VecDestroy(&xv); REFCHK(ierr);
  return 0;
}

/* MatSetValuesBatch (src/mat/interface/matrix.c:1698) on a fresh SeqAIJ matrix through the reference's host path: the
 * type has no setvaluesbatch op, so matrix.c:1712-1716 issues one MatSetValues(..., ADD_VALUES) per element.
 * nnz_hint[m] preallocates the rows (an upper bound keeps the run free of reallocations; the assembled result does not
 * depend on it).  The assembled matrix is returned as a handle for ref_mat_get_csr / ref_mat_mult. */
int ref_mat_setvaluesbatch(int m, int n, int ne, int nl, int *rows, const double *vals, const int *nnz_hint, void **out)
{
  Mat            A;
  PetscErrorCode ierr;
  ierr = ref_initialize(); REFCHK(ierr);
  ierr = MatCreate(PETSC_COMM_SELF, &A); REFCHK(ierr);
  ierr = MatSetSizes(A, m, n, m, n); REFCHK(ierr);
  ierr = MatSetType(A, MATSEQAIJ); REFCHK(ierr);
  ierr = MatSeqAIJSetPreallocation(A, 0, nnz_hint); REFCHK(ierr);
  ierr = MatSetOption(A, MAT_NEW_NONZERO_ALLOCATION_ERR, PETSC_FALSE); REFCHK(ierr);
  ierr = MatSetValuesBatch(A, ne, nl, rows, vals); REFCHK(ierr);
  ierr = MatAssemblyBegin(A, MAT_FINAL_ASSEMBLY); REFCHK(ierr);
  ierr = MatAssemblyEnd(A, MAT_FINAL_ASSEMBLY); REFCHK(ierr);
  *out = (void *)A;
  return 0;
}
