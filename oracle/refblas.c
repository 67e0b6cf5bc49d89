/* oracle/refblas.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * The reference calls BLAS level-1 through Fortran-ABI symbols (BLASdot_ -> ddot_, BLASaxpy_ -> daxpy_, ...;
 * name map: reference include/petscblaslapack_uscore.h:52-73; call sites src/vec/vec/impls/seq/bvec1.c:67,255,
 * bvec2.c:617,633).  The BLAS itself is a third-party dependency that is NOT vendored in /root/reference and is
 * not pinned by the reference (configure would accept any vendor BLAS or --download-f2cblaslapack), and neither
 * a Fortran compiler nor a system BLAS exists on this box.  This file restates the published Netlib
 * reference-BLAS semantics as plain C loops:
 *   ddot   : dtemp = dtemp + dx(i)*dy(i), i ascending (Netlib unrolls by 5 but evaluates left to right, i.e. the
 *            same association as the sequential loop below)
 *   daxpy  : dy(i) = dy(i) + da*dx(i)
 *   dscal / dcopy / dswap / dasum : the obvious loops, ascending
 *   dnrm2  : scaled sum of squares (Netlib's scale/ssq recurrence)
 *   dgemv / dgemm : column-oriented reference loops (only Dense Mat / GMRES eigen-estimates reach them)
 * Because the reference does not pin its BLAS, the summation ORDER of dot products is "parity unpinned" by the
 * reference itself at the 1e-16 level; we pin it here as left-to-right.  Compiled with -ffp-contract=off.
 *
 * LAPACK entry points the reference's off-path code links against (dense factorisation, eigen-estimates) are
 * stubs that report failure through `info`; nothing on the AIJ MatMult / CG / GMRES / Jacobi path calls them.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

typedef int blasint;

double ddot_(const blasint *n, const double *dx, const blasint *incx, const double *dy, const blasint *incy)
{
  double  dtemp = 0.0;
  blasint i, ix, iy;
  if (*n <= 0) return 0.0;
  if (*incx == 1 && *incy == 1) {
    for (i = 0; i < *n; i++) dtemp = dtemp + dx[i] * dy[i];
    return dtemp;
  }
  ix = (*incx < 0) ? (1 - *n) * *incx : 0;
  iy = (*incy < 0) ? (1 - *n) * *incy : 0;
  for (i = 0; i < *n; i++, ix += *incx, iy += *incy) dtemp = dtemp + dx[ix] * dy[iy];
  return dtemp;
}

void daxpy_(const blasint *n, const double *da, const double *dx, const blasint *incx, double *dy, const blasint *incy)
{
  blasint i, ix, iy;
  if (*n <= 0 || *da == 0.0) return;
  if (*incx == 1 && *incy == 1) {
    for (i = 0; i < *n; i++) dy[i] = dy[i] + *da * dx[i];
    return;
  }
  ix = (*incx < 0) ? (1 - *n) * *incx : 0;
  iy = (*incy < 0) ? (1 - *n) * *incy : 0;
  for (i = 0; i < *n; i++, ix += *incx, iy += *incy) dy[iy] = dy[iy] + *da * dx[ix];
}

void dscal_(const blasint *n, const double *da, double *dx, const blasint *incx)
{
  blasint i;
  if (*n <= 0 || *incx <= 0) return;
  for (i = 0; i < *n; i++) dx[i * *incx] = *da * dx[i * *incx];
}

void dcopy_(const blasint *n, const double *dx, const blasint *incx, double *dy, const blasint *incy)
{
  blasint i, ix, iy;
  if (*n <= 0) return;
  ix = (*incx < 0) ? (1 - *n) * *incx : 0;
  iy = (*incy < 0) ? (1 - *n) * *incy : 0;
  for (i = 0; i < *n; i++, ix += *incx, iy += *incy) dy[iy] = dx[ix];
}

void dswap_(const blasint *n, double *dx, const blasint *incx, double *dy, const blasint *incy)
{
  blasint i, ix, iy;
  if (*n <= 0) return;
  ix = (*incx < 0) ? (1 - *n) * *incx : 0;
  iy = (*incy < 0) ? (1 - *n) * *incy : 0;
  for (i = 0; i < *n; i++, ix += *incx, iy += *incy) { double t = dx[ix]; dx[ix] = dy[iy]; dy[iy] = t; }
}

double dasum_(const blasint *n, const double *dx, const blasint *incx)
{
  double  dtemp = 0.0;
  blasint i;
  if (*n <= 0 || *incx <= 0) return 0.0;
  for (i = 0; i < *n; i++) dtemp = dtemp + fabs(dx[i * *incx]);
  return dtemp;
}

double dnrm2_(const blasint *n, const double *x, const blasint *incx)
{
  double  scale = 0.0, ssq = 1.0;
  blasint i;
  if (*n < 1 || *incx < 1) return 0.0;
  if (*n == 1) return fabs(x[0]);
  for (i = 0; i < *n; i++) {
    double v = x[i * *incx];
    if (v != 0.0) {
      double a = fabs(v);
      if (scale < a) { ssq = 1.0 + ssq * (scale / a) * (scale / a); scale = a; }
      else ssq += (a / scale) * (a / scale);
    }
  }
  return scale * sqrt(ssq);
}

/* y := alpha*op(A)*x + beta*y, column-major A(lda,*) */
void dgemv_(const char *trans, const blasint *m, const blasint *n, const double *alpha, const double *a, const blasint *lda,
            const double *x, const blasint *incx, const double *beta, double *y, const blasint *incy)
{
  blasint i, j, leny, lenx;
  int     tr = (*trans == 'T' || *trans == 't' || *trans == 'C' || *trans == 'c');
  if (*m <= 0 || *n <= 0) return;
  lenx = tr ? *m : *n;
  leny = tr ? *n : *m;
  (void)lenx;
  if (*beta != 1.0) {
    for (i = 0; i < leny; i++) y[i * *incy] = (*beta == 0.0) ? 0.0 : *beta * y[i * *incy];
  }
  if (*alpha == 0.0) return;
  if (!tr) {
    for (j = 0; j < *n; j++) {
      double t = *alpha * x[j * *incx];
      for (i = 0; i < *m; i++) y[i * *incy] += t * a[i + (size_t)j * *lda];
    }
  } else {
    for (j = 0; j < *n; j++) {
      double t = 0.0;
      for (i = 0; i < *m; i++) t += a[i + (size_t)j * *lda] * x[i * *incx];
      y[j * *incy] += *alpha * t;
    }
  }
}

/* C := alpha*op(A)*op(B) + beta*C */
void dgemm_(const char *ta, const char *tb, const blasint *m, const blasint *n, const blasint *k, const double *alpha,
            const double *a, const blasint *lda, const double *b, const blasint *ldb, const double *beta, double *c,
            const blasint *ldc)
{
  blasint i, j, l;
  int     nota = (*ta == 'N' || *ta == 'n'), notb = (*tb == 'N' || *tb == 'n');
  for (j = 0; j < *n; j++) {
    for (i = 0; i < *m; i++) {
      double t = 0.0;
      for (l = 0; l < *k; l++) {
        double av = nota ? a[i + (size_t)l * *lda] : a[l + (size_t)i * *lda];
        double bv = notb ? b[l + (size_t)j * *ldb] : b[j + (size_t)l * *ldb];
        t += av * bv;
      }
      c[i + (size_t)j * *ldc] = *alpha * t + ((*beta == 0.0) ? 0.0 : *beta * c[i + (size_t)j * *ldc]);
    }
  }
}

/* ---- LAPACK stubs: off the hot path; report failure loudly ---- */
#define LAPACK_STUB(name)                                                                                  \
  void name(void)                                                                                          \
  {                                                                                                        \
    fprintf(stderr, "oracle/refblas.c: LAPACK routine %s is not available in the oracle build\n", #name);  \
    abort();                                                                                               \
  }
LAPACK_STUB(dgetrf_) LAPACK_STUB(dgetrs_) LAPACK_STUB(dpotrf_) LAPACK_STUB(dpotrs_) LAPACK_STUB(dpttrf_)
LAPACK_STUB(dpttrs_) LAPACK_STUB(dgesvd_) LAPACK_STUB(dorgqr_) LAPACK_STUB(dgeqrf_) LAPACK_STUB(dgeev_)
LAPACK_STUB(dtrsen_) LAPACK_STUB(dtgsen_) LAPACK_STUB(dsteqr_) LAPACK_STUB(dstein_) LAPACK_STUB(dstebz_)
LAPACK_STUB(dhseqr_) LAPACK_STUB(dgges_) LAPACK_STUB(dgesv_) LAPACK_STUB(dgerfs_) LAPACK_STUB(dsyev_)
LAPACK_STUB(dsyevx_) LAPACK_STUB(dsygv_) LAPACK_STUB(dsygvx_) LAPACK_STUB(dormqr_) LAPACK_STUB(dhgeqz_)
LAPACK_STUB(dtrtrs_) LAPACK_STUB(dgelss_) LAPACK_STUB(dtrmv_) LAPACK_STUB(dtrsm_)
