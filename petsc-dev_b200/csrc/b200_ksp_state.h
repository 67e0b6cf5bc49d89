/* b200_ksp_state.h — the device-resident scalar state of the fused CG drivers and the reference's scalar logic on it
 * (shared by b200_ksp.cu and b200_ksp_persist.cu). */
#ifndef B200_KSP_STATE_H
#define B200_KSP_STATE_H
#include "b200_internal.h"
#include <math.h>

#define KSP_ERR_FP_REASON (-100)   /* internal: NaN/Inf in a CG dot product -> PETSC_ERR_FP (cg.c:152,213,278) */

struct CgState {
  double beta, betaold, dpi, dpiold, a, b, dp, ttol, rnorm0, abstol, dtol;
  double red[4];            /* [0] z.z  [1] z.r  [2] r.r  [3] p.w   (local, then all-reduced in place) */
  int    i, its, reason, max_it, normtype, nhist, histmax, pad;
};

/* an allreduce over peer memory issued from inside a kernel (b200_p2p_allreduce_cta); size <= 1: none */
struct CgAr { unsigned long long seq; unsigned long long *const *inbox; int rank, size; };

static __device__ __forceinline__ bool bad(double v) { return isnan(v) || isinf(v); }

/* after p.w is known, red[3] (cg.c:207-220): one thread; s->reason == 0 on entry */
static __device__ __forceinline__ void cg_after_dot_body(CgState *s)
{
  s->dpi = s->red[3];
  s->betaold = s->beta;
  if (bad(s->dpi)) { s->reason = KSP_ERR_FP_REASON; return; }
  if (s->dpi == 0.0 || (s->i > 0 && s->dpi * s->dpiold <= 0.0)) { s->reason = B200_KSP_DIVERGED_INDEFINITE_MAT; return; }
  s->a = s->beta / s->dpi;
}

/* top of loop iteration i (cg.c:181-199): decides whether the iteration runs and fixes b */
static __device__ __forceinline__ void cg_top(CgState *s)
{
  s->its = s->i + 1;
  if (s->beta == 0.0) { s->reason = B200_KSP_CONVERGED_ATOL; return; }
  if (s->i > 0 && s->beta * s->betaold < 0.0) { s->reason = B200_KSP_DIVERGED_INDEFINITE_PC; return; }
  s->b = (s->i == 0) ? 0.0 : s->beta / s->betaold;
  s->dpiold = s->dpi;
}

/* after the update kernel's reductions (cg.c:227-281): norm, history, KSPDefaultConverged, new beta, next top */
static __device__ __forceinline__ void cg_after_update_body(CgState *s, double *hist)   /* one thread; s->reason == 0 on entry */
{
  double dp, beta = s->beta;
  if (s->normtype == B200_KSP_NORM_PRECONDITIONED) dp = sqrt(s->red[0]);
  else if (s->normtype == B200_KSP_NORM_UNPRECONDITIONED) dp = sqrt(s->red[2]);
  else if (s->normtype == B200_KSP_NORM_NATURAL) {
    beta = s->red[1];
    if (bad(beta)) { s->reason = KSP_ERR_FP_REASON; return; }
    dp = sqrt(fabs(beta));
  } else dp = 0.0;
  s->dp = dp;
  if (hist && s->histmax > s->nhist) hist[s->nhist++] = dp;
  if (s->normtype != B200_KSP_NORM_NONE) {
    if (bad(dp)) s->reason = B200_KSP_DIVERGED_NAN;
    else if (dp <= s->ttol) s->reason = (dp < s->abstol) ? B200_KSP_CONVERGED_ATOL : B200_KSP_CONVERGED_RTOL;
    else if (dp >= s->dtol * s->rnorm0) s->reason = B200_KSP_DIVERGED_DTOL;
    if (s->reason) return;
  }
  if (s->normtype != B200_KSP_NORM_NATURAL) {
    beta = s->red[1];
    if (bad(beta)) { s->reason = KSP_ERR_FP_REASON; return; }
  }
  s->beta = beta;
  s->i++;
  if (s->i >= s->max_it) { s->reason = B200_KSP_DIVERGED_ITS; return; }
  cg_top(s);
}


#endif
