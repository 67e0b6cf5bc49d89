/* b200_ksp.cu — KSPSolve for CG and GMRES(m) with PCJACOBI / PCNONE on device-resident vectors.
 *
 * Host-side restatement of the reference drivers; the scalar recurrences are the reference's, the vector work goes
 * through the kernels of b200_vec.cu / b200_spmv.cu:
 *   KSPSolve_CG                       src/ksp/ksp/impls/cg/cg.c:92-286
 *   KSPSolve_GMRES / KSPGMRESCycle    src/ksp/ksp/impls/gmres/gmres.c:118-238
 *   classical Gram-Schmidt            src/ksp/ksp/impls/gmres/borthog2.c:35-118
 *   Hessenberg update / back-solve    gmres.c:309-409
 *   KSPInitialResidual                src/ksp/ksp/interface/itres.c:39-72
 *   KSPDefaultConverged               src/ksp/ksp/interface/iterativ.c:702-777
 *   PCSetUp_Jacobi / PCApply_Jacobi   src/ksp/pc/impls/jacobi/jacobi.c:125-206,266-278
 *
 * Two CG forms produce the same iterates:
 *   fused = 0  op by op, exactly the reference's call sequence (3 host-visible reductions per iteration);
 *   fused = 1  the iteration's scalars (beta, dpi, a, b, dp, reason) live on the device: SpMV with p.w folded in,
 *              one update kernel doing x += a p, r -= a w, z = r.*d and the three dot products of the step, and
 *              1-thread kernels evaluating the reference's scalar logic (breakdown checks, KSPDefaultConverged,
 *              residual history).  The host only polls the reason every `check_every` iterations.
 */
#include "b200_internal.h"
#include "b200_p2p_dev.cuh"
#include "b200_ksp_state.h"
#include <math.h>
#include <vector>


/* ------------------------------------------------------------------------------------------------ operator */
static int op_mult(b200_mat A, const double *x, double *y) { return b200_mat_mult(A, x, y); }

struct Pc {
  int     type;      /* 0 none, 1 jacobi */
  int     n;
  double *d;         /* jacobi: 1/diag with zeros -> 1 */
};

__global__ void jacobi_fix_kernel(int n, double *d)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double v = d[i];
    v = (v != 0.0) ? 1.0 / v : v;      /* VecReciprocal_Default */
    if (v == 0.0) v = 1.0;             /* jacobi.c:175-180 */
    d[i] = v;
  }
}

static int pc_setup(Pc *pc, b200_mat A, int type, const double *inverse_diag_dev)
{
  pc->type = type; pc->n = A->A.m; pc->d = NULL;
  if (type == 1) {
    B200_CHK(b200_malloc(sizeof(double) * (size_t)(pc->n + 2), (void **)&pc->d));
    if (inverse_diag_dev) {                                  /* the caller's PCJACOBI has been set up already (any variant) */
      if (pc->n) B200_CHK(b200_vec_copy(pc->n, inverse_diag_dev, pc->d));
      return 0;
    }
    B200_CHK(b200_mat_get_diagonal(A, pc->d));
    if (pc->n) {
      int g = (pc->n + 255) / 256; if (g > B200_NSM * 8) g = B200_NSM * 8;
      jacobi_fix_kernel<<<g, 256, 0, g_b200.stream>>>(pc->n, pc->d);
      B200_LAUNCHED(1);
    }
  } else if (type != 0) return b200_seterr(B200_ERR_SUP, "pc_type %d: only none (0) and jacobi (1) are on this path", type);
  return 0;
}

static int pc_apply(const Pc *pc, const double *x, double *y)
{
  if (pc->type == 1) return b200_vec_pointwise_mult(pc->n, y, x, pc->d);   /* jacobi.c:275 */
  return b200_vec_copy(pc->n, x, y);                                        /* PCApply_None */
}

/* ------------------------------------------------------------------------------------------------ convergence */
struct Conv {
  double rtol, abstol, dtol, rnorm0, ttol;
  int    normtype, guess_nonzero;
};

/* KSPDefaultConverged (iterativ.c:702-777); at n == 0 needs the norm of the (preconditioned) right-hand side when
 * the guess is non-zero */
static int converged_host(Conv *cv, int n, double rnorm, int *reason, b200_mat A, const Pc *pc, const double *b, double *tmp)
{
  *reason = B200_KSP_CONVERGED_ITERATING;
  if (!n) {
    if (cv->guess_nonzero) {
      double snorm = 0.0;
      const int nl = A->A.m;
      if (cv->normtype == B200_KSP_NORM_UNPRECONDITIONED) { B200_CHK(b200_vec_norm(nl, b, B200_NORM_2, &snorm)); }
      else {
        B200_CHK(pc_apply(pc, b, tmp));
        if (cv->normtype == B200_KSP_NORM_PRECONDITIONED) { B200_CHK(b200_vec_norm(nl, tmp, B200_NORM_2, &snorm)); }
        else { double nrm; B200_CHK(b200_vec_dot(nl, b, tmp, &nrm)); snorm = sqrt(fabs(nrm)); }
      }
      if (!snorm) snorm = rnorm;
      cv->rnorm0 = snorm;
    } else cv->rnorm0 = rnorm;
    cv->ttol = fmax(cv->rtol * cv->rnorm0, cv->abstol);
  }
  if (isnan(rnorm) || isinf(rnorm)) *reason = B200_KSP_DIVERGED_NAN;
  else if (rnorm <= cv->ttol) *reason = (rnorm < cv->abstol) ? B200_KSP_CONVERGED_ATOL : B200_KSP_CONVERGED_RTOL;
  else if (rnorm >= cv->dtol * cv->rnorm0) *reason = B200_KSP_DIVERGED_DTOL;
  return 0;
}

struct Hist {
  double *h; int max, len;
  void log(double v) { if (h && max > len) h[len++] = v; }     /* KSPLogResidualHistory, kspimpl.h:131 */
};

/* ------------------------------------------------------------------------------------------------ CG, op by op */
static int cg_reference_sequence(b200_mat A, const Pc *pc, Conv *cv, int max_it, const double *B, double *X,
                                 double *R, double *Z, double *P, Hist *hist, b200_ksp_result *res)
{
  const int n = A->A.m;
  double dpi = 0.0, a = 1.0, beta = 0.0, betaold = 1.0, b = 0.0, dpiold, dp = 0.0;
  double *W = Z;                                                /* cg.c:121 */
  int reason = 0, its = 0;
  if (cv->guess_nonzero) { B200_CHK(op_mult(A, X, R)); B200_CHK(b200_vec_aypx(n, R, -1.0, B)); }   /* r = b - Ax */
  else { B200_CHK(b200_vec_copy(n, B, R)); }
  switch (cv->normtype) {
  case B200_KSP_NORM_PRECONDITIONED:
    B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_norm(n, Z, B200_NORM_2, &dp)); break;
  case B200_KSP_NORM_UNPRECONDITIONED:
    B200_CHK(b200_vec_norm(n, R, B200_NORM_2, &dp)); break;
  case B200_KSP_NORM_NATURAL:
    B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_dot(n, Z, R, &beta));
    if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
    dp = sqrt(fabs(beta)); break;
  case B200_KSP_NORM_NONE: dp = 0.0; break;
  default: return b200_seterr(B200_ERR_SUP, "norm type %d", cv->normtype);
  }
  hist->log(dp);
  res->rnorm = dp;
  B200_CHK(converged_host(cv, 0, dp, &reason, A, pc, B, P));
  if (reason) { res->its = 0; res->reason = reason; return 0; }
  if (cv->normtype != B200_KSP_NORM_PRECONDITIONED && cv->normtype != B200_KSP_NORM_NATURAL) { B200_CHK(pc_apply(pc, R, Z)); }
  if (cv->normtype != B200_KSP_NORM_NATURAL) {
    B200_CHK(b200_vec_dot(n, Z, R, &beta));
    if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
  }
  int i = 0;
  do {
    its = i + 1;
    if (beta == 0.0) { reason = B200_KSP_CONVERGED_ATOL; break; }
    else if (i > 0 && beta * betaold < 0.0) { reason = B200_KSP_DIVERGED_INDEFINITE_PC; break; }
    if (!i) { B200_CHK(b200_vec_copy(n, Z, P)); b = 0.0; }
    else { b = beta / betaold; B200_CHK(b200_vec_aypx(n, P, b, Z)); }
    dpiold = dpi;
    B200_CHK(op_mult(A, P, W));
    B200_CHK(b200_vec_dot(n, P, W, &dpi));
    betaold = beta;
    if (isnan(dpi) || isinf(dpi)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
    if (dpi == 0.0 || (i > 0 && dpi * dpiold <= 0.0)) { reason = B200_KSP_DIVERGED_INDEFINITE_MAT; break; }
    a = beta / dpi;
    B200_CHK(b200_vec_axpy(n, X, a, P));
    B200_CHK(b200_vec_axpy(n, R, -a, W));
    if (cv->normtype == B200_KSP_NORM_PRECONDITIONED) { B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_norm(n, Z, B200_NORM_2, &dp)); }
    else if (cv->normtype == B200_KSP_NORM_UNPRECONDITIONED) { B200_CHK(b200_vec_norm(n, R, B200_NORM_2, &dp)); }
    else if (cv->normtype == B200_KSP_NORM_NATURAL) {
      B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_dot(n, Z, R, &beta));
      if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
      dp = sqrt(fabs(beta));
    } else dp = 0.0;
    res->rnorm = dp;
    hist->log(dp);
    if (cv->normtype != B200_KSP_NORM_NONE) { B200_CHK(converged_host(cv, i + 1, dp, &reason, A, pc, B, NULL)); }
    if (reason) break;
    if (cv->normtype != B200_KSP_NORM_PRECONDITIONED && cv->normtype != B200_KSP_NORM_NATURAL) { B200_CHK(pc_apply(pc, R, Z)); }
    if (cv->normtype != B200_KSP_NORM_NATURAL) {
      B200_CHK(b200_vec_dot(n, Z, R, &beta));
      if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
    }
    i++;
  } while (i < max_it);
  if (i >= max_it) reason = B200_KSP_DIVERGED_ITS;
  res->its = its; res->reason = reason;
  return 0;
}

/* ------------------------------------------------------------------------------------------------ CG, fused */
__global__ void cg_begin_kernel(CgState *s) { if (threadIdx.x == 0 && s->reason == 0) cg_top(s); }

/* after p.w is known (cg.c:207-220) */
struct CgParts { const double *a; int na; const double *b; int nb; };   /* per-CTA partials of p.w left by the SpMV kernels */

__global__ void __launch_bounds__(256) cg_after_dot_kernel(CgState *s, const CgAr ar, const CgParts parts)
{
  if (parts.na + parts.nb > 0) {
    /* final, fixed-order reduction of the fused p.w partials (was a kernel of its own) */
    __shared__ double s_w[8];
    double t = 0.0;
    for (int i = threadIdx.x; i < parts.na; i += 256) t += parts.a[i];
    for (int i = threadIdx.x; i < parts.nb; i += 256) t += parts.b[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      double u = 0.0;
#pragma unroll
      for (int q = 0; q < 8; q++) u += s_w[q];
      s->red[3] = u;
    }
  }
  /* the allreduce is unconditional (collective): every rank runs the same number of them whatever s->reason says */
  if (ar.size > 1) b200_p2p_allreduce_cta(&s->red[3], 1, 0, ar.seq, ar.rank, ar.size, ar.inbox);
  else __syncthreads();
  if (threadIdx.x != 0 || s->reason != 0) return;
  cg_after_dot_body(s);
}

__global__ void __launch_bounds__(256) cg_parts_kernel(CgState *s, const CgParts parts)   /* NCCL transport: local p.w only */
{
  __shared__ double s_w[8];
  double t = 0.0;
  for (int i = threadIdx.x; i < parts.na; i += 256) t += parts.a[i];
  for (int i = threadIdx.x; i < parts.nb; i += 256) t += parts.b[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
  __syncthreads();
  if (threadIdx.x == 0) { double u = 0.0; for (int q = 0; q < 8; q++) u += s_w[q]; s->red[3] = u; }
}

__global__ void __launch_bounds__(64) cg_after_update_kernel(CgState *s, double *hist, const CgAr ar)
{
  if (ar.size > 1) b200_p2p_allreduce_cta(&s->red[0], 3, 0, ar.seq, ar.rank, ar.size, ar.inbox);
  if (threadIdx.x != 0 || s->reason != 0) return;
  cg_after_update_body(s, hist);
}

/* p = z + b p  (VecCopy at i == 0, VecAYPX after: cg.c:194,202; b == 0 cannot occur for i > 0 unless beta == 0,
 * which ends the iteration before) */
__global__ void __launch_bounds__(256) cg_pupdate_kernel(int n, const CgState *__restrict__ s, double *__restrict__ p, const double *__restrict__ z)
{
  if (s->reason != 0) return;
  const bool first = (s->i == 0);
  const double b = s->b;
  const int stride = gridDim.x * blockDim.x;
  const int n2 = n >> 1;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    const double2 zv = reinterpret_cast<const double2 *>(z)[i];
    double2 pv;
    if (first) pv = zv;
    else {
      pv = reinterpret_cast<double2 *>(p)[i];
      if (b == 1.0) { pv.x = pv.x + zv.x; pv.y = pv.y + zv.y; }            /* VecAYPX alpha == 1 -> daxpy */
      else if (b == -1.0) { pv.x = zv.x - pv.x; pv.y = zv.y - pv.y; }
      else { pv.x = zv.x + b * pv.x; pv.y = zv.y + b * pv.y; }
    }
    reinterpret_cast<double2 *>(p)[i] = pv;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int i = n - 1;
    if (first) p[i] = z[i];
    else if (b == 1.0) p[i] = p[i] + z[i];
    else if (b == -1.0) p[i] = z[i] - p[i];
    else p[i] = z[i] + b * p[i];
  }
}

/* x += a p ; r -= a w ; z = r .* d (or z = r) ; partial sums of z.z, z.r, r.r  — 5 reads + 3 writes per row */
template <bool JACOBI, bool INLINE_AFTER>
__global__ void __launch_bounds__(512) cg_update_kernel(int n, CgState *s, double *__restrict__ x, const double *__restrict__ p,
                                                        double *__restrict__ r, const double *w, double *z,   /* W == Z (cg.c:121): no restrict */
                                                        const double *__restrict__ d, double *partials, unsigned *ticket, double *hist,
                                                        const CgAr ar)
{
  /* ar.size > 1 (INLINE_AFTER only): the three sums are all-reduced over peer memory by the last CTA, which then runs the
   * scalar step.  That allreduce is collective, so a finished solve (reason != 0, the same on every rank) still goes
   * through it with zero sums; without it the kernel just returns. */
  const bool active = (s->reason == 0);
  if (!active && ar.size <= 1) return;
  const double a = s->a, ma = -a;
  double zz = 0.0, zr = 0.0, rr = 0.0;
  const int stride = gridDim.x * blockDim.x;
  const int n2 = active ? (n >> 1) : 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
    double2 xv = reinterpret_cast<double2 *>(x)[i];
    const double2 pv = reinterpret_cast<const double2 *>(p)[i];
    double2 rv = reinterpret_cast<double2 *>(r)[i];
    const double2 wv = reinterpret_cast<const double2 *>(w)[i];
    double2 zv;
    xv.x = xv.x + a * pv.x; xv.y = xv.y + a * pv.y;                      /* VecAXPY(X,a,P)   */
    rv.x = rv.x + ma * wv.x; rv.y = rv.y + ma * wv.y;                    /* VecAXPY(R,-a,W)  */
    if (JACOBI) { const double2 dv = reinterpret_cast<const double2 *>(d)[i]; zv.x = rv.x * dv.x; zv.y = rv.y * dv.y; }
    else zv = rv;
    reinterpret_cast<double2 *>(x)[i] = xv;
    reinterpret_cast<double2 *>(r)[i] = rv;
    reinterpret_cast<double2 *>(z)[i] = zv;
    zz += zv.x * zv.x; zz += zv.y * zv.y;
    zr += zv.x * rv.x; zr += zv.y * rv.y;
    rr += rv.x * rv.x; rr += rv.y * rv.y;
  }
  if (active && (n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    const int i = n - 1;
    const double xv = x[i] + a * p[i], rv = r[i] + ma * w[i], zv = JACOBI ? rv * d[i] : rv;
    x[i] = xv; r[i] = rv; z[i] = zv;
    zz += zv * zv; zr += zv * rv; rr += rv * rv;
  }
  /* block reduce, per-block partials, last block finalises in fixed order */
  __shared__ double s_w[3][32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
  double v[3] = {zz, zr, rr};
#pragma unroll
  for (int q = 0; q < 3; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
    if (lane == 0) s_w[q][wid] = v[q];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int q = 0; q < 3; q++) {
      double t = (lane < nw) ? s_w[q][lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      if (lane == 0) partials[(size_t)blockIdx.x * 3 + q] = t;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
#pragma unroll
  for (int q = 0; q < 3; q++) {
    double t = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) t += __ldcg(&partials[(size_t)b * 3 + q]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    __syncthreads();
    if (lane == 0) s_w[q][wid] = t;
    __syncthreads();
    if (wid == 0) {
      double u = (lane < nw) ? s_w[q][lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
      if (lane == 0) s->red[q] = u;
    }
  }
  if (threadIdx.x == 0) *ticket = 0;
  if (INLINE_AFTER) {
    /* the scalar recurrences that follow (norm, convergence test, new beta) run here, in the last CTA, instead of in a
     * kernel of their own (thread 0 wrote red[] itself) */
    if (ar.size > 1) b200_p2p_allreduce_cta(&s->red[0], 3, 0, ar.seq, ar.rank, ar.size, ar.inbox);
    if (threadIdx.x == 0 && active) cg_after_update_body(s, hist);
  }
}

/* generic dot into a device scalar (MPIAIJ operator: p.w cannot ride on the two-part product) */
__global__ void __launch_bounds__(512) cg_dot_kernel(int n, const CgState *s, const double *__restrict__ x, const double *__restrict__ y,
                                                     double *partials, unsigned *ticket, double *out)
{
  double acc = 0.0;
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) acc += x[i] * y[i];
  __shared__ double s_w[32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) s_w[wid] = acc;
  __syncthreads();
  if (wid == 0) {
    double t = (lane < nw) ? s_w[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) partials[blockIdx.x] = t;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double t = 0.0;
  for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) t += __ldcg(&partials[b]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  __syncthreads();
  if (lane == 0) s_w[wid] = t;
  __syncthreads();
  if (wid == 0) {
    double u = (lane < nw) ? s_w[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
    if (lane == 0) *out = u;
  }
  if (threadIdx.x == 0) *ticket = 0;
}

static int cg_fused(b200_mat A, const Pc *pc, Conv *cv, int max_it, int check_every, const double *B, double *X,
                    double *R, double *Z, double *P, Hist *hist, b200_ksp_result *res)
{
  B200Ctx &c = g_b200;
  cudaStream_t st = c.stream;
  const int n = A->A.m;
  double *W = Z;
  double dp = 0.0, beta = 0.0;
  int reason = 0;
  /* the start-up phase (cg.c:129-176) is not on the per-iteration path: op by op */
  if (cv->guess_nonzero) { B200_CHK(op_mult(A, X, R)); B200_CHK(b200_vec_aypx(n, R, -1.0, B)); }
  else { B200_CHK(b200_vec_copy(n, B, R)); }
  switch (cv->normtype) {
  case B200_KSP_NORM_PRECONDITIONED: B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_norm(n, Z, B200_NORM_2, &dp)); break;
  case B200_KSP_NORM_UNPRECONDITIONED: B200_CHK(b200_vec_norm(n, R, B200_NORM_2, &dp)); break;
  case B200_KSP_NORM_NATURAL:
    B200_CHK(pc_apply(pc, R, Z)); B200_CHK(b200_vec_dot(n, Z, R, &beta));
    if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
    dp = sqrt(fabs(beta)); break;
  case B200_KSP_NORM_NONE: dp = 0.0; break;
  default: return b200_seterr(B200_ERR_SUP, "norm type %d", cv->normtype);
  }
  hist->log(dp);
  res->rnorm = dp;
  B200_CHK(converged_host(cv, 0, dp, &reason, A, pc, B, P));
  if (reason) { res->its = 0; res->reason = reason; return 0; }
  if (cv->normtype != B200_KSP_NORM_PRECONDITIONED && cv->normtype != B200_KSP_NORM_NATURAL) { B200_CHK(pc_apply(pc, R, Z)); }
  if (cv->normtype != B200_KSP_NORM_NATURAL) {
    B200_CHK(b200_vec_dot(n, Z, R, &beta));
    if (isnan(beta) || isinf(beta)) return b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product");
  }
  if (max_it <= 0) { res->its = 0; res->reason = B200_KSP_DIVERGED_ITS; return 0; }

  /* device-resident state */
  CgState hs;
  memset(&hs, 0, sizeof(hs));
  hs.beta = beta; hs.betaold = 1.0; hs.dpi = 0.0; hs.a = 1.0; hs.b = 0.0; hs.dp = dp;
  hs.ttol = cv->ttol; hs.rnorm0 = cv->rnorm0; hs.abstol = cv->abstol; hs.dtol = cv->dtol;
  hs.i = 0; hs.its = 0; hs.reason = 0; hs.max_it = max_it; hs.normtype = cv->normtype;
  const int hist_off = hist->len;
  hs.nhist = 0; hs.histmax = hist->h ? (hist->max - hist->len) : 0;
  CgState *ds = NULL, *ps = NULL;
  double *dhist = NULL;
  B200_CHK(b200_malloc(sizeof(CgState), (void **)&ds));
  B200_CHK(b200_host_alloc(sizeof(CgState), (void **)&ps));
  if (hs.histmax > 0) { B200_CHK(b200_malloc(sizeof(double) * (size_t)hs.histmax, (void **)&dhist)); }
  B200_CUDA(cudaMemcpyAsync(ds, &hs, sizeof(hs), cudaMemcpyHostToDevice, st));
  cg_begin_kernel<<<1, 32, 0, st>>>(ds);
  B200_LAUNCHED(1);

  const bool seq = (A->kind == 0) && !A->A.compressed && (A->A.m == A->A.n) && A->A.nstrips <= 1;
  const bool multi = c.size > 1;
  const bool mpi_dot = (A->kind == 1) && !A->A.compressed && (A->A.m == A->A.n) && A->A.m > 0 && A->A.nstrips <= 1 && A->B.nstrips <= 1;
  int gp = ((n + 1) / 2 + 255) / 256; if (gp > B200_NSM * 16) gp = B200_NSM * 16; if (gp < 1) gp = 1;
  const int gu = b200_red_grid((n + 1) / 2, 512, 2);
  if (check_every < 1) check_every = 1;
  /* short iterations: the whole block of iterations as ONE persistent cooperative kernel (b200_ksp_persist.cu); it leaves
   * early when the device-side convergence test fires, so the host need not look often */
  const bool persist = !multi && seq && b200_cg_persistent_ok(A);
  if (persist && check_every < 256) check_every = 256;
  /* short iterations on one GPU: the host cannot launch four kernels as fast as the GPU finishes them (measured floor:
   * 22 us per iteration however small the matrix).  A block of iterations is therefore captured ONCE as a CUDA graph
   * (every kernel argument is the same from iteration to iteration: the scalars live in the CgState on the device) and
   * replayed; the first block runs directly (it also triggers the kernels' one-time attribute calls). */
  bool use_graph = !multi && seq && !persist && A->A.nz <= (64ll << 20) && A->A.nlong == 0;
  { const char *eg = getenv("B200_CG_GRAPH"); if (eg) use_graph = use_graph && atoi(eg) != 0; }
  const int gblock = 64;                                    /* iterations per graph; the host looks every check_every */
  if (use_graph) check_every = check_every < gblock ? gblock : (check_every / gblock) * gblock;
  const int block = use_graph ? gblock : check_every;       /* iterations issued per pass of the loop below */
  cudaGraphExec_t gexec = NULL;
  long long graph_launches = 0;
  bool capturing = false;
  int launched = 0, rc = 0, nbatch = 0, since_check = 0;
  while (launched < max_it) {
    int batch = block;
    if (launched + batch > max_it) batch = max_it - launched;
    const long long l0 = c.launches;
    if (persist) {
      rc = b200_cg_persistent_run(A, pc->type == 1, pc->d, X, R, Z, P, ds, dhist, batch); if (rc) goto done;
    } else if (gexec && batch == block) {
      if (cudaGraphLaunch(gexec, st) != cudaSuccess) { rc = b200_cuda_err(cudaGetLastError(), "cudaGraphLaunch", __FILE__, __LINE__); goto done; }
      c.launches += graph_launches;
    } else {
    if (use_graph && !gexec && nbatch >= 1 && batch == block) {
      if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) == cudaSuccess) capturing = true;
      else { cudaGetLastError(); use_graph = false; }
    }
    for (int k = 0; k < batch; k++) {
      cg_pupdate_kernel<<<gp, 256, 0, st>>>(n, ds, P, Z);
      B200_LAUNCHED(1);
      CgAr ar1 = {0, NULL, c.rank, 1}, ar2 = {0, NULL, c.rank, 1};
      CgParts parts = {NULL, 0, NULL, 0};
      bool dot_done = false;                      /* p.w, its allreduce and the scalar step after it already ran */
      if (seq && !multi && A->A.nlong == 0 && A->A.ntiles > 0) {
        /* one GPU: the product's last CTA finishes p.w and takes the scalar step — three launches per iteration */
        rc = b200_seq_mult_ex(&A->A, P, NULL, W, false, true, P, NULL, ds); if (rc) goto done;
        dot_done = true;
      } else if (seq) {
        rc = b200_seq_mult_ex(&A->A, P, NULL, W, false, true, P, NULL); if (rc) goto done;
        b200_seq_dotparts(&A->A, &parts.a, &parts.na);
      } else if (mpi_dot && b200_mpiaij_fused(A)) {
        /* one launch: ghost exchange + both blocks + p.w; over peer memory its last CTA also all-reduces p.w and takes the
         * scalar step */
        if (!multi || b200_p2p_next(&ar1.seq, &ar1.inbox)) {
          if (multi) ar1.size = c.size;
          rc = b200_mpiaij_mult_fused_cg(A, P, NULL, W, false, true, ds, &ar1); if (rc) goto done;
          dot_done = true;
        } else {
          rc = b200_mpiaij_mult_fused(A, P, NULL, W, false, true); if (rc) goto done;
          b200_mpiaij_fused_dotparts(A, &parts.a, &parts.na);
        }
      } else if (mpi_dot) {
        rc = b200_mpiaij_mult_dot(A, P, W, NULL); if (rc) goto done;
        b200_seq_dotparts(&A->A, &parts.a, &parts.na);
        b200_seq_dotparts(&A->B, &parts.b, &parts.nb);
      } else {
        rc = op_mult(A, P, W); if (rc) goto done;
        cg_dot_kernel<<<b200_red_grid(n, 512, 4), 512, 0, st>>>(n, ds, P, W, c.d_partials, c.d_ticket, &ds->red[3]);
        B200_LAUNCHED(1);
      }
      /* p.w: final reduction of the SpMV partials + allreduce (peer memory) + the scalar step, one kernel */
      if (!dot_done) {
        if (multi) {
          if (b200_p2p_next(&ar1.seq, &ar1.inbox)) ar1.size = c.size;
          else {
            if (parts.na + parts.nb > 0) {          /* NCCL needs the local sum in memory first */
              cg_parts_kernel<<<1, 256, 0, st>>>(ds, parts); B200_LAUNCHED(1);
              parts.na = parts.nb = 0;
            }
            rc = b200_comm_allreduce_dev(&ds->red[3], 1, 0); if (rc) goto done;
          }
        }
        cg_after_dot_kernel<<<1, 256, 0, st>>>(ds, ar1, parts);
        B200_LAUNCHED(1);
      }
      if (multi && !b200_p2p_next(&ar2.seq, &ar2.inbox)) {
        /* NCCL transport: update, allreduce, scalar step as three launches */
        if (pc->type == 1) cg_update_kernel<true, false><<<gu, 512, 0, st>>>(n, ds, X, P, R, W, Z, pc->d, c.d_partials, c.d_ticket, dhist, ar2);
        else               cg_update_kernel<false, false><<<gu, 512, 0, st>>>(n, ds, X, P, R, W, Z, NULL, c.d_partials, c.d_ticket, dhist, ar2);
        B200_LAUNCHED(1);
        rc = b200_comm_allreduce_dev(&ds->red[0], 3, 0); if (rc) goto done;
        cg_after_update_kernel<<<1, 64, 0, st>>>(ds, dhist, ar2);
        B200_LAUNCHED(1);
      } else {
        /* the last CTA all-reduces z.z, z.r, r.r over peer memory (one rank: nothing to reduce) and takes the scalar step */
        if (multi) ar2.size = c.size;
        if (pc->type == 1) cg_update_kernel<true, true><<<gu, 512, 0, st>>>(n, ds, X, P, R, W, Z, pc->d, c.d_partials, c.d_ticket, dhist, ar2);
        else               cg_update_kernel<false, true><<<gu, 512, 0, st>>>(n, ds, X, P, R, W, Z, NULL, c.d_partials, c.d_ticket, dhist, ar2);
        B200_LAUNCHED(1);
      }
    }
    if (capturing) {
      cudaGraph_t g = NULL;
      capturing = false;
      if (cudaStreamEndCapture(st, &g) != cudaSuccess || !g) { rc = b200_cuda_err(cudaGetLastError(), "cudaStreamEndCapture", __FILE__, __LINE__); goto done; }
      const cudaError_t ei = cudaGraphInstantiate(&gexec, g, 0);
      cudaGraphDestroy(g);
      if (ei != cudaSuccess) { gexec = NULL; rc = b200_cuda_err(ei, "cudaGraphInstantiate", __FILE__, __LINE__); goto done; }
      graph_launches = c.launches - l0;                 /* the capture recorded this block; now it runs */
      if (cudaGraphLaunch(gexec, st) != cudaSuccess) { rc = b200_cuda_err(cudaGetLastError(), "cudaGraphLaunch", __FILE__, __LINE__); goto done; }
    }
    }
    nbatch++;
    launched += batch;
    since_check += batch;
    if (since_check < check_every && launched < max_it) continue;
    since_check = 0;
    if (cudaMemcpyAsync(ps, ds, sizeof(CgState), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
      rc = b200_cuda_err(cudaGetLastError(), "CG state readback", __FILE__, __LINE__); goto done;
    }
    if (ps->reason != 0) break;
  }
  if (ps->reason == 0) ps->reason = B200_KSP_DIVERGED_ITS;
  if (ps->reason == KSP_ERR_FP_REASON) { rc = b200_seterr(B200_ERR_FP, "Infinite or not-a-number generated in dot product"); goto done; }
  res->its = ps->its; res->reason = ps->reason; res->rnorm = ps->dp;
  if (ps->nhist > 0) {
    if (cudaMemcpy(hist->h + hist_off, dhist, sizeof(double) * (size_t)ps->nhist, cudaMemcpyDeviceToHost) != cudaSuccess) {
      rc = b200_cuda_err(cudaGetLastError(), "history readback", __FILE__, __LINE__); goto done;
    }
    hist->len = hist_off + ps->nhist;
  }
done:
  if (capturing) { cudaGraph_t g = NULL; cudaStreamEndCapture(st, &g); if (g) cudaGraphDestroy(g); cudaGetLastError(); }
  if (gexec) cudaGraphExecDestroy(gexec);
  b200_free(ds); b200_host_free(ps); b200_free(dhist);
  return rc;
}

/* ------------------------------------------------------------------------------------------------ GMRES(m) */
struct Gmres {
  int     max_k;
  std::vector<double> hh, hes, rs, cc, ss, lhh, nrs;    /* hh: (max_k+2) x (max_k+1), column major as gmresimpl.h:73 */
  double *HH(int a, int b) { return &hh[(size_t)b * (max_k + 2) + a]; }
  double *HES(int a, int b) { return &hes[(size_t)b * (max_k + 1) + a]; }
};

/* VecNormalize (rvector.c:299-320) */
static int vec_normalize(int n, double *x, double *val)
{
  double nrm;
  B200_CHK(b200_vec_norm(n, x, B200_NORM_2, &nrm));
  if (nrm != 0.0 && nrm != 1.0) { B200_CHK(b200_vec_scale(n, x, 1.0 / nrm)); }
  *val = nrm;
  return 0;
}

/* KSPGMRESUpdateHessenberg (gmres.c:360-409): apply the stored rotations to column `it`, make the new one */
static void gmres_update_hessenberg(Gmres *g, int it, bool hapend, double *res, int *reason)
{
  double *hh = g->HH(0, it);
  for (int j = 1; j <= it; j++) {
    const double c = g->cc[j - 1], s = g->ss[j - 1], t = hh[j - 1];
    hh[j - 1] = c * t + s * hh[j];
    hh[j]     = c * hh[j] - s * t;
  }
  if (!hapend) {
    const double t = sqrt(hh[it] * hh[it] + hh[it + 1] * hh[it + 1]);
    if (t == 0.0) { *reason = B200_KSP_DIVERGED_NULL; return; }
    g->cc[it] = hh[it] / t;
    g->ss[it] = hh[it + 1] / t;
    g->rs[it + 1] = -(g->ss[it] * g->rs[it]);
    g->rs[it]     = g->cc[it] * g->rs[it];
    hh[it]        = g->cc[it] * hh[it] + g->ss[it] * hh[it + 1];
    *res          = fabs(g->rs[it + 1]);
  } else *res = 0.0;
}

static int gmres_solve(b200_mat A, const Pc *pc, Conv *cv, const b200_ksp_options *o, const double *B, double *X,
                       Hist *hist, b200_ksp_result *res)
{
  const int n = A->A.m, max_k = o->gmres_restart;
  if (max_k < 1) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "restart must be positive");
  if (cv->normtype != B200_KSP_NORM_PRECONDITIONED && cv->normtype != B200_KSP_NORM_NONE)
    return b200_seterr(B200_ERR_SUP, "GMRES (left preconditioning) supports the preconditioned norm only");   /* gmres.c:943 */
  Gmres g;
  g.max_k = max_k;
  g.hh.assign((size_t)(max_k + 2) * (max_k + 1), 0.0); g.hes.assign((size_t)(max_k + 1) * (max_k + 1), 0.0);
  g.rs.assign(max_k + 2, 0.0); g.cc.assign(max_k + 2, 0.0); g.ss.assign(max_k + 2, 0.0); g.lhh.assign(max_k + 2, 0.0); g.nrs.assign(max_k + 2, 0.0);
  /* VEC_TEMP, VEC_TEMP_MATOP, VV(0..max_k): separate 16-byte aligned vectors carved from one slab */
  const size_t ld = ((size_t)n + 1) & ~(size_t)1;
  double *slab = NULL;
  B200_CHK(b200_malloc(sizeof(double) * ld * (size_t)(max_k + 3) + 16, (void **)&slab));
  double *TEMP = slab, *MATOP = slab + ld;
  std::vector<const double *> vv(max_k + 1);
  for (int k = 0; k <= max_k; k++) vv[k] = slab + ld * (size_t)(k + 2);
  auto VV = [&](int k) { return const_cast<double *>(vv[k]); };
  int its_total = 0, reason = 0, rc = 0, itcount = 0;
  bool guess_zero = !cv->guess_nonzero;
  double resn = 0.0;
  while (!reason) {
    /* KSPInitialResidual (itres.c:52-67), left preconditioning */
    if (!guess_zero) {
      if ((rc = op_mult(A, X, TEMP))) break;
      if ((rc = b200_vec_copy(n, B, MATOP))) break;
      if ((rc = b200_vec_axpy(n, MATOP, -1.0, TEMP))) break;
      if ((rc = pc_apply(pc, MATOP, VV(0)))) break;
    } else {
      if ((rc = b200_vec_copy(n, B, MATOP))) break;
      if ((rc = pc_apply(pc, B, VV(0)))) break;
    }
    /* KSPGMRESCycle (gmres.c:118-209) */
    int it = 0;
    bool hapend = false;
    double res_norm, tt;
    if ((rc = vec_normalize(n, VV(0), &res_norm))) break;
    resn = res_norm;
    g.rs[0] = res_norm;
    res->rnorm = resn;
    hist->log(resn);
    if (!resn) { reason = B200_KSP_CONVERGED_ATOL; break; }
    if (cv->normtype != B200_KSP_NORM_NONE) { if ((rc = converged_host(cv, its_total, resn, &reason, A, pc, B, TEMP))) break; }
    while (!reason && it < max_k && its_total < o->max_it) {
      if (it) hist->log(resn);
      /* KSP_PCApplyBAorAB, left: VV(it+1) = B A VV(it) */
      if ((rc = op_mult(A, VV(it), MATOP))) break;
      if ((rc = pc_apply(pc, MATOP, VV(it + 1)))) break;
      double *hh = g.HH(0, it), *hes = g.HES(0, it), *lhh = g.lhh.data();
      if (o->gmres_modified_gs) {
        /* modified Gram-Schmidt (borthog.c:41-49): hh[j] = (vv(it+1), vv(j)); vv(it+1) -= hh[j] vv(j), one j at a time */
        for (int j = 0; j <= it; j++) {
          if ((rc = b200_vec_dot(n, VV(it + 1), vv[j], &hh[j]))) break;
          hes[j] = hh[j];
          if ((rc = b200_vec_axpy(n, VV(it + 1), -hh[j], vv[j]))) break;
        }
      } else {
      /* classical Gram-Schmidt (borthog2.c:52-82) */
      for (int j = 0; j <= it; j++) { hh[j] = 0.0; hes[j] = 0.0; }
      bool refine = (o->gmres_cgs_refine == 2);
      for (int pass = 0; pass < 2; pass++) {
        if ((rc = b200_vec_mdot(n, VV(it + 1), it + 1, vv.data(), lhh))) break;
        for (int j = 0; j <= it; j++) lhh[j] = -lhh[j];
        if ((rc = b200_vec_maxpy(n, VV(it + 1), it + 1, lhh, vv.data()))) break;
        for (int j = 0; j <= it; j++) { hh[j] -= lhh[j]; hes[j] -= lhh[j]; }
        if (pass == 0 && o->gmres_cgs_refine == 1) {
          double hnrm = 0.0, wnrm;
          for (int j = 0; j <= it; j++) hnrm += lhh[j] * lhh[j];
          hnrm = sqrt(hnrm);
          if ((rc = b200_vec_norm(n, VV(it + 1), B200_NORM_2, &wnrm))) break;
          if (wnrm < 1.0286 * hnrm) refine = true;
        }
        if (!refine) break;
      }
      }
      if (rc) break;
      if ((rc = vec_normalize(n, VV(it + 1), &tt))) break;
      *g.HH(it + 1, it) = tt;
      *g.HES(it + 1, it) = tt;
      double hapbnd = fabs(tt / g.rs[it]);
      if (hapbnd > o->gmres_haptol) hapbnd = o->gmres_haptol;
      if (tt < hapbnd) hapend = true;
      gmres_update_hessenberg(&g, it, hapend, &resn, &reason);
      it++;
      its_total++;
      res->rnorm = resn;
      if (reason) break;
      if (cv->normtype != B200_KSP_NORM_NONE) { if ((rc = converged_host(cv, its_total, resn, &reason, A, pc, B, TEMP))) break; }
      if (hapend) {
        if (!reason) { rc = b200_seterr(B200_ERR_PLIB, "You reached the happy break down, but convergence was not indicated. Residual norm = %g", resn); }
        break;
      }
    }
    if (rc) break;
    if (it && (reason || its_total >= o->max_it)) hist->log(resn);
    /* KSPGMRESBuildSoln (gmres.c:309-354) with it-1 */
    {
      const int k1 = it - 1;
      if (k1 >= 0) {
        double *nrs = g.nrs.data();
        if (*g.HH(k1, k1) != 0.0) nrs[k1] = g.rs[k1] / *g.HH(k1, k1);
        else { reason = B200_KSP_DIVERGED_BREAKDOWN; goto cycle_done; }
        for (int ii = 1; ii <= k1; ii++) {
          const int k = k1 - ii;
          double t = g.rs[k];
          for (int j = k + 1; j <= k1; j++) t = t - *g.HH(k, j) * nrs[j];
          if (*g.HH(k, k) == 0.0) { reason = B200_KSP_DIVERGED_BREAKDOWN; goto cycle_done; }
          nrs[k] = t / *g.HH(k, k);
        }
        if ((rc = b200_vec_set(n, TEMP, 0.0))) break;
        if ((rc = b200_vec_maxpy(n, TEMP, k1 + 1, nrs, vv.data()))) break;
        if ((rc = b200_vec_axpy(n, X, 1.0, TEMP))) break;        /* left PC: nothing to unwind */
      }
    }
  cycle_done:
    itcount += it;
    if (itcount >= o->max_it) { if (!reason) reason = B200_KSP_DIVERGED_ITS; break; }
    guess_zero = false;
  }
  b200_free(slab);
  if (rc) return rc;
  res->its = its_total; res->reason = reason;
  return 0;
}

/* ------------------------------------------------------------------------------------------------ entry */
extern "C" int b200_ksp_default_options(b200_ksp_options *o)
{
  if (!o) return b200_seterr(B200_ERR_ARG_NULL, "null options");
  memset(o, 0, sizeof(*o));
  o->ksp_type = 0; o->pc_type = 1;
  o->rtol = 1e-5; o->abstol = 1e-50; o->dtol = 1e5; o->max_it = 10000;     /* itcreate.c defaults */
  o->norm_type = B200_KSP_NORM_PRECONDITIONED;
  o->guess_nonzero = 0;
  o->gmres_restart = 30; o->gmres_cgs_refine = 0; o->gmres_haptol = 1e-30; /* gmres.c:947-956 */
  o->gmres_modified_gs = 0;
  o->fused = 1; o->check_every = 16;
  return 0;
}

extern "C" int b200_ksp_solve(b200_mat A, const b200_ksp_options *o, const double *b, double *x,
                              double *hist_host, int histmax, b200_ksp_result *result)
{
  B200_NEED_INIT();
  if (!A || !o || !b || !x || !result) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (A->A.m != A->A.n) return b200_seterr(B200_ERR_ARG_SIZ, "KSP needs a square (local diagonal block) operator, got %d x %d", A->A.m, A->A.n);
  if (b == x) return b200_seterr(B200_ERR_ARG_IDN, "b and x must be different vectors");
  const int n = A->A.m;
  memset(result, 0, sizeof(*result));
  Pc pc;
  B200_CHK(pc_setup(&pc, A, o->pc_type, o->pc_jacobi_inverse_diag));                      /* KSPSetUp -> PCSetUp (itfunc.c:267) */
  Conv cv;
  cv.rtol = o->rtol; cv.abstol = o->abstol; cv.dtol = o->dtol; cv.rnorm0 = 0; cv.ttol = 0;
  cv.normtype = o->norm_type; cv.guess_nonzero = o->guess_nonzero;
  Hist hist = {hist_host, hist_host ? histmax : 0, 0};
  int rc = 0;
  if (!o->guess_nonzero) { rc = b200_vec_set(n, x, 0.0); }      /* itfunc.c:403 */
  double *work = NULL;
  const size_t ld = ((size_t)n + 1) & ~(size_t)1;
  if (!rc) rc = b200_synchronize();
  if (!rc) rc = b200_timer_start();
  if (!rc) {
    if (o->ksp_type == 0) {
      rc = b200_malloc(sizeof(double) * ld * 3 + 16, (void **)&work);      /* KSPSetUp_CG: 3 work vectors (cg.c:74) */
      if (!rc) {
        double *R = work, *Z = work + ld, *P = work + 2 * ld;
        rc = b200_timer_start();
        if (!rc) rc = o->fused ? cg_fused(A, &pc, &cv, o->max_it, o->check_every, b, x, R, Z, P, &hist, result)
                               : cg_reference_sequence(A, &pc, &cv, o->max_it, b, x, R, Z, P, &hist, result);
      }
    } else if (o->ksp_type == 1) {
      rc = gmres_solve(A, &pc, &cv, o, b, x, &hist, result);
    } else rc = b200_seterr(B200_ERR_SUP, "ksp_type %d: only cg (0) and gmres (1) are on this path", o->ksp_type);
  }
  double ms = 0.0;
  if (!rc) rc = b200_timer_stop(&ms);
  result->solve_ms = ms;
  result->nhist = hist.len;
  b200_free(work);
  b200_free(pc.d);
  return rc;
}
