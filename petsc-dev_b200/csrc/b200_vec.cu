/* b200_vec.cu — the Krylov vector kernels for sm_100a.
 *
 * Elementwise family: one streaming kernel template, 128-bit loads/stores, each expression written with the SAME
 * association the reference's C loop uses (src/vec/vec/impls/seq/{bvec1,bvec2,dvec2}.c) and compiled with
 * -fmad=false, so every elementwise result is bit-identical to the reference built without FMA contraction.
 * Reductions: grid-stride accumulate -> warp shuffle -> block -> last-block final pass in a FIXED order
 * (deterministic run to run; differs from the reference BLAS only in summation order).
 */
#include "b200_internal.h"
#include <mutex>

#define EW_THREADS 256
#define RED_THREADS 512

/* ------------------------------------------------------------------------------------------------ elementwise */
enum { OP_SET, OP_COPY, OP_SCALE, OP_AXPY, OP_AYPX, OP_XMY /* y = x - y */, OP_WAXPY, OP_WXPY, OP_WYMX, OP_AX /* y = a x */,
       OP_AXPBY, OP_AXPBYPCZ_A1, OP_AXPBYPCZ_G1, OP_AXPBYPCZ_G0, OP_AXPBYPCZ, OP_PMULT, OP_PDIV, OP_RECIP };

/* out = f(x, y, out_old); which inputs are read is a compile-time property of OP */
template <int OP> struct Ew {
  static constexpr bool RX = !(OP == OP_SET || OP == OP_SCALE || OP == OP_RECIP);
  static constexpr bool RY = (OP == OP_WAXPY || OP == OP_WXPY || OP == OP_WYMX || OP == OP_AXPBYPCZ_A1 || OP == OP_AXPBYPCZ_G1 ||
                              OP == OP_AXPBYPCZ_G0 || OP == OP_AXPBYPCZ || OP == OP_PMULT || OP == OP_PDIV);
  static constexpr bool RO = (OP == OP_SCALE || OP == OP_AXPY || OP == OP_AYPX || OP == OP_XMY || OP == OP_AXPBY ||
                              OP == OP_AXPBYPCZ_A1 || OP == OP_AXPBYPCZ_G1 || OP == OP_AXPBYPCZ || OP == OP_RECIP);
  __device__ static __forceinline__ double f(double x, double y, double o, double a, double b, double c)
  {
    switch (OP) {
    case OP_SET:         return a;
    case OP_COPY:        return x;
    case OP_SCALE:       return a * o;                 /* dscal: da*dx(i)                                    */
    case OP_AXPY:        return o + a * x;             /* daxpy: dy(i) + da*dx(i)                            */
    case OP_AYPX:        return x + a * o;             /* dvec2.c:1008                                       */
    case OP_XMY:         return x - o;                 /* dvec2.c:991                                        */
    case OP_WAXPY:       return y + a * x;             /* dvec2.c:1108                                       */
    case OP_WXPY:        return y + x;                 /* dvec2.c:1096                                       */
    case OP_WYMX:        return y - x;                 /* dvec2.c:1099                                       */
    case OP_AX:          return a * x;                 /* bvec1.c:338                                        */
    case OP_AXPBY:       return a * x + b * o;         /* bvec1.c:347                                        */
    case OP_AXPBYPCZ_A1: return (x + b * y) + c * o;   /* bvec1.c:432                                        */
    case OP_AXPBYPCZ_G1: return (a * x + b * y) + o;   /* bvec1.c:437                                        */
    case OP_AXPBYPCZ_G0: return a * x + b * y;         /* bvec1.c:442                                        */
    case OP_AXPBYPCZ:    return (a * x + b * y) + c * o; /* bvec1.c:447                                      */
    case OP_PMULT:       return x * y;                 /* bvec2.c:251                                        */
    case OP_PDIV:        return x / y;                 /* bvec2.c:311                                        */
    case OP_RECIP:       return (o != 0.0) ? 1.0 / o : o; /* VecReciprocal_Default                           */
    }
    return 0.0;
  }
};

template <int OP, bool VEC2>
__global__ void __launch_bounds__(EW_THREADS) ew_kernel(int n, double *__restrict__ out, const double *__restrict__ x,
                                                        const double *__restrict__ y, double a, double b, double c)
{
  typedef Ew<OP> E;
  if (VEC2) {
    const int n2 = n >> 1;
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
      double2 xv = make_double2(0, 0), yv = make_double2(0, 0), ov = make_double2(0, 0);
      if (E::RX) xv = reinterpret_cast<const double2 *>(x)[i];
      if (E::RY) yv = reinterpret_cast<const double2 *>(y)[i];
      if (E::RO) ov = reinterpret_cast<const double2 *>(out)[i];
      double2 r;
      r.x = E::f(xv.x, yv.x, ov.x, a, b, c);
      r.y = E::f(xv.y, yv.y, ov.y, a, b, c);
      reinterpret_cast<double2 *>(out)[i] = r;
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
      const int i = n - 1;
      out[i] = E::f(E::RX ? x[i] : 0.0, E::RY ? y[i] : 0.0, E::RO ? out[i] : 0.0, a, b, c);
    }
  } else {
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
      out[i] = E::f(E::RX ? x[i] : 0.0, E::RY ? y[i] : 0.0, E::RO ? out[i] : 0.0, a, b, c);
  }
}

static inline bool aligned16(const void *p) { return (((uintptr_t)p) & 15) == 0; }

template <int OP>
static int ew_launch(int n, double *out, const double *x, const double *y, double a, double b, double c)
{
  B200_NEED_INIT();
  if (n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative vector length %d", n);
  if (n == 0) return 0;
  if (!out || (Ew<OP>::RX && !x) || (Ew<OP>::RY && !y)) return b200_seterr(B200_ERR_ARG_NULL, "null device vector");
  const bool v2 = aligned16(out) && (!Ew<OP>::RX || aligned16(x)) && (!Ew<OP>::RY || aligned16(y));
  long long work = v2 ? (n + 1) / 2 : n;
  long long g = (work + EW_THREADS * 2 - 1) / (EW_THREADS * 2);
  const long long gmax = (long long)B200_NSM * 16;
  if (g > gmax) g = gmax;
  if (g < 1) g = 1;
  if (v2) ew_kernel<OP, true><<<(int)g, EW_THREADS, 0, g_b200.stream>>>(n, out, x, y, a, b, c);
  else    ew_kernel<OP, false><<<(int)g, EW_THREADS, 0, g_b200.stream>>>(n, out, x, y, a, b, c);
  B200_LAUNCHED(1);
  return 0;
}

extern "C" int b200_vec_set(int n, double *x, double alpha) { return ew_launch<OP_SET>(n, x, 0, 0, alpha, 0, 0); }

extern "C" int b200_vec_copy(int n, const double *x, double *y)
{
  if (x == y) return 0;                       /* VecCopy exits when x == y (vector.c:1667) */
  return ew_launch<OP_COPY>(n, y, x, 0, 0, 0, 0);
}

__global__ void __launch_bounds__(EW_THREADS) swap_kernel(int n, double *__restrict__ x, double *__restrict__ y)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) { double t = x[i]; x[i] = y[i]; y[i] = t; }
}

extern "C" int b200_vec_swap(int n, double *x, double *y)
{
  B200_NEED_INIT();
  if (n <= 0 || x == y) return 0;
  long long g = ((long long)n + EW_THREADS * 4 - 1) / (EW_THREADS * 4);
  if (g > B200_NSM * 16) g = B200_NSM * 16;
  swap_kernel<<<(int)g, EW_THREADS, 0, g_b200.stream>>>(n, x, y);
  B200_LAUNCHED(1);
  return 0;
}

extern "C" int b200_vec_scale(int n, double *x, double alpha)
{
  if (alpha == 0.0) return b200_vec_set(n, x, 0.0);          /* bvec1.c:189 */
  if (alpha == 1.0) return 0;                                /* bvec1.c:191 */
  return ew_launch<OP_SCALE>(n, x, 0, 0, alpha, 0, 0);
}

extern "C" int b200_vec_axpy(int n, double *y, double alpha, const double *x)
{
  if (alpha == 0.0) return 0;                                /* bvec1.c:253: no-op */
  return ew_launch<OP_AXPY>(n, y, x, 0, alpha, 0, 0);
}

extern "C" int b200_vec_aypx(int n, double *y, double alpha, const double *x)
{
  if (alpha == 0.0) return b200_vec_copy(n, x, y);           /* dvec2.c:979 */
  if (alpha == 1.0) return b200_vec_axpy(n, y, 1.0, x);      /* dvec2.c:981 */
  if (alpha == -1.0) return ew_launch<OP_XMY>(n, y, x, 0, 0, 0, 0);
  return ew_launch<OP_AYPX>(n, y, x, 0, alpha, 0, 0);
}

extern "C" int b200_vec_axpby(int n, double *y, double alpha, double beta, const double *x)
{
  if (alpha == 0.0) return b200_vec_scale(n, y, beta);       /* bvec1.c:328-335 */
  if (beta == 1.0) return b200_vec_axpy(n, y, alpha, x);
  if (alpha == 1.0) return b200_vec_aypx(n, y, beta, x);
  if (beta == 0.0) return ew_launch<OP_AX>(n, y, x, 0, alpha, 0, 0);
  return ew_launch<OP_AXPBY>(n, y, x, 0, alpha, beta, 0);
}

extern "C" int b200_vec_waxpy(int n, double *w, double alpha, const double *x, const double *y)
{
  if (alpha == 1.0) return ew_launch<OP_WXPY>(n, w, x, y, 0, 0, 0);
  if (alpha == -1.0) return ew_launch<OP_WYMX>(n, w, x, y, 0, 0, 0);
  if (alpha == 0.0) return b200_vec_copy(n, y, w);
  return ew_launch<OP_WAXPY>(n, w, x, y, alpha, 0, 0);
}

extern "C" int b200_vec_axpbypcz(int n, double *z, double alpha, double beta, double gamma, const double *x, const double *y)
{
  if (alpha == 1.0) return ew_launch<OP_AXPBYPCZ_A1>(n, z, x, y, alpha, beta, gamma);
  if (gamma == 1.0) return ew_launch<OP_AXPBYPCZ_G1>(n, z, x, y, alpha, beta, gamma);
  if (gamma == 0.0) return ew_launch<OP_AXPBYPCZ_G0>(n, z, x, y, alpha, beta, gamma);
  return ew_launch<OP_AXPBYPCZ>(n, z, x, y, alpha, beta, gamma);
}

extern "C" int b200_vec_pointwise_mult(int n, double *w, const double *x, const double *y) { return ew_launch<OP_PMULT>(n, w, x, y, 0, 0, 0); }
extern "C" int b200_vec_pointwise_divide(int n, double *w, const double *x, const double *y) { return ew_launch<OP_PDIV>(n, w, x, y, 0, 0, 0); }
extern "C" int b200_vec_reciprocal(int n, double *x) { return ew_launch<OP_RECIP>(n, x, 0, 0, 0, 0, 0); }

/* ------------------------------------------------------------------------------------------------ VecMAXPY
 * y += sum_j alpha_j x_j in ONE pass over y.  The reference (dvec2.c:836-906) works in groups: first nv&3 vectors,
 * then fours, each group evaluated as  y += ((a0 x0 + a1 x1) + a2 x2) + a3 x3 ; we keep exactly that association
 * per group, so the result is bit-identical, but y is read and written once instead of once per group. */
#define MAXPY_CHUNK 16
struct MaxpyArgs { const double *x[MAXPY_CHUNK]; double a[MAXPY_CHUNK]; };

template <int NV, int G0, bool VEC2>
__global__ void __launch_bounds__(EW_THREADS) maxpy_kernel(int n, double *__restrict__ y, const MaxpyArgs args)
{
  const int stride = gridDim.x * blockDim.x;
  if (VEC2) {
    const int n2 = n >> 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
      double2 xv[NV];
#pragma unroll
      for (int v = 0; v < NV; v++) xv[v] = reinterpret_cast<const double2 *>(args.x[v])[i];
      double2 yv = reinterpret_cast<double2 *>(y)[i];
      int v = 0;
#pragma unroll
      for (int g = 0; g < (NV - G0) / 4 + 1; g++) {
        const int gs = (g == 0) ? G0 : 4;
        double tx = args.a[v] * xv[v].x, ty = args.a[v] * xv[v].y;
#pragma unroll
        for (int k = 1; k < gs; k++) { tx = tx + args.a[v + k] * xv[v + k].x; ty = ty + args.a[v + k] * xv[v + k].y; }
        yv.x = yv.x + tx; yv.y = yv.y + ty;
        v += gs;
      }
      reinterpret_cast<double2 *>(y)[i] = yv;
    }
  }
  /* scalar path, and the odd tail of the vector path */
  const int i0 = VEC2 ? (n & ~1) : 0;
  for (int i = i0 + blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double yv = y[i];
    int v = 0;
#pragma unroll
    for (int g = 0; g < (NV - G0) / 4 + 1; g++) {
      const int gs = (g == 0) ? G0 : 4;
      double t = args.a[v] * args.x[v][i];
#pragma unroll
      for (int k = 1; k < gs; k++) t = t + args.a[v + k] * args.x[v + k][i];
      yv = yv + t;
      v += gs;
    }
    y[i] = yv;
  }
}

template <int NV, int G0>
static void maxpy_launch(int n, double *y, const MaxpyArgs &args, bool v2)
{
  long long work = v2 ? (n + 1) / 2 : n;
  long long g = (work + EW_THREADS - 1) / EW_THREADS;
  if (g > B200_NSM * 8) g = B200_NSM * 8;
  if (v2) maxpy_kernel<NV, G0, true><<<(int)g, EW_THREADS, 0, g_b200.stream>>>(n, y, args);
  else    maxpy_kernel<NV, G0, false><<<(int)g, EW_THREADS, 0, g_b200.stream>>>(n, y, args);
}

template <int NV>
static void maxpy_dispatch(int n, double *y, const MaxpyArgs &args, bool v2, bool first)
{
  /* first chunk: leading group is nv&3 (or 4); later chunks: all groups are fours */
  if (first && (NV & 3)) maxpy_launch<NV, (NV & 3) ? (NV & 3) : 4>(n, y, args, v2);
  else maxpy_launch<NV, ((NV & 3) == 0) ? 4 : (NV & 3)>(n, y, args, v2);
}

extern "C" int b200_vec_maxpy(int n, double *y, int nv, const double *alpha, const double *const *x)
{
  B200_NEED_INIT();
  if (nv < 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "nv = %d", nv);
  if (n <= 0 || nv == 0) return 0;
  if (!y || !alpha || !x) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  bool v2 = aligned16(y);
  for (int v = 0; v < nv; v++) v2 = v2 && aligned16(x[v]);
  int done = 0;
  bool first = true;
  while (done < nv) {
    int rem = nv - done, take;
    if (first) {
      take = (rem <= MAXPY_CHUNK) ? rem : ((nv & 3) ? (nv & 3) + 12 : 16);
    } else take = rem < MAXPY_CHUNK ? rem : MAXPY_CHUNK;
    MaxpyArgs args;
    for (int v = 0; v < take; v++) { args.x[v] = x[done + v]; args.a[v] = alpha[done + v]; }
    switch (take) {
#define MX(N) case N: maxpy_dispatch<N>(n, y, args, v2, first); break;
      MX(1) MX(2) MX(3) MX(4) MX(5) MX(6) MX(7) MX(8) MX(9) MX(10) MX(11) MX(12) MX(13) MX(14) MX(15) MX(16)
#undef MX
    }
    B200_LAUNCHED(1);
    done += take;
    first = false;
  }
  return 0;
}

/* ------------------------------------------------------------------------------------------------ reductions */
__device__ __forceinline__ double comb(double a, double b, int op_max)
{
  if (!op_max) return a + b;
  return (a != a || b != b) ? (a + b) : (a > b ? a : b);   /* NaN propagates like bvec2.c:626 */
}

__device__ __forceinline__ double warp_red(double v, int op_max)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = comb(v, __shfl_xor_sync(0xffffffffu, v, o), op_max);
  return v;
}

/* block-level reduce of NV values, then last-block finalisation in fixed block order */
template <int NV>
__device__ __forceinline__ void block_finish(double (&acc)[NV], double *partials, unsigned *ticket, double *out, int op_max)
{
  __shared__ double s_w[NV][32];
  __shared__ bool s_last;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double r = warp_red(acc[v], op_max);
    if (lane == 0) s_w[v][wid] = r;
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int v = 0; v < NV; v++) {
      double r = (lane < nw) ? s_w[v][lane] : (op_max ? 0.0 : 0.0);
      r = warp_red(r, op_max);
      if (lane == 0) partials[(size_t)blockIdx.x * NV + v] = r;
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned t = atomicAdd(ticket, 1u);
    s_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_last) return;
  /* every CTA has taken its ticket: reset it now, by the thread that writes the results below, so that the reset is
     ordered before them (b200_reduce_finish lets the host go on as soon as it sees the results) */
  if (threadIdx.x == 0) *ticket = 0;
  __threadfence();
#pragma unroll
  for (int v = 0; v < NV; v++) {
    double r = 0.0;
    for (int b = threadIdx.x; b < (int)gridDim.x; b += blockDim.x) r = comb(r, __ldcg(&partials[(size_t)b * NV + v]), op_max);
    r = warp_red(r, op_max);
    __syncthreads();
    if (lane == 0) s_w[v][wid] = r;
    __syncthreads();
    if (wid == 0) {
      double q = (lane < nw) ? s_w[v][lane] : 0.0;
      q = warp_red(q, op_max);
      if (lane == 0) out[v] = q;
    }
  }
}

enum { R_DOT, R_SUMSQ, R_ASUM, R_AMAX, R_ASUM_SUMSQ, R_DOTNORM2 };
template <int K> struct RedN { static constexpr int NV = (K == R_ASUM_SUMSQ || K == R_DOTNORM2) ? 2 : 1; };

template <int K>
__device__ __forceinline__ void red_acc(double (&acc)[RedN<K>::NV], double x, double y)
{
  switch (K) {
  case R_DOT:        acc[0] += x * y; break;
  case R_SUMSQ:      acc[0] += x * x; break;
  case R_ASUM:       acc[0] += fabs(x); break;
  case R_AMAX:       acc[0] = comb(acc[0], fabs(x), 1); break;
  case R_ASUM_SUMSQ: acc[0] += fabs(x); acc[1] += x * x; break;
  case R_DOTNORM2:   acc[0] += x * y; acc[1] += y * y; break;     /* (s.t, t.t): veccusp.cu:1836 */
  }
}

template <int K, bool VEC2>
__global__ void __launch_bounds__(RED_THREADS) red_kernel(int n, const double *__restrict__ x, const double *__restrict__ y,
                                                         double *partials, unsigned *ticket, double *out)
{
  constexpr int NV = RedN<K>::NV;
  constexpr bool RY = (K == R_DOT || K == R_DOTNORM2);
  double acc[NV];
#pragma unroll
  for (int v = 0; v < NV; v++) acc[v] = 0.0;
  const int stride = gridDim.x * blockDim.x;
  if (VEC2) {
    const int n2 = n >> 1;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    /* two independent 128-bit loads per array per trip keep enough bytes in flight */
    for (; i + stride < n2; i += 2 * stride) {
      double2 xa = reinterpret_cast<const double2 *>(x)[i], xb = reinterpret_cast<const double2 *>(x)[i + stride];
      double2 ya = xa, yb = xb;
      if (RY) { ya = reinterpret_cast<const double2 *>(y)[i]; yb = reinterpret_cast<const double2 *>(y)[i + stride]; }
      red_acc<K>(acc, xa.x, ya.x); red_acc<K>(acc, xa.y, ya.y);
      red_acc<K>(acc, xb.x, yb.x); red_acc<K>(acc, xb.y, yb.y);
    }
    if (i < n2) {
      double2 xa = reinterpret_cast<const double2 *>(x)[i];
      double2 ya = xa;
      if (RY) ya = reinterpret_cast<const double2 *>(y)[i];
      red_acc<K>(acc, xa.x, ya.x); red_acc<K>(acc, xa.y, ya.y);
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) red_acc<K>(acc, x[n - 1], RY ? y[n - 1] : x[n - 1]);
  } else {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) red_acc<K>(acc, x[i], RY ? y[i] : x[i]);
  }
  block_finish<NV>(acc, partials, ticket, out, K == R_AMAX);
}

/* Bring nvals reduced values (at g_b200.d_res when single rank, g_b200.d_scal when they still need the cross-rank
 * reduction) to the host.  This is the one synchronisation point of VecDot/VecNorm/VecMDot, as in the reference
 * where the result must be valid on return (cg.c:207-220).
 *
 * Single rank: the last CTA writes the results straight into mapped pinned memory, so the host does not need the
 * stream's completion signal, only the values.  b200_reduce_arm() stamps the slots with a NaN payload no reduction
 * produces BEFORE the launch; b200_reduce_finish() spins on the slots and returns as soon as every one has changed
 * (the kernel's few remaining instructions — the ticket reset — finish behind it in stream order).  Every 1024 spins
 * the stream is queried: a kernel that faulted is reported, and a stream that has completed with a slot still
 * stamped (input that carries the very same payload) falls through to the plain synchronise below, as does
 * everything with B200_REDUCE_POLL=0. */
#define B200_RES_STAMP 0x7ff4b200dead5eedULL           /* signalling-NaN pattern: arithmetic results are quiet NaNs */
static int g_res_armed = 0;
static int g_res_poll  = -1;

void b200_reduce_arm(int nvals)
{
  B200Ctx &c = g_b200;
  if (g_res_poll < 0) { const char *e = getenv("B200_REDUCE_POLL"); g_res_poll = (e && !atoi(e)) ? 0 : 1; }
  g_res_armed = 0;
  if (!g_res_poll || !c.inited || nvals < 1 || nvals > B200_RES_SLOTS) return;
  volatile unsigned long long *h = reinterpret_cast<volatile unsigned long long *>(c.h_res);
  for (int v = 0; v < nvals; v++) h[v] = B200_RES_STAMP;
  g_res_armed = nvals;
}

int b200_reduce_finish(int nvals, double *result_host, bool allreduce, int op_max)
{
  B200Ctx &c = g_b200;
  const int armed = g_res_armed;
  g_res_armed = 0;
  if (allreduce && c.size > 1) {
    B200_CHK(b200_comm_allreduce_dev(c.d_scal, nvals, op_max));
    B200_CUDA(cudaMemcpyAsync(c.h_res, c.d_scal, sizeof(double) * nvals, cudaMemcpyDeviceToHost, c.stream));
  } else if (armed >= nvals && nvals > 0) {
    volatile unsigned long long *h = reinterpret_cast<volatile unsigned long long *>(c.h_res);
    unsigned spins = 0;
    int v = nvals - 1;                                   /* the highest slot is written last */
    while (v >= 0) {
      if (h[v] != B200_RES_STAMP) { v--; continue; }
      if ((++spins & 1023u) == 0) {
        cudaError_t q = cudaStreamQuery(c.stream);
        if (q == cudaSuccess) break;
        if (q != cudaErrorNotReady) return b200_cuda_err(q, "reduction kernel", __FILE__, __LINE__);
      }
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    }
    if (v < 0) {
      for (int k = 0; k < nvals; k++) result_host[k] = c.h_res[k];
      return 0;
    }
  }
  B200_CUDA(cudaStreamSynchronize(c.stream));
  for (int v = 0; v < nvals; v++) result_host[v] = c.h_res[v];
  return 0;
}

/* where a reduction launch writes; call it BEFORE the launch: it stamps the nvals host-visible slots (see above) */
static inline double *red_out(bool allreduce, int nvals)
{
  if (allreduce && g_b200.size > 1) return g_b200.d_scal;
  b200_reduce_arm(nvals);
  return g_b200.d_res;
}

template <int K>
static int red_launch(int n, const double *x, const double *y, double *out)
{
  const bool v2 = aligned16(x) && (!y || aligned16(y));
  /* (capping the grid at ONE wave of resident CTAs — 4 x 148 of 512 lanes — instead of two was measured: VecNorm 37.3 us
     against 38.4 us, VecDot 56.6 against 57.3 us at n = 256^3: within the noise, left alone) */
  const int g = b200_red_grid(v2 ? (n + 1) / 2 : n, RED_THREADS, 4);
  if (v2) red_kernel<K, true><<<g, RED_THREADS, 0, g_b200.stream>>>(n, x, y, g_b200.d_partials, g_b200.d_ticket, out);
  else    red_kernel<K, false><<<g, RED_THREADS, 0, g_b200.stream>>>(n, x, y, g_b200.d_partials, g_b200.d_ticket, out);
  B200_LAUNCHED(1);
  return 0;
}

static int dot_impl(int n, const double *x, const double *y, double *r, bool allreduce)
{
  B200_NEED_INIT();
  if (!r) return b200_seterr(B200_ERR_ARG_NULL, "null result pointer");
  if (n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative vector length");
  if (n > 0 && (!x || !y)) return b200_seterr(B200_ERR_ARG_NULL, "null device vector");
  B200_CHK(red_launch<R_DOT>(n, x, y, red_out(allreduce, 1)));
  return b200_reduce_finish(1, r, allreduce, 0);
}
extern "C" int b200_vec_dot(int n, const double *x, const double *y, double *r) { return dot_impl(n, x, y, r, true); }
extern "C" int b200_vec_dot_local(int n, const double *x, const double *y, double *r) { return dot_impl(n, x, y, r, false); }

static int norm_impl(int n, const double *x, int type, double *r, bool allreduce)
{
  B200_NEED_INIT();
  if (!r) return b200_seterr(B200_ERR_ARG_NULL, "null result pointer");
  if (n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative vector length");
  if (n > 0 && !x) return b200_seterr(B200_ERR_ARG_NULL, "null device vector");
  double *out = red_out(allreduce, 2);
  switch (type) {
  case B200_NORM_2:
  case B200_NORM_FROBENIUS:                       /* sqrt(ddot(x,x)): bvec2.c:617-618; MPI: sum of squares, then sqrt pvec2.c:59-62 */
    B200_CHK(red_launch<R_SUMSQ>(n, x, NULL, out));
    B200_CHK(b200_reduce_finish(1, r, allreduce, 0));
    r[0] = sqrt(r[0]);
    return 0;
  case B200_NORM_1:
    B200_CHK(red_launch<R_ASUM>(n, x, NULL, out));
    return b200_reduce_finish(1, r, allreduce, 0);
  case B200_NORM_INFINITY:
    B200_CHK(red_launch<R_AMAX>(n, x, NULL, out));
    return b200_reduce_finish(1, r, allreduce, 1);
  case B200_NORM_1_AND_2:
    B200_CHK(red_launch<R_ASUM_SUMSQ>(n, x, NULL, out));
    B200_CHK(b200_reduce_finish(2, r, allreduce, 0));
    r[1] = sqrt(r[1]);
    return 0;
  }
  return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "unknown NormType %d", type);
}
extern "C" int b200_vec_norm(int n, const double *x, int type, double *r) { return norm_impl(n, x, type, r, true); }
extern "C" int b200_vec_norm_local(int n, const double *x, int type, double *r) { return norm_impl(n, x, type, r, false); }

extern "C" int b200_vec_dotnorm2(int n, const double *s, const double *t, double *dp, double *nm)
{
  B200_NEED_INIT();
  if (!dp || !nm) return b200_seterr(B200_ERR_ARG_NULL, "null result pointer");
  double r[2];
  B200_CHK(red_launch<R_DOTNORM2>(n, s, t, red_out(true, 2)));
  B200_CHK(b200_reduce_finish(2, r, true, 0));
  *dp = r[0]; *nm = r[1];
  return 0;
}

/* ------------------------------------------------------------------------------------------------ VecMDot
 * z_j = x . y_j for up to 16 vectors per launch: x is read once per launch, each y_j streams from its own buffer
 * (GMRES basis vectors are separate allocations, gmres.c:415-439), NV accumulators per thread. */
#define MDOT_CHUNK 16
struct MdotArgs { const double *y[MDOT_CHUNK]; };

/* 256 threads and at least 2 CTAs per SM: up to 128 registers per thread, so that ALL NV + 1 loads of an iteration are
 * in flight together.  (With 512-thread CTAs ptxas capped the kernel at 56-64 registers and, for some NV, issued the loads
 * one at a time through a single register pair — NV = 14 ran at 4.6 TB/s against 6.6 TB/s for NV = 8.) */
#define MDOT_THREADS 256
template <int NV, bool VEC2>
__global__ void __launch_bounds__(MDOT_THREADS, 2) mdot_kernel(int n, const double *__restrict__ x, const MdotArgs args,
                                                              double *partials, unsigned *ticket, double *out)
{
  double acc[NV];
#pragma unroll
  for (int v = 0; v < NV; v++) acc[v] = 0.0;
  const int stride = gridDim.x * blockDim.x;
  if (VEC2) {
    const int n2 = n >> 1;
#pragma unroll 1
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) {
      double2 yv[NV];
#pragma unroll
      for (int v = 0; v < NV; v++) yv[v] = __ldcs(reinterpret_cast<const double2 *>(args.y[v]) + i);
      const double2 xv = __ldcs(reinterpret_cast<const double2 *>(x) + i);
#pragma unroll
      for (int v = 0; v < NV; v++) { acc[v] += xv.x * yv[v].x; acc[v] += xv.y * yv[v].y; }
    }
    if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
#pragma unroll
      for (int v = 0; v < NV; v++) acc[v] += x[n - 1] * args.y[v][n - 1];
    }
  } else {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      const double xv = x[i];
#pragma unroll
      for (int v = 0; v < NV; v++) acc[v] += xv * args.y[v][i];
    }
  }
  block_finish<NV>(acc, partials, ticket, out, 0);
}

template <int NV>
static void mdot_launch(int n, const double *x, const MdotArgs &args, bool v2, double *out)
{
  const int g = b200_red_grid(v2 ? (n + 1) / 2 : n, MDOT_THREADS, 2);
  if (v2) mdot_kernel<NV, true><<<g, MDOT_THREADS, 0, g_b200.stream>>>(n, x, args, g_b200.d_partials, g_b200.d_ticket, out);
  else    mdot_kernel<NV, false><<<g, MDOT_THREADS, 0, g_b200.stream>>>(n, x, args, g_b200.d_partials, g_b200.d_ticket, out);
}

static int mdot_impl(int n, const double *x, int nv, const double *const *y, double *z, bool allreduce)
{
  B200_NEED_INIT();
  if (nv < 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "nv = %d", nv);
  if (nv == 0) return 0;
  if (nv > B200_RES_SLOTS) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "nv = %d exceeds %d", nv, B200_RES_SLOTS);
  if (!z || !y || (n > 0 && !x)) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  bool v2 = aligned16(x);
  for (int v = 0; v < nv; v++) v2 = v2 && aligned16(y[v]);
  double *out = red_out(allreduce, nv);
  for (int done = 0; done < nv;) {
    const int take = (nv - done < MDOT_CHUNK) ? nv - done : MDOT_CHUNK;
    MdotArgs args;
    for (int v = 0; v < take; v++) args.y[v] = y[done + v];
    switch (take) {
#define MD(N) case N: mdot_launch<N>(n, x, args, v2, out + done); break;
      MD(1) MD(2) MD(3) MD(4) MD(5) MD(6) MD(7) MD(8) MD(9) MD(10) MD(11) MD(12) MD(13) MD(14) MD(15) MD(16)
#undef MD
    }
    B200_LAUNCHED(1);
    done += take;
  }
  return b200_reduce_finish(nv, z, allreduce, 0);
}
extern "C" int b200_vec_mdot(int n, const double *x, int nv, const double *const *y, double *z) { return mdot_impl(n, x, nv, y, z, true); }
extern "C" int b200_vec_mdot_local(int n, const double *x, int nv, const double *const *y, double *z) { return mdot_impl(n, x, nv, y, z, false); }

/* ------------------------------------------------------------------------------------------------ VecMDot, no allocation
 * b200_vec_mdot_noalloc: the mdot_kernel above for callers that may not allocate (CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * It does not need b200_init (which allocates) and calls no cudaMalloc/cudaFree or allocating library routine:
 *  - the per-CTA partial sums and the last-CTA ticket live in the file-scope device array below, which every GPU
 *    that loads the library holds until exit: 256 B header + B200_NOALLOC_MAX_CTAS x B200_NOALLOC_MAX_NV doubles
 *    (148.25 KB);
 *  - the y pointers travel in the kernel's argument block (MdotArgs), the nv results are written to a DEVICE buffer
 *    the caller passes, on the stream the caller passes; nothing is read back and nothing synchronises.
 * Limits: nv <= B200_NOALLOC_MAX_NV (one launch), n <= B200_NOALLOC_MAX_N, reduction grid <= B200_NOALLOC_MAX_CTAS.
 * Every call on a device shares the array: the ticket word is cleared on the stream before each launch (a call that
 * died half way would otherwise leave it counting), the partials are fully rewritten by each launch before they are
 * read.  Two calls that overlap on the GPU (different streams, no ordering between them) race on it. */
#define B200_NOALLOC_MAX_NV    MDOT_CHUNK
#define B200_NOALLOC_MAX_CTAS  B200_RED_MAXBLK
#define B200_NOALLOC_MAX_N     (1 << 30)
#define B200_NOALLOC_HEADER    32                 /* doubles: 256 B holding the ticket word                                */
#define B200_NOALLOC_MAXDEV    64

static __device__ __align__(256) double b200_noalloc_scratch[B200_NOALLOC_HEADER + B200_NOALLOC_MAX_CTAS * B200_NOALLOC_MAX_NV];

/* In host code the array's name is only a stand-in for the symbol, not a device address: each device's address is
 * looked up once, with that device current, and that pointer is what the kernel is given. */
static double *g_noalloc_addr[B200_NOALLOC_MAXDEV];
static std::mutex g_noalloc_mutex;

static int noalloc_scratch(double **ptr)
{
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= B200_NOALLOC_MAXDEV) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "device ordinal %d outside 0..%d", dev, B200_NOALLOC_MAXDEV - 1);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (!g_noalloc_addr[dev]) {
    double *p = NULL;
    B200_CUDA(cudaGetSymbolAddress((void **)&p, b200_noalloc_scratch));
    if (!p) return b200_seterr(B200_ERR_LIB, "cudaGetSymbolAddress gave a null address on device %d", dev);
    g_noalloc_addr[dev] = p;
  }
  *ptr = g_noalloc_addr[dev];
  return 0;
}

template <int NV>
static void mdot_launch_noalloc(int g, int n, const double *x, const MdotArgs &args, bool v2, double *scratch, double *out, cudaStream_t st)
{
  double   *partials = scratch + B200_NOALLOC_HEADER;
  unsigned *ticket   = reinterpret_cast<unsigned *>(scratch);
  if (v2) mdot_kernel<NV, true><<<g, MDOT_THREADS, 0, st>>>(n, x, args, partials, ticket, out);
  else    mdot_kernel<NV, false><<<g, MDOT_THREADS, 0, st>>>(n, x, args, partials, ticket, out);
}

extern "C" int b200_vec_mdot_noalloc(int n, const double *x, int nv, const double *const *y, double *z_dev, void *cuda_stream)
{
  if (nv < 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "nv = %d", nv);
  if (nv > B200_NOALLOC_MAX_NV)
    return b200_seterr(B200_ERR_ARG_SIZ, "b200_vec_mdot_noalloc: nv = %d exceeds the %d vectors its static scratch holds (use b200_vec_mdot)", nv, B200_NOALLOC_MAX_NV);
  if (n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative vector length");
  if (n > B200_NOALLOC_MAX_N)
    return b200_seterr(B200_ERR_ARG_SIZ, "b200_vec_mdot_noalloc: n = %d exceeds the maximum %d (use b200_vec_mdot)", n, B200_NOALLOC_MAX_N);
  if (nv == 0) return 0;
  if (!z_dev || !y || (n > 0 && !x)) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  bool v2 = aligned16(x);
  MdotArgs args;
  for (int v = 0; v < nv; v++) {
    if (n > 0 && !y[v]) return b200_seterr(B200_ERR_ARG_NULL, "null device vector y[%d]", v);
    args.y[v] = y[v];
    v2 = v2 && aligned16(y[v]);
  }
  for (int v = nv; v < MDOT_CHUNK; v++) args.y[v] = NULL;
  const int g = b200_red_grid(v2 ? (n + 1) / 2 : n, MDOT_THREADS, 2);       /* the grid b200_vec_mdot uses: same sums, same order */
  if (g > B200_NOALLOC_MAX_CTAS)
    return b200_seterr(B200_ERR_ARG_SIZ, "b200_vec_mdot_noalloc: %d CTAs exceed the %d its static scratch holds", g, B200_NOALLOC_MAX_CTAS);
  double *scratch = NULL;
  B200_CHK(noalloc_scratch(&scratch));
  cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : (g_b200.inited ? g_b200.stream : (cudaStream_t)0);
  B200_CUDA(cudaMemsetAsync(scratch, 0, B200_NOALLOC_HEADER * sizeof(double), st));
  switch (nv) {
#define MD(N) case N: mdot_launch_noalloc<N>(g, n, x, args, v2, scratch, z_dev, st); break;
    MD(1) MD(2) MD(3) MD(4) MD(5) MD(6) MD(7) MD(8) MD(9) MD(10) MD(11) MD(12) MD(13) MD(14) MD(15) MD(16)
#undef MD
  }
  B200_LAUNCHED(1);
  return 0;
}
