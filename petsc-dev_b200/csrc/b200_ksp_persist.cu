/* b200_ksp_persist.cu — KSPSolve_CG (src/ksp/ksp/impls/cg/cg.c:181-281) as ONE persistent cooperative kernel per block of
 * iterations, for problems whose iteration is shorter than the kernel boundaries around it.
 *
 * An experiment kept as an option (B200_CG_PERSISTENT=1), not the default: see b200_cg_persistent_ok for the numbers.
 * The grid stays resident for `iters` iterations and the phases are separated by grid barriers:
 *
 *   A  p = z + b p                                   | barrier
 *   B  w = A p (TMA-fed row tiles, one lane group per row, partial p.w per CTA)
 *                                                    | barrier; the LAST CTA to arrive sums the partials in fixed order and
 *                                                    |          runs cg.c:207-220 (dpi, breakdown tests, a = beta/dpi)
 *   C  x += a p ; r -= a w ; z = r .* d ; partial z.z, z.r, r.r
 *                                                    | barrier; the last CTA sums the partials and runs cg.c:227-281 + the top of
 *                                                    |          the next iteration (norm, history, KSPDefaultConverged, beta, b)
 *
 * Same arithmetic, per element and per row, as the multi-kernel fused driver (b200_ksp.cu): one lane per row sums left to
 * right like PetscSparseDensePlusDot (aij.h:384-386); the reductions are fixed-order trees.  The scalar recurrences live in
 * the CgState struct in global memory; every CTA re-reads what it needs after each barrier (ld.global.cg).
 */
#include "b200_internal.h"
#include "b200_spmv_kernels.cuh"
#include "b200_ksp_state.h"
#include <math.h>

struct CgPersistArgs {
  int           n, ntiles, iters, jacobi, keep_l2;
  const int4   *tiles;
  const int    *ai, *aj;
  const double *aa;
  double       *x, *r, *zw, *p;      /* zw: Z and W share one vector (cg.c:121) */
  const double *d;
  CgState      *s;
  double       *dotpart, *partials, *hist;
  unsigned     *bar;                 /* [0] arrival count, [1] generation */
};

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned *p)
{
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(unsigned *p, unsigned v)
{
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

/* Grid barrier with a serial section: returns true in the CTA that arrived last, BEFORE the others are released; that
 * CTA must call grid_release() when its serial work is done.  All CTAs are resident (cooperative launch). */
__device__ __forceinline__ bool grid_arrive(unsigned *bar, unsigned nblocks, unsigned *gen_out)
{
  __shared__ unsigned s_gen, s_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned g = ld_acquire_gpu_u32(&bar[1]);
    __threadfence();
    const unsigned t = atomicAdd(&bar[0], 1u);
    s_gen = g; s_last = (t == nblocks - 1u);
    if (s_last) __threadfence();
  }
  __syncthreads();
  *gen_out = s_gen;
  return s_last != 0u;
}
__device__ __forceinline__ void grid_release(unsigned *bar, unsigned gen)
{
  __syncthreads();
  if (threadIdx.x == 0) { bar[0] = 0u; __threadfence(); st_release_gpu_u32(&bar[1], gen + 1u); }
}
__device__ __forceinline__ void grid_wait(const unsigned *bar, unsigned gen)
{
  if (threadIdx.x == 0) { while (ld_acquire_gpu_u32(&bar[1]) == gen) { } }
  __syncthreads();
}

/* fixed-order sum of v[0], v[stride], ... (n values) by the whole CTA; result valid in thread 0 */
template <int THREADS>
__device__ __forceinline__ double cta_sum_strided(const double *v, int n, int stride, double *s_w)
{
  double t = 0.0;
  for (int i = threadIdx.x; i < n; i += THREADS) t += __ldcg(&v[(size_t)i * stride]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = t;
  __syncthreads();
  double u = 0.0;
  if (threadIdx.x == 0) for (int q = 0; q < THREADS / 32; q++) u += s_w[q];
  return u;
}

/* three interleaved sums v[3 i + q] at once (same fixed order per sum as cta_sum_strided); results valid in thread 0 */
template <int THREADS>
__device__ __forceinline__ void cta_sum3(const double *v, int n, double (*s_w)[THREADS / 32], double (&out)[3])
{
  double t[3] = {0.0, 0.0, 0.0};
  for (int i = threadIdx.x; i < n; i += THREADS) {
#pragma unroll
    for (int q = 0; q < 3; q++) t[q] += __ldcg(&v[(size_t)i * 3 + q]);
  }
#pragma unroll
  for (int q = 0; q < 3; q++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t[q] += __shfl_xor_sync(0xffffffffu, t[q], o);
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { s_w[0][threadIdx.x >> 5] = t[0]; s_w[1][threadIdx.x >> 5] = t[1]; s_w[2][threadIdx.x >> 5] = t[2]; }
  __syncthreads();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int q = 0; q < 3; q++) { double u = 0.0; for (int w = 0; w < THREADS / 32; w++) u += s_w[q][w]; out[q] = u; }
  }
}

template <int G, int THREADS, int TILE_NNZ, int STAGES>
__global__ void __launch_bounds__(THREADS)
cg_persistent_kernel(const CgPersistArgs a)
{
  constexpr int SE = TILE_NNZ + SPMV_ALIGN_PAD;
  extern __shared__ __align__(128) unsigned char spmv_smem[];
  double   *s_a  = reinterpret_cast<double *>(spmv_smem);
  int      *s_j  = reinterpret_cast<int *>(spmv_smem + (size_t)STAGES * SE * 8);
  uint64_t *bars = reinterpret_cast<uint64_t *>(spmv_smem + (size_t)STAGES * SE * 12);
  __shared__ double s_w[3][THREADS / 32];

  const int tid = threadIdx.x, r = tid / G, g = tid % G;
  const int nblk = gridDim.x, stride = gridDim.x;
  const int n = a.n, n2 = n >> 1;
  uint64_t pol = 0;
  uint32_t nwait[STAGES];                            /* tiles consumed so far from each stage: mbarrier phase parity */
#pragma unroll
  for (int q = 0; q < STAGES; q++) nwait[q] = 0;
  if (tid == 0) {
#pragma unroll
    for (int q = 0; q < STAGES; q++) b200_mbar_init(&bars[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    /* a matrix that fits in L2 next to the vectors is re-read from there every iteration: do not mark it evict-first */
    if (a.keep_l2) asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol));
    else pol = b200_policy_evict_first();
  }
  __syncthreads();
  auto issue = [&](int q, int4 d) {                 /* thread 0 only */
    const int ka = d.z & ~3, kb = (d.w + 3) & ~3;
    const uint32_t nn = (uint32_t)(kb - ka);
    if (nn == 0) { b200_mbar_expect_tx(&bars[q], 0); return; }
    b200_mbar_expect_tx(&bars[q], nn * 12u);
    b200_bulk_g2s(s_a + (size_t)q * SE, a.aa + ka, nn * 8u, &bars[q], pol);
    b200_bulk_g2s(s_j + (size_t)q * SE, a.aj + ka, nn * 4u, &bars[q], pol);
  };
  CgState *s = a.s;
  unsigned gen;

  for (int iter = 0; iter < a.iters; iter++) {
    if (__ldcg(&s->reason) != 0) break;              /* uniform: every CTA reads the state the last barrier published */
    /* ---- A: p = z + b p (VecCopy at i == 0, VecAYPX after: cg.c:194,202) ---- */
    {
      const bool first = (__ldcg(&s->i) == 0);
      const double b = __ldcg(&s->b);
      const double2 *z2 = reinterpret_cast<const double2 *>(a.zw);
      double2 *p2 = reinterpret_cast<double2 *>(a.p);
      for (int i = blockIdx.x * THREADS + tid; i < n2; i += nblk * THREADS) {
        const double2 zv = __ldcg(&z2[i]);
        double2 pv;
        if (first) pv = zv;
        else {
          pv = p2[i];
          if (b == 1.0) { pv.x = pv.x + zv.x; pv.y = pv.y + zv.y; }
          else if (b == -1.0) { pv.x = zv.x - pv.x; pv.y = zv.y - pv.y; }
          else { pv.x = zv.x + b * pv.x; pv.y = zv.y + b * pv.y; }
        }
        p2[i] = pv;
      }
      if ((n & 1) && blockIdx.x == 0 && tid == 0) {
        const int i = n - 1;
        if (first) a.p[i] = a.zw[i];
        else if (b == 1.0) a.p[i] = a.p[i] + a.zw[i];
        else if (b == -1.0) a.p[i] = a.zw[i] - a.p[i];
        else a.p[i] = a.zw[i] + b * a.p[i];
      }
    }
    if (grid_arrive(a.bar, nblk, &gen)) grid_release(a.bar, gen); else grid_wait(a.bar, gen);

    /* ---- B: w = A p, partial p.w ---- */
    {
      double dot = 0.0;
      int t = blockIdx.x;
      if (tid == 0) {
#pragma unroll
        for (int q = 0; q < STAGES; q++) if (t + q * stride < a.ntiles) issue(q, __ldg(&a.tiles[t + q * stride]));
      }
      for (int it = 0; t < a.ntiles; t += stride, it++) {
        const int q = it % STAGES;
        const int4 dsc = __ldg(&a.tiles[t]);
        int b0 = 0, e0 = 0, row = -1;
        if (r < dsc.y - dsc.x) { row = dsc.x + r; b0 = __ldg(&a.ai[row]); e0 = __ldg(&a.ai[row + 1]); }
        uint32_t par = 0;
#pragma unroll
        for (int u = 0; u < STAGES; u++) if (u == q) { par = nwait[u] & 1u; nwait[u]++; }
        b200_mbar_wait(&bars[q], par);
        const int ka = dsc.z & ~3;
        const double *sa = s_a + (size_t)q * SE - ka;
        const int    *sj = s_j + (size_t)q * SE - ka;
        double sum = 0.0;
        if (G == 1) {
#pragma unroll 4
          for (int k = b0; k < e0; k++) sum += sa[k] * __ldcg(&a.p[sj[k]]);
        } else {
#pragma unroll 4
          for (int k = b0 + g; k < e0; k += G) sum += sa[k] * __ldcg(&a.p[sj[k]]);
          __syncwarp();
#pragma unroll
          for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
        }
        if (g == 0 && row >= 0) { a.zw[row] = sum; dot += __ldcg(&a.p[row]) * sum; }
        __syncthreads();
        if (tid == 0 && t + STAGES * stride < a.ntiles) issue(q, __ldg(&a.tiles[t + STAGES * stride]));
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
      if ((tid & 31) == 0) s_w[0][tid >> 5] = dot;
      __syncthreads();
      if (tid == 0) { double u = 0.0; for (int q = 0; q < THREADS / 32; q++) u += s_w[0][q]; a.dotpart[blockIdx.x] = u; }
    }
    if (grid_arrive(a.bar, nblk, &gen)) {
      const double pw = cta_sum_strided<THREADS>(a.dotpart, nblk, 1, s_w[0]);
      if (tid == 0) {                                 /* cg.c:207-220 */
        s->red[3] = pw;
        s->dpi = pw;
        s->betaold = s->beta;
        if (isnan(pw) || isinf(pw)) s->reason = KSP_ERR_FP_REASON;
        else if (pw == 0.0 || (s->i > 0 && pw * s->dpiold <= 0.0)) s->reason = B200_KSP_DIVERGED_INDEFINITE_MAT;
        else s->a = s->beta / pw;
      }
      grid_release(a.bar, gen);
    } else grid_wait(a.bar, gen);
    if (__ldcg(&s->reason) != 0) break;

    /* ---- C: x += a p ; r -= a w ; z = r .* d ; partial z.z, z.r, r.r ---- */
    {
      const double al = __ldcg(&s->a), ma = -al;
      double zz = 0.0, zr = 0.0, rr = 0.0;
      double2 *x2 = reinterpret_cast<double2 *>(a.x), *r2 = reinterpret_cast<double2 *>(a.r), *w2 = reinterpret_cast<double2 *>(a.zw);
      const double2 *p2 = reinterpret_cast<const double2 *>(a.p), *d2 = reinterpret_cast<const double2 *>(a.d);
      for (int i = blockIdx.x * THREADS + tid; i < n2; i += nblk * THREADS) {
        double2 xv = x2[i], rv = r2[i], zv;
        const double2 pv = p2[i], wv = __ldcg(&w2[i]);   /* w was written by other CTAs in phase B */
        xv.x = xv.x + al * pv.x; xv.y = xv.y + al * pv.y;
        rv.x = rv.x + ma * wv.x; rv.y = rv.y + ma * wv.y;
        if (a.jacobi) { const double2 dv = d2[i]; zv.x = rv.x * dv.x; zv.y = rv.y * dv.y; } else zv = rv;
        x2[i] = xv; r2[i] = rv; w2[i] = zv;
        zz += zv.x * zv.x; zz += zv.y * zv.y;
        zr += zv.x * rv.x; zr += zv.y * rv.y;
        rr += rv.x * rv.x; rr += rv.y * rv.y;
      }
      if ((n & 1) && blockIdx.x == 0 && tid == 0) {
        const int i = n - 1;
        const double xv = a.x[i] + al * a.p[i], rv = a.r[i] + ma * a.zw[i], zv = a.jacobi ? rv * a.d[i] : rv;
        a.x[i] = xv; a.r[i] = rv; a.zw[i] = zv;
        zz += zv * zv; zr += zv * rv; rr += rv * rv;
      }
      double v[3] = {zz, zr, rr};
#pragma unroll
      for (int q = 0; q < 3; q++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
        if ((tid & 31) == 0) s_w[q][tid >> 5] = v[q];
      }
      __syncthreads();
      if (tid < 3) { double u = 0.0; for (int q = 0; q < THREADS / 32; q++) u += s_w[tid][q]; a.partials[(size_t)blockIdx.x * 3 + tid] = u; }
    }
    if (grid_arrive(a.bar, nblk, &gen)) {
      double red[3];
      cta_sum3<THREADS>(a.partials, nblk, s_w, red);     /* one pass for the three sums: the other CTAs are waiting */
      if (tid == 0) {
        s->red[0] = red[0]; s->red[1] = red[1]; s->red[2] = red[2];
        cg_after_update_body(s, a.hist);              /* cg.c:227-281 and the top of the next iteration */
      }
      grid_release(a.bar, gen);
    } else grid_wait(a.bar, gen);
  }
}

/* ------------------------------------------------------------------------------------------------ host */
typedef void (*persist_kern_t)(const CgPersistArgs);
#define PERSIST_STAGES 2                               /* the next tile streams in while this one is summed */
static persist_kern_t persist_kernel(int G)
{
  switch (G) {
    case 1:  return cg_persistent_kernel<1, 256, 2048, PERSIST_STAGES>;
    case 2:  return cg_persistent_kernel<2, 256, 2048, PERSIST_STAGES>;
    case 4:  return cg_persistent_kernel<4, 256, 2048, PERSIST_STAGES>;
    case 8:  return cg_persistent_kernel<8, 256, 2048, PERSIST_STAGES>;
    case 16: return cg_persistent_kernel<16, 256, 2048, PERSIST_STAGES>;
    default: return cg_persistent_kernel<32, 256, 2048, PERSIST_STAGES>;
  }
}

/* may this solve run as the persistent kernel?  (one rank, one sequential block with the default tile shape, no long rows) */
bool b200_cg_persistent_ok(b200_mat A)
{
  const char *e = getenv("B200_CG_PERSISTENT");
  if (e && !atoi(e)) return false;
  const B200Seq *S = &A->A;
  if (g_b200.size > 1 || A->kind != 0 || S->m != S->n || S->m == 0 || S->ntiles == 0) return false;
  if (S->compressed || S->nstrips > 1 || S->nlong || S->cfg != 0 || S->park) return false;
  int coop = 0;
  if (cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, g_b200.device) != cudaSuccess || !coop) return false;
  if (e && atoi(e)) return true;
  /* Opt-in only (B200_CG_PERSISTENT=1).  Measured on ex2's 5-point Laplacian, m = n = 100 ... 1000 (profiles/
   * r02_c1_sweep.txt): 23.6 / 24.0 / 27.3 / 42.0 us per iteration at 10^4 / 9 10^4 / 2.5 10^5 / 10^6 rows against 22.7 / 22.8 /
   * 25.4 / 41.1 us with four launches — three grid-wide barriers with a serial scalar section each cost what the four
   * launches cost, at every size.  What the small problems were short of was host launch rate, and the CUDA-graph replay
   * of the four-launch iteration (b200_ksp.cu) fixes that without the barriers. */
  return false;
}

static unsigned *g_bar = NULL;

/* run up to `iters` iterations; ds = device CgState already primed by cg_top (b200_ksp.cu) */
int b200_cg_persistent_run(b200_mat A, int jacobi, const double *d, double *X, double *R, double *ZW, double *P, CgState *ds, double *dhist, int iters)
{
  B200Ctx &c = g_b200;
  const B200Seq *S = &A->A;
  persist_kern_t k = persist_kernel(S->G);
  const size_t smem = SpmvSmem<256, 2048, PERSIST_STAGES>::BYTES;
  static int occ_cache[33] = {0};
  if (!occ_cache[S->G]) {
    B200_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int o = 0;
    B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, k, 256, smem));
    if (o < 1) return b200_seterr(B200_ERR_LIB, "persistent CG kernel does not fit on an SM");
    int cap = 4;                                      /* 4 x 256 lanes per SM: fewer CTAs make cheaper barriers */
    const char *eo = getenv("B200_CG_PERSIST_OCC");
    if (eo && atoi(eo) > 0) cap = atoi(eo);
    occ_cache[S->G] = o > cap ? cap : o;
  }
  int grid = c.nsm * occ_cache[S->G];
  if (!g_bar) {
    B200_CUDA(cudaMalloc((void **)&g_bar, 64));
    B200_CUDA(cudaMemset(g_bar, 0, 64));
  }
  CgPersistArgs a;
  memset(&a, 0, sizeof(a));
  a.n = S->m; a.ntiles = S->ntiles; a.iters = iters; a.jacobi = jacobi;
  a.keep_l2 = (12.0 * (double)S->nz + 48.0 * (double)S->m) <= 100e6;
  { const char *ek = getenv("B200_CG_PERSIST_KEEPL2"); if (ek) a.keep_l2 = atoi(ek) != 0; }
  a.tiles = S->d_tiles; a.ai = S->d_i; a.aj = S->d_j; a.aa = S->d_a;
  a.x = X; a.r = R; a.zw = ZW; a.p = P; a.d = d; a.s = ds;
  a.dotpart = c.d_partials; a.partials = c.d_partials + B200_RED_MAXBLK; a.hist = dhist; a.bar = g_bar;
  if (grid > B200_RED_MAXBLK) grid = B200_RED_MAXBLK;
  void *params[] = {(void *)&a};
  B200_CUDA(cudaLaunchCooperativeKernel((const void *)k, dim3(grid), dim3(256), params, smem, c.stream));
  B200_LAUNCHED(1);
  return 0;
}
