/* b200_spmv.cu — AIJ (CSR) MatMult / MatMultAdd for sm_100a.
 *
 * Replaces MatMult_SeqAIJ / MatMultAdd_SeqAIJ (src/mat/impls/aij/seq/aij.c:1225-1358) and the cusp::multiply call of
 * MatMult_SeqAIJCUSP (src/mat/impls/aij/seq/seqcusp/aijcusp.cu:349-390).
 *
 * Row binning (done once at upload, the device analogue of MatAssemblyEnd_SeqAIJ's compressed-row/inode checks):
 *   - "row tiles": maximal runs of consecutive rows holding <= tile_nnz entries and <= threads/G rows, walked by the
 *     persistent TMA-fed tile kernel (b200_spmv_kernels.cuh).  G lanes per row, chosen from the mean row length: G = 1
 *     (short rows, e.g. the 5/7-point stencils) sums each row left to right exactly like the reference, so with
 *     -fmad=false those rows are bit-identical to the CPU result; G > 1 uses a fixed shuffle tree.
 *   - "long rows" (> long_thresh entries): one CTA per row, strided loads, fixed-shape tree reduction
 *     (deterministic; differs from the reference in association only).
 *   - blocks that are mostly empty rows (the MPIAIJ off-diagonal block) are visited through the list of non-empty
 *     rows only (what Mat_CompressedRow does on the CPU: src/mat/utils/compressedrow.c:28-80).
 * Algorithmic traffic per MatMult: 12 B/nnz + 4 B/row (offsets) + 8 B/row (y) + 8 B/col (x once).
 */
#include "b200_internal.h"
#include <vector>
#include <thread>
#include <functional>
#include <algorithm>
#include <math.h>

#include "b200_spmv_kernels.cuh"


/* launch shapes of the tile kernel; the active one is chosen at upload (default 0, B200_SPMV_CFG overrides — a tuning
 * aid, every shape computes the same numbers) */
struct SpmvCfg { int threads, tile_nnz, stages, ctas_per_sm; };
static const SpmvCfg g_cfgs[] = {
  {256, 2048, 1, 8},      /* default: 2048 resident lanes per SM; 8 single-stage CTAs overlap each other's loads     */
  {256, 2048, 2, 4},
  {512, 4096, 2, 2},
  {128, 1024, 1, 16},
  {512, 4096, 1, 4},
  {256, 1792, 2, 5},
  {256, 4096, 2, 2},
  {384, 3072, 1, 5},
};
#define SPMV_NCFG ((int)(sizeof(g_cfgs) / sizeof(g_cfgs[0])))
static int g_tune_cfg = -1, g_tune_G = -1, g_tune_long = -1, g_tune_park = -1;

/* deterministic final pass over the per-CTA partials of the fused dot */
__global__ void __launch_bounds__(1024) dot_final_kernel(int n, const double *__restrict__ part, int n2, const double *__restrict__ part2, double *out)
{
  __shared__ double s_red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) s += part[i];
  for (int i = threadIdx.x; i < n2; i += 1024) s += part2[i];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) s_red[wid] = s;
  __syncthreads();
  if (wid == 0) {
    double v = s_red[lane];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) out[0] = v;
  }
}

__global__ void zero_rows_kernel(int n, double *y)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) y[i] = 0.0;
}

/* later calls: d[i] = a[diagpos[i]] (MatGetDiagonal_SeqAIJ with the marks of MatMarkDiagonal_SeqAIJ, aij.c:1040-1073,1486) */
__global__ void __launch_bounds__(256) diag_gather_kernel(int r0, int m, const int *__restrict__ pos, const double *__restrict__ aa, double *__restrict__ d)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = r0 + blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) { const int p = __ldg(&pos[i]); d[i] = p >= 0 ? __ldg(&aa[p]) : 0.0; }
}

/* d[i] = a_ii or 0 (MatGetDiagonal_SeqAIJ aij.c:1040-1073); columns ascending -> bisection; the position found is kept */
__global__ void diag_kernel(int r0, int m, const int *__restrict__ ai, const int *__restrict__ aj, const double *__restrict__ aa,
                            int coloff, double *__restrict__ d, int *__restrict__ pos)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = r0 + blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
    int lo = ai[i], hi = ai[i + 1];
    const int want = i + coloff;
    double v = 0.0;
    int at = -1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const int c = aj[mid];
      if (c == want) { v = aa[mid]; at = mid; break; }
      if (c < want) lo = mid + 1; else hi = mid;
    }
    d[i] = v;
    pos[i] = at;
  }
}

/* ------------------------------------------------------------------------------------------------ host side */
void b200_seq_free(B200Seq *S)
{
  if (!S) return;
  if (S->T) { b200_seq_free(S->T); free(S->T); S->T = NULL; }
  if (S->strips) {
    for (int k = 0; k < S->nstrips; k++) b200_seq_free(&S->strips[k]);
    free(S->strips); free(S->strip_nzoff);
  }
  if (S->d_ci && S->d_ci != S->d_i) cudaFree(S->d_ci);
  cudaFree(S->d_i); cudaFree(S->d_j); cudaFree(S->d_a); cudaFree(S->d_rindex);
  cudaFree(S->d_wtiles); cudaFree(S->d_wlong); cudaFree(S->d_wlongstart); cudaFree(S->d_wlongpart);
  cudaFree(S->d_ticket); cudaFree(S->d_diagpos); cudaFree(S->d_tiles); cudaFree(S->d_long); cudaFree(S->d_longstart); cudaFree(S->d_longpart); cudaFree(S->d_dotpart);
  memset(S, 0, sizeof(*S));
}

/* i is a host array.  j / a are host arrays that are copied, unless adopt: then they are device arrays of nz + 8
 * entries (cudaMalloc'ed by the caller, e.g. the device assembly) whose ownership passes to S, and d_i_adopt — the
 * device copy of i the caller already has, m + 1 entries — is adopted as well instead of being uploaded again. */
static int g_park_cap = 32;

extern "C" int b200_spmv_set_park_cap(int cap)
{
  g_park_cap = cap;
  return 0;
}

static int seq_upload_one(B200Seq *S, int m, int n, const int *i, const int *j, const double *a, bool adopt = false, int *d_i_adopt = NULL)
{
  memset(S, 0, sizeof(*S));
  {
    static bool gather_mode_set = false;
    if (!gather_mode_set) {
      const char *e = getenv("B200_SPMV_GATHER_CG");
      const int v = e ? atoi(e) : 1;
      B200_CUDA(cudaMemcpyToSymbol(b200_c_gather_cg, &v, sizeof(int)));
      gather_mode_set = true;
    }
  }
  if (adopt) { S->d_j = const_cast<int *>(j); S->d_a = const_cast<double *>(a); S->d_i = d_i_adopt; }   /* owned from here on, also on failure */
  if (m < 0 || n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative matrix size %d x %d", m, n);
  if (!i || i[0] != 0) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets must start at 0");
  const long long nz = i[m];
  if (nz < 0) return b200_seterr(B200_ERR_ARG_WRONG, "negative nonzero count");
  if ((nz > 0 || adopt) && (!j || !a)) return b200_seterr(B200_ERR_ARG_NULL, "null column/value arrays");
  S->m = m; S->n = n; S->nz = nz; S->req_G = g_tune_G;
  /* row statistics + the visiting order */
  /* one pass: validity, non-empty rows, longest row, sum of squared lengths (integer accumulators: a floating-point
   * running sum makes this loop latency-bound, and it runs at every upload) */
  int nonempty = 0, maxlen = 0, bad = 0;
  unsigned long long sumsq = 0;
  for (int r = 0; r < m; r++) {
    const int l = i[r + 1] - i[r];
    bad |= l;                                              /* sign bit set <=> some length negative */
    nonempty += (l > 0);
    maxlen = l > maxlen ? l : maxlen;
    sumsq += (unsigned long long)((long long)l * l);
  }
  if (bad < 0) {
    for (int r = 0; r < m; r++) if (i[r + 1] < i[r]) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets decrease at row %d", r);
  }
  S->nonzerorow = nonempty;
  /* compressed-row rule: reference uses it when the fraction of EMPTY rows exceeds 0.6 (aij.c:818,859) */
  S->compressed = (m > 0) && ((double)(m - nonempty) > 0.6 * (double)m);
  std::vector<int> ci, rindex;
  const int *vi = i;
  int nrows = m;
  if (S->compressed) {
    ci.reserve(nonempty + 1); rindex.reserve(nonempty);
    ci.push_back(0);
    for (int r = 0; r < m; r++) if (i[r + 1] > i[r]) { rindex.push_back(r); ci.push_back(i[r + 1]); }
    /* offsets of consecutive non-empty rows are contiguous in j/a, so ci[k] == i[rindex[k]] */
    vi = ci.data(); nrows = nonempty;
  }
  S->ncrows = nrows;
  /* lanes per row from the mean length of the non-empty rows (B200_SPMV_G forces a value: 1 = reference order) */
  {
    int cfg = 0;
    const char *ec = getenv("B200_SPMV_CFG");
    if (ec && atoi(ec) >= 0 && atoi(ec) < SPMV_NCFG) cfg = atoi(ec);
    if (g_tune_cfg >= 0 && g_tune_cfg < SPMV_NCFG) cfg = g_tune_cfg;
    S->cfg = cfg;
    /* statistics of the non-empty rows: mean and coefficient of variation of the lengths */
    const double mean = nonempty ? (double)nz / (double)nonempty : 0.0;
    double ss;
    if (maxlen <= (1 << 26)) {
      ss = (double)sumsq - (double)nonempty * mean * mean;  /* sum (l - mean)^2 over the non-empty rows; sumsq <= nz * maxlen < 2^57 */
      if (ss < 0.0) ss = 0.0;
    } else {                                                /* gigantic rows: the integer sum could wrap */
      ss = 0.0;
      for (int r = 0; r < m; r++) { const double l = (double)(i[r + 1] - i[r]); if (l > 0) ss += (l - mean) * (l - mean); }
    }
    const double cv = (nonempty && mean > 0) ? sqrt(ss / (double)nonempty) / mean : 0.0;
    /* lanes per row: one lane (reference summation order) up to 32 entries per row, else about 8-16 entries per lane */
    int G = 1;
    if (mean > 32.0) while (G < 32 && mean >= 16.0 * G) G *= 2;
    const char *eg = getenv("B200_SPMV_G");
    if (eg) { const int v = atoi(eg); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) G = v; }
    if (g_tune_G > 0) G = g_tune_G;
    S->G = G;
    S->park = cv > 0.5;
    const char *ep = getenv("B200_SPMV_PARK");
    if (ep) S->park = atoi(ep) != 0;
    if (g_tune_park >= 0) S->park = g_tune_park != 0;
    S->long_thresh = g_cfgs[cfg].tile_nnz < 64 * G ? g_cfgs[cfg].tile_nnz : 64 * G;
    if (S->long_thresh < 256) S->long_thresh = 256 < g_cfgs[cfg].tile_nnz ? 256 : g_cfgs[cfg].tile_nnz;
    /* parked variant with one lane per row: a long row no longer holds its tile up (the row's warp sums what exceeds the
     * cap, b200_spmv_set_park_cap), so only rows that do not fit a tile are cut into chunk tiles — C4: 3565 such rows
     * instead of 84 746, 3.81 -> 3.63 ms per product (profiles/r02_plaw_sweep_cap_10M.txt) */
    S->park_cap = (S->park && S->req_G <= 0 && g_park_cap > 0) ? g_park_cap : 0;   /* an explicit request for a number of lanes per row (1 = the reference order) is honoured as it is */
    if (S->park_cap > 0 && G == 1) S->long_thresh = g_cfgs[cfg].tile_nnz;  /* the explicit settings below still override */
    const char *el = getenv("B200_SPMV_LONG");
    if (el && atoi(el) > 0 && atoi(el) <= g_cfgs[cfg].tile_nnz) S->long_thresh = atoi(el);
    if (g_tune_long > 0 && g_tune_long <= g_cfgs[cfg].tile_nnz) S->long_thresh = g_tune_long;
  }
  const int tile_nnz = g_cfgs[S->cfg].tile_nnz, tile_rows = g_cfgs[S->cfg].threads / S->G;
  /* binning: row tiles, and chunk tiles {row position, -(slot+1), k_begin, k_end} for the long rows */
  std::vector<int4> tiles;
  std::vector<int>  longrows, longstart;
  int ts = 0, nchunks = 0;
  auto close_tile = [&](int re) { if (re > ts) tiles.push_back(make_int4(ts, re, vi[ts], vi[re])); };
  if (maxlen <= S->long_thresh) {
    /* no long rows: the greedy packing below reduces to "as many rows as fit", found per tile by a bounded binary
     * search instead of a walk over every row (same tiles) */
    tiles.reserve((size_t)(nz / tile_nnz) + (size_t)(nrows / tile_rows) + 2);
    while (ts < nrows) {
      const int hi = ts + tile_rows < nrows ? ts + tile_rows : nrows;
      int re = hi;
      if (vi[hi] - vi[ts] > tile_nnz)
        re = (int)(std::upper_bound(vi + ts + 1, vi + hi + 1, (long long)vi[ts] + tile_nnz, [](long long v, int e) { return v < (long long)e; }) - vi) - 1;
      if (re <= ts) re = ts + 1;                            /* cannot happen (a row is at most long_thresh <= tile_nnz long) */
      close_tile(re);
      ts = re;
    }
  } else
  for (int r = 0; r < nrows; r++) {
    const int len = vi[r + 1] - vi[r];
    if (len > S->long_thresh) {
      close_tile(r);
      longrows.push_back(r); longstart.push_back(nchunks);
      for (int kb = vi[r]; kb < vi[r + 1]; kb += tile_nnz) {
        const int ke = kb + tile_nnz < vi[r + 1] ? kb + tile_nnz : vi[r + 1];
        tiles.push_back(make_int4(r, -(nchunks + 1), kb, ke));
        nchunks++;
      }
      ts = r + 1;
      continue;
    }
    if (vi[r + 1] - vi[ts] > tile_nnz || r - ts >= tile_rows) { close_tile(r); ts = r; }
  }
  close_tile(nrows);
  longstart.push_back(nchunks);
  S->nchunks = nchunks;
  S->ntiles = (int)tiles.size();
  S->nlong  = (int)longrows.size();
  S->grid   = g_b200.nsm * g_cfgs[S->cfg].ctas_per_sm;
  { const char *eg = getenv("B200_SPMV_CTAS"); if (eg && atoi(eg) > 0) S->grid = g_b200.nsm * atoi(eg); }   /* experiment: fewer resident CTAs per SM */
  if (S->grid > S->ntiles) S->grid = S->ntiles;

  const size_t pad = 8;
  if (!S->d_i) {
    B200_CUDA(cudaMalloc((void **)&S->d_i, sizeof(int) * ((size_t)m + 1)));
    B200_CHK(b200_upload_pageable(S->d_i, i, sizeof(int) * ((size_t)m + 1)));
  }
  if (!adopt) {
    B200_CUDA(cudaMalloc((void **)&S->d_j, sizeof(int) * ((size_t)nz + pad)));
    B200_CUDA(cudaMalloc((void **)&S->d_a, sizeof(double) * ((size_t)nz + pad)));
    if (nz) {
      B200_CHK(b200_upload_pageable(S->d_j, j, sizeof(int) * (size_t)nz));
      B200_CHK(b200_upload_pageable(S->d_a, a, sizeof(double) * (size_t)nz));
    }
  }
  B200_CUDA(cudaMemset(S->d_j + nz, 0, sizeof(int) * pad));
  B200_CUDA(cudaMemset(S->d_a + nz, 0, sizeof(double) * pad));
  if (S->compressed) {
    B200_CUDA(cudaMalloc((void **)&S->d_ci, sizeof(int) * ((size_t)nrows + 1)));
    B200_CUDA(cudaMemcpy(S->d_ci, ci.data(), sizeof(int) * ((size_t)nrows + 1), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMalloc((void **)&S->d_rindex, sizeof(int) * ((size_t)nrows + 1)));
    B200_CUDA(cudaMemcpy(S->d_rindex, rindex.data(), sizeof(int) * (size_t)nrows, cudaMemcpyHostToDevice));
  } else {
    S->d_ci = S->d_i;
  }
  if (S->ntiles) {
    B200_CUDA(cudaMalloc((void **)&S->d_tiles, sizeof(int4) * tiles.size()));
    B200_CUDA(cudaMemcpy(S->d_tiles, tiles.data(), sizeof(int4) * tiles.size(), cudaMemcpyHostToDevice));
  }
  if (S->nlong) {
    B200_CUDA(cudaMalloc((void **)&S->d_long, sizeof(int) * longrows.size()));
    B200_CUDA(cudaMemcpy(S->d_long, longrows.data(), sizeof(int) * longrows.size(), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMalloc((void **)&S->d_longstart, sizeof(int) * longstart.size()));
    B200_CUDA(cudaMemcpy(S->d_longstart, longstart.data(), sizeof(int) * longstart.size(), cudaMemcpyHostToDevice));
    B200_CUDA(cudaMalloc((void **)&S->d_longpart, sizeof(double) * (size_t)nchunks));
  }
  B200_CUDA(cudaMalloc((void **)&S->d_dotpart, sizeof(double) * ((size_t)S->grid + S->nlong + 1)));
  B200_CUDA(cudaMalloc((void **)&S->d_ticket, sizeof(unsigned)));
  B200_CUDA(cudaMemset(S->d_ticket, 0, sizeof(unsigned)));
  /* the warp-tile form for irregular blocks with one lane per row (B200_SPMV_WARP=0 keeps the CTA-tile form) */
  {
    const char *ew = getenv("B200_SPMV_WARP");
    const bool want = S->park_cap > 0 && nz > 0 && !(ew && !atoi(ew));
    if (want) {
      std::vector<int4> wt;
      std::vector<int>  wl, wls;
      int wts = 0, wch = 0;
      wt.reserve((size_t)(nz / SPMV_WARP_TILE) * 2 + 16);
      auto wclose = [&](int re) { if (re > wts) wt.push_back(make_int4(wts, re, vi[wts], vi[re])); };
      for (int r = 0; r < nrows; r++) {
        const int len = vi[r + 1] - vi[r];
        if (len > SPMV_WARP_TILE) {
          wclose(r);
          wl.push_back(r); wls.push_back(wch);
          for (int kb = vi[r]; kb < vi[r + 1]; kb += SPMV_WARP_TILE) {
            const int ke = kb + SPMV_WARP_TILE < vi[r + 1] ? kb + SPMV_WARP_TILE : vi[r + 1];
            wt.push_back(make_int4(r, -(wch + 1), kb, ke));
            wch++;
          }
          wts = r + 1;
          continue;
        }
        if (vi[r + 1] - vi[wts] > SPMV_WARP_TILE || r - wts >= 32) { wclose(r); wts = r; }
      }
      wclose(nrows);
      wls.push_back(wch);
      S->nwtiles = (int)wt.size(); S->nwlong = (int)wl.size(); S->nwchunks = wch;
      B200_CUDA(cudaMalloc((void **)&S->d_wtiles, sizeof(int4) * (wt.size() + 1)));
      B200_CUDA(cudaMemcpy(S->d_wtiles, wt.data(), sizeof(int4) * wt.size(), cudaMemcpyHostToDevice));
      if (S->nwlong) {
        B200_CUDA(cudaMalloc((void **)&S->d_wlong, sizeof(int) * wl.size()));
        B200_CUDA(cudaMemcpy(S->d_wlong, wl.data(), sizeof(int) * wl.size(), cudaMemcpyHostToDevice));
        B200_CUDA(cudaMalloc((void **)&S->d_wlongstart, sizeof(int) * wls.size()));
        B200_CUDA(cudaMemcpy(S->d_wlongstart, wls.data(), sizeof(int) * wls.size(), cudaMemcpyHostToDevice));
        B200_CUDA(cudaMalloc((void **)&S->d_wlongpart, sizeof(double) * (size_t)(wch + 1)));
      }
    }
  }
  return 0;
}

static int g_tune_strips = -1;         /* -1 automatic, 0/1 off, k > 1 forced */

/* How many column strips?  Stencil-like matrices gather from a narrow window of x that lives in L2 by itself; a matrix
 * whose rows span (almost) all columns drags the whole of x through L2 against the matrix stream, and once x is larger
 * than about half of the 126 MB L2 every gather goes to DRAM (measured: 10 M-row power-law matrix, x = 80 MB: 9.7 ms
 * per product unblocked).  Strips of ~24 MB of x keep the active window resident. */
static int choose_strips(int m, int n, const int *i, const int *j)
{
  int forced = g_tune_strips;
  const char *e = getenv("B200_SPMV_STRIPS");
  if (e) forced = atoi(e);
  if (forced >= 0) return forced < 1 ? 1 : (forced > 64 ? 64 : forced);
  if (8.0 * (double)n <= 48e6 || m == 0) return 1;
  double span = 0.0; long long rows = 0;
  const int step = m > (1 << 20) ? m >> 20 : 1;          /* sample about a million rows */
  for (int r = 0; r < m; r += step) if (i[r + 1] > i[r]) { span += (double)(j[i[r + 1] - 1] - j[i[r]]); rows++; }
  if (!rows || 8.0 * span / (double)rows <= 24e6) return 1;
  int S = (int)((8.0 * (double)n + 24e6 - 1) / 24e6);
  return S > 64 ? 64 : S;
}

int b200_seq_upload(B200Seq *S, int m, int n, const int *i, const int *j, const double *a)
{
  memset(S, 0, sizeof(*S));
  if (m < 0 || n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative matrix size %d x %d", m, n);
  if (!i || i[0] != 0) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets must start at 0");
  const long long nz = i[m];
  const int ns = (nz > 0 && j) ? choose_strips(m, n, i, j) : 1;
  if (ns <= 1) return seq_upload_one(S, m, n, i, j, a);
  /* split by column range; within a row the entries of a strip are a contiguous run because columns ascend */
  S->m = m; S->n = n; S->nz = nz; S->nstrips = ns; S->req_G = g_tune_G;
  S->strips = (B200Seq *)calloc((size_t)ns, sizeof(B200Seq));
  S->strip_nzoff = (long long *)calloc((size_t)ns + 1, sizeof(long long));
  if (!S->strips || !S->strip_nzoff) return b200_seterr(B200_ERR_MEM, "out of host memory");
  const int width = (int)(((long long)n + ns - 1) / ns);
  /* one pass: validity, non-empty rows, longest row, sum of squared lengths (integer accumulators: a floating-point
   * running sum makes this loop latency-bound, and it runs at every upload) */
  int nonempty = 0, maxlen = 0, bad = 0;
  unsigned long long sumsq = 0;
  for (int r = 0; r < m; r++) {
    const int l = i[r + 1] - i[r];
    bad |= l;                                              /* sign bit set <=> some length negative */
    nonempty += (l > 0);
    maxlen = l > maxlen ? l : maxlen;
    sumsq += (unsigned long long)((long long)l * l);
  }
  if (bad < 0) {
    for (int r = 0; r < m; r++) if (i[r + 1] < i[r]) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets decrease at row %d", r);
  }
  S->nonzerorow = nonempty;
  /* per strip: count every row's entries below the strip's upper column bound, prefix-sum, copy — the counting and the
   * copying on row ranges in parallel (a 500 M-entry matrix took 4 s on one thread); the result is the same whatever
   * the number of threads */
  int T = (int)std::thread::hardware_concurrency();
  if (T < 1 || nz < (1 << 22)) T = 1;
  { const char *e = getenv("B200_HOST_THREADS"); if (e && atoi(e) > 0) T = atoi(e); }
  if (T > 16) T = 16;
  if (T > m) T = m > 0 ? m : 1;
  std::vector<int> cut((size_t)T + 1, m);
  cut[0] = 0;
  for (int t = 1; t < T; t++) cut[t] = (int)(std::lower_bound(i, i + m + 1, (int)(nz * t / T)) - i);
  for (int t = 1; t <= T; t++) if (cut[t] < cut[t - 1]) cut[t] = cut[t - 1];
  auto par = [&](const std::function<void(int)> &f) {
    if (T == 1) { f(0); return; }
    std::vector<std::thread> th;
    for (int t = 1; t < T; t++) th.emplace_back(f, t);
    f(0);
    for (auto &x : th) x.join();
  };
  std::vector<int> si((size_t)m + 1), sj, cur(i, i + m), nxt((size_t)m);   /* cur: next unread entry of every row */
  std::vector<double> sa;
  for (int k = 0; k < ns; k++) {
    const long long chi = (long long)(k + 1) * width;
    par([&](int t) {
      for (int r = cut[t]; r < cut[t + 1]; r++) {
        int c = cur[r];
        const int e = i[r + 1];
        while (c < e && (long long)j[c] < chi) c++;
        nxt[r] = c;
      }
    });
    si[0] = 0;
    for (int r = 0; r < m; r++) si[(size_t)r + 1] = si[r] + (nxt[r] - cur[r]);
    const size_t snz = (size_t)si[m];
    sj.resize(snz); sa.resize(snz);
    par([&](int t) {
      for (int r = cut[t]; r < cut[t + 1]; r++) {
        const int b = cur[r], len = nxt[r] - b;
        if (len > 0) {
          memcpy(sj.data() + si[r], j + b, sizeof(int) * (size_t)len);
          memcpy(sa.data() + si[r], a + b, sizeof(double) * (size_t)len);
        }
        cur[r] = nxt[r];
      }
    });
    S->strip_nzoff[k + 1] = S->strip_nzoff[k] + (long long)snz;
    int rc = seq_upload_one(&S->strips[k], m, n, si.data(), sj.data(), sa.data());
    if (rc) return rc;
  }
  /* the descriptive fields report the first strip's binning */
  S->cfg = S->strips[0].cfg; S->G = S->strips[0].G; S->park = S->strips[0].park; S->long_thresh = S->strips[0].long_thresh;
  S->grid = S->strips[0].grid;
  for (int k = 0; k < ns; k++) { S->ntiles += S->strips[k].ntiles; S->nlong += S->strips[k].nlong; }
  S->ncrows = m;
  return 0;
}

/* Block from a CSR whose column ids and values already live on the device (device assembly, b200_assembly.cu): only
 * the row offsets pass through the host (the row binning is computed from them); d_i (m + 1 entries, may be NULL: then
 * i_host is uploaded) and d_j / d_a (nz + 8 entries each) are adopted.  Column strips are not formed on this path. */
int b200_seq_adopt(B200Seq *S, int m, int n, const int *i_host, int *d_i, int *d_j, double *d_a)
{
  return seq_upload_one(S, m, n, i_host, d_j, d_a, true, d_i);
}

extern "C" int b200_spmv_set_strips(int nstrips)
{
  g_tune_strips = nstrips;
  return 0;
}

/* The explicit transpose, built once (lazily) from the device copy: MatMultTranspose_SeqAIJ (aij.c:1078-1135) walks
 * the rows in order and does y[j] += x[i] a_ij, so every y[j] receives its terms in ascending source-row order; a CSR
 * of A^T whose rows list the source rows ascending, multiplied by the one-lane-per-row kernel starting from y (or 0),
 * performs exactly the same additions in the same order — bit-identical, and it streams like any other MatMult
 * instead of scattering atomics. */
int b200_seq_transpose(B200Seq *S)
{
  if (S->T) return 0;
  const int m = S->m, n = S->n;
  const int nblk = S->nstrips > 1 ? S->nstrips : 1;
  std::vector<int> ti((size_t)n + 1, 0);
  std::vector<std::vector<int>>    bi(nblk), bj(nblk);
  std::vector<std::vector<double>> ba(nblk);
  for (int k = 0; k < nblk; k++) {
    const B200Seq *C = S->nstrips > 1 ? &S->strips[k] : S;
    bi[k].resize((size_t)m + 1); bj[k].resize((size_t)C->nz); ba[k].resize((size_t)C->nz);
    B200_CUDA(cudaMemcpy(bi[k].data(), C->d_i, sizeof(int) * ((size_t)m + 1), cudaMemcpyDeviceToHost));
    if (C->nz) {
      B200_CUDA(cudaMemcpy(bj[k].data(), C->d_j, sizeof(int) * (size_t)C->nz, cudaMemcpyDeviceToHost));
      B200_CUDA(cudaMemcpy(ba[k].data(), C->d_a, sizeof(double) * (size_t)C->nz, cudaMemcpyDeviceToHost));
    }
    for (long long e = 0; e < C->nz; e++) ti[(size_t)bj[k][(size_t)e] + 1]++;
  }
  for (int c = 0; c < n; c++) ti[(size_t)c + 1] += ti[c];
  std::vector<int> tj((size_t)S->nz), pos(ti.begin(), ti.end() - 1);
  std::vector<double> ta((size_t)S->nz);
  for (int k = 0; k < nblk; k++)                                  /* a column lives in exactly one strip */
    for (int r = 0; r < m; r++)
      for (int e = bi[k][r]; e < bi[k][(size_t)r + 1]; e++) {
        const int c = bj[k][e], p = pos[c]++;
        tj[p] = r; ta[p] = ba[k][e];
      }
  S->T = (B200Seq *)calloc(1, sizeof(B200Seq));
  if (!S->T) return b200_seterr(B200_ERR_MEM, "out of host memory");
  const int keepG = g_tune_G;
  g_tune_G = S->req_G;                                            /* the transpose follows its parent's lanes-per-row request */
  int rc = b200_seq_upload(S->T, n, m, ti.data(), tj.data(), ta.data());
  g_tune_G = keepG;
  if (rc) { b200_seq_free(S->T); free(S->T); S->T = NULL; }
  return rc;
}

/* ---- tile-kernel dispatch: (launch shape) x (lanes per row) x (add, dot) */
template <int G, bool PARK, int THREADS, int TILE_NNZ, int STAGES>
static cudaError_t launch_tile5(const B200Seq *S, bool add, bool dot, const double *x, const double *yin, double *z, const double *dotx, const SpmvCgTail &tail, cudaStream_t st)
{
  const size_t smem = SpmvSmem<THREADS, TILE_NNZ, STAGES>::BYTES;
#define GO(ADD, DOT) do { \
    auto kern = spmv_tile_kernel<G, PARK, ADD, DOT, THREADS, TILE_NNZ, STAGES>; \
    static bool attr_done = false; \
    if (!attr_done) { cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); if (e != cudaSuccess) return e; attr_done = true; } \
    kern<<<S->grid, THREADS, smem, st>>>(S->ntiles, S->d_tiles, S->d_ci, S->d_rindex, S->d_j, S->d_a, x, yin, z, S->d_dotpart, S->d_longpart, dotx, tail); \
  } while (0)
  if (add) { if (dot) GO(true, true); else GO(true, false); }
  else     { if (dot) GO(false, true); else GO(false, false); }
#undef GO
  return cudaGetLastError();
}

template <int G, int THREADS, int TILE_NNZ, int STAGES>
static cudaError_t launch_tile4(const B200Seq *S, bool add, bool dot, const double *x, const double *yin, double *z, const double *dotx, const SpmvCgTail &tail, cudaStream_t st)
{
  return S->park ? launch_tile5<G, true, THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st)
                 : launch_tile5<G, false, THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
}

template <int THREADS, int TILE_NNZ, int STAGES>
static cudaError_t launch_tile3(const B200Seq *S, bool add, bool dot, const double *x, const double *yin, double *z, const double *dotx, const SpmvCgTail &tail, cudaStream_t st)
{
  switch (S->G) {
    case 1:  return launch_tile4<1,  THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
    case 2:  return launch_tile4<2,  THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
    case 4:  return launch_tile4<4,  THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
    case 8:  return launch_tile4<8,  THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
    case 16: return launch_tile4<16, THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
    default: return launch_tile4<32, THREADS, TILE_NNZ, STAGES>(S, add, dot, x, yin, z, dotx, tail, st);
  }
}

static cudaError_t launch_tile(const B200Seq *S, bool add, bool dot, const double *x, const double *yin, double *z, const double *dotx, const SpmvCgTail &tail, cudaStream_t st)
{
  switch (S->cfg) {                       /* must mirror g_cfgs[] */
    case 0:  return launch_tile3<256, 2048, 1>(S, add, dot, x, yin, z, dotx, tail, st);
    case 1:  return launch_tile3<256, 2048, 2>(S, add, dot, x, yin, z, dotx, tail, st);
    case 2:  return launch_tile3<512, 4096, 2>(S, add, dot, x, yin, z, dotx, tail, st);
    case 3:  return launch_tile3<128, 1024, 1>(S, add, dot, x, yin, z, dotx, tail, st);
    case 4:  return launch_tile3<512, 4096, 1>(S, add, dot, x, yin, z, dotx, tail, st);
    case 5:  return launch_tile3<256, 1792, 2>(S, add, dot, x, yin, z, dotx, tail, st);
    case 6:  return launch_tile3<256, 4096, 2>(S, add, dot, x, yin, z, dotx, tail, st);
    default: return launch_tile3<384, 3072, 1>(S, add, dot, x, yin, z, dotx, tail, st);
  }
}

/* z = (add ? yin : 0) + S x.  want_dot: the CTAs leave partials of dotx . (z - yin) in S->d_dotpart (dotx NULL = x);
 * dot_out (device scalar), when non-NULL, receives their sum (one more tiny kernel). */
int b200_seq_mult_ex(const B200Seq *S, const double *x, const double *yin, double *z, bool add, bool want_dot, const double *dotx, double *dot_out, CgState *cg)
{
  cudaStream_t st = g_b200.stream;
  if (S->m == 0) return 0;
  if (cg && (!want_dot || S->nlong || S->nstrips > 1 || !S->ntiles)) return b200_seterr(B200_ERR_SUP, "the CG tail needs the fused dot on a block without long rows or strips");
  const SpmvCgTail tail = {cg, S->d_ticket, S->park_cap};
  if (S->nstrips > 1) {
    if (want_dot) return b200_seterr(B200_ERR_SUP, "fused dot is not available for a column-stripped block");
    for (int k = 0; k < S->nstrips; k++) {
      /* the first strip starts the row sums (from 0 or from yin), every later one continues them in place */
      B200_CHK(b200_seq_mult_ex(&S->strips[k], x, k == 0 ? yin : z, z, k == 0 ? add : true, false, NULL, NULL, NULL));
    }
    return 0;
  }
  if (!dotx) dotx = x;
  if (add && z != yin && S->compressed) {
    /* rows the compressed list skips keep y (aij.c:1313-1315) */
    B200_CUDA(cudaMemcpyAsync(z, yin, sizeof(double) * (size_t)S->m, cudaMemcpyDeviceToDevice, st));
  }
  if (!add && S->compressed) {
    zero_rows_kernel<<<B200_NSM * 4, 256, 0, st>>>(S->m, z);
    B200_LAUNCHED(1);
  }
  if (want_dot && S->compressed && !add) return b200_seterr(B200_ERR_SUP, "fused dot needs an uncompressed block (or the MatMultAdd form)");
  if (S->nwtiles && !want_dot) {
    /* irregular block, plain product: one warp per 256-entry tile, no CTA barriers (spmv_warp_kernel) */
    int g = (S->nwtiles + 7) / 8;
    if (g > g_b200.nsm * 6) g = g_b200.nsm * 6;
    if (add) spmv_warp_kernel<true><<<g, 256, 0, st>>>(S->nwtiles, S->d_wtiles, S->d_ci, S->d_rindex, S->d_j, S->d_a, x, yin, z, S->d_wlongpart, S->park_cap);
    else     spmv_warp_kernel<false><<<g, 256, 0, st>>>(S->nwtiles, S->d_wtiles, S->d_ci, S->d_rindex, S->d_j, S->d_a, x, yin, z, S->d_wlongpart, S->park_cap);
    B200_LAUNCHED(1);
    if (S->nwlong) {
      const int fg = (S->nwlong + 127) / 128;
      if (add) spmv_long_finish_kernel<true, false><<<fg, 128, 0, st>>>(S->nwlong, S->d_wlong, S->d_wlongstart, S->d_rindex, S->d_wlongpart, dotx, yin, z, NULL);
      else     spmv_long_finish_kernel<false, false><<<fg, 128, 0, st>>>(S->nwlong, S->d_wlong, S->d_wlongstart, S->d_rindex, S->d_wlongpart, dotx, yin, z, NULL);
      B200_LAUNCHED(1);
    }
    return 0;
  }
  if (S->ntiles) {
    B200_CUDA(launch_tile(S, add, want_dot, x, yin, z, dotx, tail, st));
    B200_LAUNCHED(1);
  }
  if (S->nlong) {
    double *dp = S->d_dotpart + S->grid;
    const int fg = (S->nlong + 127) / 128;
#define FIN(ADD, DOT) spmv_long_finish_kernel<ADD, DOT><<<fg, 128, 0, st>>>(S->nlong, S->d_long, S->d_longstart, S->d_rindex, S->d_longpart, dotx, yin, z, dp)
    if (add) { if (want_dot) FIN(true, true); else FIN(true, false); }
    else     { if (want_dot) FIN(false, true); else FIN(false, false); }
#undef FIN
    B200_LAUNCHED(1);
  }
  if (want_dot && dot_out) {
    dot_final_kernel<<<1, 1024, 0, st>>>((S->ntiles ? S->grid : 0) + S->nlong, S->ntiles ? S->d_dotpart : S->d_dotpart + S->grid, 0, NULL, dot_out);
    B200_LAUNCHED(1);
  }
  return 0;
}

/* where the last want_dot product left its per-CTA partials */
void b200_seq_dotparts(const B200Seq *S, const double **p, int *n)
{
  *n = (S->m == 0) ? 0 : (S->ntiles ? S->grid : 0) + S->nlong;
  *p = S->ntiles ? S->d_dotpart : S->d_dotpart + S->grid;
}

int b200_seq_mult(const B200Seq *S, const double *x, const double *yin, double *z, bool add, double *dot_out)
{
  return b200_seq_mult_ex(S, x, yin, z, add, dot_out != NULL, NULL, dot_out, NULL);
}

/* y = A x on the row-partitioned operator with the LOCAL part of x . y left in dot_out (device scalar): the diagonal
 * block contributes x . (A_d x), the off-diagonal MatMultAdd x . (B lvec); one final kernel adds both partial lists. */
int b200_mpiaij_mult_dot(b200_mat A, const double *x, double *y, double *dot_out)
{
  const B200Seq *Sa = &A->A, *Sb = &A->B;
  if (A->fusedplan) {
    B200_CHK(b200_mpiaij_mult_fused(A, x, NULL, y, false, true));
    if (!dot_out) return 0;
    const double *p; int n;
    b200_mpiaij_fused_dotparts(A, &p, &n);
    dot_final_kernel<<<1, 1024, 0, g_b200.stream>>>(n, p, 0, NULL, dot_out);
    B200_LAUNCHED(1);
    return 0;
  }
  if (Sa->compressed || Sa->m != Sa->n || Sa->nstrips > 1 || Sb->nstrips > 1) return b200_seterr(B200_ERR_SUP, "fused MatMult+dot needs a square, uncompressed, unstripped diagonal block");
  B200_CHK(b200_mpiaij_exchange_begin(A, x));
  B200_CHK(b200_seq_mult_ex(Sa, x, NULL, y, false, true, x, NULL, NULL));
  B200_CHK(b200_mpiaij_exchange_end(A));
  B200_CHK(b200_seq_mult_ex(Sb, A->d_lvec[A->epoch & 1], y, y, true, true, x, NULL, NULL));
  if (Sa->m == 0 || !dot_out) return 0;                          /* dot_out NULL: the caller reduces the partial lists itself */
  const int na = (Sa->ntiles ? Sa->grid : 0) + Sa->nlong, nb = (Sb->ntiles ? Sb->grid : 0) + Sb->nlong;
  dot_final_kernel<<<1, 1024, 0, g_b200.stream>>>(na, Sa->ntiles ? Sa->d_dotpart : Sa->d_dotpart + Sa->grid,
                                                  nb, Sb->ntiles ? Sb->d_dotpart : Sb->d_dotpart + Sb->grid, dot_out);
  B200_LAUNCHED(1);
  return 0;
}

/* ------------------------------------------------------------------------------------------------ C ABI */
extern "C" int b200_mat_create_seqaij(int m, int n, const int *i, const int *j, const double *a, b200_mat *A)
{
  B200_NEED_INIT();
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null output handle");
  *A = NULL;
  b200_mat M = (b200_mat)calloc(1, sizeof(*M));
  if (!M) return b200_seterr(B200_ERR_MEM, "out of host memory");
  M->kind = 0;
  int rc = b200_seq_upload(&M->A, m, n, i, j, a);
  if (rc) { b200_seq_free(&M->A); free(M); return rc; }
  *A = M;
  return 0;
}

extern "C" int b200_mat_update_values(b200_mat A, const double *a)
{
  B200_NEED_INIT();
  if (!A || !a) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (A->kind != 0) return b200_seterr(B200_ERR_SUP, "value refresh is implemented for the sequential type");
  if (A->A.nstrips > 1) return b200_seterr(B200_ERR_SUP, "value refresh is not implemented for a column-stripped matrix (recreate it)");
  if (A->A.T) { b200_seq_free(A->A.T); free(A->A.T); A->A.T = NULL; }        /* rebuilt from the new values on demand */
  return b200_upload_pageable(A->A.d_a, a, sizeof(double) * (size_t)A->A.nz);
}

extern "C" int b200_mat_update_values_mpiaij(b200_mat A, const double *ad, const double *ao)
{
  B200_NEED_INIT();
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  if (A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not a parallel matrix");
  if ((A->A.nz && !ad) || (A->B.nz && !ao)) return b200_seterr(B200_ERR_ARG_NULL, "null value array");
  if (A->A.nstrips > 1 || A->B.nstrips > 1) return b200_seterr(B200_ERR_SUP, "value refresh is not implemented for a column-stripped block (recreate the matrix)");
  B200Seq *blk[2] = {&A->A, &A->B};
  const double *src[2] = {ad, ao};
  for (int k = 0; k < 2; k++) {
    if (blk[k]->T) { b200_seq_free(blk[k]->T); free(blk[k]->T); blk[k]->T = NULL; }   /* rebuilt from the new values on demand */
    if (blk[k]->nz) B200_CHK(b200_upload_pageable(blk[k]->d_a, src[k], sizeof(double) * (size_t)blk[k]->nz));
  }
  return 0;
}

extern "C" int b200_mat_get_size(b200_mat A, int *m, int *n, long long *nz)
{
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  if (m) *m = A->A.m;
  if (n) *n = A->A.n;
  if (nz) *nz = A->A.nz + (A->kind ? A->B.nz : 0);
  return 0;
}

extern "C" int b200_mat_flops(b200_mat A, double *f)
{
  if (!A || !f) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *f = 2.0 * (double)A->A.nz - (double)A->A.nonzerorow;          /* aij.c:1281 */
  if (A->kind) *f += 2.0 * (double)A->B.nz;                      /* MatMultAdd: aij.c:1343 */
  return 0;
}

extern "C" int b200_mat_bytes(b200_mat A, double *b)
{
  if (!A || !b) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  double v = 12.0 * (double)A->A.nz + 4.0 * ((double)A->A.m + 1) + 8.0 * (double)A->A.n + 8.0 * (double)A->A.m;
  if (A->kind) {
    /* off-diagonal MatMultAdd: only the rows it visits count (the non-empty ones when stored compressed: offsets, row
     * ids, y read + write), the ghost vector once, and the exchange itself (gather read + ghost-vector write) */
    const double rows = (double)A->B.ncrows;
    v += 12.0 * (double)A->B.nz + 4.0 * (rows + 1) + (A->B.compressed ? 4.0 * rows : 0.0) + 16.0 * rows + 8.0 * (double)A->nghost
       + 16.0 * (double)A->nghost;
  }
  *b = v;
  return 0;
}

extern "C" int b200_mat_bin_info(b200_mat A, int *ntiles, int *nlong, int *compressed)
{
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  const B200Seq *S = &A->A;
  if (ntiles) *ntiles = S->ntiles;
  if (nlong) *nlong = S->nlong;
  if (compressed) *compressed = S->compressed;
  return 0;
}

extern "C" int b200_mat_bin_info2(b200_mat A, int *lanes, int *long_thresh, int *grid, int *shape, int *park)
{
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  if (lanes) *lanes = A->A.G;
  if (long_thresh) *long_thresh = A->A.long_thresh;
  if (grid) *grid = A->A.grid;
  if (shape) *shape = A->A.cfg;
  if (park) *park = A->A.park;
  return 0;
}

extern "C" int b200_spmv_set_tuning(int shape, int lanes, int long_thresh, int park)
{
  if (shape >= SPMV_NCFG) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "launch shape %d out of range (0..%d)", shape, SPMV_NCFG - 1);
  if (lanes > 0 && (lanes > 32 || (lanes & (lanes - 1)))) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "lanes per row must be a power of two <= 32");
  g_tune_cfg = shape; g_tune_G = lanes; g_tune_long = long_thresh; g_tune_park = park;
  return 0;
}

extern "C" int b200_mat_mult(b200_mat A, const double *x, double *y)
{
  B200_NEED_INIT();
  if (!A || !x || !y) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == y) return b200_seterr(B200_ERR_ARG_IDN, "x and y must be different vectors");   /* matrix.c:2143 */
  if (A->kind == 0) return b200_seq_mult(&A->A, x, NULL, y, false, NULL);
  /* MatMult_MPIAIJ mpiaij.c:1102-1116: scatter begin | diagonal block | scatter end | off-diagonal multadd —
   * one kernel when the operator has a one-launch plan, else the four steps (b200_comm.cu) */
  return b200_mpiaij_mult(A, x, NULL, y, false);
}

/* diagnostic: one MPIAIJ MatMult with CUDA events between its phases (exchange begin | diagonal block | exchange end |
 * off-diagonal block); each phase includes the launch gap in front of it */
extern "C" int b200_mat_mult_phases(b200_mat A, const double *x, double *y, double *ms4)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !ms4) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (A->kind != 1) return b200_seterr(B200_ERR_ARG_WRONG, "not an MPIAIJ matrix");
  static cudaEvent_t ev[5];
  static bool made = false;
  if (!made) { for (int k = 0; k < 5; k++) B200_CUDA(cudaEventCreate(&ev[k])); made = true; }
  cudaStream_t st = g_b200.stream;
  B200_CUDA(cudaEventRecord(ev[0], st));
  B200_CHK(b200_mpiaij_exchange_begin(A, x));
  B200_CUDA(cudaEventRecord(ev[1], st));
  B200_CHK(b200_seq_mult(&A->A, x, NULL, y, false, NULL));
  B200_CUDA(cudaEventRecord(ev[2], st));
  B200_CHK(b200_mpiaij_exchange_end(A));
  B200_CUDA(cudaEventRecord(ev[3], st));
  B200_CHK(b200_seq_mult(&A->B, A->d_lvec[A->epoch & 1], y, y, true, NULL));
  B200_CUDA(cudaEventRecord(ev[4], st));
  B200_CUDA(cudaStreamSynchronize(st));
  for (int k = 0; k < 4; k++) { float f = 0; B200_CUDA(cudaEventElapsedTime(&f, ev[k], ev[k + 1])); ms4[k] = f; }
  return 0;
}

extern "C" int b200_mat_multadd(b200_mat A, const double *x, const double *y, double *z)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !z) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == z) return b200_seterr(B200_ERR_ARG_IDN, "x and z must be different vectors");
  if (A->kind == 0) return b200_seq_mult(&A->A, x, y, z, true, NULL);
  /* MatMultAdd_MPIAIJ mpiaij.c:1132-1143 */
  return b200_mpiaij_mult(A, x, y, z, true);
}

extern "C" int b200_mat_multtranspose(b200_mat A, const double *x, double *y)
{
  B200_NEED_INIT();
  if (!A || !x || !y) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == y) return b200_seterr(B200_ERR_ARG_IDN, "x and y must be different vectors");   /* matrix.c:2207 */
  if (A->kind != 0) return b200_mpiaij_multtranspose(A, x, NULL, y);
  B200_CHK(b200_seq_transpose(&A->A));
  if (A->A.n > 0 && A->A.m == 0) return b200_vec_set(A->A.n, y, 0.0);
  return b200_seq_mult(A->A.T, x, NULL, y, false, NULL);
}

extern "C" int b200_mat_multtransposeadd(b200_mat A, const double *x, const double *y, double *z)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !z) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == z) return b200_seterr(B200_ERR_ARG_IDN, "x and z must be different vectors");
  if (A->kind != 0) return b200_mpiaij_multtranspose(A, x, y, z);
  B200_CHK(b200_seq_transpose(&A->A));
  if (A->A.m == 0) return (z != y && A->A.n > 0) ? b200_vec_copy(A->A.n, y, z) : 0;
  return b200_seq_mult(A->A.T, x, y, z, true, NULL);
}

extern "C" int b200_mat_mult_dot(b200_mat A, const double *x, double *y, double *xdoty)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !xdoty) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (A->kind != 0 || A->A.m != A->A.n || A->A.compressed || A->A.nstrips > 1) return b200_seterr(B200_ERR_SUP, "fused MatMult+dot needs a square sequential matrix stored as one block");
  b200_reduce_arm(0);          /* not stamped: the product's dot may reach its slot in two steps, so wait for the stream */
  B200_CHK(b200_seq_mult(&A->A, x, NULL, y, false, g_b200.d_res));
  return b200_reduce_finish(1, xdoty, false, 0);
}

extern "C" int b200_mat_get_diagonal(b200_mat A, double *d)
{
  B200_NEED_INIT();
  if (!A || !d) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  const B200Seq *S = &A->A;
  if (S->m == 0) return 0;
  /* the diagonal block stores local column ids, so the diagonal entry of local row i is column i in both types
   * (MatGetDiagonal_MPIAIJ forwards to the diagonal block, mpiaij.c:1246-1256) */
  if (S->nstrips > 1) {
    /* the diagonal entry of row i sits in the strip that owns column i */
    const int width = (int)(((long long)S->n + S->nstrips - 1) / S->nstrips);
    B200_CUDA(cudaMemsetAsync(d, 0, sizeof(double) * (size_t)S->m, g_b200.stream));
    for (int k = 0; k < S->nstrips; k++) {
      const int r0 = k * width, r1 = (k + 1) * width < S->m ? (k + 1) * width : S->m;
      if (r0 >= r1) continue;
      const B200Seq *C = &S->strips[k];
      int g = (r1 - r0 + 255) / 256; if (g > B200_NSM * 8) g = B200_NSM * 8;
      B200Seq *Cw = &A->A.strips[k];
      /* the strip keeps positions for its own rows r0..r1 only (indexed by the global row: pointer shifted by r0) */
      if (Cw->d_diagpos) diag_gather_kernel<<<g, 256, 0, g_b200.stream>>>(r0, r1, C->d_diagpos - r0, C->d_a, d);
      else {
        B200_CUDA(cudaMalloc((void **)&Cw->d_diagpos, sizeof(int) * (size_t)(r1 - r0 + 1)));
        diag_kernel<<<g, 256, 0, g_b200.stream>>>(r0, r1, C->d_i, C->d_j, C->d_a, 0, d, Cw->d_diagpos - r0);
      }
      B200_LAUNCHED(1);
    }
    return 0;
  }
  int g = (S->m + 255) / 256; if (g > B200_NSM * 8) g = B200_NSM * 8;
  if (S->d_diagpos) diag_gather_kernel<<<g, 256, 0, g_b200.stream>>>(0, S->m, S->d_diagpos, S->d_a, d);
  else {
    B200_CUDA(cudaMalloc((void **)&A->A.d_diagpos, sizeof(int) * (size_t)(S->m + 1)));
    diag_kernel<<<g, 256, 0, g_b200.stream>>>(0, S->m, S->d_i, S->d_j, S->d_a, 0, d, A->A.d_diagpos);
  }
  B200_LAUNCHED(1);
  return 0;
}

extern "C" int b200_mat_strips(b200_mat A, int *diag_strips, int *offdiag_strips)
{
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  if (diag_strips) *diag_strips = A->A.nstrips > 1 ? A->A.nstrips : 1;
  if (offdiag_strips) *offdiag_strips = A->kind ? (A->B.nstrips > 1 ? A->B.nstrips : 1) : 0;
  return 0;
}

extern "C" int b200_mat_destroy(b200_mat A)
{
  if (!A) return 0;
  if (g_b200.inited) cudaStreamSynchronize(g_b200.stream);
  b200_seq_free(&A->A);
  if (A->batchplan) b200_batchplan_free(A->batchplan);
  if (A->kind) {
    b200_seq_free(&A->B);
    b200_mpiaij_fused_free(A);
    if (A->ev_done) cudaEventDestroy(A->ev_done);
    if (A->peerplan) b200_mpiaij_peer_free(A);
    cudaFree(A->d_ghost_t);
    cudaFree(A->d_garray); cudaFree(A->d_lvec[0]); cudaFree(A->d_lvec[1]); cudaFree(A->d_to_indices); cudaFree(A->d_sendbuf);
    free(A->to_procs); free(A->to_starts); free(A->from_procs); free(A->from_starts);
  }
  free(A);
  return 0;
}
