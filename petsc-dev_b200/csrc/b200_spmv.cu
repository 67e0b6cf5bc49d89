/* b200_spmv.cu — AIJ (CSR) MatMult / MatMultAdd for sm_100a.
 *
 * Replaces MatMult_SeqAIJ / MatMultAdd_SeqAIJ (src/mat/impls/aij/seq/aij.c:1225-1358) and the cusp::multiply call of
 * MatMult_SeqAIJCUSP (src/mat/impls/aij/seq/seqcusp/aijcusp.cu:349-390).
 *
 * Row binning (done once at upload, the device analogue of MatAssemblyEnd_SeqAIJ's compressed-row/inode checks):
 *   - "stream tiles": maximal runs of consecutive short rows holding <= TILE_NNZ entries.  One CTA per tile: the
 *     tile's values and column ids are streamed with 128-bit loads, multiplied by the gathered x (read-only path,
 *     L1/L2 resident) and parked in shared memory; then one thread per row adds its products LEFT TO RIGHT starting
 *     from 0 (MatMult) or y[i] (MatMultAdd) — the reference's exact association (aij.h:384-386), so with -fmad=false
 *     these rows are bit-identical to the CPU result.  Row offsets of the tile are staged in shared memory.
 *   - "long rows" (> LONG_THRESH entries): one CTA per row, strided loads, fixed-shape tree reduction
 *     (deterministic; differs from the reference in association only).
 *   - blocks that are mostly empty rows (the MPIAIJ off-diagonal block) are visited through the list of non-empty
 *     rows only (what Mat_CompressedRow does on the CPU: src/mat/utils/compressedrow.c:28-80).
 * Algorithmic traffic per MatMult: 12 B/nnz + 4 B/row (offsets) + 8 B/row (y) + 8 B/col (x once).
 */
#include "b200_internal.h"
#include <vector>

#define SPMV_THREADS 256
#define TILE_NNZ     2048
#define TILE_ROWS    512
#define LONG_THRESH  256
#define LONG_THREADS 128
#define TILE_GROUPS  (TILE_NNZ / 4 + 1)                               /* 16-byte groups a tile can touch          */
#define TILE_ITERS   ((TILE_GROUPS + SPMV_THREADS - 1) / SPMV_THREADS)

template <bool ADD, bool DOT>
__global__ void __launch_bounds__(SPMV_THREADS)
spmv_stream_kernel(const int4 *__restrict__ tiles, const int *__restrict__ ci, const int *__restrict__ rindex,
                   const int *__restrict__ aj, const double *__restrict__ aa, const double *__restrict__ x,
                   const double *yin, double *z, double *__restrict__ dotpart)
{
  __shared__ double s_prod[TILE_NNZ];
  __shared__ int    s_rp[TILE_ROWS + 1];
  __shared__ double s_red[SPMV_THREADS / 32];
  const int4 t  = __ldg(&tiles[blockIdx.x]);
  const int  rb = t.x, nrows = t.y - t.x, k0 = t.z, k1 = t.w;
  const int  tid = threadIdx.x;

  /* stage the tile's row offsets (relative to k0) */
  for (int r = tid; r <= nrows; r += SPMV_THREADS) s_rp[r] = __ldg(&ci[rb + r]) - k0;

  /* phase 1: stream a/j in aligned 16-byte groups, gather x, park the products */
  const int g0 = k0 >> 2, g1 = (k1 + 3) >> 2;
  int4    jv[TILE_ITERS];
  double2 av[TILE_ITERS][2];
#pragma unroll
  for (int it = 0; it < TILE_ITERS; it++) {
    const int g = g0 + tid + it * SPMV_THREADS;
    if (g < g1) {
      jv[it]    = __ldcs(reinterpret_cast<const int4 *>(aj) + g);
      av[it][0] = __ldcs(reinterpret_cast<const double2 *>(aa) + 2 * (size_t)g);
      av[it][1] = __ldcs(reinterpret_cast<const double2 *>(aa) + 2 * (size_t)g + 1);
    }
  }
#pragma unroll
  for (int it = 0; it < TILE_ITERS; it++) {
    const int g = g0 + tid + it * SPMV_THREADS;
    if (g < g1) {
      const int k = 4 * g - k0;                          /* position of the group's first entry in the tile */
      const int n1 = k1 - k0;
      if (k >= 0 && k + 3 < n1) {
        const double p0 = av[it][0].x * __ldg(&x[jv[it].x]);
        const double p1 = av[it][0].y * __ldg(&x[jv[it].y]);
        const double p2 = av[it][1].x * __ldg(&x[jv[it].z]);
        const double p3 = av[it][1].y * __ldg(&x[jv[it].w]);
        s_prod[k] = p0; s_prod[k + 1] = p1; s_prod[k + 2] = p2; s_prod[k + 3] = p3;
      } else {
        if (k     >= 0 && k     < n1) s_prod[k]     = av[it][0].x * __ldg(&x[jv[it].x]);
        if (k + 1 >= 0 && k + 1 < n1) s_prod[k + 1] = av[it][0].y * __ldg(&x[jv[it].y]);
        if (k + 2 >= 0 && k + 2 < n1) s_prod[k + 2] = av[it][1].x * __ldg(&x[jv[it].z]);
        if (k + 3 >= 0 && k + 3 < n1) s_prod[k + 3] = av[it][1].y * __ldg(&x[jv[it].w]);
      }
    }
  }
  __syncthreads();

  /* phase 2: one thread per row, left-to-right sum */
  double dot = 0.0;
  for (int r = tid; r < nrows; r += SPMV_THREADS) {
    const int row = rindex ? __ldg(&rindex[rb + r]) : rb + r;
    const int b = s_rp[r], e = s_rp[r + 1];
    double sum = ADD ? yin[row] : 0.0;
    for (int k = b; k < e; k++) sum += s_prod[k];
    z[row] = sum;
    if (DOT) dot += __ldg(&x[row]) * sum;
  }
  if (DOT) {
    const int lane = tid & 31, wid = tid >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) s_red[wid] = dot;
    __syncthreads();
    if (wid == 0) {
      double v = (lane < SPMV_THREADS / 32) ? s_red[lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) dotpart[blockIdx.x] = v;
    }
  }
}

/* one CTA per long row */
template <bool ADD, bool DOT>
__global__ void __launch_bounds__(LONG_THREADS)
spmv_long_kernel(const int *__restrict__ longrows, const int *__restrict__ ci, const int *__restrict__ rindex,
                 const int *__restrict__ aj, const double *__restrict__ aa, const double *__restrict__ x,
                 const double *yin, double *z, double *__restrict__ dotpart)
{
  __shared__ double s_red[LONG_THREADS / 32];
  const int r = __ldg(&longrows[blockIdx.x]);
  const int row = rindex ? __ldg(&rindex[r]) : r;
  const int k0 = __ldg(&ci[r]), k1 = __ldg(&ci[r + 1]);
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  int k = k0 + threadIdx.x;
  for (; k + 3 * LONG_THREADS < k1; k += 4 * LONG_THREADS) {
    const int j0 = __ldcs(&aj[k]), j1 = __ldcs(&aj[k + LONG_THREADS]), j2 = __ldcs(&aj[k + 2 * LONG_THREADS]), j3 = __ldcs(&aj[k + 3 * LONG_THREADS]);
    const double a0 = __ldcs(&aa[k]), a1 = __ldcs(&aa[k + LONG_THREADS]), a2 = __ldcs(&aa[k + 2 * LONG_THREADS]), a3 = __ldcs(&aa[k + 3 * LONG_THREADS]);
    s0 += a0 * __ldg(&x[j0]); s1 += a1 * __ldg(&x[j1]); s2 += a2 * __ldg(&x[j2]); s3 += a3 * __ldg(&x[j3]);
  }
  for (; k < k1; k += LONG_THREADS) s0 += __ldcs(&aa[k]) * __ldg(&x[__ldcs(&aj[k])]);
  double s = (s0 + s1) + (s2 + s3);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) s_red[wid] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < LONG_THREADS / 32; w++) v += s_red[w];
    const double out = ADD ? yin[row] + v : v;
    z[row] = out;
    if (DOT) dotpart[blockIdx.x] = __ldg(&x[row]) * out;
  }
}

/* deterministic final pass over the per-CTA partials of the fused dot */
__global__ void __launch_bounds__(1024) dot_final_kernel(int n, const double *__restrict__ part, double *out)
{
  __shared__ double s_red[32];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) s += part[i];
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

/* d[i] = a_ii or 0 (MatGetDiagonal_SeqAIJ aij.c:1040-1073); columns ascending -> bisection */
__global__ void diag_kernel(int m, const int *__restrict__ ai, const int *__restrict__ aj, const double *__restrict__ aa,
                            int coloff, double *__restrict__ d)
{
  const int stride = gridDim.x * blockDim.x;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += stride) {
    int lo = ai[i], hi = ai[i + 1];
    const int want = i + coloff;
    double v = 0.0;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      const int c = aj[mid];
      if (c == want) { v = aa[mid]; break; }
      if (c < want) lo = mid + 1; else hi = mid;
    }
    d[i] = v;
  }
}

/* ------------------------------------------------------------------------------------------------ host side */
void b200_seq_free(B200Seq *S)
{
  if (!S) return;
  if (S->d_ci && S->d_ci != S->d_i) cudaFree(S->d_ci);
  cudaFree(S->d_i); cudaFree(S->d_j); cudaFree(S->d_a); cudaFree(S->d_rindex);
  cudaFree(S->d_tiles); cudaFree(S->d_long); cudaFree(S->d_dotpart);
  memset(S, 0, sizeof(*S));
}

int b200_seq_upload(B200Seq *S, int m, int n, const int *i, const int *j, const double *a)
{
  memset(S, 0, sizeof(*S));
  if (m < 0 || n < 0) return b200_seterr(B200_ERR_ARG_SIZ, "negative matrix size %d x %d", m, n);
  if (!i || i[0] != 0) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets must start at 0");
  const long long nz = i[m];
  if (nz < 0) return b200_seterr(B200_ERR_ARG_WRONG, "negative nonzero count");
  if (nz > 0 && (!j || !a)) return b200_seterr(B200_ERR_ARG_NULL, "null column/value arrays");
  S->m = m; S->n = n; S->nz = nz;
  /* row statistics + the visiting order */
  int nonempty = 0;
  for (int r = 0; r < m; r++) {
    if (i[r + 1] < i[r]) return b200_seterr(B200_ERR_ARG_WRONG, "row offsets decrease at row %d", r);
    nonempty += (i[r + 1] > i[r]);
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
  /* binning */
  std::vector<int4> tiles;
  std::vector<int>  longrows;
  int ts = 0;
  auto close_tile = [&](int re) { if (re > ts) tiles.push_back(make_int4(ts, re, vi[ts], vi[re])); };
  for (int r = 0; r < nrows; r++) {
    const int len = vi[r + 1] - vi[r];
    if (len > LONG_THRESH) { close_tile(r); longrows.push_back(r); ts = r + 1; continue; }
    if (vi[r + 1] - vi[ts] > TILE_NNZ || r - ts >= TILE_ROWS) { close_tile(r); ts = r; }
  }
  close_tile(nrows);
  S->ntiles = (int)tiles.size();
  S->nlong  = (int)longrows.size();

  const size_t pad = 8;
  B200_CUDA(cudaMalloc((void **)&S->d_i, sizeof(int) * ((size_t)m + 1)));
  B200_CUDA(cudaMalloc((void **)&S->d_j, sizeof(int) * ((size_t)nz + pad)));
  B200_CUDA(cudaMalloc((void **)&S->d_a, sizeof(double) * ((size_t)nz + pad)));
  B200_CUDA(cudaMemcpy(S->d_i, i, sizeof(int) * ((size_t)m + 1), cudaMemcpyHostToDevice));
  if (nz) {
    B200_CUDA(cudaMemcpy(S->d_j, j, sizeof(int) * (size_t)nz, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(S->d_a, a, sizeof(double) * (size_t)nz, cudaMemcpyHostToDevice));
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
  }
  B200_CUDA(cudaMalloc((void **)&S->d_dotpart, sizeof(double) * ((size_t)S->ntiles + S->nlong + 1)));
  return 0;
}

/* z = (add ? yin : 0) + S x ; dot_out (device scalar) receives x.z when non-NULL (needs m == n, all rows visited) */
int b200_seq_mult(const B200Seq *S, const double *x, const double *yin, double *z, bool add, double *dot_out)
{
  cudaStream_t st = g_b200.stream;
  if (S->m == 0) return 0;
  if (add && z != yin && S->compressed) {
    /* rows the compressed list skips keep y (aij.c:1313-1315) */
    B200_CUDA(cudaMemcpyAsync(z, yin, sizeof(double) * (size_t)S->m, cudaMemcpyDeviceToDevice, st));
  }
  if (!add && S->compressed) {
    zero_rows_kernel<<<B200_NSM * 4, 256, 0, st>>>(S->m, z);
    B200_LAUNCHED(1);
  }
  if (dot_out && S->compressed) return b200_seterr(B200_ERR_SUP, "fused dot needs an uncompressed square block");
#define SPMV_GO(ADD, DOT) do { \
    if (S->ntiles) { spmv_stream_kernel<ADD, DOT><<<S->ntiles, SPMV_THREADS, 0, st>>>(S->d_tiles, S->d_ci, S->d_rindex, S->d_j, S->d_a, x, yin, z, S->d_dotpart); B200_LAUNCHED(1); } \
    if (S->nlong)  { spmv_long_kernel<ADD, DOT><<<S->nlong, LONG_THREADS, 0, st>>>(S->d_long, S->d_ci, S->d_rindex, S->d_j, S->d_a, x, yin, z, S->d_dotpart + S->ntiles); B200_LAUNCHED(1); } \
  } while (0)
  if (add) { if (dot_out) SPMV_GO(true, true); else SPMV_GO(true, false); }
  else     { if (dot_out) SPMV_GO(false, true); else SPMV_GO(false, false); }
#undef SPMV_GO
  if (dot_out) {
    dot_final_kernel<<<1, 1024, 0, st>>>(S->ntiles + S->nlong, S->d_dotpart, dot_out);
    B200_LAUNCHED(1);
  }
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
  B200_CUDA(cudaMemcpyAsync(A->A.d_a, a, sizeof(double) * (size_t)A->A.nz, cudaMemcpyHostToDevice, g_b200.stream));
  B200_CUDA(cudaStreamSynchronize(g_b200.stream));
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
  if (A->kind) v += 12.0 * (double)A->B.nz + 4.0 * ((double)A->B.m + 1) + 8.0 * (double)A->nghost + 16.0 * (double)A->B.m + 16.0 * (double)A->nghost;
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

extern "C" int b200_mat_mult(b200_mat A, const double *x, double *y)
{
  B200_NEED_INIT();
  if (!A || !x || !y) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == y) return b200_seterr(B200_ERR_ARG_IDN, "x and y must be different vectors");   /* matrix.c:2143 */
  if (A->kind == 0) return b200_seq_mult(&A->A, x, NULL, y, false, NULL);
  /* MatMult_MPIAIJ mpiaij.c:1102-1116: scatter begin | diagonal block | scatter end | off-diagonal multadd */
  B200_CHK(b200_mpiaij_exchange_begin(A, x));
  B200_CHK(b200_seq_mult(&A->A, x, NULL, y, false, NULL));
  B200_CHK(b200_mpiaij_exchange_end(A));
  return b200_seq_mult(&A->B, A->d_lvec[A->epoch & 1], y, y, true, NULL);
}

extern "C" int b200_mat_multadd(b200_mat A, const double *x, const double *y, double *z)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !z) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (x == z) return b200_seterr(B200_ERR_ARG_IDN, "x and z must be different vectors");
  if (A->kind == 0) return b200_seq_mult(&A->A, x, y, z, true, NULL);
  /* MatMultAdd_MPIAIJ mpiaij.c:1132-1143 */
  B200_CHK(b200_mpiaij_exchange_begin(A, x));
  B200_CHK(b200_seq_mult(&A->A, x, y, z, true, NULL));
  B200_CHK(b200_mpiaij_exchange_end(A));
  return b200_seq_mult(&A->B, A->d_lvec[A->epoch & 1], z, z, true, NULL);
}

extern "C" int b200_mat_mult_dot(b200_mat A, const double *x, double *y, double *xdoty)
{
  B200_NEED_INIT();
  if (!A || !x || !y || !xdoty) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (A->kind != 0 || A->A.m != A->A.n || A->A.compressed) return b200_seterr(B200_ERR_SUP, "fused MatMult+dot needs a square sequential matrix");
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
  int g = (S->m + 255) / 256; if (g > B200_NSM * 8) g = B200_NSM * 8;
  diag_kernel<<<g, 256, 0, g_b200.stream>>>(S->m, S->d_i, S->d_j, S->d_a, 0, d);
  B200_LAUNCHED(1);
  return 0;
}

extern "C" int b200_mat_destroy(b200_mat A)
{
  if (!A) return 0;
  if (g_b200.inited) cudaStreamSynchronize(g_b200.stream);
  b200_seq_free(&A->A);
  if (A->kind) {
    b200_seq_free(&A->B);
    cudaFree(A->d_garray); cudaFree(A->d_lvec[0]); cudaFree(A->d_lvec[1]); cudaFree(A->d_to_indices); cudaFree(A->d_sendbuf);
    free(A->to_procs); free(A->to_starts); free(A->from_procs); free(A->from_starts);
  }
  free(A);
  return 0;
}
