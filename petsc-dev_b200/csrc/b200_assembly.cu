/* Device assembly of a SeqAIJ matrix from element matrices: MatSetValuesBatch (src/mat/interface/matrix.c:1698).
 *
 * What the reference does.  Host path (matrix.c:1712-1716): one MatSetValues(A, nl, rows_e, nl, rows_e, v_e, ADD_VALUES)
 * per element; MatSetValues_SeqAIJ (aij.c:293-368) walks the nl rows of the block in order, and within a row the nl
 * columns in order, inserting a new nonzero (sorted position, value = v) or adding to the existing one.  So a nonzero
 * (r, c) ends up holding  ((v_0 + v_1) + v_2) + ...  with its contributions ordered by (element, row-in-element,
 * column-in-element), explicit zeros stay in the pattern, negative ids are skipped.  Device path of the reference
 * (MatSetValuesBatch_SeqAIJCUSP, aijAssemble.cu:157-290): expand to a fat COO, global sort by (i, j), reduce by key,
 * COO -> CSR, copy back.
 *
 * What this file does instead (no global sort, no fat COO in memory).  Every use of a row id in elem_rows is an
 * "occurrence" o = e*nl + a; it contributes the nl fat entries (elem_rows[e*nl + b], elem_mats[o*nl + b]), b = 0..nl-1.
 *   1. count the occurrences of every row (atomics), exclusive scan -> occurrence offsets            [4 B / occurrence]
 *   2. fill the per-row occurrence lists (atomic cursors: arbitrary order inside a row)              [8 B / occurrence]
 *   3. one warp per row (one CTA for rows with more than 512 fat entries): sort the row's occurrences ascending
 *      (this restores the element order, so the result does not depend on the atomics), expand them to fat keys
 *      (column << 32 | position) in shared memory, bitonic sort, mark the first entry of every distinct column,
 *      write the sorted source positions ("perm", 4 B / fat entry) and the number of distinct columns of the row
 *   4. exclusive scan of the row lengths -> CSR row offsets; second sweep writes the column ids and the start of every
 *      nonzero's run inside perm
 *   5. values: nonzero k = elem_mats[perm[run]] summed left to right over its run — the same additions in the same
 *      order as the host path.  Steps 1-4 depend on elem_rows only: with keep_plan the runs stay on the device and a
 *      new set of element matrices costs one pass of step 5 (b200_mat_setvaluesbatch_values).
 * Step 5 is HBM-bound: 8 B (value) + 4 B (perm) per fat entry, 4 B (run start) + 8 B (result) per nonzero. */
#include "b200_internal.h"
#include <vector>
#include <chrono>
#include <cstdio>
#include <cstdlib>

typedef unsigned long long u64;
typedef unsigned int       u32;

#define ASM_SENTINEL 0x7fffffffu      /* slot of the row's slice that holds no entry (negative column id)            */
#define ASM_HEAD     0x80000000u      /* first entry of a run of equal columns                                        */
#define ASM_WARP_CAP 512              /* fat entries a warp sorts in its shared-memory slab                           */
#define ASM_CTA_CAP  4096             /* fat entries a CTA sorts in shared memory; longer rows sort in global scratch */
#define ASM_WARPS    8
#define ASM_FULL     0xffffffffu

struct AsmStat { int bad_row, bad_col, nbig, maxfat; };

struct BatchPlan {
  long long nvals;                    /* ne*nl*nl: length of elem_mats                                                */
  long long nfat;                     /* slots in perm                                                                */
  long long nz;
  u32      *d_perm;                   /* [nfat] source position in elem_mats | ASM_HEAD, rows in order, columns ascending */
  int      *d_segoff;                 /* [nz+1] first slot of every nonzero's run                                     */
};

void b200_batchplan_free(void *plan)
{
  BatchPlan *P = (BatchPlan *)plan;
  if (!P) return;
  cudaFree(P->d_perm); cudaFree(P->d_segoff);
  free(P);
}

namespace {
/* Device buffers of one assembly; whatever is still listed is released on return, what the matrix keeps (column ids,
 * values, the kept plan) is taken off the list with release().  Everything is a stream-ordered allocation from the
 * device's default pool, whose release threshold is raised to 8 GiB once: after the first assembly an allocation costs
 * microseconds (plain cudaMalloc/cudaFree of these half-gigabyte buffers was 2/3 of the wall time).  The kept buffers
 * are later freed with cudaFree, which accepts pool allocations and hands them back to the pool. */
struct DevBufs {
  std::vector<void *> items;
  cudaStream_t st;
  explicit DevBufs(cudaStream_t s) : st(s)
  {
    static bool pool_ready = false;
    if (!pool_ready) {
      int dev = 0;
      cudaMemPool_t pool;
      if (cudaGetDevice(&dev) == cudaSuccess && cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
        unsigned long long keep = 8ull << 30;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
      }
      pool_ready = true;
    }
  }
  ~DevBufs() { for (void *q : items) cudaFreeAsync(q, st); }
  template <class T> cudaError_t get(T **out, size_t count)
  {
    void *q = NULL;
    cudaError_t e = cudaMallocAsync(&q, sizeof(T) * (count ? count : 1), st);
    if (e == cudaSuccess) items.push_back(q);
    *out = (T *)q;
    return e;
  }
  void release(void *q) { for (size_t k = 0; k < items.size(); k++) if (items[k] == q) { items.erase(items.begin() + k); return; } }
};

struct AsmTrace {                     /* B200_ASM_TRACE=1: wall-clock milliseconds of the phases on stderr */
  bool on;
  std::chrono::steady_clock::time_point t0;
  AsmTrace() : on(getenv("B200_ASM_TRACE") != NULL), t0(std::chrono::steady_clock::now()) {}
  void mark(const char *what)
  {
    if (!on) return;
    cudaDeviceSynchronize();
    auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[b200 assembly] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};
}

/* ---- exclusive scan of n ints into out[0..n] (out[n] = total) ---------------------------------------------------- */
#define SCAN_T 256
#define SCAN_E 8

__device__ __forceinline__ int cta_exclusive(int s, int *wsum, int *total)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  int inc = s;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(ASM_FULL, inc, d); if (lane >= d) inc += t; }
  if (lane == 31) wsum[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int w = lane < nwarp ? wsum[lane] : 0;
    int wi = w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(ASM_FULL, wi, d); if (lane >= d) wi += t; }
    if (lane < nwarp) wsum[lane] = wi - w;
    if (lane == 31) wsum[32] = wi;
  }
  __syncthreads();
  const int excl = inc - s + wsum[warp];
  *total = wsum[32];
  __syncthreads();
  return excl;
}

__global__ void __launch_bounds__(SCAN_T) scan_block_kernel(int n, const int *__restrict__ in, int *__restrict__ out, int *__restrict__ bsum)
{
  __shared__ int wsum[33];
  const long long base = (long long)blockIdx.x * (SCAN_T * SCAN_E) + (long long)threadIdx.x * SCAN_E;
  int v[SCAN_E], s = 0;
#pragma unroll
  for (int e = 0; e < SCAN_E; e++) { v[e] = base + e < n ? in[base + e] : 0; s += v[e]; }
  int total;
  int excl = cta_exclusive(s, wsum, &total);
#pragma unroll
  for (int e = 0; e < SCAN_E; e++) { if (base + e < n) out[base + e] = excl; excl += v[e]; }
  if (threadIdx.x == 0) bsum[blockIdx.x] = total;
}

__global__ void __launch_bounds__(1024) scan_sums_kernel(int nb, int *bsum)
{
  __shared__ int wsum[33];
  int carry = 0;
  for (int base = 0; base < nb; base += 1024) {
    const int idx = base + threadIdx.x;
    const int v = idx < nb ? bsum[idx] : 0;
    int total;
    const int excl = cta_exclusive(v, wsum, &total);
    if (idx < nb) bsum[idx] = carry + excl;
    carry += total;
  }
  if (threadIdx.x == 0) bsum[nb] = carry;
}

__global__ void scan_add_kernel(int n, int *__restrict__ out, const int *__restrict__ bsum, int nb)
{
  for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += (long long)gridDim.x * blockDim.x)
    out[idx] += bsum[idx / (SCAN_T * SCAN_E)];
  if (blockIdx.x == 0 && threadIdx.x == 0) out[n] = bsum[nb];
}

static int dev_exclusive_scan(int n, const int *in, int *out, DevBufs &tmp, cudaStream_t st)
{
  const int nb = n > 0 ? (n + SCAN_T * SCAN_E - 1) / (SCAN_T * SCAN_E) : 0;
  int *bsum;
  B200_CUDA(tmp.get(&bsum, (size_t)nb + 1));
  if (nb) { scan_block_kernel<<<nb, SCAN_T, 0, st>>>(n, in, out, bsum); B200_LAUNCHED(1); }
  scan_sums_kernel<<<1, 1024, 0, st>>>(nb, bsum); B200_LAUNCHED(1);
  int g = nb ? (n + 255) / 256 : 1;
  if (g > B200_NSM * 16) g = B200_NSM * 16;
  scan_add_kernel<<<g, 256, 0, st>>>(n, out, bsum, nb); B200_LAUNCHED(1);
  return 0;
}

/* ---- steps 1, 2: occurrences of every row ------------------------------------------------------------------------- */
/* Row window [rstart, rstart + m): the sequential matrix has rstart = 0 and row_limit = m (a larger id is an error);
 * the local rows of a parallel matrix have row_limit = global size, and ids of rows other ranks own are passed over
 * here — they still serve as column ids in step 3 (MatSetValues_MPIAIJ keeps the owned rows and stashes the others,
 * mpiaij.c:496-570). */
__global__ void asm_count_kernel(long long nocc, const int *__restrict__ rows, int rstart, int m, int row_limit, int *__restrict__ cnt, AsmStat *st)
{
  for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < nocc; o += (long long)gridDim.x * blockDim.x) {
    const int g = rows[o];
    if (g < 0) continue;                                   /* aij.c:311: negative rows are ignored */
    if (g >= row_limit) { atomicMax(&st->bad_row, g); continue; }
    const int r = g - rstart;
    if (r < 0 || r >= m) continue;                          /* another rank's row */
    atomicAdd(&cnt[r], 1);
  }
}

/* cnt counts down to zero while the slots fill from the back; the order inside a row is fixed later by the sort */
__global__ void asm_fill_kernel(long long nocc, const int *__restrict__ rows, int rstart, int m, const int *__restrict__ off, int *__restrict__ cnt, u32 *__restrict__ occ)
{
  for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < nocc; o += (long long)gridDim.x * blockDim.x) {
    const int g = rows[o];
    if (g < 0) continue;
    const int r = g - rstart;
    if (r < 0 || r >= m) continue;
    const int pos = atomicSub(&cnt[r], 1) - 1;
    occ[off[r] + pos] = (u32)o;
  }
}

__global__ void asm_classify_kernel(int m, int nl, const int *__restrict__ off, int *__restrict__ big, AsmStat *st)
{
  for (long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x; r < m; r += (long long)gridDim.x * blockDim.x) {
    const long long fat = (long long)(off[r + 1] - off[r]) * nl;
    if (fat > ASM_WARP_CAP) {
      big[atomicAdd(&st->nbig, 1)] = (int)r;
      atomicMax(&st->maxfat, (int)fat);
    }
  }
}

/* ---- step 3: per-row sort ----------------------------------------------------------------------------------------- */
template <bool WARP> __device__ __forceinline__ void grp_sync() { if (WARP) __syncwarp(); else __syncthreads(); }

/* Tried and dropped: the same network for rows of at most 64 keys held in registers (two per lane, partners by
 * 64-bit shuffles): as many instructions per round as this shared-memory form, no gain (per-row sort 2.95 -> 3.1 ms). */
template <bool WARP>
__device__ void bitonic_sort(u64 *k, int npad, int tid, int nt)
{
  for (int size = 2; size <= npad; size <<= 1)
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < (npad >> 1); t += nt) {
        const int i = 2 * t - (t & (stride - 1));
        const int j = i + stride;
        const bool up = (i & size) == 0;
        const u64 a = k[i], b = k[j];
        if ((a > b) == up) { k[i] = b; k[j] = a; }
      }
      grp_sync<WARP>();
    }
}

/* One row, by one warp (WARP) or one CTA.  keys: room for the padded number of fat entries.  Returns (to every thread
 * of a warp; through *cta_heads for a CTA) the number of distinct columns. */
template <bool WARP>
__device__ void asm_sort_row(int r, int nl, int n, const int *__restrict__ off, u32 *occ, const int *__restrict__ rows,
                             u64 *keys, u32 *__restrict__ perm, int *__restrict__ rowlen, AsmStat *st, int tid, int nt, int *cta_heads)
{
  const int o0 = off[r], c = off[r + 1] - o0;
  if (c == 0) { if (tid == 0) rowlen[r] = 0; return; }
  /* occurrences ascending = element order, then row-in-element order */
  int cpad = 1;
  while (cpad < c) cpad <<= 1;
  for (int t = tid; t < cpad; t += nt) keys[t] = t < c ? (u64)occ[o0 + t] : ~0ull;
  grp_sync<WARP>();
  bitonic_sort<WARP>(keys, cpad, tid, nt);
  for (int t = tid; t < c; t += nt) occ[o0 + t] = (u32)keys[t];
  grp_sync<WARP>();
  /* fat keys: (column, position in contribution order); skipped entries sort to the end */
  const int L = c * nl;
  const size_t f0 = (size_t)o0 * nl;
  int lpad = 1;
  while (lpad < L) lpad <<= 1;
  for (int p = tid; p < lpad; p += nt) {
    u64 key = ~0ull;
    if (p < L) {
      const u32 o = occ[o0 + p / nl];
      const int b = p % nl;
      const int col = rows[(size_t)(o / (u32)nl) * nl + b];
      if (col >= n) atomicMax(&st->bad_col, col);
      else if (col >= 0) key = ((u64)(u32)col << 32) | (u32)p;   /* aij.c:325: negative columns are ignored */
    }
    keys[p] = key;
  }
  grp_sync<WARP>();
  bitonic_sort<WARP>(keys, lpad, tid, nt);
  int heads = 0;
  for (int q = tid; q < L; q += nt) {
    const u64 key = keys[q];
    u32 out = ASM_SENTINEL;
    if (key != ~0ull) {
      const u32 p = (u32)key;
      const bool head = q == 0 || (u32)(keys[q - 1] >> 32) != (u32)(key >> 32);
      out = (occ[o0 + p / nl] * (u32)nl + p % nl) | (head ? ASM_HEAD : 0u);
      heads += head;
    }
    perm[f0 + q] = out;
  }
  if (WARP) {
    heads = __reduce_add_sync(ASM_FULL, heads);
    if (tid == 0) rowlen[r] = heads;
    __syncwarp();
  } else {
    heads = __reduce_add_sync(ASM_FULL, heads);
    if ((tid & 31) == 0 && heads) atomicAdd(cta_heads, heads);
    __syncthreads();
    if (tid == 0) { rowlen[r] = *cta_heads; *cta_heads = 0; }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(ASM_WARPS * 32) asm_sort_warp_kernel(int m, int nl, int n, const int *__restrict__ off, u32 *occ, const int *__restrict__ rows,
                                                                        u32 *__restrict__ perm, int *__restrict__ rowlen, AsmStat *st)
{
  __shared__ u64 slab[ASM_WARPS][ASM_WARP_CAP];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (long long r = (long long)blockIdx.x * ASM_WARPS + warp; r < m; r += (long long)gridDim.x * ASM_WARPS) {
    const long long fat = (long long)(off[r + 1] - off[r]) * nl;
    if (fat > ASM_WARP_CAP) continue;                       /* the CTA kernel's rows */
    asm_sort_row<true>((int)r, nl, n, off, occ, rows, slab[warp], perm, rowlen, st, lane, 32, NULL);
  }
}

__global__ void __launch_bounds__(512) asm_sort_cta_kernel(int nbig, const int *__restrict__ big, int nl, int n, const int *__restrict__ off, u32 *occ, const int *__restrict__ rows,
                                                          u32 *__restrict__ perm, int *__restrict__ rowlen, AsmStat *st, u64 *scratch, size_t scratch_per_cta)
{
  __shared__ u64 slab[ASM_CTA_CAP];
  __shared__ int heads;
  if (threadIdx.x == 0) heads = 0;
  __syncthreads();
  for (int b = blockIdx.x; b < nbig; b += gridDim.x) {
    const int r = big[b];
    const long long fat = (long long)(off[r + 1] - off[r]) * nl;
    u64 *keys = fat <= ASM_CTA_CAP ? slab : scratch + (size_t)blockIdx.x * scratch_per_cta;
    asm_sort_row<false>(r, nl, n, off, occ, rows, keys, perm, rowlen, st, threadIdx.x, blockDim.x, &heads);
  }
}

/* ---- step 4: column ids and run starts ------------------------------------------------------------------------------ */
__global__ void __launch_bounds__(256) asm_build_kernel(int m, int nl, const int *__restrict__ off, const u32 *__restrict__ perm, const int *__restrict__ ai,
                                                        const int *__restrict__ rows, int *__restrict__ aj, int *__restrict__ segoff)
{
  const int lane = threadIdx.x & 31;
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long r = gw; r < m; r += nw) {
    const size_t f0 = (size_t)off[r] * nl;
    const int L = (off[r + 1] - off[r]) * nl;
    int base = ai[r];
    for (int q0 = 0; q0 < L; q0 += 32) {
      const int q = q0 + lane;
      const u32 v = q < L ? perm[f0 + q] : ASM_SENTINEL;
      const bool head = (v & ASM_HEAD) != 0;
      const u32 ballot = __ballot_sync(ASM_FULL, head);
      if (head) {
        const int k = base + __popc(ballot & ((1u << lane) - 1u));
        const u32 seq = v & ~ASM_HEAD;
        const u32 o = seq / (u32)nl, b = seq % (u32)nl;
        segoff[k] = (int)(f0 + q);
        aj[k] = rows[(size_t)(o / (u32)nl) * nl + b];
      }
      base += __popc(ballot);
    }
  }
  if (gw == 0 && lane == 0) segoff[ai[m]] = (int)((size_t)off[m] * nl);
}

/* ---- step 5: values = contributions summed in order ---------------------------------------------------------------- */
/* Nonzero k = its run of perm, summed in slot order.  The first ASM_VAL_UNROLL slots of the run are loaded together
 * (all index loads, then all value loads in flight), the additions stay in slot order; slots past the run or holding
 * the sentinel (tail of a row's slice) add nothing.
 * Tried and dropped: one warp per row reading the element values in their own order (whole sectors) and dropping them
 * into sorted slots in shared memory through an inverse map — same bytes, but the per-row latency chain left the
 * SMs idle: 1.31 ms against 0.49 ms for this kernel on the 128^3 hexahedral mesh. */
#define ASM_VAL_UNROLL 8
__global__ void __launch_bounds__(256) asm_values_kernel(long long nz, const int *__restrict__ segoff, const u32 *__restrict__ perm, const double *__restrict__ em, double *__restrict__ a)
{
  for (long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x; k < nz; k += (long long)gridDim.x * blockDim.x) {
    const int p0 = segoff[k], p1 = segoff[k + 1];
    u32 idx[ASM_VAL_UNROLL];
    double v[ASM_VAL_UNROLL];
#pragma unroll
    for (int t = 0; t < ASM_VAL_UNROLL; t++) idx[t] = p0 + t < p1 ? (__ldcs(perm + p0 + t) & ~ASM_HEAD) : ASM_SENTINEL;
#pragma unroll
    for (int t = 0; t < ASM_VAL_UNROLL; t++) v[t] = idx[t] != ASM_SENTINEL ? __ldcs(em + idx[t]) : 0.0;
    double s = v[0];                                        /* a new nonzero starts as its first value (aij.c:350-366) */
#pragma unroll
    for (int t = 1; t < ASM_VAL_UNROLL; t++) if (idx[t] != ASM_SENTINEL) s += v[t];
    for (int p = p0 + ASM_VAL_UNROLL; p < p1; p++) {
      const u32 i = perm[p];
      if (i != ASM_SENTINEL) s += em[i];
    }
    a[k] = s;
  }
}

static int grid_for(long long n, int threads)
{
  long long g = (n + threads - 1) / threads;
  if (g < 1) g = 1;
  if (g > B200_NSM * 16) g = B200_NSM * 16;
  return (int)g;
}

static int setvaluesbatch_window(int m, int rstart, int row_limit, int n, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                                 int on_device, int keep_plan, b200_mat *A);

extern "C" int b200_mat_setvaluesbatch(int m, int n, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                                       int on_device, int keep_plan, b200_mat *A)
{
  return setvaluesbatch_window(m, 0, m, n, ne, nl, elem_rows, elem_mats, on_device, keep_plan, A);
}

extern "C" int b200_mat_setvaluesbatch_rows(int m_local, int rstart, int N, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                                            int on_device, int keep_plan, b200_mat *A)
{
  if (rstart < 0 || (long long)rstart + m_local > N) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "rows %d..%d are not inside 0..%d", rstart, rstart + m_local, N);
  return setvaluesbatch_window(m_local, rstart, N, N, ne, nl, elem_rows, elem_mats, on_device, keep_plan, A);
}

static int setvaluesbatch_window(int m, int rstart, int row_limit, int n, long long ne, int nl, const int *elem_rows, const double *elem_mats,
                                 int on_device, int keep_plan, b200_mat *A)
{
  B200_NEED_INIT();
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null output handle");
  *A = NULL;
  if (m < 0 || n < 0 || ne < 0 || nl < 1) return b200_seterr(B200_ERR_ARG_SIZ, "bad sizes m %d n %d elements %lld rows per element %d", m, n, ne, nl);
  const long long nocc = ne * nl, nvals = nocc * nl;
  if (nvals >= 0x7fffffffLL) return b200_seterr(B200_ERR_ARG_SIZ, "element matrices hold %lld values; the limit is 2^31 - 2 per matrix", nvals);
  if (ne > 0 && (!elem_rows || !elem_mats)) return b200_seterr(B200_ERR_ARG_NULL, "null element arrays");
  cudaStream_t st = g_b200.stream;
  DevBufs tmp(st);
  AsmTrace trace;
  const int *d_rows = elem_rows;
  const double *d_em = elem_mats;
  if (!on_device) {
    int *r; double *v;
    B200_CUDA(tmp.get(&r, (size_t)nocc));
    B200_CUDA(tmp.get(&v, (size_t)nvals));
    if (ne) {
      B200_CUDA(cudaMemcpyAsync(r, elem_rows, sizeof(int) * (size_t)nocc, cudaMemcpyHostToDevice, st));
      B200_CUDA(cudaMemcpyAsync(v, elem_mats, sizeof(double) * (size_t)nvals, cudaMemcpyHostToDevice, st));
    }
    d_rows = r; d_em = v;
  }
  int *cnt, *off, *rowlen, *ai, *big;
  u32 *occ, *perm;
  AsmStat *dstat, hstat = {-1, -1, 0, 0};
  B200_CUDA(tmp.get(&cnt, (size_t)m + 1));
  B200_CUDA(tmp.get(&off, (size_t)m + 1));
  B200_CUDA(tmp.get(&rowlen, (size_t)m + 1));
  B200_CUDA(tmp.get(&ai, (size_t)m + 1));
  B200_CUDA(tmp.get(&occ, (size_t)nocc));
  B200_CUDA(tmp.get(&perm, (size_t)nvals));
  B200_CUDA(tmp.get(&big, (size_t)(nvals / ASM_WARP_CAP) + 1));
  B200_CUDA(tmp.get(&dstat, 1));
  trace.mark("allocations");
  B200_CUDA(cudaMemcpyAsync(dstat, &hstat, sizeof(hstat), cudaMemcpyHostToDevice, st));
  B200_CUDA(cudaMemsetAsync(cnt, 0, sizeof(int) * ((size_t)m + 1), st));
  /* 1, 2: occurrence lists */
  asm_count_kernel<<<grid_for(nocc, 256), 256, 0, st>>>(nocc, d_rows, rstart, m, row_limit, cnt, dstat); B200_LAUNCHED(1);
  B200_CHK(dev_exclusive_scan(m, cnt, off, tmp, st));
  asm_fill_kernel<<<grid_for(nocc, 256), 256, 0, st>>>(nocc, d_rows, rstart, m, off, cnt, occ); B200_LAUNCHED(1);
  asm_classify_kernel<<<grid_for(m, 256), 256, 0, st>>>(m, nl, off, big, dstat); B200_LAUNCHED(1);
  B200_CUDA(cudaMemcpyAsync(&hstat, dstat, sizeof(hstat), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  if (hstat.bad_row >= 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "Row too large: row %d max %d", hstat.bad_row, row_limit - 1);
  trace.mark("occurrence lists");
  /* 3: per-row sort */
  asm_sort_warp_kernel<<<grid_for(m, ASM_WARPS), ASM_WARPS * 32, 0, st>>>(m, nl, n, off, occ, d_rows, perm, rowlen, dstat); B200_LAUNCHED(1);
  if (hstat.nbig) {
    int g = hstat.nbig < B200_NSM * 2 ? hstat.nbig : B200_NSM * 2;
    u64 *scratch = NULL;
    size_t per = 0;
    if (hstat.maxfat > ASM_CTA_CAP) {
      per = 1;
      while (per < (size_t)hstat.maxfat) per <<= 1;
      if (g > B200_NSM) g = B200_NSM;
      const size_t room = (size_t)1 << 27;                  /* at most 1 GiB of scratch keys: fewer CTAs for gigantic rows */
      if (per * (size_t)g > room) g = per >= room ? 1 : (int)(room / per);
      B200_CUDA(tmp.get(&scratch, per * (size_t)g));
    }
    asm_sort_cta_kernel<<<g, 512, 0, st>>>(hstat.nbig, big, nl, n, off, occ, d_rows, perm, rowlen, dstat, scratch, per); B200_LAUNCHED(1);
  }
  trace.mark("per-row sort");
  /* 4: row offsets, column ids, run starts */
  B200_CHK(dev_exclusive_scan(m, rowlen, ai, tmp, st));
  /* the row offsets go to the host (pinned scratch: a pageable target made this copy 1.4 ms for 2 M rows) for the row
   * binning of the SpMV kernel, which then runs on the host WHILE the two kernels below fill column ids and values */
  int *hi = (int *)b200_pinned_scratch(sizeof(int) * ((size_t)m + 1));
  if (!hi) return b200_seterr(B200_ERR_MEM, "out of pinned host memory");
  B200_CUDA(cudaMemcpyAsync(hi, ai, sizeof(int) * ((size_t)m + 1), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaMemcpyAsync(&hstat, dstat, sizeof(hstat), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  if (hstat.bad_col >= 0) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "Column too large: col %d max %d", hstat.bad_col, n - 1);
  const long long nz = hi[m];
  int *d_j, *segoff;
  double *d_a;
  trace.mark("row offsets to host");
  B200_CUDA(tmp.get(&d_j, (size_t)nz + 8));
  B200_CUDA(tmp.get(&d_a, (size_t)nz + 8));
  B200_CUDA(tmp.get(&segoff, (size_t)nz + 1));
  asm_build_kernel<<<grid_for(m, 8), 256, 0, st>>>(m, nl, off, perm, ai, d_rows, d_j, segoff); B200_LAUNCHED(1);
  /* 5: values */
  if (nz) { asm_values_kernel<<<grid_for(nz, 256), 256, 0, st>>>(nz, segoff, perm, d_em, d_a); B200_LAUNCHED(1); }

  b200_mat M = (b200_mat)calloc(1, sizeof(*M));
  if (!M) return b200_seterr(B200_ERR_MEM, "out of host memory");
  M->kind = 0;
  tmp.release(ai); tmp.release(d_j); tmp.release(d_a);      /* adopted by the block, also when the call below fails */
  int rc = b200_seq_adopt(&M->A, m, n, hi, ai, d_j, d_a);
  if (rc) { cudaStreamSynchronize(st); b200_seq_free(&M->A); free(M); return rc; }
  B200_CUDA(cudaStreamSynchronize(st));
  trace.mark("column ids + values | row binning (host)");
  if (keep_plan) {
    BatchPlan *P = (BatchPlan *)calloc(1, sizeof(*P));
    if (!P) { b200_mat_destroy(M); return b200_seterr(B200_ERR_MEM, "out of host memory"); }
    P->nvals = nvals; P->nfat = nvals; P->nz = nz; P->d_perm = perm; P->d_segoff = segoff;
    tmp.release(perm); tmp.release(segoff);
    M->batchplan = P;
  }
  *A = M;
  return 0;
}

extern "C" int b200_mat_setvaluesbatch_values(b200_mat A, const double *elem_mats, int on_device)
{
  B200_NEED_INIT();
  if (!A || !elem_mats) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  BatchPlan *P = (BatchPlan *)A->batchplan;
  if (A->kind != 0 || !P) return b200_seterr(B200_ERR_ARG_WRONGSTATE, "matrix was not built by b200_mat_setvaluesbatch with keep_plan");
  cudaStream_t st = g_b200.stream;
  DevBufs tmp(st);
  const double *d_em = elem_mats;
  if (!on_device) {
    double *v;
    B200_CUDA(tmp.get(&v, (size_t)P->nvals));
    B200_CUDA(cudaMemcpyAsync(v, elem_mats, sizeof(double) * (size_t)P->nvals, cudaMemcpyHostToDevice, st));
    d_em = v;
  }
  if (A->A.T) { b200_seq_free(A->A.T); free(A->A.T); A->A.T = NULL; }        /* rebuilt from the new values on demand */
  if (P->nz) { asm_values_kernel<<<grid_for(P->nz, 256), 256, 0, st>>>(P->nz, P->d_segoff, P->d_perm, d_em, A->A.d_a); B200_LAUNCHED(1); }
  if (!on_device) B200_CUDA(cudaStreamSynchronize(st));    /* the staging copy is freed on return */
  return 0;
}

extern "C" int b200_mat_get_csr(b200_mat A, int *i, int *j, double *a)
{
  B200_NEED_INIT();
  if (!A) return b200_seterr(B200_ERR_ARG_NULL, "null matrix");
  if (A->kind != 0 || A->A.nstrips > 1) return b200_seterr(B200_ERR_SUP, "CSR read-back is for sequential, unstripped matrices");
  cudaStream_t st = g_b200.stream;
  const B200Seq *S = &A->A;
  if (i) B200_CUDA(cudaMemcpyAsync(i, S->d_i, sizeof(int) * ((size_t)S->m + 1), cudaMemcpyDeviceToHost, st));
  if (j && S->nz) B200_CUDA(cudaMemcpyAsync(j, S->d_j, sizeof(int) * (size_t)S->nz, cudaMemcpyDeviceToHost, st));
  if (a && S->nz) B200_CUDA(cudaMemcpyAsync(a, S->d_a, sizeof(double) * (size_t)S->nz, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  return 0;
}

/* ================================================================================================ device-side split
 * The local rows of a parallel matrix (m rows x N GLOBAL columns, ascending per row — what b200_mat_setvaluesbatch_rows
 * assembles) split into the diagonal block (columns [cstart, cend) -> local ids) and the compacted off-diagonal block
 * (columns -> positions in the sorted list of distinct off-process columns, `garray`) WITHOUT a trip through the host:
 * MatSetValues_MPIAIJ's routing (mpiaij.c:507-565) + MatSetUpMultiply_MPIAIJ's compaction (mmaij.c:28-64), the step the
 * reference's device assembly does with two COO sorts (mpiaijAssemble.cu:200-460).  No sort here: columns ascend inside a
 * row, so a row's diagonal-block entries are ONE contiguous run found by two bisections; the distinct off-process columns
 * are the set bits of an N-bit map (atomicOr), garray is that map read out in order, and an entry's ghost id is the
 * number of set bits below its column (word prefix + popc) — garray ascending exactly as mmaij.c:50 sorts it.
 * Only the m+1 row offsets of the two blocks and garray (nghost ints) go to the host: the SpMV row binning and the
 * scatter plan are built from them. */
struct b200_split_s {
  int        m, ncols_local, nghost;
  long long  nzA, nzB;
  int       *d_Ai, *d_Aj, *d_Bi, *d_Bj;
  double    *d_Aa, *d_Ba;
  int       *h_Ai, *h_Bi, *h_garray;
};

__device__ __forceinline__ int split_lower_bound(const int *__restrict__ j, int b, int e, int key)
{
  while (b < e) { const int mid = (b + e) >> 1; if (__ldg(&j[mid]) < key) b = mid + 1; else e = mid; }
  return b;
}

/* one warp per row: run bounds, counts, and the off-process columns marked in the bit map */
__global__ void __launch_bounds__(256) split_count_kernel(int m, const int *__restrict__ li, const int *__restrict__ lj, int cstart, int cend,
                                                          int *__restrict__ acnt, int *__restrict__ bcnt, int *__restrict__ runlo, unsigned *bitmap)
{
  const int lane = threadIdx.x & 31, nw = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += nw) {
    const int b = __ldg(&li[r]), e = __ldg(&li[r + 1]);
    const int lo = split_lower_bound(lj, b, e, cstart), hi = split_lower_bound(lj, lo, e, cend);
    if (lane == 0) { acnt[r] = hi - lo; bcnt[r] = (e - b) - (hi - lo); runlo[r] = lo; }
    for (int k = b + lane; k < lo; k += 32) { const int c = __ldg(&lj[k]); atomicOr(&bitmap[c >> 5], 1u << (c & 31)); }
    for (int k = hi + lane; k < e; k += 32) { const int c = __ldg(&lj[k]); atomicOr(&bitmap[c >> 5], 1u << (c & 31)); }
  }
}

__global__ void split_popc_kernel(int nwords, const unsigned *__restrict__ bitmap, int *__restrict__ wcnt)
{
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += gridDim.x * blockDim.x) wcnt[w] = __popc(bitmap[w]);
}

__global__ void split_garray_kernel(int nwords, const unsigned *__restrict__ bitmap, const int *__restrict__ wpre, int *__restrict__ garray)
{
  for (int w = blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += gridDim.x * blockDim.x) {
    unsigned bits = bitmap[w];
    int p = wpre[w];
    while (bits) { const int t = __ffs(bits) - 1; bits &= bits - 1; garray[p++] = w * 32 + t; }
  }
}

/* one warp per row: the diagonal run with local column ids, the two off-diagonal runs with ghost ids, values alongside */
__global__ void __launch_bounds__(256) split_fill_kernel(int m, const int *__restrict__ li, const int *__restrict__ lj, const double *__restrict__ la,
                                                         int cstart, const int *__restrict__ runlo, const int *__restrict__ Ai, const int *__restrict__ Bi,
                                                         const unsigned *__restrict__ bitmap, const int *__restrict__ wpre,
                                                         int *__restrict__ Aj, double *__restrict__ Aa, int *__restrict__ Bj, double *__restrict__ Ba)
{
  const int lane = threadIdx.x & 31, nw = (gridDim.x * blockDim.x) >> 5;
  for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < m; r += nw) {
    const int b = __ldg(&li[r]), e = __ldg(&li[r + 1]), lo = __ldg(&runlo[r]);
    const int pa = __ldg(&Ai[r]), na = __ldg(&Ai[r + 1]) - pa, pb = __ldg(&Bi[r]), hi = lo + na;
    for (int k = lane; k < na; k += 32) { Aj[pa + k] = __ldg(&lj[lo + k]) - cstart; Aa[pa + k] = __ldg(&la[lo + k]); }
    const int nb = (e - b) - na, nleft = lo - b;
    for (int k = lane; k < nb; k += 32) {
      const int src = k < nleft ? b + k : hi + (k - nleft);
      const int c = __ldg(&lj[src]);
      Bj[pb + k] = __ldg(&wpre[c >> 5]) + __popc(__ldg(&bitmap[c >> 5]) & ((1u << (c & 31)) - 1u));
      Ba[pb + k] = __ldg(&la[src]);
    }
  }
}

extern "C" int b200_split_destroy(b200_split S)
{
  if (!S) return 0;
  cudaFree(S->d_Ai); cudaFree(S->d_Aj); cudaFree(S->d_Aa); cudaFree(S->d_Bi); cudaFree(S->d_Bj); cudaFree(S->d_Ba);
  free(S->h_Ai); free(S->h_Bi); free(S->h_garray);
  free(S);
  return 0;
}

extern "C" int b200_mat_split_device(b200_mat L, int cstart, int cend, int *nghost_out, b200_split *out)
{
  B200_NEED_INIT();
  if (!L || !out) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *out = NULL;
  if (L->kind != 0 || L->A.nstrips > 1 || L->A.compressed) return b200_seterr(B200_ERR_ARG_WRONG, "the local rows must be one uncompressed sequential block");
  const B200Seq *A = &L->A;
  const int m = A->m, N = A->n;
  if (cstart < 0 || cend < cstart || cend > N) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "column range [%d,%d) outside [0,%d)", cstart, cend, N);
  cudaStream_t st = g_b200.stream;
  DevBufs tmp(st);
  const int nwords = (N + 31) / 32;
  int *acnt, *bcnt, *runlo, *wcnt, *wpre;
  unsigned *bitmap;
  b200_split S = (b200_split)calloc(1, sizeof(*S));
  if (!S) return b200_seterr(B200_ERR_MEM, "out of host memory");
  S->m = m; S->ncols_local = cend - cstart;
#define SPLIT_FAIL(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { b200_split_destroy(S); return b200_cuda_err(e__, #call, __FILE__, __LINE__); } } while (0)
  SPLIT_FAIL(tmp.get(&acnt, (size_t)m + 1));
  SPLIT_FAIL(tmp.get(&bcnt, (size_t)m + 1));
  SPLIT_FAIL(tmp.get(&runlo, (size_t)m + 1));
  SPLIT_FAIL(tmp.get(&bitmap, (size_t)nwords + 1));
  SPLIT_FAIL(tmp.get(&wcnt, (size_t)nwords + 1));
  SPLIT_FAIL(tmp.get(&wpre, (size_t)nwords + 1));
  SPLIT_FAIL(tmp.get(&S->d_Ai, (size_t)m + 1)); tmp.release(S->d_Ai);
  SPLIT_FAIL(tmp.get(&S->d_Bi, (size_t)m + 1)); tmp.release(S->d_Bi);
  SPLIT_FAIL(cudaMemsetAsync(bitmap, 0, sizeof(unsigned) * ((size_t)nwords + 1), st));
  const int gw = m > 0 ? (m + 7) / 8 : 1;                 /* 8 warps per CTA */
  split_count_kernel<<<gw > B200_NSM * 16 ? B200_NSM * 16 : gw, 256, 0, st>>>(m, A->d_i, A->d_j, cstart, cend, acnt, bcnt, runlo, bitmap);
  split_popc_kernel<<<grid_for(nwords, 256), 256, 0, st>>>(nwords, bitmap, wcnt);
  g_b200.launches += 2;
  int rc = dev_exclusive_scan(m, acnt, S->d_Ai, tmp, st);
  if (!rc) rc = dev_exclusive_scan(m, bcnt, S->d_Bi, tmp, st);
  if (!rc) rc = dev_exclusive_scan(nwords, wcnt, wpre, tmp, st);
  if (rc) { b200_split_destroy(S); return rc; }
  S->h_Ai = (int *)malloc(sizeof(int) * ((size_t)m + 1));
  S->h_Bi = (int *)malloc(sizeof(int) * ((size_t)m + 1));
  if (!S->h_Ai || !S->h_Bi) { b200_split_destroy(S); return b200_seterr(B200_ERR_MEM, "out of host memory"); }
  int ng = 0;
  SPLIT_FAIL(cudaMemcpyAsync(S->h_Ai, S->d_Ai, sizeof(int) * ((size_t)m + 1), cudaMemcpyDeviceToHost, st));
  SPLIT_FAIL(cudaMemcpyAsync(S->h_Bi, S->d_Bi, sizeof(int) * ((size_t)m + 1), cudaMemcpyDeviceToHost, st));
  SPLIT_FAIL(cudaMemcpyAsync(&ng, wpre + nwords, sizeof(int), cudaMemcpyDeviceToHost, st));
  SPLIT_FAIL(cudaStreamSynchronize(st));
  S->nzA = S->h_Ai[m]; S->nzB = S->h_Bi[m]; S->nghost = ng;
  int *d_garray;
  SPLIT_FAIL(tmp.get(&d_garray, (size_t)ng + 1));
  SPLIT_FAIL(tmp.get(&S->d_Aj, (size_t)S->nzA + 8)); tmp.release(S->d_Aj);
  SPLIT_FAIL(tmp.get(&S->d_Aa, (size_t)S->nzA + 8)); tmp.release(S->d_Aa);
  SPLIT_FAIL(tmp.get(&S->d_Bj, (size_t)S->nzB + 8)); tmp.release(S->d_Bj);
  SPLIT_FAIL(tmp.get(&S->d_Ba, (size_t)S->nzB + 8)); tmp.release(S->d_Ba);
  split_garray_kernel<<<grid_for(nwords, 256), 256, 0, st>>>(nwords, bitmap, wpre, d_garray);
  split_fill_kernel<<<gw > B200_NSM * 16 ? B200_NSM * 16 : gw, 256, 0, st>>>(m, A->d_i, A->d_j, A->d_a, cstart, runlo, S->d_Ai, S->d_Bi, bitmap, wpre,
                                                                             S->d_Aj, S->d_Aa, S->d_Bj, S->d_Ba);
  g_b200.launches += 2;
  S->h_garray = (int *)malloc(sizeof(int) * ((size_t)ng + 1));
  if (!S->h_garray) { b200_split_destroy(S); return b200_seterr(B200_ERR_MEM, "out of host memory"); }
  SPLIT_FAIL(cudaMemcpyAsync(S->h_garray, d_garray, sizeof(int) * (size_t)ng, cudaMemcpyDeviceToHost, st));
  SPLIT_FAIL(cudaStreamSynchronize(st));
  SPLIT_FAIL(cudaGetLastError());
#undef SPLIT_FAIL
  if (nghost_out) *nghost_out = ng;
  *out = S;
  return 0;
}

extern "C" int b200_split_get(b200_split S, long long *nzA, long long *nzB, int *nghost, int *garray_host)
{
  if (!S) return b200_seterr(B200_ERR_ARG_NULL, "null split");
  if (nzA) *nzA = S->nzA;
  if (nzB) *nzB = S->nzB;
  if (nghost) *nghost = S->nghost;
  if (garray_host && S->nghost) memcpy(garray_host, S->h_garray, sizeof(int) * (size_t)S->nghost);
  return 0;
}

/* b200_mat_create_mpiaij with the two blocks taken from the split (consumed: S is destroyed, also on failure) */
extern "C" int b200_mat_create_mpiaij_split(b200_split S, int N,
                                            int nto, const int *to_procs, const int *to_starts, const int *to_indices,
                                            int nfrom, const int *from_procs, const int *from_starts, b200_mat *Aout)
{
  B200_NEED_INIT();
  if (!S || !Aout) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *Aout = NULL;
  b200_mat M = (b200_mat)calloc(1, sizeof(*M));
  if (!M) { b200_split_destroy(S); return b200_seterr(B200_ERR_MEM, "out of host memory"); }
  M->kind = 1;
  int rc = b200_seq_adopt(&M->A, S->m, S->ncols_local, S->h_Ai, S->d_Ai, S->d_Aj, S->d_Aa);
  S->d_Ai = NULL; S->d_Aj = NULL; S->d_Aa = NULL;          /* owned by the block from here on, also on failure */
  if (!rc) {
    rc = b200_seq_adopt(&M->B, S->m, S->nghost, S->h_Bi, S->d_Bi, S->d_Bj, S->d_Ba);
    S->d_Bi = NULL; S->d_Bj = NULL; S->d_Ba = NULL;
  }
  if (rc) { b200_mat_destroy(M); b200_split_destroy(S); return rc; }
  rc = b200_mpiaij_finish(M, N, S->nghost, S->h_Bi, S->h_garray, nto, to_procs, to_starts, to_indices, nfrom, from_procs, from_starts, Aout);
  b200_split_destroy(S);
  return rc;
}
