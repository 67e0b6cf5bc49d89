/* b200_spmv_kernels.cuh — the sm_100a CSR SpMV kernels (device code only; hosts: b200_spmv.cu, tools/).
 *
 * spmv_tile_kernel: persistent CTAs walk "row tiles" (contiguous runs of rows, <= TILE_NNZ entries, <= THREADS/G rows).
 *   - the tile's values and column ids are brought into shared memory by the TMA engine (1-D cp.async.bulk, completion
 *     on an mbarrier, L2 evict-first policy so the matrix stream does not push x out of L2), STAGES tiles deep, so the
 *     HBM stream is continuous and never touches the LSU/L1 data pipe (ncu on the first version of this kernel showed
 *     that pipe, not DRAM, was the limiter: 69 % l1tex wavefronts at 53 % of HBM peak);
 *   - G lanes own one row.  G = 1: the lane walks its row LEFT TO RIGHT starting from 0 (MatMult) or y[i] (MatMultAdd) —
 *     the reference's association exactly (PetscSparseDensePlusDot, src/mat/impls/aij/seq/aij.h:384-386; aij.c:1338),
 *     so with -fmad=false the result is bit-identical to MatMult_SeqAIJ.  Consecutive lanes own consecutive rows: the
 *     x gathers of a warp are (for banded matrices) one contiguous 256-byte run, the shared-memory reads are strided by
 *     the row length (conflict-free for odd lengths).  G > 1: lanes stride the row, fixed xor-shuffle tree at the end
 *     (deterministic; association differs from the reference, well inside the 1e-12 bound);
 *   - row offsets / y of the NEXT tile and the tile descriptor after it are prefetched into registers while the
 *     current tile is computed.
 *   - PARK variant (irregular matrices): a first pass in which the lanes stride the tile's ENTRIES (perfectly balanced,
 *     8 independent x gathers in flight per lane) parks the products in shared memory; the row sums then only read
 *     shared memory, so a long row in the tile delays its lanes by LDS+DADD latencies, not by L2 gathers;
 *   - rows longer than the threshold are cut into chunks of one tile each (descriptor .y = -(slot+1)); the whole CTA
 *     strides a chunk and leaves one partial per chunk; spmv_long_finish_kernel adds a row's partials in chunk order.
 *     The chunks travel through the same TMA pipeline and are dealt round-robin to the CTAs like any other tile. */
#ifndef B200_SPMV_KERNELS_CUH
#define B200_SPMV_KERNELS_CUH
#include <cuda_runtime.h>
#include <stdint.h>
#include "b200_ksp_state.h"     /* CgState + the scalar step the DOT variants can take in their last CTA */

#define SPMV_ALIGN_PAD 8      /* entries of slack per stage: the bulk copies start/end on 4-entry boundaries */

__device__ __forceinline__ uint32_t b200_smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void b200_mbar_init(uint64_t *bar, int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b200_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void b200_mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b200_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void b200_mbar_wait(uint64_t *bar, uint32_t parity)
{
  asm volatile("{\n"
               ".reg .pred P1;\n"
               "LAB_WAIT:\n"
               "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
               "@P1 bra DONE;\n"
               "bra LAB_WAIT;\n"
               "DONE:\n"
               "}" ::"r"(b200_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t b200_policy_evict_first()
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
/* 1-D TMA bulk copy global -> shared; bytes % 16 == 0, both addresses 16-byte aligned */
__device__ __forceinline__ void b200_bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar, uint64_t pol)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(b200_smem_u32(dst)), "l"(src), "r"(bytes), "r"(b200_smem_u32(bar)), "l"(pol) : "memory");
}

template <int THREADS, int TILE_NNZ, int STAGES>
struct SpmvSmem {
  static constexpr int    STAGE_ELEMS = TILE_NNZ + SPMV_ALIGN_PAD;
  static constexpr size_t BYTES = (size_t)STAGES * STAGE_ELEMS * 12 + STAGES * 8 + 64;
};

/* The parked gather pass of irregular matrices loads x with ld.global.cg (L2 only): its L1 hit rate is 2-3 % (ncu,
 * profiles/r02_spmv_plaw_10M_full.txt), so allocating the lines in L1 only costs tag work — measured 0.744 ms against
 * 0.799 ms with ld.global.nc on the 2 M-row power-law sample.  B200_SPMV_GATHER_CG=0 (read once in b200_spmv.cu) restores
 * the ld.global.nc loads for comparison. */
static __constant__ int b200_c_gather_cg;

/* per-launch extras.  cg/ticket: optional tail of the DOT variants — the CG state whose scalar step the kernel's last CTA
 * takes (NULL: none).  park_cap: parked (irregular-matrix) variant with one lane per row — how many entries of a row its own
 * lane sums before the row's warp takes over; <= 0: the lane sums the whole row, i.e. the reference's left-to-right order for
 * every row (what a caller who ASKED for one lane per row gets: b200_spmv_set_tuning(lanes_per_row = 1)). */
struct SpmvCgTail { CgState *cg; unsigned *ticket; int park_cap; };

/* tiles[t] = {row_begin, row_end, k_begin, k_end} in visiting order (ci = offsets in visiting order; rindex maps a
 * visiting position to the row id when the block is stored compressed, else NULL). */
/* the default shape (256 lanes, one stage) is held to 32 registers = 8 resident CTAs per SM in every variant: at 40
 * registers (6 CTAs) the same walk streams 5 % slower (0.262 ms against 0.250 ms per 256^3 product) */
template <int G, bool PARK, bool ADD, bool DOT, int THREADS, int TILE_NNZ, int STAGES>
__global__ void __launch_bounds__(THREADS, (THREADS == 256 && STAGES == 1) ? ((ADD && DOT) ? 6 : 8) : 1)   /* ADD+DOT would spill at 32 */
spmv_tile_kernel(int ntiles, const int4 *__restrict__ tiles, const int *__restrict__ ci, const int *__restrict__ rindex,
                 const int *__restrict__ aj, const double *__restrict__ aa, const double *__restrict__ x,
                 const double *yin, double *z, double *__restrict__ dotpart, double *__restrict__ longpart,
                 const double *__restrict__ dotx, const SpmvCgTail tail)
{
  /* DOT: the CTA leaves its share of dotx . (z - yin) in dotpart (dotx == x, yin == 0 for the plain MatMult; for the
   * off-diagonal MatMultAdd dotx is the LOCAL vector, so that diag + off-diag partials add up to x . (A x)) */
  constexpr int SE = TILE_NNZ + SPMV_ALIGN_PAD;
  extern __shared__ __align__(128) unsigned char spmv_smem[];
  double   *s_a  = reinterpret_cast<double *>(spmv_smem);
  int      *s_j  = reinterpret_cast<int *>(spmv_smem + (size_t)STAGES * SE * 8);
  uint64_t *bars = reinterpret_cast<uint64_t *>(spmv_smem + (size_t)STAGES * SE * 12);
  __shared__ double s_red[THREADS / 32];

  const int tid = threadIdx.x, r = tid / G, g = tid % G;
  const int stride = gridDim.x;
  int t = blockIdx.x;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) b200_mbar_init(&bars[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](int s, int4 d) {                 /* thread 0 only */
    const int ka = d.z & ~3, kb = (d.w + 3) & ~3;
    const uint32_t n = (uint32_t)(kb - ka);
    if (n == 0) { b200_mbar_expect_tx(&bars[s], 0); return; }
    const uint64_t pol = b200_policy_evict_first();  /* made where it is used: two registers less across the tile loop */
    b200_mbar_expect_tx(&bars[s], n * 12u);
    b200_bulk_g2s(s_a + (size_t)s * SE, aa + ka, n * 8u, &bars[s], pol);
    b200_bulk_g2s(s_j + (size_t)s * SE, aj + ka, n * 4u, &bars[s], pol);
  };

  /* prologue: fill the pipeline */
  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) {
      const int tt = t + s * stride;
      if (tt < ntiles) issue(s, __ldg(&tiles[tt]));
    }
  }

  /* register pipeline of descriptors and row bounds */
  int4 d_cur = make_int4(0, 0, 0, 0), d_nxt = make_int4(0, 0, 0, 0);
  if (t < ntiles) d_cur = __ldg(&tiles[t]);
  if (t + stride < ntiles) d_nxt = __ldg(&tiles[t + stride]);
  int b_cur = 0, e_cur = 0, row_cur = -1;
  double y_cur = 0.0;
  /* b, e are positions INSIDE the tile's stage (entry number minus the tile's aligned first entry): the row sums index
     the stage directly, no per-tile base pointers live in registers (this is what lets the ADD variants fit 32 registers) */
  auto load_row = [&](const int4 &d, int &b, int &e, int &row, double &y0) {
    row = -1; b = e = 0;
    if (r < d.y - d.x) {
      const int p = d.x + r, ka = d.z & ~3;
      b = __ldg(&ci[p]) - ka; e = __ldg(&ci[p + 1]) - ka;
      row = rindex ? __ldg(&rindex[p]) : p;
      if (ADD && g == 0) y0 = yin[row];
    }
  };
  if (t < ntiles) load_row(d_cur, b_cur, e_cur, row_cur, y_cur);

  double dot = 0.0;
  for (int it = 0; t < ntiles; it++, t += stride) {
    const int s = it % STAGES;
    const uint32_t parity = (uint32_t)((it / STAGES) & 1);
    /* prefetch: bounds of the next tile, descriptor of the one after */
    int b_nxt = 0, e_nxt = 0, row_nxt = -1;
    double y_nxt = 0.0;
    int4 d_nn = make_int4(0, 0, 0, 0);
    if (t + stride < ntiles) load_row(d_nxt, b_nxt, e_nxt, row_nxt, y_nxt);
    if (t + 2 * stride < ntiles) d_nn = __ldg(&tiles[t + 2 * stride]);

    b200_mbar_wait(&bars[s], parity);
    {
      double    *sa = s_a + (size_t)s * SE;                  /* indexed by the position inside the stage */
      const int *sj = s_j + (size_t)s * SE;
      const int k0 = d_cur.z - (d_cur.z & ~3), k1 = d_cur.w - (d_cur.z & ~3);   /* the tile's entries inside the stage */
      if (d_cur.y < 0) {
        /* one chunk of a long row: every lane strides the chunk, fixed-shape block tree, partial -> longpart[slot] */
        double p0 = 0.0, p1 = 0.0;
        int k = k0 + tid;
        for (; k + THREADS < k1; k += 2 * THREADS) {
          p0 += sa[k] * __ldg(&x[sj[k]]);
          p1 += sa[k + THREADS] * __ldg(&x[sj[k + THREADS]]);
        }
        if (k < k1) p0 += sa[k] * __ldg(&x[sj[k]]);
        double v = p0 + p1;
        const int lane = tid & 31, wid = tid >> 5;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[wid] = v;
        __syncthreads();
        if (tid == 0) {
          double w = 0.0;
#pragma unroll
          for (int q = 0; q < THREADS / 32; q++) w += s_red[q];
          longpart[-d_cur.y - 1] = w;
        }
      } else {
        if (PARK) {
          /* balanced gather pass: products parked in place of the values, then the row sums read shared memory only */
          if (b200_c_gather_cg) {
#pragma unroll 8
            for (int k = k0 + tid; k < k1; k += THREADS) sa[k] = sa[k] * __ldcg(&x[sj[k]]);
          } else {
#pragma unroll 8
            for (int k = k0 + tid; k < k1; k += THREADS) sa[k] = sa[k] * __ldg(&x[sj[k]]);
          }
          __syncthreads();
        }
        if (G == 1 && PARK) {
          /* Row sums from the parked products.  One lane per row sums left to right — but only the first PARK_CAP entries:
           * with power-law row lengths a tile's longest row (up to the long-row threshold, 256) used to keep 255 lanes
           * waiting at the barrier below for one serial chain of shared-memory loads (ncu, C4: 37 % of the warp samples
           * at the two barriers, 11 % in this loop, 40 % in the gathers).  What a row has beyond PARK_CAP entries its WARP
           * sums together: lanes stride the rest (consecutive addresses, no bank conflicts), a shuffle tree, the row's lane
           * adds it on.  Rows of up to PARK_CAP entries keep the reference's order exactly; so does every row when the
           * caller asked for one lane per row (tail.park_cap == 0). */
          const int PARK_CAP = tail.park_cap > 0 ? tail.park_cap : 0x7fffffff;
          const int lane = tid & 31;
          const int len = e_cur - b_cur;                    /* 0 for lanes without a row */
          double sum = ADD ? y_cur : 0.0;
          {
            const int e1 = b_cur + (len < PARK_CAP ? len : PARK_CAP);
#pragma unroll 4
            for (int k = b_cur; k < e1; k++) sum += sa[k];
          }
          unsigned more = __ballot_sync(0xffffffffu, len > PARK_CAP);
          while (more) {
            const int src = __ffs(more) - 1;
            more &= more - 1;
            const int bb = __shfl_sync(0xffffffffu, b_cur, src) + PARK_CAP, ee = __shfl_sync(0xffffffffu, e_cur, src);
            double part = 0.0;
            for (int k = bb + lane; k < ee; k += 32) part += sa[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
            if (lane == src) sum += part;
          }
          if (row_cur >= 0) {
            z[row_cur] = sum;
            if (DOT) {
              const double xd = __ldg(&dotx[row_cur]);
              dot += xd * sum;
              if (ADD) dot -= xd * y_cur;
            }
          }
        } else if (G == 1) {
          if (row_cur >= 0) {
            double sum = ADD ? y_cur : 0.0;
#pragma unroll 4
            for (int k = b_cur; k < e_cur; k++) sum += sa[k] * __ldg(&x[sj[k]]);
            z[row_cur] = sum;
            if (DOT) {
              const double xd = __ldg(&dotx[row_cur]);
              dot += xd * sum;
              if (ADD) dot -= xd * y_cur;
            }
          }
        } else {
          /* lanes without a row have b == e == 0 and fall through to the (warp-converged) shuffles */
          double sum = 0.0;
          if (PARK) {
#pragma unroll 4
            for (int k = b_cur + g; k < e_cur; k += G) sum += sa[k];
          } else {
#pragma unroll 4
            for (int k = b_cur + g; k < e_cur; k += G) sum += sa[k] * __ldg(&x[sj[k]]);
          }
          __syncwarp();
#pragma unroll
          for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
          if (g == 0 && row_cur >= 0) {
            if (DOT) dot += __ldg(&dotx[row_cur]) * sum;
            if (ADD) sum = y_cur + sum;
            z[row_cur] = sum;
          }
        }
      }
    }
    __syncthreads();                                          /* everyone is done with stage s */
    if (tid == 0) {
      const int tt = t + STAGES * stride;
      if (tt < ntiles) issue(s, __ldg(&tiles[tt]));
    }
    d_cur = d_nxt; d_nxt = d_nn;
    b_cur = b_nxt; e_cur = e_nxt; row_cur = row_nxt; y_cur = y_nxt;
  }

  if (DOT) {
    const int lane = tid & 31, wid = tid >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
    if (lane == 0) s_red[wid] = dot;
    __syncthreads();
    if (wid == 0) {
      double v = (lane < THREADS / 32) ? s_red[lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) dotpart[blockIdx.x] = v;
    }
    if (tail.cg) {
      /* fused CG on one GPU: the last CTA to get here sums the partials of p.w (the order cg_after_dot_kernel uses) and
       * takes the scalar step that follows the product (cg.c:207-220) — one launch less per iteration */
      __shared__ bool s_last;
      if (tid == 0) { __threadfence(); s_last = (atomicAdd(tail.ticket, 1u) == gridDim.x - 1u); }
      __syncthreads();
      if (s_last) {                                          /* CTA-uniform */
        __threadfence();
        double u = 0.0;
        for (int i = tid; i < (int)gridDim.x; i += THREADS) u += __ldcg(&dotpart[i]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
        __syncthreads();
        if (lane == 0) s_red[wid] = u;
        __syncthreads();
        if (tid == 0) {
          double w = 0.0;
#pragma unroll
          for (int q = 0; q < THREADS / 32; q++) w += s_red[q];
          tail.cg->red[3] = w;
          *tail.ticket = 0;
          if (tail.cg->reason == 0) cg_after_dot_body(tail.cg);
        }
      }
    }
  }
}

/* ---- irregular matrices: one WARP per tile, no CTA-wide barrier ------------------------------------------------------
 * The CTA-tile kernel above meets two barriers per tile in its parked form, and a microbenchmark of exactly that form with
 * perfectly balanced rows (tools/gather_peak.cu) tops out at 105-120 G gathers/s against 275 G/s for a kernel that only
 * gathers: the phases serialise inside a CTA.  Here a tile is at most 256 entries / 32 whole rows and belongs to one warp:
 * lanes load ids and values coalesced straight from global memory (streaming loads), keep 8 gathers of x each in flight,
 * leave the products in the warp's 2 KB of shared memory, and after a __syncwarp every row's lane sums its row left to right
 * (up to `cap` entries; what a row has beyond that the warp sums together, as in the parked CTA form).  The 64 warps of an
 * SM drift apart, so the load/store path always has gathers to do.  Rows longer than a tile are cut into 256-entry chunks,
 * one warp and one partial each, summed in chunk order by spmv_long_finish_kernel. */
#define SPMV_WARP_TILE 256
template <bool ADD>
__global__ void __launch_bounds__(256, 6)
spmv_warp_kernel(int ntiles, const int4 *__restrict__ tiles, const int *__restrict__ ci, const int *__restrict__ rindex,
                 const int *__restrict__ aj, const double *__restrict__ aa, const double *__restrict__ x,
                 const double *yin, double *z, double *__restrict__ longpart, int cap)
{
  __shared__ double s_p[8][SPMV_WARP_TILE];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  double *sp = s_p[wid];
  const int nw = gridDim.x * 8;
  const int CAP = cap > 0 ? cap : 0x7fffffff;
  for (int t = blockIdx.x * 8 + wid; t < ntiles; t += nw) {
    const int4 d = __ldg(&tiles[t]);                         /* warp-uniform */
    const int k0 = d.z, n = d.w - d.z;
    /* this lane's row (row tiles only): bounds relative to the tile, fetched before the gathers so that everything is in flight together */
    int b = 0, e = 0, row = -1;
    double y0 = 0.0;
    if (d.y >= 0 && lane < d.y - d.x) {
      const int pos = d.x + lane;
      b = __ldg(&ci[pos]) - k0; e = __ldg(&ci[pos + 1]) - k0;
      row = rindex ? __ldg(&rindex[pos]) : pos;
      if (ADD) y0 = yin[row];
    }
    int c[8];
    double v[8];
#pragma unroll
    for (int u = 0; u < 8; u++) { const int k = lane + 32 * u; c[u] = k < n ? __ldcs(&aj[k0 + k]) : -1; }
#pragma unroll
    for (int u = 0; u < 8; u++) { const int k = lane + 32 * u; v[u] = k < n ? __ldcs(&aa[k0 + k]) : 0.0; }
#pragma unroll
    for (int u = 0; u < 8; u++) v[u] = c[u] >= 0 ? v[u] * __ldcg(&x[c[u]]) : 0.0;
    if (d.y < 0) {
      /* chunk of a long row: fixed-shape tree over the 256 products, partial -> longpart[slot] */
      double s = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) longpart[-d.y - 1] = s;
      continue;
    }
#pragma unroll
    for (int u = 0; u < 8; u++) sp[lane + 32 * u] = v[u];
    __syncwarp();
    const int len = e - b;
    double sum = ADD ? y0 : 0.0;
    {
      const int e1 = b + (len < CAP ? len : CAP);
#pragma unroll 4
      for (int k = b; k < e1; k++) sum += sp[k];
    }
    unsigned more = __ballot_sync(0xffffffffu, len > CAP);
    while (more) {
      const int src = __ffs(more) - 1;
      more &= more - 1;
      const int bb = __shfl_sync(0xffffffffu, b, src) + CAP, ee = __shfl_sync(0xffffffffu, e, src);
      double part = 0.0;
      for (int k = bb + lane; k < ee; k += 32) part += sp[k];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (lane == src) sum += part;
    }
    if (row >= 0) z[row] = sum;
    __syncwarp();                                            /* the next tile overwrites sp */
  }
}

/* long rows: sum the chunk partials of each row in chunk order (fixed association), add y, write z (+ the x.z term) */
template <bool ADD, bool DOT>
__global__ void __launch_bounds__(128)
spmv_long_finish_kernel(int nlong, const int *__restrict__ longrows, const int *__restrict__ longstart,
                        const int *__restrict__ rindex, const double *__restrict__ longpart,
                        const double *__restrict__ dotx, const double *yin, double *z, double *__restrict__ dotpart)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlong) return;
  const int r = __ldg(&longrows[i]);
  const int row = rindex ? __ldg(&rindex[r]) : r;
  double v = 0.0;
  for (int c = __ldg(&longstart[i]); c < __ldg(&longstart[i + 1]); c++) v += longpart[c];
  const double out = ADD ? yin[row] + v : v;
  z[row] = out;
  if (DOT) dotpart[i] = __ldg(&dotx[row]) * v;
}

#endif /* B200_SPMV_KERNELS_CUH */
