/* b200_spmv_mpi_kernels.cuh — MatMult_MPIAIJ (src/mat/impls/aij/mpi/mpiaij.c:1102-1116) as ONE kernel:
 *
 *     VecScatterBegin | y = A_d x_local | VecScatterEnd | y += B lvec
 *
 * The persistent TMA-fed tile walk of spmv_tile_kernel (b200_spmv_kernels.cuh), plus
 *   phase 0  the first CTAs gather x[to_indices] and store it straight into the neighbours' ghost vectors over NVLink
 *            (Pack_1 + send + UnPack_1 fused, as pack_put_kernel does stand-alone), one system-scope fence per CTA, and the
 *            last of them raises the arrival flag on every neighbour.  Nothing in phase 0 waits on the current exchange of
 *            any rank (only on credits of exchanges two steps back), so no cycle of waits can form between ranks;
 *   tiles    interior tiles first.  Tiles holding rows with off-diagonal entries ("boundary tiles") come last: the first
 *            time a CTA reaches one it waits (ld.acquire.sys) for the arrival flags of this exchange — by then the ghost
 *            values have normally landed long ago — and each row's lane CONTINUES its register sum over the row's
 *            off-diagonal entries (ghost values read with ld.global.cg, they were written by another GPU).  That is the
 *            association of MatMultAdd_SeqAIJ started from y_i (aij.c:1338): with one lane per row the result is
 *            bit-identical to the reference's two-step product, and y is written once instead of written, re-read and
 *            re-written.
 * No second launch, no event join, no wait kernel.  Used when the off-diagonal block is thin (stencil-like operators);
 * matrices whose off-diagonal block carries most of the entries keep the two-product path of b200_comm.cu. */
#ifndef B200_SPMV_MPI_KERNELS_CUH
#define B200_SPMV_MPI_KERNELS_CUH
#include "b200_spmv_kernels.cuh"
#include "b200_p2p_dev.cuh"
#include "b200_ksp_state.h"

struct MpiSpmvArgs {
  int           ntiles;
  const int4   *tiles;        /* {row_begin, row_end, k_begin, k_end} of the diagonal block, interior tiles first; the sign
                                 bit of k_begin marks a tile with off-diagonal entries in some row ("boundary tile")     */
  const int    *ai, *aj;  const double *aa;       /* diagonal block                                                     */
  const int    *bi, *bj;  const double *ba;       /* off-diagonal block, full row offsets [m+1], columns index lvec     */
  const double *x, *yin;  double *z;
  const double *lvec;         /* this exchange's ghost vector (local arena)                                              */
  double       *dotpart;  const double *dotx;
  /* exchange */
  int           nsend, npack; /* entries to send; CTAs that take part in phase 0 (>= 1 when this rank has any neighbour) */
  const int    *to_idx;
  PeerPlanDev   plan;
  int           parity;
  unsigned long long epoch;
  unsigned     *ticket;
  /* CG (DOT only, cg != NULL): the last CTA to finish sums the partials of p.w in a fixed order, all-reduces the sum over
   * peer memory and runs the scalar step that follows it (cg.c:207-220) — no kernel of its own for that */
  CgState      *cg;
  CgAr          ar;
  unsigned     *ticket_dot;
};

/* The off-diagonal entries of one row, lane g of G striding them, added to `sum` in order (kept out of line: it runs for
 * the boundary tiles only and must not cost the streaming loop registers). */
template <int G>
__device__ __noinline__ double b200_offdiag_row(const int *__restrict__ bi, const int *__restrict__ bj, const double *__restrict__ ba,
                                                const double *lvec, int row, int g, double sum)
{
  const int bb = __ldg(&bi[row]), be = __ldg(&bi[row + 1]);
  for (int k = bb + g; k < be; k += G) sum += __ldg(&ba[k]) * __ldcg(&lvec[__ldg(&bj[k])]);   /* aij.c:1338, continued */
  return sum;
}

/* 6 resident CTAs per SM = the 40 registers the kernel wants.  Holding it to the sequential kernel's 8 CTAs (32 registers)
 * spills ~50 bytes inside the tile loop and was measured 1.7x SLOWER on 2 GPUs (0.441 ms against 0.258 ms per 256^3
 * product, profiles/r02_bench_n2_*.json), so that variant is gone. */
template <int G, bool ADD, bool DOT, int THREADS, int TILE_NNZ>
__global__ void __launch_bounds__(THREADS, 6)
spmv_mpiaij_kernel(const MpiSpmvArgs a)
{
  constexpr int SE = TILE_NNZ + SPMV_ALIGN_PAD;
  extern __shared__ __align__(128) unsigned char spmv_smem[];
  double   *s_a  = reinterpret_cast<double *>(spmv_smem);
  int      *s_j  = reinterpret_cast<int *>(spmv_smem + (size_t)SE * 8);
  uint64_t *bars = reinterpret_cast<uint64_t *>(spmv_smem + (size_t)SE * 12);
  __shared__ double s_red[THREADS / 32];
  __shared__ int    s_starts[B200_MAX_RANKS + 1];
  __shared__ double *s_dst[B200_MAX_RANKS];
  __shared__ bool   s_last;

  const int tid = threadIdx.x, r = tid / G, g = tid % G;
  const int stride = gridDim.x;
  const int ntiles = a.ntiles;
  int t = blockIdx.x;
  uint64_t pol = 0;

  if (tid == 0) {
    b200_mbar_init(&bars[0], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    pol = b200_policy_evict_first();
  }
  __syncthreads();

  auto issue = [&](int4 d) {                        /* thread 0 only */
    const int ka = (d.z & 0x7fffffff) & ~3, kb = (d.w + 3) & ~3;
    const uint32_t n = (uint32_t)(kb - ka);
    if (n == 0) { b200_mbar_expect_tx(&bars[0], 0); return; }
    b200_mbar_expect_tx(&bars[0], n * 12u);
    b200_bulk_g2s(s_a, a.aa + ka, n * 8u, &bars[0], pol);
    b200_bulk_g2s(s_j, a.aj + ka, n * 4u, &bars[0], pol);
  };
  if (tid == 0 && t < ntiles) issue(__ldg(&a.tiles[t]));        /* the first tile streams in while phase 0 runs */

  /* ---- phase 0: VecScatterBegin ---- */
  if ((int)blockIdx.x < a.npack) {
    const PeerPlanDev &plan = a.plan;
    /* credits first (stream order: this rank's previous product, which read the other parity's buffer, is complete) */
    if (blockIdx.x == 0 && tid < plan.nfrom) st_release_sys(plan.rcredit[tid], a.epoch - 1);
    if (tid <= plan.nto) s_starts[tid] = plan.to_starts[tid];
    if (tid < plan.nto) {
      s_dst[tid] = plan.dst[a.parity * plan.nto + tid];
      { const long long t0 = clock64(); while (ld_acquire_sys(plan.lcredit[tid]) + 2 < a.epoch) b200_spin_guard(t0); }
    }
    __syncthreads();
    int k = 0;
    for (int i = blockIdx.x * THREADS + tid; i < a.nsend; i += a.npack * THREADS) {
      while (i >= s_starts[k + 1]) k++;
      s_dst[k][i - s_starts[k]] = __ldg(&a.x[__ldg(&a.to_idx[i])]);
    }
    __syncthreads();
    if (tid == 0) {
      __threadfence_system();                       /* cumulative over the CTA's stores through the barrier */
      s_last = (atomicAdd(a.ticket, 1u) == (unsigned)a.npack - 1u);
    }
    __syncthreads();
    if (s_last) {
      if (tid < plan.nto) { __threadfence_system(); st_release_sys(plan.rflag[tid], a.epoch); }
      if (tid == 0) *a.ticket = 0;
    }
  }

  /* ---- tiles ---- */
  int4 d_cur = make_int4(0, 0, 0, 0), d_nxt = make_int4(0, 0, 0, 0);
  if (t < ntiles) d_cur = __ldg(&a.tiles[t]);
  if (t + stride < ntiles) d_nxt = __ldg(&a.tiles[t + stride]);
  int b_cur = 0, e_cur = 0, row_cur = -1;
  double y_cur = 0.0;
  auto load_row = [&](const int4 &d, int &b, int &e, int &row, double &y0) {
    row = -1; b = e = 0;
    if (r < d.y - d.x) {
      row = d.x + r;
      b = __ldg(&a.ai[row]); e = __ldg(&a.ai[row + 1]);
      if (ADD && g == 0) y0 = a.yin[row];
    }
  };
  if (t < ntiles) load_row(d_cur, b_cur, e_cur, row_cur, y_cur);

  bool arrived = false;
  double dot = 0.0;
  for (int it = 0; t < ntiles; it++, t += stride) {
    const uint32_t parity = (uint32_t)(it & 1);
    int b_nxt = 0, e_nxt = 0, row_nxt = -1;
    double y_nxt = 0.0;
    int4 d_nn = make_int4(0, 0, 0, 0);
    if (t + stride < ntiles) load_row(d_nxt, b_nxt, e_nxt, row_nxt, y_nxt);
    if (t + 2 * stride < ntiles) d_nn = __ldg(&a.tiles[t + 2 * stride]);
    const bool f_cur = d_cur.z < 0;

    if (f_cur && !arrived) {
      /* VecScatterEnd: the ghost values of this exchange are in lvec once every sender's flag shows the epoch */
      if (tid < a.plan.nfrom) { const long long t0 = clock64(); while (ld_acquire_sys(a.plan.lflag[tid]) < a.epoch) b200_spin_guard(t0); }
      __syncthreads();
      arrived = true;
    }
    b200_mbar_wait(&bars[0], parity);
    {
      const int ka = (d_cur.z & 0x7fffffff) & ~3;
      const double *sa = s_a - ka;                   /* indexed by the global entry number */
      const int    *sj = s_j - ka;
      if (G == 1) {
        if (row_cur >= 0) {
          double sum = ADD ? y_cur : 0.0;
#pragma unroll 4
          for (int k = b_cur; k < e_cur; k++) sum += sa[k] * __ldg(&a.x[sj[k]]);
          if (f_cur) sum = b200_offdiag_row<1>(a.bi, a.bj, a.ba, a.lvec, row_cur, 0, sum);   /* boundary tiles only */
          a.z[row_cur] = sum;
          if (DOT) {
            const double xd = __ldg(&a.dotx[row_cur]);
            dot += xd * sum;
            if (ADD) dot -= xd * y_cur;
          }
        }
      } else {
        double sum = 0.0, sb = 0.0;
#pragma unroll 4
        for (int k = b_cur + g; k < e_cur; k += G) sum += sa[k] * __ldg(&a.x[sj[k]]);
        if (f_cur && row_cur >= 0) sb = b200_offdiag_row<G>(a.bi, a.bj, a.ba, a.lvec, row_cur, g, 0.0);
        __syncwarp();
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) { sum += __shfl_xor_sync(0xffffffffu, sum, o, G); sb += __shfl_xor_sync(0xffffffffu, sb, o, G); }
        if (g == 0 && row_cur >= 0) {
          if (ADD) sum = y_cur + sum;
          sum = sum + sb;                             /* diagonal block first, then the off-diagonal block */
          if (DOT) { const double xd = __ldg(&a.dotx[row_cur]); dot += xd * sum; if (ADD) dot -= xd * y_cur; }
          a.z[row_cur] = sum;
        }
      }
    }
    __syncthreads();                                  /* everyone is done with the stage */
    if (tid == 0) {
      const int tt = t + stride;
      if (tt < ntiles) issue(d_nxt);
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
      if (lane == 0) a.dotpart[blockIdx.x] = v;
    }
    if (a.cg) {
      if (tid == 0) { __threadfence(); s_last = (atomicAdd(a.ticket_dot, 1u) == gridDim.x - 1u); }
      __syncthreads();
      if (s_last) {                                   /* CTA-uniform */
        __threadfence();
        double u = 0.0;
        for (int i = tid; i < (int)gridDim.x; i += THREADS) u += __ldcg(&a.dotpart[i]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
        if (lane == 0) s_red[wid] = u;
        __syncthreads();
        if (tid == 0) {
          double v = 0.0;
#pragma unroll
          for (int q = 0; q < THREADS / 32; q++) v += s_red[q];
          a.cg->red[3] = v;
          *a.ticket_dot = 0;
        }
        /* collective: every rank runs it whatever cg->reason says */
        if (a.ar.size > 1) b200_p2p_allreduce_cta(&a.cg->red[3], 1, 0, a.ar.seq, a.ar.rank, a.ar.size, a.ar.inbox);
        else __syncthreads();
        if (tid == 0 && a.cg->reason == 0) cg_after_dot_body(a.cg);
      }
    }
  }
}

#endif
