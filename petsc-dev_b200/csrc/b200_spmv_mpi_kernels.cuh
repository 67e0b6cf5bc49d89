/* b200_spmv_mpi_kernels.cuh — MatMult_MPIAIJ (src/mat/impls/aij/mpi/mpiaij.c:1102-1116) as ONE kernel:
 *
 *     VecScatterBegin | y = A_d x_local | VecScatterEnd | y += B lvec
 *
 * The persistent TMA-fed tile walk of spmv_tile_kernel (b200_spmv_kernels.cuh), plus
 *   phase 0  the first CTAs gather x[to_indices] and store it straight into the neighbours' ghost vectors over NVLink
 *            (Pack_1 + send + UnPack_1 fused, as pack_put_kernel does stand-alone), one system-scope fence per CTA, and the
 *            last of them raises the arrival flag on every neighbour.  Nothing in phase 0 waits on the current exchange of
 *            any rank (only on credits of exchanges two steps back), so no cycle of waits can form between ranks;
 *   tiles    the sequential kernel's walk over the diagonal block, unchanged (same registers, same 8 CTAs per SM);
 *   phase 2  the CTA then waits (ld.acquire.sys) for the arrival flags of this exchange — the ghost values landed while
 *            the tiles streamed — and revisits those of ITS tiles that hold rows with off-diagonal entries: each such
 *            row's lane re-reads the z it wrote and CONTINUES the sum over the row's off-diagonal entries (ghost values
 *            read with ld.global.cg, they were written by another GPU).  That is the association of MatMultAdd_SeqAIJ
 *            started from y_i (aij.c:1338): with one lane per row the result is bit-identical to the reference's
 *            two-step product.  Only the boundary rows (under 1 % for a slab of a 3-D grid) are touched twice.
 * No second launch, no event join, no wait kernel.  Used when the off-diagonal block is thin (stencil-like operators);
 * matrices whose off-diagonal block carries most of the entries keep the two-product path of b200_comm.cu. */
#ifndef B200_SPMV_MPI_KERNELS_CUH
#define B200_SPMV_MPI_KERNELS_CUH
#include "b200_spmv_kernels.cuh"
#include "b200_p2p_dev.cuh"
#include "b200_ksp_state.h"

struct MpiSpmvArgs {
  int           ntiles;
  const int4   *tiles;        /* {row_begin, row_end, k_begin, k_end} of the diagonal block (the sequential kernel's table)    */
  const int    *bt_start, *bt_list;   /* per CTA of the launch grid: which tiles of its walk hold rows with off-diagonal
                                         entries ("boundary tiles"), bt_list[bt_start[c] .. bt_start[c+1])                */
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

#ifndef FUSED_CTAS_PER_SM
#define FUSED_CTAS_PER_SM 8        /* 32 registers, like the sequential kernel: 2048 resident lanes per SM */
#endif
template <int G, bool ADD, bool DOT, int THREADS, int TILE_NNZ>
__global__ void __launch_bounds__(THREADS, FUSED_CTAS_PER_SM)
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

  if (tid == 0) {
    b200_mbar_init(&bars[0], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto issue = [&](int4 d) {                        /* thread 0 only */
    const int ka = d.z & ~3, kb = (d.w + 3) & ~3;
    const uint32_t n = (uint32_t)(kb - ka);
    if (n == 0) { b200_mbar_expect_tx(&bars[0], 0); return; }
    const uint64_t pol = b200_policy_evict_first();  /* made where it is used: two registers less across the tile loop */
    b200_mbar_expect_tx(&bars[0], n * 12u);
    b200_bulk_g2s(s_a, a.aa + ka, n * 8u, &bars[0], pol);
    b200_bulk_g2s(s_j, a.aj + ka, n * 4u, &bars[0], pol);
  };
  if (tid == 0 && t < ntiles) issue(__ldg(&a.tiles[t]));        /* the first tile streams in while phase 0 runs */

  /* ---- phase 0: VecScatterBegin ----
   * (which CTAs pack makes no measurable difference: the first npack — which carry one tile more when the tile count is not
   * a multiple of the grid — 0.2530 ms per 256^3 product on 2 GPUs, the last npack 0.2540 ms) */
  const int pk = (int)blockIdx.x < a.npack ? (int)blockIdx.x : -1;
  if (pk >= 0) {
    const PeerPlanDev &plan = a.plan;
    /* credits first (stream order: this rank's previous product, which read the other parity's buffer, is complete) */
    if (pk == 0 && tid < plan.nfrom) st_release_sys(plan.rcredit[tid], a.epoch - 1);
    if (tid <= plan.nto) s_starts[tid] = plan.to_starts[tid];
    if (tid < plan.nto) {
      s_dst[tid] = plan.dst[a.parity * plan.nto + tid];
      { const long long t0 = clock64(); while (ld_acquire_sys(plan.lcredit[tid]) + 2 < a.epoch) b200_spin_guard(t0); }
    }
    __syncthreads();
    int k = 0;
    for (int i = pk * THREADS + tid; i < a.nsend; i += a.npack * THREADS) {
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
  int4 d_nxt = make_int4(0, 0, 0, 0);
  if (t + stride < ntiles) d_nxt = __ldg(&a.tiles[t + stride]);
  int b_cur = 0, e_cur = 0, row_cur = -1;
  double y_cur = 0.0;
  /* b, e are positions INSIDE the tile's stage (entry number minus the tile's aligned first entry), so the sums below
     index the two shared arrays directly: no per-tile base pointers to keep in registers */
  auto load_row = [&](const int4 &d, int &b, int &e, int &row, double &y0) {
    row = -1; b = e = 0;
    if (r < d.y - d.x) {
      row = d.x + r;
      const int ka = d.z & ~3;
      b = __ldg(&a.ai[row]) - ka; e = __ldg(&a.ai[row + 1]) - ka;
      if (ADD && g == 0) y0 = a.yin[row];
    }
  };
  if (t < ntiles) load_row(__ldg(&a.tiles[t]), b_cur, e_cur, row_cur, y_cur);

  double dot = 0.0;
  for (int it = 0; t < ntiles; it++, t += stride) {
    const uint32_t parity = (uint32_t)(it & 1);
    int b_nxt = 0, e_nxt = 0, row_nxt = -1;
    double y_nxt = 0.0;
    int4 d_nn = make_int4(0, 0, 0, 0);
    if (t + stride < ntiles) load_row(d_nxt, b_nxt, e_nxt, row_nxt, y_nxt);
    if (t + 2 * stride < ntiles) d_nn = __ldg(&a.tiles[t + 2 * stride]);

    b200_mbar_wait(&bars[0], parity);
    if (G == 1) {
      if (row_cur >= 0) {
        double sum = ADD ? y_cur : 0.0;
#pragma unroll 4
        for (int k = b_cur; k < e_cur; k++) sum += s_a[k] * __ldg(&a.x[s_j[k]]);
        a.z[row_cur] = sum;
        if (DOT) {
          const double xd = __ldg(&a.dotx[row_cur]);
          dot += xd * sum;
          if (ADD) dot -= xd * y_cur;
        }
      }
    } else {
      double sum = 0.0;
#pragma unroll 4
      for (int k = b_cur + g; k < e_cur; k += G) sum += s_a[k] * __ldg(&a.x[s_j[k]]);
      __syncwarp();
#pragma unroll
      for (int o = G / 2; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o, G);
      if (g == 0 && row_cur >= 0) {
        if (ADD) sum = y_cur + sum;
        if (DOT) { const double xd = __ldg(&a.dotx[row_cur]); dot += xd * sum; if (ADD) dot -= xd * y_cur; }
        a.z[row_cur] = sum;
      }
    }
    __syncthreads();                                  /* everyone is done with the stage */
    if (tid == 0) {
      const int tt = t + stride;
      if (tt < ntiles) issue(d_nxt);
    }
    d_nxt = d_nn;
    b_cur = b_nxt; e_cur = e_nxt; row_cur = row_nxt; y_cur = y_nxt;
  }

  /* ---- VecScatterEnd + y += B lvec on the rows that have off-diagonal entries ----
   * bt_list[bt_start[c] .. bt_start[c+1]) are the tiles of THIS CTA's walk that hold such rows (built with the plan for
   * exactly this grid), so every row is finished by the thread that wrote its diagonal part above: z[row] is re-read and
   * the row's lane CONTINUES the sum over the off-diagonal entries — the association of MatMultAdd_SeqAIJ started from
   * y_i (aij.c:1338), bit-identical to the reference's two-step product with one lane per row.  Kept out of the tile loop:
   * in it, the call cost the loop 8 registers, i.e. two of the eight resident CTAs per SM (0.262 ms against 0.250 ms per
   * 256^3 product, profiles/r02_ctas_per_sm.txt). */
  {
    const int q0 = __ldg(&a.bt_start[blockIdx.x]), q1 = __ldg(&a.bt_start[blockIdx.x + 1]);
    if (q1 > q0) {
      /* the ghost values of this exchange are in lvec once every sender's flag shows the epoch */
      if (tid < a.plan.nfrom) { const long long t0 = clock64(); while (ld_acquire_sys(a.plan.lflag[tid]) < a.epoch) b200_spin_guard(t0); }
      __syncthreads();
    }
    for (int q = q0; q < q1; q++) {
      const int4 d = __ldg(&a.tiles[__ldg(&a.bt_list[q])]);
      int bb = 0, be = 0, row = -1;
      if (r < d.y - d.x) { row = d.x + r; bb = __ldg(&a.bi[row]); be = __ldg(&a.bi[row + 1]); }
      if (G == 1) {
        if (be > bb) {
          const double z0 = a.z[row];
          double sum = z0;
          for (int k = bb; k < be; k++) sum += __ldg(&a.ba[k]) * __ldcg(&a.lvec[__ldg(&a.bj[k])]);   /* ghost values: written by another GPU */
          a.z[row] = sum;
          if (DOT) { const double xd = __ldg(&a.dotx[row]); dot += xd * sum; dot -= xd * z0; }
        }
      } else {
        double sb = 0.0;
        for (int k = bb + g; k < be; k += G) sb += __ldg(&a.ba[k]) * __ldcg(&a.lvec[__ldg(&a.bj[k])]);
        __syncwarp();
#pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) sb += __shfl_xor_sync(0xffffffffu, sb, o, G);
        if (g == 0 && be > bb) {
          const double z0 = a.z[row], sum = z0 + sb;   /* diagonal block first, then the off-diagonal block */
          a.z[row] = sum;
          if (DOT) { const double xd = __ldg(&a.dotx[row]); dot += xd * sum; dot -= xd * z0; }
        }
      }
    }
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
