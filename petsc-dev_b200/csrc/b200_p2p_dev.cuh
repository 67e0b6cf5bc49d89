/* b200_p2p_dev.cuh — device side of the NVLink peer-memory transport (system-scope loads/stores, the inbox allreduce).
 * Shared by b200_comm.cu (stand-alone allreduce kernel) and b200_ksp.cu (allreduce folded into the CG scalar kernels). */
#ifndef B200_P2P_DEV_CUH
#define B200_P2P_DEV_CUH
#include "b200_internal.h"

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v)
{
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys_f64(double *p, double v)
{
  asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ double ld_relaxed_sys_f64(const double *p)
{
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

/* per-matrix peer plan: where this rank's packed entries go on each neighbour, and the flags/credits around them */
struct PeerPlanDev {
  int nto, nfrom;
  const int *to_starts;                          /* [nto+1]                                                          */
  double *const *dst;                            /* [2][nto]: neighbour k's ghost-vector window for me, per parity   */
  unsigned long long *const *rflag;              /* [nto]  arrival flag on neighbour k (its flag[my rank])           */
  const unsigned long long *const *lcredit;      /* [nto]  local credit written by neighbour k (consumed exchange)   */
  unsigned long long *const *rcredit;            /* [nfrom] credit slot on sender k (its credit[my rank])            */
  const unsigned long long *const *lflag;        /* [nfrom] local arrival flag written by sender k                   */
};

/* Every wait on a peer is bounded: a rank that died (or a collective called out of order) must end in a loud failure
 * (the kernel traps, the next CUDA call reports it), never in a GPU that spins for ever.  ~60 s of SM cycles. */
#define B200_SPIN_LIMIT_CYCLES (120ll * 1000ll * 1000ll * 1000ll)
__device__ __forceinline__ void b200_spin_guard(long long t0)
{
  if (clock64() - t0 > B200_SPIN_LIMIT_CYCLES) __trap();
}

/* Called by ALL threads of one CTA (>= 64 threads).  Thread t < size stores this rank's n values into rank t's inbox
 * slot [seq parity][rank], fences, stores the sequence number, then waits for rank t's slot here; the values are then
 * summed (or maxed) in rank order — bitwise the same on every rank — back into vals[0..n).  n <= P2P_AR_MAXV.
 * Buffer reuse: a rank can be at sequence q+2 only after it has seen every peer's q+1, which a peer stores only after
 * its kernel for q has finished reading — two parities suffice. */
__device__ __forceinline__ void b200_p2p_allreduce_cta(double *vals, int n, int op_max, unsigned long long seq, int rank, int size,
                                                       unsigned long long *const *__restrict__ inbox)
{
  const int t = threadIdx.x;
  const size_t par = (size_t)(seq & 1ull) * B200_MAX_RANKS * P2P_AR_SLOT;
  __syncthreads();                                            /* vals[] may have been written by another thread      */
  if (t < size) {
    unsigned long long *slot = inbox[t] + par + (size_t)rank * P2P_AR_SLOT;
    for (int v = 0; v < n; v++) st_relaxed_sys_f64(reinterpret_cast<double *>(slot) + 1 + v, vals[v]);
    __threadfence_system();
    st_release_sys(slot, seq);
    const unsigned long long *mine = inbox[rank] + par + (size_t)t * P2P_AR_SLOT;
    { const long long t0 = clock64(); while (ld_acquire_sys(mine) != seq) b200_spin_guard(t0); }
  }
  __syncthreads();
  if (t < n) {
    const double *base = reinterpret_cast<const double *>(inbox[rank] + par) + 1 + t;
    double acc = ld_relaxed_sys_f64(base);
    for (int r = 1; r < size; r++) {
      const double w = ld_relaxed_sys_f64(base + (size_t)r * P2P_AR_SLOT);
      acc = op_max ? ((w > acc || w != w) ? w : acc) : acc + w;                    /* fixed rank order             */
    }
    vals[t] = acc;
  }
  __syncthreads();
}

#endif
