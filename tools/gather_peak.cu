// Development tool / evidence: the ceiling of fully divergent 8-byte gathers on this GPU, the bound of the power-law SpMV (C4).
// Every lane of a warp loads a double from its own random 32-byte sector of a window of x (the SpMV's access pattern: hashed
// column ids), column ids streamed coalesced like the matrix' j array.  Reported: gathers per second for windows that are
// L2-resident (one column strip of C4: 20 MB) and not (80 MB), with ld.global.nc and ld.global.cg.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/gather_peak tools/gather_peak.cu && tools/gather_peak
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

template <int MODE, int UNROLL>
__global__ void __launch_bounds__(256) gather_kernel(long long n, const int *__restrict__ idx, const double *__restrict__ x, double *out)
{
  double acc = 0.0;
  const long long stride = (long long)gridDim.x * blockDim.x * UNROLL;
  for (long long i = (long long)blockIdx.x * blockDim.x * UNROLL + threadIdx.x; i < n; i += stride) {
    int c[UNROLL];
    double v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; u++) c[u] = (i + u * 256 < n) ? __ldcs(&idx[i + u * 256]) : 0;
#pragma unroll
    for (int u = 0; u < UNROLL; u++) v[u] = MODE ? __ldcg(&x[c[u]]) : __ldg(&x[c[u]]);
#pragma unroll
    for (int u = 0; u < UNROLL; u++) acc += v[u];
  }
  if (acc == 1.2345e300) out[0] = acc;                  // keep the loads
}

// The SpMV's parked form without its row structure: per 2048-entry tile, column ids and values are staged in shared memory
// (coalesced loads standing in for the TMA copy), barrier, the gather pass sa[k] = sa[k] * x[sj[k]] exactly as in
// spmv_tile_kernel, barrier, then every lane sums 8 parked products (perfectly balanced "rows").  What this loses against
// the pure gather kernel is the price of the shared-memory traffic and the two barriers per tile, not of irregular rows.
template <int MODE>
__global__ void __launch_bounds__(256) parked_kernel(long long n, const int *__restrict__ idx, const double *__restrict__ val, const double *__restrict__ x, double *out)
{
  __shared__ double sa[2048];
  __shared__ int sj[2048];
  double acc = 0.0;
  const long long ntiles = n / 2048;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const long long base = t * 2048;
#pragma unroll
    for (int u = 0; u < 8; u++) { sj[threadIdx.x + u * 256] = __ldcs(&idx[base + threadIdx.x + u * 256]); sa[threadIdx.x + u * 256] = __ldcs(&val[base + threadIdx.x + u * 256]); }
    __syncthreads();
#pragma unroll 8
    for (int k = threadIdx.x; k < 2048; k += 256) sa[k] = sa[k] * (MODE ? __ldcg(&x[sj[k]]) : __ldg(&x[sj[k]]));
    __syncthreads();
    if (MODE == 2) {                                      // conflict-free balanced sums: lane-strided
#pragma unroll
      for (int u = 0; u < 8; u++) acc += sa[threadIdx.x + u * 256];
    } else {                                              // 8 consecutive products per lane (row-like, bank conflicts)
#pragma unroll
      for (int u = 0; u < 8; u++) acc += sa[threadIdx.x * 8 + u];
    }
    __syncthreads();
  }
  if (acc == 1.2345e300) out[0] = acc;
}

template <int MODE>
static double run_parked(long long n, const int *d_idx, const double *d_val, const double *d_x, double *d_out, int ctas_per_sm)
{
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = 148 * ctas_per_sm;
  for (int w = 0; w < 2; w++) parked_kernel<MODE><<<grid, 256>>>(n, d_idx, d_val, d_x, d_out);
  CK(cudaEventRecord(e0));
  const int reps = 5;
  for (int r = 0; r < reps; r++) parked_kernel<MODE><<<grid, 256>>>(n, d_idx, d_val, d_x, d_out);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)(n / 2048 * 2048) * reps / (ms * 1e-3) / 1e9;
}

static uint64_t splitmix(uint64_t &s) { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }

template <int MODE, int UNROLL>
static double run(long long n, const int *d_idx, const double *d_x, double *d_out, int ctas_per_sm)
{
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int grid = 148 * ctas_per_sm;
  for (int w = 0; w < 2; w++) gather_kernel<MODE, UNROLL><<<grid, 256>>>(n, d_idx, d_x, d_out);
  CK(cudaEventRecord(e0));
  const int reps = 5;
  for (int r = 0; r < reps; r++) gather_kernel<MODE, UNROLL><<<grid, 256>>>(n, d_idx, d_x, d_out);
  CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
  float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
  return (double)n * reps / (ms * 1e-3) / 1e9;          // G gathers/s
}

int main()
{
  const long long n = 200ll * 1000 * 1000;              // gathers per launch (C4: 125 M per strip)
  int *h_idx = (int *)malloc(sizeof(int) * n), *d_idx;
  double *d_x, *d_out, *d_val;
  CK(cudaMalloc(&d_idx, sizeof(int) * n)); CK(cudaMalloc(&d_out, 64));
  CK(cudaMalloc(&d_val, sizeof(double) * n)); CK(cudaMemset(d_val, 0, sizeof(double) * n));
  const long long windows[3] = {2500000, 10000000, 40000000};       // doubles: 20 MB (one C4 strip), 80 MB (all of C4's x), 320 MB
  CK(cudaMalloc(&d_x, sizeof(double) * windows[2])); CK(cudaMemset(d_x, 0, sizeof(double) * windows[2]));
  int dev = 0, clk = 0, sms = 0;
  CK(cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, dev)); CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  printf("# %d SMs, max SM clock %d MHz; %lld gathers per launch, index stream coalesced (4 B per gather)\n", sms, clk / 1000, n);
  for (int w = 0; w < 3; w++) {
    uint64_t s = 12345 + w;
    for (long long i = 0; i < n; i++) h_idx[i] = (int)(splitmix(s) % (uint64_t)windows[w]);
    CK(cudaMemcpy(d_idx, h_idx, sizeof(int) * n, cudaMemcpyHostToDevice));
    const double a = run<0, 8>(n, d_idx, d_x, d_out, 8), b = run<1, 8>(n, d_idx, d_x, d_out, 8);
    const double c = run<1, 4>(n, d_idx, d_x, d_out, 8), d = run<1, 16>(n, d_idx, d_x, d_out, 4);
    printf("window %4lld MB: ld.global.nc %.1f G gathers/s | ld.global.cg %.1f (8 in flight per lane) %.1f (4) %.1f (16, 4 CTAs/SM)  -> %.2f cycles per gather per SM at the max clock\n",
           windows[w] * 8 / 1000000, a, b, c, d, (double)sms * clk * 1e3 / (b * 1e9));
    if (w < 2) {
      const double p8 = run_parked<1>(n, d_idx, d_val, d_x, d_out, 8), p6 = run_parked<1>(n, d_idx, d_val, d_x, d_out, 6), q8 = run_parked<2>(n, d_idx, d_val, d_x, d_out, 8);
      printf("               parked form (ids+values staged in shared memory, products stored back, 2 barriers per 2048-entry tile, 12 B/entry streamed):\n"
             "               %.1f G gathers/s at 8 CTAs/SM, %.1f at 6; with conflict-free sums %.1f\n", p8, p6, q8);
    }
  }
  printf("# C4 SpMV (10 M rows, 499.7 M entries, 4 strips of 20 MB of x): 3.76 ms per product = %.1f G gathers/s WITH the 6 GB matrix stream\n", 499.7e6 / 3.76e-3 / 1e9);
  return 0;
}
