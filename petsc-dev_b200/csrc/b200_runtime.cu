/* b200_runtime.cu — device selection, streams, scratch, memory, timers, error reporting.
 * Plays the role of the reference's GPU bring-up in PetscInitialize (src/sys/objects/pinit.c:838-846) and of
 * WaitForGPU / CHKERRCUSP (src/vec/vec/impls/seq/seqcusp/cuspvecimpl.h:84-88). */
#include "b200_internal.h"

B200Ctx g_b200 = {};
static char g_errbuf[1024] = "";

int b200_seterr(int code, const char *fmt, ...)
{
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_errbuf, sizeof(g_errbuf), fmt, ap);
  va_end(ap);
  return code;
}

int b200_cuda_err(cudaError_t e, const char *what, const char *file, int line)
{
  return b200_seterr(B200_ERR_LIB, "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
}

extern "C" const char *b200_last_error(void) { return g_errbuf; }
extern "C" int b200_version(void) { return 100; }

extern "C" int b200_device_count(int *n)
{
  if (!n) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  B200_CUDA(cudaGetDeviceCount(n));
  return 0;
}

extern "C" int b200_init(int device)
{
  B200Ctx &c = g_b200;
  if (c.inited) {
    if (device >= 0 && device != c.device) return b200_seterr(B200_ERR_ARG_WRONGSTATE, "already initialised on device %d", c.device);
    return 0;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1)
    return b200_seterr(B200_ERR_LIB, "no CUDA device available (%s): libpetscb200 has no CPU fallback", cudaGetErrorString(e));
  if (device < 0) { B200_CUDA(cudaGetDevice(&device)); }
  if (device >= ndev) return b200_seterr(B200_ERR_ARG_OUTOFRANGE, "device %d of %d", device, ndev);
  B200_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) return b200_seterr(B200_ERR_SUP, "device %s is sm_%d%d; this library is built for sm_100a only", prop.name, prop.major, prop.minor);
  c.device = device;
  c.nsm    = prop.multiProcessorCount;
  B200_CUDA(cudaStreamCreateWithFlags(&c.own_stream, cudaStreamNonBlocking));
  B200_CUDA(cudaStreamCreateWithFlags(&c.comm_stream, cudaStreamNonBlocking));
  c.stream = c.own_stream;
  B200_CUDA(cudaEventCreate(&c.ev_t0));
  B200_CUDA(cudaEventCreate(&c.ev_t1));
  B200_CUDA(cudaEventCreateWithFlags(&c.ev_fork, cudaEventDisableTiming));
  B200_CUDA(cudaEventCreateWithFlags(&c.ev_join, cudaEventDisableTiming));
  B200_CUDA(cudaHostAlloc((void **)&c.h_res, B200_RES_SLOTS * sizeof(double), cudaHostAllocMapped));
  B200_CUDA(cudaHostGetDevicePointer((void **)&c.d_res, c.h_res, 0));
  B200_CUDA(cudaMalloc((void **)&c.d_partials, sizeof(double) * B200_RED_MAXBLK * B200_RED_MAXV));
  B200_CUDA(cudaMalloc((void **)&c.d_ticket, sizeof(unsigned) * 16));
  B200_CUDA(cudaMemset(c.d_ticket, 0, sizeof(unsigned) * 16));
  B200_CUDA(cudaMalloc((void **)&c.d_scal, sizeof(double) * B200_RES_SLOTS));
  B200_CUDA(cudaMemset(c.d_scal, 0, sizeof(double) * B200_RES_SLOTS));
  B200_CUDA(cudaMalloc(&c.d_ptrtab, 4096));
  B200_CUDA(cudaHostAlloc(&c.h_ptrtab, 4096, cudaHostAllocDefault));
  c.l2bytes = (size_t)256 << 20;   /* > 126 MB L2 */
  c.l2flush = NULL;
  c.launches = 0;
  c.rank = 0; c.size = 1; c.nccl = NULL; c.p2p = false; c.peer = NULL;
  c.inited = true;
  return 0;
}

bool b200_ready(void) { return g_b200.inited; }

extern "C" int b200_finalize(void)
{
  B200Ctx &c = g_b200;
  if (!c.inited) return 0;
  cudaDeviceSynchronize();
  b200_comm_destroy();
  cudaFree(c.d_partials); cudaFree(c.d_ticket); cudaFree(c.d_scal); cudaFree(c.d_ptrtab);
  cudaFreeHost(c.h_res); cudaFreeHost(c.h_ptrtab);
  if (c.l2flush) cudaFree(c.l2flush);
  cudaEventDestroy(c.ev_t0); cudaEventDestroy(c.ev_t1); cudaEventDestroy(c.ev_fork); cudaEventDestroy(c.ev_join);
  cudaStreamDestroy(c.own_stream); cudaStreamDestroy(c.comm_stream);
  memset(&c, 0, sizeof(c));
  return 0;
}

extern "C" int b200_set_stream(void *s)
{
  B200_NEED_INIT();
  if (s) g_b200.multi_stream = true;
  else if (g_b200.multi_stream) { B200_CUDA(cudaDeviceSynchronize()); g_b200.multi_stream = false; }   /* back on one stream: nothing in flight elsewhere */
  g_b200.stream = s ? (cudaStream_t)s : g_b200.own_stream;
  return 0;
}

extern "C" int b200_get_stream(void **s)
{
  B200_NEED_INIT();
  *s = (void *)g_b200.stream;
  return 0;
}

/* caller-owned streams (for b200_set_stream): lets a host program double-buffer, e.g. upload the next vector while the
 * previous result is still on its way back */
extern "C" int b200_stream_create(void **s)
{
  B200_NEED_INIT();
  if (!s) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  cudaStream_t st;
  B200_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
  *s = (void *)st;
  return 0;
}

extern "C" int b200_stream_destroy(void *s)
{
  if (!s) return 0;
  if (g_b200.inited && g_b200.stream == (cudaStream_t)s) g_b200.stream = g_b200.own_stream;
  B200_CUDA(cudaStreamDestroy((cudaStream_t)s));
  return 0;
}

extern "C" int b200_stream_synchronize(void *s)
{
  B200_NEED_INIT();
  B200_CUDA(cudaStreamSynchronize(s ? (cudaStream_t)s : g_b200.own_stream));
  return 0;
}

extern "C" int b200_synchronize(void)
{
  B200_NEED_INIT();
  B200_CUDA(cudaStreamSynchronize(g_b200.stream));
  return 0;
}

extern "C" int b200_launch_count(long long *n)
{
  if (!n) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *n = g_b200.launches;
  return 0;
}

extern "C" int b200_timer_start(void)
{
  B200_NEED_INIT();
  B200_CUDA(cudaEventRecord(g_b200.ev_t0, g_b200.stream));
  return 0;
}

extern "C" int b200_timer_stop(double *ms)
{
  B200_NEED_INIT();
  float f = 0.f;
  B200_CUDA(cudaEventRecord(g_b200.ev_t1, g_b200.stream));
  B200_CUDA(cudaEventSynchronize(g_b200.ev_t1));
  B200_CUDA(cudaEventElapsedTime(&f, g_b200.ev_t0, g_b200.ev_t1));
  if (ms) *ms = (double)f;
  return 0;
}

extern "C" int b200_flush_l2(void)
{
  B200_NEED_INIT();
  if (!g_b200.l2flush) { B200_CUDA(cudaMalloc(&g_b200.l2flush, g_b200.l2bytes)); }
  B200_CUDA(cudaMemsetAsync(g_b200.l2flush, 1, g_b200.l2bytes, g_b200.stream));
  return 0;
}

extern "C" int b200_malloc(size_t bytes, void **dptr)
{
  B200_NEED_INIT();
  if (!dptr) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  *dptr = NULL;
  if (!bytes) bytes = 16;
  cudaError_t e = cudaMalloc(dptr, bytes);
  if (e == cudaErrorMemoryAllocation) { cudaGetLastError(); return b200_seterr(B200_ERR_MEM, "cudaMalloc of %zu bytes failed", bytes); }
  B200_CUDA(e);
  return 0;
}

extern "C" int b200_free(void *dptr)
{
  if (dptr) { B200_CUDA(cudaFree(dptr)); }
  return 0;
}

extern "C" int b200_host_alloc(size_t bytes, void **hptr)
{
  B200_NEED_INIT();
  if (!hptr) return b200_seterr(B200_ERR_ARG_NULL, "null argument");
  if (!bytes) bytes = 16;
  B200_CUDA(cudaHostAlloc(hptr, bytes, cudaHostAllocDefault));
  return 0;
}

extern "C" int b200_host_free(void *hptr)
{
  if (hptr) { B200_CUDA(cudaFreeHost(hptr)); }
  return 0;
}

extern "C" int b200_memcpy_h2d_async(void *dst, const void *src, size_t bytes)
{
  B200_NEED_INIT();
  if (bytes) { B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, g_b200.stream)); }
  return 0;
}

extern "C" int b200_memcpy_d2h_async(void *dst, const void *src, size_t bytes)
{
  B200_NEED_INIT();
  if (bytes) { B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, g_b200.stream)); }
  return 0;
}

extern "C" int b200_memcpy_h2d(void *dst, const void *src, size_t bytes)
{
  B200_CHK(b200_memcpy_h2d_async(dst, src, bytes));
  B200_CUDA(cudaStreamSynchronize(g_b200.stream));
  return 0;
}

extern "C" int b200_memcpy_d2h(void *dst, const void *src, size_t bytes)
{
  B200_CHK(b200_memcpy_d2h_async(dst, src, bytes));
  B200_CUDA(cudaStreamSynchronize(g_b200.stream));
  return 0;
}

extern "C" int b200_memset(void *dst, int byte, size_t bytes)
{
  B200_NEED_INIT();
  if (bytes) { B200_CUDA(cudaMemsetAsync(dst, byte, bytes, g_b200.stream)); }
  return 0;
}
