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
  b200_upload_release();
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

/* ------------------------------------------------------------------------------------------------ staged uploads
 * Host -> device copy of a large PAGEABLE array (the CSR arrays a caller hands to b200_mat_create_*: the reference's
 * a->i/a->j/a->a come from PetscMalloc).  A plain cudaMemcpy stages such memory through one small driver buffer on one
 * thread (measured 7.5 GB/s for the 1.4 GB of configs[1]); here W host threads copy alternating 1 MB chunks into their
 * own pinned buffers and queue the DMA of each chunk on their own stream, so the host-side copies run in parallel and
 * overlap the transfers (what MatCUSPCopyToGPU's `assign` from a host array cannot do, aijcusp.cu:126-253).
 * Synchronous like the cudaMemcpy it replaces. */
#include <thread>
#include <vector>
#include <atomic>

#define B200_UP_CHUNK   ((size_t)1 << 20)     /* pinning costs ~1 ms per MB: 16 MB in all for 8 workers */
#define B200_UP_MAXW    8
#define B200_UP_MIN     ((size_t)64 << 20)     /* below this one ordinary copy is as fast as setting the pipeline up */
struct UpWorker { cudaStream_t st; void *buf[2]; cudaEvent_t ev[2]; bool ready; };
static UpWorker g_up[B200_UP_MAXW];
static int      g_up_max = 0;
static char    *g_up_pinned = NULL;        /* ONE pinned block for all workers' buffers: every cudaHostAlloc call has a
                                              fixed cost of milliseconds (16 calls were ~70 ms of the first big upload) */

/* workers (stream + two events) are set up as uploads need them */
static int up_worker_init(int k)
{
  if (g_up[k].ready) return 0;
  if (!g_up_pinned) B200_CUDA(cudaHostAlloc((void **)&g_up_pinned, (size_t)B200_UP_MAXW * 2 * B200_UP_CHUNK, cudaHostAllocDefault));
  B200_CUDA(cudaStreamCreateWithFlags(&g_up[k].st, cudaStreamNonBlocking));
  for (int b = 0; b < 2; b++) {
    g_up[k].buf[b] = g_up_pinned + ((size_t)k * 2 + b) * B200_UP_CHUNK;
    B200_CUDA(cudaEventCreateWithFlags(&g_up[k].ev[b], cudaEventDisableTiming));
  }
  g_up[k].ready = true;
  return 0;
}

static int up_workers_for(size_t bytes)
{
  if (!g_up_max) {
    int w = (int)std::thread::hardware_concurrency() / 2;
    const char *e = getenv("B200_UPLOAD_THREADS");
    if (e && atoi(e) > 0) w = atoi(e);
    g_up_max = w < 1 ? 1 : (w > B200_UP_MAXW ? B200_UP_MAXW : w);
  }
  size_t w = bytes / ((size_t)32 << 20);                   /* one worker per 32 MB, so a 100 MB matrix pins little */
  if (w < 2) w = 2;
  return w > (size_t)g_up_max ? g_up_max : (int)w;
}

static void  *g_pin_scratch = NULL;
static size_t g_pin_scratch_bytes = 0;
void *b200_pinned_scratch(size_t bytes)
{
  if (bytes > g_pin_scratch_bytes) {
    if (g_pin_scratch) cudaFreeHost(g_pin_scratch);
    g_pin_scratch = NULL; g_pin_scratch_bytes = 0;
    const size_t want = bytes + (bytes >> 2) + 4096;
    if (cudaHostAlloc(&g_pin_scratch, want, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); g_pin_scratch = NULL; return NULL; }
    g_pin_scratch_bytes = want;
  }
  return g_pin_scratch;
}

void b200_upload_release(void)
{
  if (g_pin_scratch) { cudaFreeHost(g_pin_scratch); g_pin_scratch = NULL; g_pin_scratch_bytes = 0; }
  for (int k = 0; k < B200_UP_MAXW; k++) {
    if (!g_up[k].ready) continue;
    cudaStreamDestroy(g_up[k].st);
    for (int b = 0; b < 2; b++) cudaEventDestroy(g_up[k].ev[b]);
    g_up[k].ready = false;
  }
  if (g_up_pinned) { cudaFreeHost(g_up_pinned); g_up_pinned = NULL; }
}

int b200_upload_pageable(void *dst, const void *src, size_t bytes)
{
  B200Ctx &c = g_b200;
  if (!bytes) return 0;
  cudaPointerAttributes at;
  const bool pinned = cudaPointerGetAttributes(&at, src) == cudaSuccess && (at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged);
  cudaGetLastError();
  if (pinned || bytes < B200_UP_MIN) {                      /* small, or already pinned: one ordinary copy */
    B200_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, c.stream));
    B200_CUDA(cudaStreamSynchronize(c.stream));
    return 0;
  }
  const size_t nchunks = (bytes + B200_UP_CHUNK - 1) / B200_UP_CHUNK;
  const int W = up_workers_for(bytes);
  for (int w = 0; w < W; w++) B200_CHK(up_worker_init(w));
  std::atomic<int> err(0);
  const int device = c.device;
  /* the destination may still be in use by work queued on the compute stream (a value refresh of a live matrix) */
  B200_CUDA(cudaEventRecord(c.ev_fork, c.stream));
  auto work = [&](int w) {
    UpWorker &u = g_up[w];
    if (cudaSetDevice(device) != cudaSuccess) { err = 1; return; }
    if (cudaStreamWaitEvent(u.st, c.ev_fork, 0) != cudaSuccess) { err = 1; return; }
    int b = 0;
    for (size_t k = (size_t)w; k < nchunks && !err; k += (size_t)W, b ^= 1) {
      const size_t off = k * B200_UP_CHUNK, len = bytes - off < B200_UP_CHUNK ? bytes - off : B200_UP_CHUNK;
      if (cudaEventSynchronize(u.ev[b]) != cudaSuccess) { err = 1; return; }        /* the buffer's previous transfer has left */
      memcpy(u.buf[b], (const char *)src + off, len);
      if (cudaMemcpyAsync((char *)dst + off, u.buf[b], len, cudaMemcpyHostToDevice, u.st) != cudaSuccess) { err = 1; return; }
      if (cudaEventRecord(u.ev[b], u.st) != cudaSuccess) { err = 1; return; }
    }
  };
  std::vector<std::thread> th;
  for (int w = 1; w < W; w++) th.emplace_back(work, w);
  work(0);
  for (auto &t : th) t.join();
  if (err) return b200_cuda_err(cudaGetLastError(), "staged upload", __FILE__, __LINE__);
  /* complete on return like the cudaMemcpy it replaces: later readers of the destination may sit on any stream */
  for (int w = 0; w < W; w++) B200_CUDA(cudaStreamSynchronize(g_up[w].st));
  return 0;
}
