/* b200mpiexec — launcher of the b200mpi ranks (see mpi.h):   b200mpiexec -n P [-env NAME VALUE]... prog args...
 *
 * Creates the shared segment (memfd, so nothing is left behind in /dev/shm), starts P copies of `prog` with
 * B200MPI_RANK / B200MPI_SIZE / B200MPI_FD in their environment, waits for all of them.  When one rank ends
 * abnormally the others are terminated; the exit status is the first non-zero status seen.  Ring size:
 * B200MPI_RING_KB (default: 4 MiB for P <= 16, shrinking so that the P*P rings stay within 1 GiB of address space). */
#define _GNU_SOURCE
#include <errno.h>
#include <sched.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>

#define B200MPI_MAGIC  0xB200A11Eu
#define B200MPI_MAXP   256
#define SEG_HDR_BYTES  8192

typedef struct {                                   /* must match SegHdr in b200mpi.c */
  uint32_t     magic;
  int          size;
  uint64_t     ring_bytes;
  volatile int abort_flag, abort_code;
  volatile int next_ctx;
  volatile int nattached;
  int          pids[B200MPI_MAXP];
} SegHdr;

static pid_t g_pids[B200MPI_MAXP];
static int   g_n = 0;

static void kill_all(int sig) { for (int i = 0; i < g_n; i++) if (g_pids[i] > 0) kill(g_pids[i], sig); }
static void on_signal(int sig) { kill_all(SIGTERM); _exit(128 + sig); }

int main(int argc, char **argv)
{
  int n = 1, a = 1;
  while (a < argc && argv[a][0] == '-') {
    if ((!strcmp(argv[a], "-n") || !strcmp(argv[a], "-np")) && a + 1 < argc) { n = atoi(argv[a + 1]); a += 2; }
    else if (!strcmp(argv[a], "-env") && a + 2 < argc) { setenv(argv[a + 1], argv[a + 2], 1); a += 3; }
    else break;
  }
  if (a >= argc || n < 1 || n > B200MPI_MAXP) {
    fprintf(stderr, "usage: b200mpiexec -n P [-env NAME VALUE]... program [args...]   (1 <= P <= %d)\n", B200MPI_MAXP);
    return 2;
  }
  uint64_t ring = 4u << 20;
  while ((uint64_t)n * n * ring > (1ull << 30) && ring > (64u << 10)) ring >>= 1;
  const char *e = getenv("B200MPI_RING_KB");
  if (e && atol(e) >= 4) { ring = 4096; while (ring < (uint64_t)atol(e) * 1024) ring <<= 1; }
  const size_t total = SEG_HDR_BYTES + (size_t)n * n * (128 + ring);
  int fd = memfd_create("b200mpi", 0);
  if (fd < 0 || ftruncate(fd, (off_t)total)) { fprintf(stderr, "b200mpiexec: cannot create the shared segment: %s\n", strerror(errno)); return 2; }
  SegHdr *seg = (SegHdr *)mmap(NULL, SEG_HDR_BYTES, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  if (seg == MAP_FAILED) { fprintf(stderr, "b200mpiexec: mmap: %s\n", strerror(errno)); return 2; }
  seg->magic = B200MPI_MAGIC; seg->size = n; seg->ring_bytes = ring; seg->next_ctx = 3;
  signal(SIGINT, on_signal); signal(SIGTERM, on_signal); signal(SIGHUP, on_signal);
  char buf[32];
  snprintf(buf, sizeof(buf), "%d", n);  setenv("B200MPI_SIZE", buf, 1);
  snprintf(buf, sizeof(buf), "%d", fd); setenv("B200MPI_FD", buf, 1);
  fflush(stdout); fflush(stderr);
  g_n = n;
  /* one core per rank when there are enough of them (B200MPI_BIND=0 switches the binding off): spinning ranks that migrate
   * between cores pay for it in every message */
  cpu_set_t allowed;
  int ncpu = 0, cpus[CPU_SETSIZE], bind = 1;
  const char *eb = getenv("B200MPI_BIND");
  if (eb && !atoi(eb)) bind = 0;
  CPU_ZERO(&allowed);
  if (sched_getaffinity(0, sizeof(allowed), &allowed) == 0)
    for (int c = 0; c < CPU_SETSIZE; c++) if (CPU_ISSET(c, &allowed)) cpus[ncpu++] = c;
  if (ncpu < n) bind = 0;
  for (int r = 0; r < n; r++) {
    pid_t p = fork();
    if (p < 0) { perror("b200mpiexec: fork"); kill_all(SIGTERM); return 2; }
    if (p == 0) {
      snprintf(buf, sizeof(buf), "%d", r); setenv("B200MPI_RANK", buf, 1);
      if (bind) { cpu_set_t one; CPU_ZERO(&one); CPU_SET(cpus[(size_t)r * ncpu / n], &one); sched_setaffinity(0, sizeof(one), &one); }
      execvp(argv[a], argv + a);
      fprintf(stderr, "b200mpiexec: cannot run %s: %s\n", argv[a], strerror(errno));
      _exit(127);
    }
    g_pids[r] = p;
  }
  int status = 0, left = n, rc = 0;
  while (left > 0) {
    pid_t p = wait(&status);
    if (p < 0) { if (errno == EINTR) continue; break; }
    left--;
    for (int i = 0; i < n; i++) if (g_pids[i] == p) g_pids[i] = 0;
    int code = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + (WIFSIGNALED(status) ? WTERMSIG(status) : 0);
    if (code && !rc) {
      rc = seg->abort_flag && seg->abort_code ? seg->abort_code : code;
      seg->abort_code = rc; seg->abort_flag = 1;         /* spinning ranks notice and leave; the rest is terminated */
      usleep(200000);
      kill_all(SIGTERM);
    }
  }
  return rc;
}
