/* b200mpi.c — the N-rank, one-node MPI subset behind petsc-dev_b200/mpishim/mpi.h (see the header for the model).
 *
 * Layout of the shared segment (created by b200mpiexec, mapped by every rank):
 *   SegHdr | RingCtl[P*P] | ring data [P*P][ring_bytes]        ring (s,d) = index s*P + d, single producer s, single consumer d
 * Wire format inside a ring: MsgHdr (32 bytes) + payload padded to 8 bytes; a message is contiguous in its ring, so
 * messages between one pair of ranks never overtake each other (MPI's ordering rule).
 * The receiver's progress engine (called from every blocking or testing call) drains its P incoming rings: a header is
 * matched against the posted receives (in posting order); unmatched messages are copied into an unexpected queue, so a
 * sender never depends on the receiver having posted anything — only on the receiver entering the library.
 */
#define _GNU_SOURCE
#include "mpi.h"
#include <errno.h>
#include <sched.h>
#include <signal.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/prctl.h>
#include <time.h>
#include <unistd.h>

#define B200MPI_MAGIC  0xB200A11Eu
#define B200MPI_MAXP   256
#define SEG_HDR_BYTES  8192
#define COLL_CTX       0x40000000
#define TAG_UB_VALUE   0x0fffffff

typedef struct {
  uint32_t     magic;
  int          size;
  uint64_t     ring_bytes;          /* power of two */
  volatile int abort_flag, abort_code;
  volatile int next_ctx;
  volatile int nattached;
  int          pids[B200MPI_MAXP];
} SegHdr;

typedef struct {
  volatile uint64_t tail; char pad0[56];       /* bytes produced (written by the sender)   */
  volatile uint64_t head; char pad1[56];       /* bytes consumed (written by the receiver) */
} RingCtl;

typedef struct { uint64_t nbytes; int32_t ctx, tag; int32_t pad[4]; } MsgHdr;   /* 32 bytes */

typedef struct Req {
  int    kind;                     /* 0 free, 1 send, 2 recv */
  int    persistent, active, done;
  char  *buf;
  size_t bytes;                    /* send: payload; recv: capacity */
  int    peer_world;               /* send: destination; recv: source or MPI_ANY_SOURCE (world numbering) */
  int    tag, ctx;
  MPI_Comm comm;
  size_t sent; int hdr_sent;
  MPI_Status st;
  struct Req *next;
  int    peer_arg, count; MPI_Datatype dt;     /* as given (persistent restarts) */
  int    handle;
} Req;

typedef struct Unexp {
  int    src_world, ctx, tag;
  size_t nbytes;
  char  *data;
  int    complete;
  Req   *claimed;
  struct Unexp *next;
} Unexp;

typedef struct { int have_hdr; MsgHdr hdr; size_t got, wire, cap; char *dst; Req *req; Unexp *ux; } InState;

typedef struct { int keyval; void *val; } Attr;
typedef struct {
  int   used, ctx, size, rank;
  int  *world;                     /* comm rank -> world rank */
  int  *inv;                       /* world rank -> comm rank, -1 */
  Attr *attr; int nattr, cattr;
  int   errh;
  unsigned coll_seq;
} Comm;

typedef struct { int used; MPI_Copy_function *copy; MPI_Delete_function *del; void *extra; } Keyval;
typedef struct { int used; size_t size; MPI_Datatype base; int mult; } Dtype;   /* derived (contiguous) types */
typedef struct { int used; MPI_User_function *fn; int commute; } UserOp;
typedef struct { int used, n; int *world; } Group;

static struct {
  int      inited, finalized;
  int      rank, size;
  SegHdr  *seg;
  RingCtl *ctl;
  char    *data;
  uint64_t ring_bytes;
  InState *in;                     /* [size] */
  Req    **sendq_head, **sendq_tail;   /* [size] */
  int      nsend_pending;
  Req     *posted_head, *posted_tail;
  Unexp   *ux_head, *ux_tail;
  Req    **reqs; int nreqs;
  Comm    *comms; int ncomms;
  Keyval  *keys;  int nkeys;
  Dtype   *dts;   int ndts;
  UserOp  *ops;   int nops;
  Group   *groups; int ngroups;
  int      tag_ub;
} G;

static const size_t builtin_size[] = {0, 1, 1, 1, 1, 2, 2, 4, 4, 8, 8, 8, 8, 4, 8, 16, 1, 8, 8, 16, 16, 8, 16, 16, 8};
#define NBUILTIN ((int)(sizeof(builtin_size) / sizeof(builtin_size[0])))
#define DT_DERIVED0 64
#define OP_USER0    32
#define KEY_USER0   8

/* ------------------------------------------------------------------------------------------------ errors */
static void die(const char *fmt, ...)
{
  va_list ap;
  fprintf(stderr, "[b200mpi rank %d] fatal: ", G.rank);
  va_start(ap, fmt); vfprintf(stderr, fmt, ap); va_end(ap);
  fprintf(stderr, "\n");
  fflush(stderr);
  MPI_Abort(MPI_COMM_WORLD, 86);
}

static void check_abort(void)
{
  if (G.seg && G.seg->abort_flag) { fflush(stdout); fflush(stderr); _exit(G.seg->abort_code ? G.seg->abort_code : 1); }
}

/* ------------------------------------------------------------------------------------------------ handles */
static Comm *comm_of(MPI_Comm c)
{
  if (c <= 0 || c >= G.ncomms || !G.comms[c].used) die("invalid communicator %d", c);
  return &G.comms[c];
}

static size_t dt_size(MPI_Datatype dt)
{
  if (dt > 0 && dt < NBUILTIN) return builtin_size[dt];
  if (dt >= DT_DERIVED0 && dt - DT_DERIVED0 < G.ndts && G.dts[dt - DT_DERIVED0].used) return G.dts[dt - DT_DERIVED0].size;
  die("invalid datatype %d", dt);
  return 0;
}

static Req *req_new(void)
{
  int h;
  for (h = 1; h < G.nreqs; h++) if (!G.reqs[h]) break;
  if (h >= G.nreqs) {
    int n = G.nreqs ? 2 * G.nreqs : 64;
    G.reqs = (Req **)realloc(G.reqs, sizeof(Req *) * (size_t)n);
    memset(G.reqs + G.nreqs, 0, sizeof(Req *) * (size_t)(n - G.nreqs));
    h = G.nreqs ? G.nreqs : 1;
    G.nreqs = n;
  }
  Req *r = (Req *)calloc(1, sizeof(Req));
  r->handle = h;
  G.reqs[h] = r;
  return r;
}

static Req *req_of(MPI_Request h)
{
  if (h <= 0 || h >= G.nreqs || !G.reqs[h]) die("invalid request %d", h);
  return G.reqs[h];
}

static void req_free(Req *r) { G.reqs[r->handle] = NULL; free(r); }

/* ------------------------------------------------------------------------------------------------ rings */
static inline RingCtl *ring_ctl(int s, int d) { return &G.ctl[(size_t)s * G.size + d]; }
static inline char    *ring_dat(int s, int d) { return G.data + ((size_t)s * G.size + d) * G.ring_bytes; }

static void ring_write(int d, const void *src, size_t n, uint64_t *tail)          /* caller checked the space */
{
  char *base = ring_dat(G.rank, d);
  const uint64_t off = *tail & (G.ring_bytes - 1);
  const size_t first = n < G.ring_bytes - off ? n : (size_t)(G.ring_bytes - off);
  memcpy(base + off, src, first);
  if (n > first) memcpy(base, (const char *)src + first, n - first);
  *tail += n;
}

static void ring_read(int s, void *dst, size_t n, uint64_t *head)                 /* dst NULL: discard */
{
  if (dst) {
    const char *base = ring_dat(s, G.rank);
    const uint64_t off = *head & (G.ring_bytes - 1);
    const size_t first = n < G.ring_bytes - off ? n : (size_t)(G.ring_bytes - off);
    memcpy(dst, base + off, first);
    if (n > first) memcpy((char *)dst + first, base, n - first);
  }
  *head += n;
}

/* ------------------------------------------------------------------------------------------------ matching */
static inline int req_matches(const Req *r, int ctx, int src_world, int tag)
{
  return r->ctx == ctx && (r->peer_world == MPI_ANY_SOURCE || r->peer_world == src_world) && (r->tag == MPI_ANY_TAG || r->tag == tag);
}

static Req *posted_take(int ctx, int src_world, int tag)
{
  Req *prev = NULL;
  for (Req *r = G.posted_head; r; prev = r, r = r->next) {
    if (!req_matches(r, ctx, src_world, tag)) continue;
    if (prev) prev->next = r->next; else G.posted_head = r->next;
    if (G.posted_tail == r) G.posted_tail = prev;
    r->next = NULL;
    return r;
  }
  return NULL;
}

static void posted_remove(Req *x)
{
  Req *prev = NULL;
  for (Req *r = G.posted_head; r; prev = r, r = r->next) {
    if (r != x) continue;
    if (prev) prev->next = r->next; else G.posted_head = r->next;
    if (G.posted_tail == r) G.posted_tail = prev;
    r->next = NULL;
    return;
  }
}

static void ux_unlink(Unexp *x)
{
  Unexp *prev = NULL;
  for (Unexp *u = G.ux_head; u; prev = u, u = u->next) {
    if (u != x) continue;
    if (prev) prev->next = u->next; else G.ux_head = u->next;
    if (G.ux_tail == u) G.ux_tail = prev;
    return;
  }
}

static void recv_finish(Req *r, int src_world, int tag, size_t nbytes)
{
  Comm *c = &G.comms[r->comm];
  r->st.MPI_SOURCE = (c->used && src_world < G.size) ? c->inv[src_world] : src_world;
  r->st.MPI_TAG    = tag;
  r->st.MPI_ERROR  = nbytes > r->bytes ? MPI_ERR_TRUNCATE : MPI_SUCCESS;
  r->st.b200_bytes = (int)(nbytes > r->bytes ? r->bytes : nbytes);
  if (nbytes > r->bytes) die("message of %zu bytes truncated by a receive of %zu bytes (tag %d)", nbytes, r->bytes, tag);
  r->done = 1;
}

/* a message whose payload is entirely at hand (self send) */
static void deliver_local(int ctx, int tag, const void *buf, size_t nbytes)
{
  Req *r = posted_take(ctx, G.rank, tag);
  if (r) {
    memcpy(r->buf, buf, nbytes < r->bytes ? nbytes : r->bytes);
    recv_finish(r, G.rank, tag, nbytes);
    return;
  }
  Unexp *u = (Unexp *)calloc(1, sizeof(Unexp));
  u->src_world = G.rank; u->ctx = ctx; u->tag = tag; u->nbytes = nbytes; u->complete = 1;
  u->data = (char *)malloc(nbytes ? nbytes : 1);
  memcpy(u->data, buf, nbytes);
  if (G.ux_tail) G.ux_tail->next = u; else G.ux_head = u;
  G.ux_tail = u;
}

static void in_complete(int s)
{
  InState *in = &G.in[s];
  if (in->req) recv_finish(in->req, s, in->hdr.tag, (size_t)in->hdr.nbytes);
  else {
    Unexp *u = in->ux;
    u->complete = 1;
    if (u->claimed) {
      Req *r = u->claimed;
      memcpy(r->buf, u->data, u->nbytes < r->bytes ? u->nbytes : r->bytes);
      recv_finish(r, s, u->tag, u->nbytes);
      ux_unlink(u);
      free(u->data); free(u);
    }
  }
  in->have_hdr = 0; in->req = NULL; in->ux = NULL;
}

static void drain(int s)
{
  RingCtl *rc = ring_ctl(s, G.rank);
  InState *in = &G.in[s];
  uint64_t head = rc->head;
  for (;;) {
    const uint64_t tail = __atomic_load_n(&rc->tail, __ATOMIC_ACQUIRE);
    uint64_t avail = tail - head;
    if (!in->have_hdr) {
      if (avail < sizeof(MsgHdr)) break;
      ring_read(s, &in->hdr, sizeof(MsgHdr), &head);
      in->have_hdr = 1; in->got = 0;
      in->wire = ((size_t)in->hdr.nbytes + 7) & ~(size_t)7;
      in->req = posted_take(in->hdr.ctx, s, in->hdr.tag);
      if (in->req) { in->dst = in->req->buf; in->cap = in->req->bytes; in->ux = NULL; }
      else {
        Unexp *u = (Unexp *)calloc(1, sizeof(Unexp));
        u->src_world = s; u->ctx = in->hdr.ctx; u->tag = in->hdr.tag; u->nbytes = (size_t)in->hdr.nbytes;
        u->data = (char *)malloc(u->nbytes ? u->nbytes : 1);
        if (!u->data) die("out of memory for an unexpected message of %zu bytes", u->nbytes);
        if (G.ux_tail) G.ux_tail->next = u; else G.ux_head = u;
        G.ux_tail = u;
        in->ux = u; in->dst = u->data; in->cap = u->nbytes;
      }
      __atomic_store_n(&rc->head, head, __ATOMIC_RELEASE);
      if (in->wire == 0) { in_complete(s); continue; }
      avail = tail - head;
    }
    if (avail == 0) break;
    size_t n = in->wire - in->got;
    if (n > avail) n = (size_t)avail;
    /* bytes [got, got+n) of the wire image; only those below min(nbytes, cap) are stored */
    size_t lim = (size_t)in->hdr.nbytes < in->cap ? (size_t)in->hdr.nbytes : in->cap;
    size_t keep = in->got < lim ? (lim - in->got < n ? lim - in->got : n) : 0;
    if (keep) ring_read(s, in->dst + in->got, keep, &head);
    if (n > keep) ring_read(s, NULL, n - keep, &head);
    in->got += n;
    __atomic_store_n(&rc->head, head, __ATOMIC_RELEASE);
    if (in->got == in->wire) in_complete(s);
  }
}

static void push(int d)
{
  RingCtl *rc = ring_ctl(G.rank, d);
  uint64_t tail = rc->tail;
  Req *r;
  while ((r = G.sendq_head[d])) {
    uint64_t space = G.ring_bytes - (tail - __atomic_load_n(&rc->head, __ATOMIC_ACQUIRE));
    if (!r->hdr_sent) {
      if (space < sizeof(MsgHdr)) break;
      MsgHdr h; memset(&h, 0, sizeof(h));
      h.nbytes = r->bytes; h.ctx = r->ctx; h.tag = r->tag;
      ring_write(d, &h, sizeof(h), &tail);
      __atomic_store_n(&rc->tail, tail, __ATOMIC_RELEASE);
      r->hdr_sent = 1; r->sent = 0;
      space -= sizeof(MsgHdr);
    }
    const size_t wire = (r->bytes + 7) & ~(size_t)7;
    while (r->sent < wire) {
      space = G.ring_bytes - (tail - __atomic_load_n(&rc->head, __ATOMIC_ACQUIRE));
      if (!space) break;
      size_t n = wire - r->sent;
      if (n > space) n = (size_t)space;
      const size_t real = r->sent < r->bytes ? (r->bytes - r->sent < n ? r->bytes - r->sent : n) : 0;
      if (real) ring_write(d, r->buf + r->sent, real, &tail);
      if (n > real) { static const char z[8] = {0}; ring_write(d, z, n - real, &tail); }   /* n - real < 8 */
      r->sent += n;
      __atomic_store_n(&rc->tail, tail, __ATOMIC_RELEASE);
    }
    if (r->sent < wire) break;
    G.sendq_head[d] = r->next;
    if (!G.sendq_head[d]) G.sendq_tail[d] = NULL;
    r->next = NULL;
    r->done = 1;
    G.nsend_pending--;
  }
}

static void progress(void)
{
  if (G.size <= 1) return;
  for (int s = 0; s < G.size; s++) if (s != G.rank) drain(s);
  if (G.nsend_pending) for (int d = 0; d < G.size; d++) if (G.sendq_head[d]) push(d);
}

static void idle(unsigned *spins)
{
  if (++*spins > 2000) { check_abort(); sched_yield(); if (*spins > 200000) { struct timespec ts = {0, 50000}; nanosleep(&ts, NULL); } }
}

/* ------------------------------------------------------------------------------------------------ raw p2p */
static void start_send(Req *r)
{
  r->done = 0; r->active = 1; r->hdr_sent = 0; r->sent = 0; r->next = NULL;
  if (r->peer_world == MPI_PROC_NULL) { r->done = 1; return; }
  if (r->peer_world == G.rank) { deliver_local(r->ctx, r->tag, r->buf, r->bytes); r->done = 1; return; }
  const int d = r->peer_world;
  if (G.sendq_tail[d]) G.sendq_tail[d]->next = r; else G.sendq_head[d] = r;
  G.sendq_tail[d] = r;
  G.nsend_pending++;
  push(d);
}

static void start_recv(Req *r)
{
  r->done = 0; r->active = 1; r->next = NULL;
  memset(&r->st, 0, sizeof(r->st));
  if (r->peer_world == MPI_PROC_NULL) { r->st.MPI_SOURCE = MPI_PROC_NULL; r->st.MPI_TAG = MPI_ANY_TAG; r->done = 1; return; }
  for (Unexp *u = G.ux_head; u; u = u->next) {
    if (u->claimed || !req_matches(r, u->ctx, u->src_world, u->tag)) continue;
    if (u->complete) {
      memcpy(r->buf, u->data, u->nbytes < r->bytes ? u->nbytes : r->bytes);
      recv_finish(r, u->src_world, u->tag, u->nbytes);
      ux_unlink(u);
      free(u->data); free(u);
    } else u->claimed = r;
    return;
  }
  if (G.posted_tail) G.posted_tail->next = r; else G.posted_head = r;
  G.posted_tail = r;
}

static void wait_req(Req *r)
{
  unsigned spins = 0;
  while (!r->done) { progress(); if (!r->done) idle(&spins); }
}

static Req *make_req(int kind, void *buf, size_t bytes, int peer_world, int tag, int ctx, MPI_Comm comm)
{
  Req *r = req_new();
  r->kind = kind; r->buf = (char *)buf; r->bytes = bytes; r->peer_world = peer_world; r->tag = tag; r->ctx = ctx; r->comm = comm;
  return r;
}

static int world_of(Comm *c, int rank, int allow_any)
{
  if (rank == MPI_PROC_NULL) return MPI_PROC_NULL;
  if (allow_any && rank == MPI_ANY_SOURCE) return MPI_ANY_SOURCE;
  if (rank < 0 || rank >= c->size) die("rank %d outside a communicator of size %d", rank, c->size);
  return c->world[rank];
}

/* internal blocking transfers for the collectives: `to` / `from` are comm ranks */
static void csend(Comm *c, MPI_Comm h, const void *buf, size_t bytes, int to, int tag)
{
  Req *r = make_req(1, (void *)buf, bytes, c->world[to], tag, c->ctx | COLL_CTX, h);
  start_send(r); wait_req(r); req_free(r);
}
static void crecv(Comm *c, MPI_Comm h, void *buf, size_t bytes, int from, int tag)
{
  Req *r = make_req(2, buf, bytes, c->world[from], tag, c->ctx | COLL_CTX, h);
  start_recv(r); wait_req(r); req_free(r);
}
static Req *cisend(Comm *c, MPI_Comm h, const void *buf, size_t bytes, int to, int tag)
{
  Req *r = make_req(1, (void *)buf, bytes, c->world[to], tag, c->ctx | COLL_CTX, h);
  start_send(r); return r;
}
static Req *cirecv(Comm *c, MPI_Comm h, void *buf, size_t bytes, int from, int tag)
{
  Req *r = make_req(2, buf, bytes, c->world[from], tag, c->ctx | COLL_CTX, h);
  start_recv(r); return r;
}
static int coll_tag(Comm *c) { return (int)(++c->coll_seq & 0x0fffffff); }

/* ------------------------------------------------------------------------------------------------ init / finalize */
int b200mpi_world_rank(void) { const char *e = getenv("B200MPI_RANK"); return G.inited ? G.rank : (e ? atoi(e) : 0); }
int b200mpi_world_size(void) { const char *e = getenv("B200MPI_SIZE"); return G.inited ? G.size : (e ? atoi(e) : 1); }

static int comm_alloc(void)
{
  int h;
  for (h = 3; h < G.ncomms; h++) if (!G.comms[h].used) break;
  if (h >= G.ncomms) {
    int n = G.ncomms ? 2 * G.ncomms : 16;
    G.comms = (Comm *)realloc(G.comms, sizeof(Comm) * (size_t)n);
    memset(G.comms + G.ncomms, 0, sizeof(Comm) * (size_t)(n - G.ncomms));
    h = G.ncomms < 3 ? 3 : G.ncomms;
    G.ncomms = n;
  }
  return h;
}

static void comm_fill(int h, int ctx, int size, int rank, const int *world)
{
  Comm *c = &G.comms[h];
  memset(c, 0, sizeof(*c));
  c->used = 1; c->ctx = ctx; c->size = size; c->rank = rank; c->errh = MPI_ERRORS_ARE_FATAL;
  c->world = (int *)malloc(sizeof(int) * (size_t)size);
  c->inv   = (int *)malloc(sizeof(int) * (size_t)(G.size > 0 ? G.size : 1));
  for (int i = 0; i < G.size; i++) c->inv[i] = -1;
  for (int i = 0; i < size; i++) { c->world[i] = world[i]; c->inv[world[i]] = i; }
}

int MPI_Init(int *argc, char ***argv)
{
  (void)argc; (void)argv;
  if (G.inited) return MPI_SUCCESS;
  const char *er = getenv("B200MPI_RANK"), *es = getenv("B200MPI_SIZE"), *ef = getenv("B200MPI_FD");
  G.rank = 0; G.size = 1;
  if (er && es && ef && atoi(es) > 1) {
    G.rank = atoi(er); G.size = atoi(es);
    if (G.size > B200MPI_MAXP || G.rank < 0 || G.rank >= G.size) { fprintf(stderr, "b200mpi: bad rank/size %d/%d\n", G.rank, G.size); exit(86); }
    const int fd = atoi(ef);
    SegHdr probe;
    if (pread(fd, &probe, sizeof(probe), 0) != (ssize_t)sizeof(probe) || probe.magic != B200MPI_MAGIC || probe.size != G.size) {
      fprintf(stderr, "b200mpi: the launcher's segment (fd %d) is not usable\n", fd); exit(86);
    }
    const size_t total = SEG_HDR_BYTES + (size_t)G.size * G.size * (sizeof(RingCtl) + probe.ring_bytes);
    void *p = mmap(NULL, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) { fprintf(stderr, "b200mpi: mmap of the shared segment failed: %s\n", strerror(errno)); exit(86); }
    G.seg = (SegHdr *)p;
    G.ctl = (RingCtl *)((char *)p + SEG_HDR_BYTES);
    G.data = (char *)p + SEG_HDR_BYTES + (size_t)G.size * G.size * sizeof(RingCtl);
    G.ring_bytes = probe.ring_bytes;
    G.seg->pids[G.rank] = (int)getpid();
    prctl(PR_SET_PDEATHSIG, SIGKILL);                  /* a rank never outlives its launcher */
    __atomic_add_fetch(&G.seg->nattached, 1, __ATOMIC_SEQ_CST);
  }
  G.in = (InState *)calloc((size_t)G.size, sizeof(InState));
  G.sendq_head = (Req **)calloc((size_t)G.size, sizeof(Req *));
  G.sendq_tail = (Req **)calloc((size_t)G.size, sizeof(Req *));
  G.tag_ub = TAG_UB_VALUE;
  G.ncomms = 0; G.comms = NULL;
  comm_alloc();                                        /* table of 16, handles 0..2 reserved */
  int *all = (int *)malloc(sizeof(int) * (size_t)G.size);
  for (int i = 0; i < G.size; i++) all[i] = i;
  comm_fill(MPI_COMM_WORLD, 1, G.size, G.rank, all);
  comm_fill(MPI_COMM_SELF, 2, 1, 0, &G.rank);
  free(all);
  G.nkeys = 0; G.keys = NULL;
  G.inited = 1;
  if (G.size > 1) {
    unsigned spins = 0;
    while (G.seg->nattached < G.size) idle(&spins);    /* everybody's rings are mapped before the first message */
  }
  return MPI_SUCCESS;
}

int MPI_Init_thread(int *argc, char ***argv, int required, int *provided)
{
  (void)required;
  if (provided) *provided = MPI_THREAD_FUNNELED;
  return MPI_Init(argc, argv);
}

int MPI_Initialized(int *flag) { *flag = G.inited; return MPI_SUCCESS; }
int MPI_Finalized(int *flag) { *flag = G.finalized; return MPI_SUCCESS; }

int MPI_Finalize(void)
{
  if (!G.inited || G.finalized) return MPI_SUCCESS;
  MPI_Barrier(MPI_COMM_WORLD);
  fflush(stdout); fflush(stderr);
  G.finalized = 1;
  return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm comm, int errorcode)
{
  (void)comm;
  fflush(stdout); fflush(stderr);
  if (G.seg) {
    G.seg->abort_code = errorcode ? errorcode : 1;
    __atomic_store_n(&G.seg->abort_flag, 1, __ATOMIC_SEQ_CST);
    for (int r = 0; r < G.size; r++) if (r != G.rank && G.seg->pids[r] > 0) kill(G.seg->pids[r], SIGTERM);
  }
  _exit(errorcode ? errorcode : 1);
}

double MPI_Wtime(void) { struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec; }
double MPI_Wtick(void) { return 1e-9; }

int MPI_Get_processor_name(char *name, int *resultlen)
{
  if (gethostname(name, MPI_MAX_PROCESSOR_NAME - 1)) strcpy(name, "localhost");
  name[MPI_MAX_PROCESSOR_NAME - 1] = 0;
  *resultlen = (int)strlen(name);
  return MPI_SUCCESS;
}

int MPI_Error_string(int errorcode, char *string, int *resultlen)
{
  snprintf(string, MPI_MAX_ERROR_STRING, "b200mpi error %d", errorcode);
  *resultlen = (int)strlen(string);
  return MPI_SUCCESS;
}
int MPI_Error_class(int errorcode, int *errorclass) { *errorclass = errorcode; return MPI_SUCCESS; }
int MPI_Errhandler_create(MPI_Handler_function *fn, MPI_Errhandler *eh) { (void)fn; *eh = MPI_ERRORS_RETURN; return MPI_SUCCESS; }
int MPI_Errhandler_set(MPI_Comm comm, MPI_Errhandler eh) { comm_of(comm)->errh = eh; return MPI_SUCCESS; }
int MPI_Errhandler_get(MPI_Comm comm, MPI_Errhandler *eh) { *eh = comm_of(comm)->errh; return MPI_SUCCESS; }
int MPI_Errhandler_free(MPI_Errhandler *eh) { *eh = MPI_ERRHANDLER_NULL; return MPI_SUCCESS; }
int MPI_Pcontrol(const int level, ...) { (void)level; return MPI_SUCCESS; }

/* ------------------------------------------------------------------------------------------------ communicators */
int MPI_Comm_size(MPI_Comm comm, int *size) { *size = comm_of(comm)->size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm comm, int *rank) { *rank = comm_of(comm)->rank; return MPI_SUCCESS; }
int MPI_Comm_test_inter(MPI_Comm comm, int *flag) { (void)comm; *flag = 0; return MPI_SUCCESS; }

static int new_ctx(void)
{
  static int local_next = 3;
  if (G.seg) return __atomic_fetch_add(&G.seg->next_ctx, 1, __ATOMIC_SEQ_CST);
  return local_next++;
}

int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm)
{
  Comm *c = comm_of(comm);
  int ctx = 0;
  if (c->rank == 0) ctx = new_ctx();
  if (c->size > 1) MPI_Bcast(&ctx, 1, MPI_INT, 0, comm);
  const int h = comm_alloc();
  c = comm_of(comm);                                   /* the table may have moved */
  comm_fill(h, ctx, c->size, c->rank, c->world);
  /* attribute copy callbacks (MPI-1 semantics: NULL_COPY_FN -> not propagated) */
  for (int i = 0; i < c->nattr; i++) {
    const int kv = c->attr[i].keyval;
    Keyval *k = &G.keys[kv - KEY_USER0];
    if (!k->used || !k->copy) continue;
    void *out = NULL; int flag = 0;
    if (k->copy == MPI_DUP_FN) { out = c->attr[i].val; flag = 1; }
    else k->copy(comm, kv, k->extra, c->attr[i].val, &out, &flag);
    if (flag) MPI_Attr_put(h, kv, out);
    c = comm_of(comm);
  }
  *newcomm = h;
  return MPI_SUCCESS;
}

typedef struct { int color, key, world, crank; } SplitRec;
static int split_cmp(const void *a, const void *b)
{
  const SplitRec *x = (const SplitRec *)a, *y = (const SplitRec *)b;
  if (x->key != y->key) return x->key < y->key ? -1 : 1;
  return x->crank - y->crank;
}

int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  SplitRec mine = {color, key, G.rank, c->rank}, *all = (SplitRec *)malloc(sizeof(SplitRec) * (size_t)n);
  MPI_Allgather(&mine, (int)sizeof(SplitRec), MPI_BYTE, all, (int)sizeof(SplitRec), MPI_BYTE, comm);
  /* the members of my colour, ordered by (key, old rank); the first one allocates the context */
  SplitRec *grp = (SplitRec *)malloc(sizeof(SplitRec) * (size_t)n);
  int m = 0;
  if (color != MPI_UNDEFINED) for (int i = 0; i < n; i++) if (all[i].color == color) grp[m++] = all[i];
  qsort(grp, (size_t)m, sizeof(SplitRec), split_cmp);
  int *ctxs = (int *)malloc(sizeof(int) * (size_t)n), myctx = 0;
  if (m && grp[0].world == G.rank) myctx = new_ctx();
  MPI_Allgather(&myctx, 1, MPI_INT, ctxs, 1, MPI_INT, comm);
  if (!m) *newcomm = MPI_COMM_NULL;
  else {
    int *world = (int *)malloc(sizeof(int) * (size_t)m), me = -1;
    for (int i = 0; i < m; i++) { world[i] = grp[i].world; if (grp[i].world == G.rank) me = i; }
    const int h = comm_alloc();
    comm_fill(h, ctxs[grp[0].crank], m, me, world);
    free(world);
    *newcomm = h;
  }
  free(ctxs); free(grp); free(all);
  return MPI_SUCCESS;
}

static Group *group_of(MPI_Group g)
{
  if (g <= 0 || g >= G.ngroups || !G.groups[g].used) die("invalid group %d", g);
  return &G.groups[g];
}

static int group_new(int n, const int *world)
{
  int h;
  for (h = 2; h < G.ngroups; h++) if (!G.groups[h].used) break;
  if (h >= G.ngroups) {
    int nn = G.ngroups ? 2 * G.ngroups : 16;
    G.groups = (Group *)realloc(G.groups, sizeof(Group) * (size_t)nn);
    memset(G.groups + G.ngroups, 0, sizeof(Group) * (size_t)(nn - G.ngroups));
    if (!G.ngroups) { G.groups[1].used = 1; G.groups[1].n = 0; G.groups[1].world = NULL; }   /* MPI_GROUP_EMPTY */
    h = G.ngroups < 2 ? 2 : G.ngroups;
    G.ngroups = nn;
  }
  G.groups[h].used = 1; G.groups[h].n = n;
  G.groups[h].world = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
  if (n > 0) memcpy(G.groups[h].world, world, sizeof(int) * (size_t)n);
  return h;
}

int MPI_Comm_group(MPI_Comm comm, MPI_Group *group) { Comm *c = comm_of(comm); *group = group_new(c->size, c->world); return MPI_SUCCESS; }
int MPI_Group_size(MPI_Group group, int *size) { *size = group_of(group)->n; return MPI_SUCCESS; }
int MPI_Group_rank(MPI_Group group, int *rank)
{
  Group *g = group_of(group);
  *rank = MPI_UNDEFINED;
  for (int i = 0; i < g->n; i++) if (g->world[i] == G.rank) *rank = i;
  return MPI_SUCCESS;
}
int MPI_Group_incl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup)
{
  Group *g = group_of(group);
  int *w = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
  for (int i = 0; i < n; i++) { if (ranks[i] < 0 || ranks[i] >= g->n) die("MPI_Group_incl: rank %d outside the group", ranks[i]); w[i] = g->world[ranks[i]]; }
  *newgroup = group_new(n, w);
  free(w);
  return MPI_SUCCESS;
}
int MPI_Group_excl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup)
{
  Group *g = group_of(group);
  int *w = (int *)malloc(sizeof(int) * (size_t)(g->n > 0 ? g->n : 1)), m = 0;
  for (int i = 0; i < g->n; i++) {
    int out = 0;
    for (int k = 0; k < n; k++) if (ranks[k] == i) out = 1;
    if (!out) w[m++] = g->world[i];
  }
  *newgroup = group_new(m, w);
  free(w);
  return MPI_SUCCESS;
}
int MPI_Group_translate_ranks(MPI_Group g1, int n, const int *ranks1, MPI_Group g2, int *ranks2)
{
  Group *a = group_of(g1), *b = group_of(g2);
  for (int i = 0; i < n; i++) {
    ranks2[i] = MPI_UNDEFINED;
    if (ranks1[i] == MPI_PROC_NULL) { ranks2[i] = MPI_PROC_NULL; continue; }
    const int w = a->world[ranks1[i]];
    for (int k = 0; k < b->n; k++) if (b->world[k] == w) ranks2[i] = k;
  }
  return MPI_SUCCESS;
}
int MPI_Group_free(MPI_Group *group)
{
  if (*group > 1) { Group *g = group_of(*group); free(g->world); g->used = 0; }
  *group = MPI_GROUP_NULL;
  return MPI_SUCCESS;
}

/* collective over `comm`; ranks outside `group` get MPI_COMM_NULL */
int MPI_Comm_create(MPI_Comm comm, MPI_Group group, MPI_Comm *newcomm)
{
  Comm *c = comm_of(comm);
  Group *g = group_of(group);
  int ctx = 0;
  if (c->rank == 0) ctx = new_ctx();
  if (c->size > 1) MPI_Bcast(&ctx, 1, MPI_INT, 0, comm);
  int me = -1;
  for (int i = 0; i < g->n; i++) if (g->world[i] == G.rank) me = i;
  if (me < 0) { *newcomm = MPI_COMM_NULL; return MPI_SUCCESS; }
  const int h = comm_alloc();
  g = group_of(group);
  comm_fill(h, ctx, g->n, me, g->world);
  *newcomm = h;
  return MPI_SUCCESS;
}

int MPI_Comm_free(MPI_Comm *comm)
{
  if (*comm == MPI_COMM_NULL) return MPI_SUCCESS;
  Comm *c = comm_of(*comm);
  /* delete callbacks; a callback may delete further attributes of this very communicator */
  while (c->nattr > 0) {
    const int kv = c->attr[c->nattr - 1].keyval;
    MPI_Attr_delete(*comm, kv);
    c = comm_of(*comm);
  }
  if (*comm > MPI_COMM_SELF) { free(c->world); free(c->inv); free(c->attr); memset(c, 0, sizeof(*c)); }
  *comm = MPI_COMM_NULL;
  return MPI_SUCCESS;
}

int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int *result)
{
  if (a == b) { *result = MPI_IDENT; return MPI_SUCCESS; }
  Comm *x = comm_of(a), *y = comm_of(b);
  if (x->size != y->size) { *result = MPI_UNEQUAL; return MPI_SUCCESS; }
  int same = 1, similar = 1;
  for (int i = 0; i < x->size; i++) { if (x->world[i] != y->world[i]) same = 0; if (y->inv[x->world[i]] < 0) similar = 0; }
  *result = same ? MPI_CONGRUENT : (similar ? MPI_SIMILAR : MPI_UNEQUAL);
  return MPI_SUCCESS;
}

/* ------------------------------------------------------------------------------------------------ attributes */
int MPI_Keyval_create(MPI_Copy_function *copy_fn, MPI_Delete_function *delete_fn, int *keyval, void *extra_state)
{
  int k;
  for (k = 0; k < G.nkeys; k++) if (!G.keys[k].used) break;
  if (k >= G.nkeys) {
    int n = G.nkeys ? 2 * G.nkeys : 16;
    G.keys = (Keyval *)realloc(G.keys, sizeof(Keyval) * (size_t)n);
    memset(G.keys + G.nkeys, 0, sizeof(Keyval) * (size_t)(n - G.nkeys));
    k = G.nkeys;
    G.nkeys = n;
  }
  G.keys[k].used = 1; G.keys[k].copy = copy_fn; G.keys[k].del = delete_fn; G.keys[k].extra = extra_state;
  *keyval = k + KEY_USER0;
  return MPI_SUCCESS;
}

int MPI_Keyval_free(int *keyval)
{
  const int k = *keyval - KEY_USER0;
  if (k >= 0 && k < G.nkeys) G.keys[k].used = 0;
  *keyval = MPI_KEYVAL_INVALID;
  return MPI_SUCCESS;
}

int MPI_Attr_put(MPI_Comm comm, int keyval, void *attr)
{
  Comm *c = comm_of(comm);
  for (int i = 0; i < c->nattr; i++) if (c->attr[i].keyval == keyval) {
    Keyval *k = &G.keys[keyval - KEY_USER0];
    void *old = c->attr[i].val;
    if (k->del && old != attr) { k->del(comm, keyval, old, k->extra); c = comm_of(comm); }
    c->attr[i].val = attr;
    return MPI_SUCCESS;
  }
  if (c->nattr == c->cattr) { c->cattr = c->cattr ? 2 * c->cattr : 8; c->attr = (Attr *)realloc(c->attr, sizeof(Attr) * (size_t)c->cattr); }
  c->attr[c->nattr].keyval = keyval; c->attr[c->nattr].val = attr; c->nattr++;
  return MPI_SUCCESS;
}

int MPI_Attr_get(MPI_Comm comm, int keyval, void *attr_ptr, int *flag)
{
  Comm *c = comm_of(comm);
  if (keyval == MPI_TAG_UB) { *(void **)attr_ptr = &G.tag_ub; *flag = 1; return MPI_SUCCESS; }
  if (keyval >= 0 && keyval < KEY_USER0) { static int zero = 0; *(void **)attr_ptr = &zero; *flag = 1; return MPI_SUCCESS; }
  *flag = 0;
  for (int i = 0; i < c->nattr; i++) if (c->attr[i].keyval == keyval) { *(void **)attr_ptr = c->attr[i].val; *flag = 1; break; }
  return MPI_SUCCESS;
}

int MPI_Attr_delete(MPI_Comm comm, int keyval)
{
  Comm *c = comm_of(comm);
  int found = 0;
  void *val = NULL;
  for (int i = 0; i < c->nattr; i++) if (c->attr[i].keyval == keyval) { found = 1; val = c->attr[i].val; break; }
  if (!found) return MPI_SUCCESS;
  /* the callback runs while the attribute is still attached (MPICH's order; PETSc's Petsc_DelComm looks the attribute
   * up again from inside the callback, src/sys/objects/pinit.c:422-441) and may itself delete attributes */
  const int k = keyval - KEY_USER0;
  if (k >= 0 && k < G.nkeys && G.keys[k].used && G.keys[k].del) G.keys[k].del(comm, keyval, val, G.keys[k].extra);
  c = comm_of(comm);
  for (int i = 0; i < c->nattr; i++) if (c->attr[i].keyval == keyval) {
    for (int j = i; j + 1 < c->nattr; j++) c->attr[j] = c->attr[j + 1];
    c->nattr--;
    break;
  }
  return MPI_SUCCESS;
}

/* ------------------------------------------------------------------------------------------------ datatypes, ops */
int MPI_Type_size(MPI_Datatype dt, int *size) { *size = (int)dt_size(dt); return MPI_SUCCESS; }
int MPI_Type_extent(MPI_Datatype dt, MPI_Aint *extent) { *extent = (MPI_Aint)dt_size(dt); return MPI_SUCCESS; }
int MPI_Type_get_extent(MPI_Datatype dt, MPI_Aint *lb, MPI_Aint *extent) { *lb = 0; *extent = (MPI_Aint)dt_size(dt); return MPI_SUCCESS; }

static int dtype_new(size_t size, MPI_Datatype base, int mult)
{
  int k;
  for (k = 0; k < G.ndts; k++) if (!G.dts[k].used) break;
  if (k >= G.ndts) {
    int n = G.ndts ? 2 * G.ndts : 16;
    G.dts = (Dtype *)realloc(G.dts, sizeof(Dtype) * (size_t)n);
    memset(G.dts + G.ndts, 0, sizeof(Dtype) * (size_t)(n - G.ndts));
    k = G.ndts;
    G.ndts = n;
  }
  G.dts[k].used = 1; G.dts[k].size = size; G.dts[k].base = base; G.dts[k].mult = mult;
  return k + DT_DERIVED0;
}

int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newtype)
{
  MPI_Datatype base = old; int mult = count;
  if (old >= DT_DERIVED0) { base = G.dts[old - DT_DERIVED0].base; mult = count * G.dts[old - DT_DERIVED0].mult; }
  *newtype = dtype_new((size_t)count * dt_size(old), base, mult);
  return MPI_SUCCESS;
}
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype old, MPI_Datatype *newtype)
{
  if (stride != blocklength && count > 1) die("MPI_Type_vector with gaps (stride %d, block %d) is outside the b200mpi subset", stride, blocklength);
  return MPI_Type_contiguous(count * blocklength, old, newtype);
}
int MPI_Type_dup(MPI_Datatype old, MPI_Datatype *newtype) { return MPI_Type_contiguous(1, old, newtype); }
int MPI_Type_commit(MPI_Datatype *dt) { (void)dt; return MPI_SUCCESS; }
int MPI_Type_free(MPI_Datatype *dt)
{
  if (*dt >= DT_DERIVED0 && *dt - DT_DERIVED0 < G.ndts) G.dts[*dt - DT_DERIVED0].used = 0;
  *dt = MPI_DATATYPE_NULL;
  return MPI_SUCCESS;
}

int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op)
{
  int k;
  for (k = 0; k < G.nops; k++) if (!G.ops[k].used) break;
  if (k >= G.nops) {
    int n = G.nops ? 2 * G.nops : 16;
    G.ops = (UserOp *)realloc(G.ops, sizeof(UserOp) * (size_t)n);
    memset(G.ops + G.nops, 0, sizeof(UserOp) * (size_t)(n - G.nops));
    k = G.nops;
    G.nops = n;
  }
  G.ops[k].used = 1; G.ops[k].fn = fn; G.ops[k].commute = commute;
  *op = k + OP_USER0;
  return MPI_SUCCESS;
}
int MPI_Op_free(MPI_Op *op)
{
  if (*op >= OP_USER0 && *op - OP_USER0 < G.nops) G.ops[*op - OP_USER0].used = 0;
  *op = MPI_OP_NULL;
  return MPI_SUCCESS;
}

int MPI_Get_count(const MPI_Status *status, MPI_Datatype dt, int *count)
{
  const size_t s = dt_size(dt);
  *count = s ? (int)((size_t)status->b200_bytes / s) : 0;
  return MPI_SUCCESS;
}
int MPI_Get_elements(const MPI_Status *status, MPI_Datatype dt, int *count) { return MPI_Get_count(status, dt, count); }
int MPI_Address(void *location, MPI_Aint *address) { *address = (MPI_Aint)(intptr_t)location; return MPI_SUCCESS; }

/* inout[i] = in[i] (op) inout[i] — `in` is the LEFT operand, as MPI defines user functions */
#define ARITH(T) do { const T *a = (const T *)in; T *b = (T *)inout; \
    switch (op) { \
    case MPI_SUM:  for (long i = 0; i < n; i++) b[i] = a[i] + b[i]; break; \
    case MPI_PROD: for (long i = 0; i < n; i++) b[i] = a[i] * b[i]; break; \
    case MPI_MAX:  for (long i = 0; i < n; i++) b[i] = a[i] > b[i] ? a[i] : b[i]; break; \
    case MPI_MIN:  for (long i = 0; i < n; i++) b[i] = a[i] < b[i] ? a[i] : b[i]; break; \
    case MPI_LAND: for (long i = 0; i < n; i++) b[i] = (T)(a[i] && b[i]); break; \
    case MPI_LOR:  for (long i = 0; i < n; i++) b[i] = (T)(a[i] || b[i]); break; \
    case MPI_LXOR: for (long i = 0; i < n; i++) b[i] = (T)((!a[i]) != (!b[i])); break; \
    default: return 1; } } while (0)
#define BITS(T) do { const T *a = (const T *)in; T *b = (T *)inout; \
    switch (op) { \
    case MPI_BAND: for (long i = 0; i < n; i++) b[i] = a[i] & b[i]; return 0; \
    case MPI_BOR:  for (long i = 0; i < n; i++) b[i] = a[i] | b[i]; return 0; \
    case MPI_BXOR: for (long i = 0; i < n; i++) b[i] = a[i] ^ b[i]; return 0; \
    default: break; } } while (0)
#define LOC(S) do { const S *a = (const S *)in; S *b = (S *)inout; \
    for (long i = 0; i < n; i++) { \
      if (op == MPI_MAXLOC) { if (a[i].v > b[i].v || (a[i].v == b[i].v && a[i].i < b[i].i)) b[i] = a[i]; } \
      else                  { if (a[i].v < b[i].v || (a[i].v == b[i].v && a[i].i < b[i].i)) b[i] = a[i]; } } } while (0)

static int builtin_apply(MPI_Op op, MPI_Datatype dt, const void *in, void *inout, long n)
{
  if (op == MPI_REPLACE) { memcpy(inout, in, (size_t)n * dt_size(dt)); return 0; }
  if (op == MPI_MAXLOC || op == MPI_MINLOC) {
    typedef struct { int v, i; } P2I; typedef struct { double v; int i; } PDI; typedef struct { float v; int i; } PFI;
    typedef struct { long v; int i; } PLI; typedef struct { short v; int i; } PSI; typedef struct { double v, i; } P2D;
    switch (dt) {
    case MPI_2INT: LOC(P2I); return 0;
    case MPI_DOUBLE_INT: LOC(PDI); return 0;
    case MPI_FLOAT_INT: LOC(PFI); return 0;
    case MPI_LONG_INT: LOC(PLI); return 0;
    case MPI_SHORT_INT: LOC(PSI); return 0;
    case MPI_2DOUBLE: LOC(P2D); return 0;
    default: return 1;
    }
  }
  switch (dt) {
  case MPI_CHAR: case MPI_SIGNED_CHAR:       BITS(signed char); ARITH(signed char); break;
  case MPI_UNSIGNED_CHAR: case MPI_BYTE:     BITS(unsigned char); ARITH(unsigned char); break;
  case MPI_SHORT:                            BITS(short); ARITH(short); break;
  case MPI_UNSIGNED_SHORT:                   BITS(unsigned short); ARITH(unsigned short); break;
  case MPI_INT:                              BITS(int); ARITH(int); break;
  case MPI_UNSIGNED:                         BITS(unsigned); ARITH(unsigned); break;
  case MPI_LONG:                             BITS(long); ARITH(long); break;
  case MPI_UNSIGNED_LONG:                    BITS(unsigned long); ARITH(unsigned long); break;
  case MPI_LONG_LONG_INT:                    BITS(long long); ARITH(long long); break;
  case MPI_UNSIGNED_LONG_LONG:               BITS(unsigned long long); ARITH(unsigned long long); break;
  case MPI_FLOAT:                            ARITH(float); break;
  case MPI_DOUBLE:                           ARITH(double); break;
  case MPI_LONG_DOUBLE:                      ARITH(long double); break;
  default: return 1;
  }
  return 0;
}

static void apply_op(MPI_Op op, MPI_Datatype dt, void *in, void *inout, int count)
{
  if (op >= OP_USER0) {
    if (op - OP_USER0 >= G.nops || !G.ops[op - OP_USER0].used) die("invalid reduction operation %d", op);
    G.ops[op - OP_USER0].fn(in, inout, &count, &dt);
    return;
  }
  MPI_Datatype base = dt; long n = count;
  if (dt >= DT_DERIVED0) { base = G.dts[dt - DT_DERIVED0].base; n = (long)count * G.dts[dt - DT_DERIVED0].mult; }
  if (builtin_apply(op, base, in, inout, n)) die("reduction operation %d is not defined for datatype %d", op, dt);
}

/* ------------------------------------------------------------------------------------------------ point to point */
static int isend_common(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req, int persistent)
{
  Comm *c = comm_of(comm);
  if (tag < 0 || tag > G.tag_ub) die("send tag %d out of range", tag);
  Req *r = make_req(1, (void *)buf, (size_t)count * dt_size(dt), world_of(c, dest, 0), tag, c->ctx, comm);
  r->persistent = persistent; r->peer_arg = dest; r->count = count; r->dt = dt;
  if (!persistent) start_send(r);
  *req = r->handle;
  return MPI_SUCCESS;
}

static int irecv_common(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req, int persistent)
{
  Comm *c = comm_of(comm);
  Req *r = make_req(2, buf, (size_t)count * dt_size(dt), world_of(c, source, 1), tag, c->ctx, comm);
  r->persistent = persistent; r->peer_arg = source; r->count = count; r->dt = dt;
  if (!persistent) start_recv(r);
  *req = r->handle;
  return MPI_SUCCESS;
}

int MPI_Isend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 0); }
int MPI_Issend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 0); }
int MPI_Irsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 0); }
int MPI_Irecv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req) { return irecv_common(buf, count, dt, source, tag, comm, req, 0); }
int MPI_Send_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 1); }
int MPI_Ssend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 1); }
int MPI_Rsend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 1); }
int MPI_Bsend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req) { return isend_common(buf, count, dt, dest, tag, comm, req, 1); }
int MPI_Recv_init(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req) { return irecv_common(buf, count, dt, source, tag, comm, req, 1); }

int MPI_Start(MPI_Request *req)
{
  Req *r = req_of(*req);
  if (!r->persistent) die("MPI_Start on a request that is not persistent");
  if (r->active && !r->done) die("MPI_Start on a request that is still active");
  if (r->kind == 1) start_send(r); else start_recv(r);
  return MPI_SUCCESS;
}
int MPI_Startall(int count, MPI_Request *reqs) { for (int i = 0; i < count; i++) MPI_Start(&reqs[i]); return MPI_SUCCESS; }

static void set_status(MPI_Status *status, const Req *r)
{
  if (!status) return;
  if (r && r->kind == 2) *status = r->st;
  else { status->MPI_SOURCE = MPI_ANY_SOURCE; status->MPI_TAG = MPI_ANY_TAG; status->MPI_ERROR = MPI_SUCCESS; status->b200_bytes = 0; status->b200_cancelled = 0; }
}

/* a completed request leaves the user's hands: freed (plain) or made inactive (persistent) */
static void retire(MPI_Request *h, Req *r)
{
  if (r->persistent) r->active = 0;
  else { req_free(r); *h = MPI_REQUEST_NULL; }
}

static inline int is_active(MPI_Request h)
{
  if (h == MPI_REQUEST_NULL) return 0;
  Req *r = req_of(h);
  return !(r->persistent && !r->active);
}

int MPI_Wait(MPI_Request *req, MPI_Status *status)
{
  if (!is_active(*req)) { set_status(status, NULL); return MPI_SUCCESS; }
  Req *r = req_of(*req);
  wait_req(r);
  set_status(status, r);
  retire(req, r);
  return MPI_SUCCESS;
}

int MPI_Test(MPI_Request *req, int *flag, MPI_Status *status)
{
  if (!is_active(*req)) { *flag = 1; set_status(status, NULL); return MPI_SUCCESS; }
  Req *r = req_of(*req);
  if (!r->done) progress();
  *flag = r->done;
  if (r->done) { set_status(status, r); retire(req, r); }
  return MPI_SUCCESS;
}

int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *statuses)
{
  for (int i = 0; i < count; i++) MPI_Wait(&reqs[i], statuses ? &statuses[i] : MPI_STATUS_IGNORE);
  return MPI_SUCCESS;
}

int MPI_Testall(int count, MPI_Request *reqs, int *flag, MPI_Status *statuses)
{
  progress();
  for (int i = 0; i < count; i++) if (is_active(reqs[i]) && !req_of(reqs[i])->done) { *flag = 0; return MPI_SUCCESS; }
  *flag = 1;
  return MPI_Waitall(count, reqs, statuses);
}

int MPI_Waitany(int count, MPI_Request *reqs, int *index, MPI_Status *status)
{
  unsigned spins = 0;
  for (;;) {
    int live = 0;
    for (int i = 0; i < count; i++) {
      if (!is_active(reqs[i])) continue;
      live = 1;
      Req *r = req_of(reqs[i]);
      if (r->done) { *index = i; set_status(status, r); retire(&reqs[i], r); return MPI_SUCCESS; }
    }
    if (!live) { *index = MPI_UNDEFINED; set_status(status, NULL); return MPI_SUCCESS; }
    progress();
    idle(&spins);
  }
}

int MPI_Testany(int count, MPI_Request *reqs, int *index, int *flag, MPI_Status *status)
{
  int live = 0;
  progress();
  for (int i = 0; i < count; i++) {
    if (!is_active(reqs[i])) continue;
    live = 1;
    Req *r = req_of(reqs[i]);
    if (r->done) { *index = i; *flag = 1; set_status(status, r); retire(&reqs[i], r); return MPI_SUCCESS; }
  }
  *index = MPI_UNDEFINED; *flag = live ? 0 : 1;
  return MPI_SUCCESS;
}

int MPI_Waitsome(int incount, MPI_Request *reqs, int *outcount, int *indices, MPI_Status *statuses)
{
  unsigned spins = 0;
  for (;;) {
    int live = 0, n = 0;
    for (int i = 0; i < incount; i++) {
      if (!is_active(reqs[i])) continue;
      live = 1;
      Req *r = req_of(reqs[i]);
      if (r->done) { indices[n] = i; set_status(statuses ? &statuses[n] : NULL, r); retire(&reqs[i], r); n++; }
    }
    if (!live) { *outcount = MPI_UNDEFINED; return MPI_SUCCESS; }
    if (n) { *outcount = n; return MPI_SUCCESS; }
    progress();
    idle(&spins);
  }
}

int MPI_Request_free(MPI_Request *req)
{
  if (*req == MPI_REQUEST_NULL) return MPI_SUCCESS;
  Req *r = req_of(*req);
  if (r->active && !r->done) {                        /* let it finish first: nobody else will ever look at it again */
    if (r->kind == 2) posted_remove(r);
    else wait_req(r);
  }
  req_free(r);
  *req = MPI_REQUEST_NULL;
  return MPI_SUCCESS;
}

int MPI_Cancel(MPI_Request *req)
{
  Req *r = req_of(*req);
  if (r->kind == 2 && r->active && !r->done) { posted_remove(r); r->done = 1; r->st.b200_cancelled = 1; }
  return MPI_SUCCESS;
}
int MPI_Test_cancelled(const MPI_Status *status, int *flag) { *flag = status->b200_cancelled; return MPI_SUCCESS; }

int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm)
{
  MPI_Request r;
  isend_common(buf, count, dt, dest, tag, comm, &r, 0);
  return MPI_Wait(&r, MPI_STATUS_IGNORE);
}
int MPI_Ssend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm) { return MPI_Send(buf, count, dt, dest, tag, comm); }
int MPI_Rsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm) { return MPI_Send(buf, count, dt, dest, tag, comm); }
int MPI_Bsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm) { return MPI_Send(buf, count, dt, dest, tag, comm); }

int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status)
{
  MPI_Request r;
  irecv_common(buf, count, dt, source, tag, comm, &r, 0);
  return MPI_Wait(&r, status);
}

int MPI_Sendrecv(const void *sbuf, int scount, MPI_Datatype sdt, int dest, int stag,
                 void *rbuf, int rcount, MPI_Datatype rdt, int source, int rtag, MPI_Comm comm, MPI_Status *status)
{
  MPI_Request r[2];
  irecv_common(rbuf, rcount, rdt, source, rtag, comm, &r[0], 0);
  isend_common(sbuf, scount, sdt, dest, stag, comm, &r[1], 0);
  MPI_Wait(&r[1], MPI_STATUS_IGNORE);
  return MPI_Wait(&r[0], status);
}

int MPI_Iprobe(int source, int tag, MPI_Comm comm, int *flag, MPI_Status *status)
{
  Comm *c = comm_of(comm);
  Req probe; memset(&probe, 0, sizeof(probe));
  probe.ctx = c->ctx; probe.tag = tag; probe.peer_world = world_of(c, source, 1);
  progress();
  *flag = 0;
  for (Unexp *u = G.ux_head; u; u = u->next) {
    if (u->claimed || !req_matches(&probe, u->ctx, u->src_world, u->tag)) continue;
    *flag = 1;
    if (status) { status->MPI_SOURCE = c->inv[u->src_world]; status->MPI_TAG = u->tag; status->MPI_ERROR = MPI_SUCCESS; status->b200_bytes = (int)u->nbytes; status->b200_cancelled = 0; }
    break;
  }
  return MPI_SUCCESS;
}

int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status)
{
  int flag = 0; unsigned spins = 0;
  for (;;) { MPI_Iprobe(source, tag, comm, &flag, status); if (flag) return MPI_SUCCESS; idle(&spins); }
}

/* ------------------------------------------------------------------------------------------------ collectives */
int MPI_Barrier(MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  for (int k = 1; k < n; k <<= 1) {                    /* dissemination barrier */
    char t = 0, u = 0;
    Req *r = cirecv(c, comm, &u, 1, (me - k + n) % n, tag);
    csend(c, comm, &t, 1, (me + k) % n, tag);
    wait_req(r); req_free(r);
  }
  return MPI_SUCCESS;
}

int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, tag = coll_tag(c);
  if (n == 1) return MPI_SUCCESS;
  const size_t bytes = (size_t)count * dt_size(dt);
  const int vr = (c->rank - root + n) % n;
  int mask = 1;
  while (mask < n) { if (vr & mask) { crecv(c, comm, buf, bytes, (vr - mask + root) % n, tag); break; } mask <<= 1; }
  mask >>= 1;
  while (mask > 0) { if (vr + mask < n) csend(c, comm, buf, bytes, (vr + mask + root) % n, tag); mask >>= 1; }
  return MPI_SUCCESS;
}

int MPI_Reduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  const size_t bytes = (size_t)count * dt_size(dt);
  const void *src = (sbuf == MPI_IN_PLACE) ? rbuf : sbuf;
  if (n == 1) { if (src != rbuf) memcpy(rbuf, src, bytes); return MPI_SUCCESS; }
  /* binomial tree over the ranks in their natural order (rank 0 collects), operands combined lower-rank (op) higher-rank:
   * the association is fixed by the communicator size alone, so floating-point sums are reproducible */
  char *acc = (char *)malloc(bytes ? bytes : 1), *tmp = (char *)malloc(bytes ? bytes : 1);
  memcpy(acc, src, bytes);
  int mask = 1;
  while (mask < n) {
    if (me & mask) { csend(c, comm, acc, bytes, me - mask, tag); break; }
    if (me + mask < n) {
      crecv(c, comm, tmp, bytes, me + mask, tag);
      apply_op(op, dt, acc, tmp, count);               /* tmp = acc (op) tmp */
      char *t = acc; acc = tmp; tmp = t;
    }
    mask <<= 1;
  }
  if (root == 0) { if (me == 0) memcpy(rbuf, acc, bytes); }
  else {
    if (me == 0) csend(c, comm, acc, bytes, root, tag);
    else if (me == root) crecv(c, comm, rbuf, bytes, 0, tag);
  }
  free(acc); free(tmp);
  return MPI_SUCCESS;
}

int MPI_Allreduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
  MPI_Reduce(sbuf, rbuf, count, dt, op, 0, comm);
  return MPI_Bcast(rbuf, count, dt, 0, comm);
}

int MPI_Scan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  const size_t bytes = (size_t)count * dt_size(dt);
  if (sbuf != MPI_IN_PLACE && sbuf != rbuf) memcpy(rbuf, sbuf, bytes);
  if (me > 0) {
    char *tmp = (char *)malloc(bytes ? bytes : 1);
    crecv(c, comm, tmp, bytes, me - 1, tag);
    apply_op(op, dt, tmp, rbuf, count);                /* rbuf = prefix(0..me-1) (op) mine */
    free(tmp);
  }
  if (me + 1 < n) csend(c, comm, rbuf, bytes, me + 1, tag);
  return MPI_SUCCESS;
}

int MPI_Exscan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  const size_t bytes = (size_t)count * dt_size(dt);
  char *mine = (char *)malloc(bytes ? bytes : 1);
  memcpy(mine, sbuf == MPI_IN_PLACE ? rbuf : sbuf, bytes);
  if (me > 0) {
    crecv(c, comm, rbuf, bytes, me - 1, tag);          /* prefix(0..me-1) */
    apply_op(op, dt, rbuf, mine, count);               /* mine = prefix (op) mine */
  }
  if (me + 1 < n) csend(c, comm, mine, bytes, me + 1, tag);
  free(mine);
  return MPI_SUCCESS;
}

int MPI_Reduce_scatter(const void *sbuf, void *rbuf, const int *rcounts, MPI_Datatype dt, MPI_Op op, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  int total = 0, *displs = (int *)malloc(sizeof(int) * (size_t)n);
  for (int i = 0; i < n; i++) { displs[i] = total; total += rcounts[i]; }
  char *all = (char *)malloc((size_t)total * dt_size(dt) + 1);
  MPI_Reduce(sbuf, all, total, dt, op, 0, comm);
  MPI_Scatterv(all, rcounts, displs, dt, rbuf, rcounts[c->rank], dt, 0, comm);
  free(all); free(displs);
  return MPI_SUCCESS;
}

int MPI_Gatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs, MPI_Datatype rdt, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  if (me != root) { csend(c, comm, sbuf, (size_t)scount * dt_size(sdt), root, tag); return MPI_SUCCESS; }
  const size_t rs = dt_size(rdt);
  Req **rq = (Req **)malloc(sizeof(Req *) * (size_t)n);
  for (int i = 0; i < n; i++) {
    rq[i] = NULL;
    if (i == me) { if (sbuf != MPI_IN_PLACE) memcpy((char *)rbuf + (size_t)displs[i] * rs, sbuf, (size_t)scount * dt_size(sdt)); }
    else rq[i] = cirecv(c, comm, (char *)rbuf + (size_t)displs[i] * rs, (size_t)rcounts[i] * rs, i, tag);
  }
  for (int i = 0; i < n; i++) if (rq[i]) { wait_req(rq[i]); req_free(rq[i]); }
  free(rq);
  return MPI_SUCCESS;
}

int MPI_Gather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  int *cnt = (int *)malloc(sizeof(int) * 2 * (size_t)n), *dis = cnt + n;
  for (int i = 0; i < n; i++) { cnt[i] = rcount; dis[i] = i * rcount; }
  const int rc = MPI_Gatherv(sbuf, scount, sdt, rbuf, cnt, dis, rdt, root, comm);
  free(cnt);
  return rc;
}

int MPI_Scatterv(const void *sbuf, const int *scounts, const int *displs, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  if (me != root) { crecv(c, comm, rbuf, (size_t)rcount * dt_size(rdt), root, tag); return MPI_SUCCESS; }
  const size_t ss = dt_size(sdt);
  Req **rq = (Req **)malloc(sizeof(Req *) * (size_t)n);
  for (int i = 0; i < n; i++) {
    rq[i] = NULL;
    if (i == me) { if (rbuf != MPI_IN_PLACE) memcpy(rbuf, (const char *)sbuf + (size_t)displs[i] * ss, (size_t)scounts[i] * ss); }
    else rq[i] = cisend(c, comm, (const char *)sbuf + (size_t)displs[i] * ss, (size_t)scounts[i] * ss, i, tag);
  }
  for (int i = 0; i < n; i++) if (rq[i]) { wait_req(rq[i]); req_free(rq[i]); }
  free(rq);
  return MPI_SUCCESS;
}

int MPI_Scatter(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  int *cnt = (int *)malloc(sizeof(int) * 2 * (size_t)n), *dis = cnt + n;
  for (int i = 0; i < n; i++) { cnt[i] = scount; dis[i] = i * scount; }
  const int rc = MPI_Scatterv(sbuf, cnt, dis, sdt, rbuf, rcount, rdt, root, comm);
  free(cnt);
  return rc;
}

int MPI_Alltoallv(const void *sbuf, const int *scounts, const int *sdispls, MPI_Datatype sdt,
                  void *rbuf, const int *rcounts, const int *rdispls, MPI_Datatype rdt, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  const size_t ss = dt_size(sdt), rs = dt_size(rdt);
  Req **rq = (Req **)calloc(2 * (size_t)n, sizeof(Req *));
  for (int i = 0; i < n; i++) if (i != me) rq[i] = cirecv(c, comm, (char *)rbuf + (size_t)rdispls[i] * rs, (size_t)rcounts[i] * rs, i, tag);
  memcpy((char *)rbuf + (size_t)rdispls[me] * rs, (const char *)sbuf + (size_t)sdispls[me] * ss, (size_t)scounts[me] * ss);
  for (int k = 1; k < n; k++) { const int i = (me + k) % n; rq[n + i] = cisend(c, comm, (const char *)sbuf + (size_t)sdispls[i] * ss, (size_t)scounts[i] * ss, i, tag); }
  for (int i = 0; i < 2 * n; i++) if (rq[i]) { wait_req(rq[i]); req_free(rq[i]); }
  free(rq);
  return MPI_SUCCESS;
}

int MPI_Alltoall(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  int *v = (int *)malloc(sizeof(int) * 4 * (size_t)n);
  for (int i = 0; i < n; i++) { v[i] = scount; v[n + i] = i * scount; v[2 * n + i] = rcount; v[3 * n + i] = i * rcount; }
  const int rc = MPI_Alltoallv(sbuf, v, v + n, sdt, rbuf, v + 2 * n, v + 3 * n, rdt, comm);
  free(v);
  return rc;
}

int MPI_Allgatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs, MPI_Datatype rdt, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size, me = c->rank, tag = coll_tag(c);
  const size_t rs = dt_size(rdt);
  const void *mine = (sbuf == MPI_IN_PLACE) ? (const char *)rbuf + (size_t)displs[me] * rs : sbuf;
  const size_t mybytes = (sbuf == MPI_IN_PLACE) ? (size_t)rcounts[me] * rs : (size_t)scount * dt_size(sdt);
  Req **rq = (Req **)calloc(2 * (size_t)n, sizeof(Req *));
  for (int i = 0; i < n; i++) if (i != me) rq[i] = cirecv(c, comm, (char *)rbuf + (size_t)displs[i] * rs, (size_t)rcounts[i] * rs, i, tag);
  if (sbuf != MPI_IN_PLACE) memcpy((char *)rbuf + (size_t)displs[me] * rs, mine, mybytes);
  for (int k = 1; k < n; k++) { const int i = (me + k) % n; rq[n + i] = cisend(c, comm, mine, mybytes, i, tag); }
  for (int i = 0; i < 2 * n; i++) if (rq[i]) { wait_req(rq[i]); req_free(rq[i]); }
  free(rq);
  return MPI_SUCCESS;
}

int MPI_Allgather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm)
{
  Comm *c = comm_of(comm);
  const int n = c->size;
  int *cnt = (int *)malloc(sizeof(int) * 2 * (size_t)n), *dis = cnt + n;
  for (int i = 0; i < n; i++) { cnt[i] = rcount; dis[i] = i * rcount; }
  const int rc = MPI_Allgatherv(sbuf, scount, sdt, rbuf, cnt, dis, rdt, comm);
  free(cnt);
  return rc;
}
