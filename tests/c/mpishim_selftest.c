/* Self-test of the b200mpi subset (petsc-dev_b200/mpishim): run as  b200mpiexec -n P mpishim_selftest.
 * Prints "b200mpi selftest ok P" from rank 0 when every check passed on every rank; any failure aborts the job. */
#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(c) do { if (!(c)) { fprintf(stderr, "[%d] CHECK failed line %d: %s\n", rank, __LINE__, #c); MPI_Abort(MPI_COMM_WORLD, 3); } } while (0)

static int ndel = 0;
static int del_fn(MPI_Comm comm, int keyval, void *attr, void *extra) { (void)comm; (void)keyval; (void)attr; (void)extra; ndel++; return MPI_SUCCESS; }
static void maxsum(void *in, void *inout, int *len, MPI_Datatype *dt)
{
  int *a = (int *)in, *b = (int *)inout;
  (void)dt;
  for (int i = 0; i < *len; i++) { b[2 * i] = a[2 * i] > b[2 * i] ? a[2 * i] : b[2 * i]; b[2 * i + 1] += a[2 * i + 1]; }
}

int main(int argc, char **argv)
{
  int rank, size, flag;
  MPI_Init(&argc, &argv);
  MPI_Comm_rank(MPI_COMM_WORLD, &rank);
  MPI_Comm_size(MPI_COMM_WORLD, &size);
  MPI_Initialized(&flag); CHECK(flag);
  const int right = (rank + 1) % size, left = (rank + size - 1) % size;

  /* ring of blocking sends (eager), big messages larger than any ring, both directions at once */
  {
    const int n = 3 * 1024 * 1024;                      /* 24 MB of doubles */
    double *s = malloc(sizeof(double) * n), *r = malloc(sizeof(double) * n);
    for (int i = 0; i < n; i++) s[i] = rank * 1e6 + i;
    MPI_Request q[2];
    MPI_Irecv(r, n, MPI_DOUBLE, left, 7, MPI_COMM_WORLD, &q[0]);
    MPI_Isend(s, n, MPI_DOUBLE, right, 7, MPI_COMM_WORLD, &q[1]);
    MPI_Status st[2];
    MPI_Waitall(2, q, st);
    int cnt; MPI_Get_count(&st[0], MPI_DOUBLE, &cnt);
    CHECK(cnt == n && st[0].MPI_SOURCE == left && st[0].MPI_TAG == 7);
    for (int i = 0; i < n; i += 4099) CHECK(r[i] == left * 1e6 + i);
    CHECK(q[0] == MPI_REQUEST_NULL && q[1] == MPI_REQUEST_NULL);
    /* blocking send before the receive is posted (unexpected queue), odd byte counts */
    char msg[24], got[24];
    snprintf(msg, sizeof(msg), "hello %05d", rank);
    MPI_Send(msg, 11, MPI_CHAR, right, 3, MPI_COMM_WORLD);
    MPI_Send(s, 5, MPI_DOUBLE, right, 4, MPI_COMM_WORLD);
    MPI_Recv(r, 5, MPI_DOUBLE, left, 4, MPI_COMM_WORLD, MPI_STATUS_IGNORE);      /* tags out of order */
    MPI_Recv(got, 11, MPI_CHAR, left, 3, MPI_COMM_WORLD, &st[0]);
    char want[24]; snprintf(want, sizeof(want), "hello %05d", left);
    CHECK(!memcmp(got, want, 11) && r[4] == left * 1e6 + 4);
    free(s); free(r);
  }
  /* MPI_ANY_SOURCE / ANY_TAG + MPI_Waitany: everybody sends its rank to rank 0 */
  {
    if (rank == 0) {
      int *v = calloc(size, sizeof(int)), seen = 0;
      MPI_Request *q = malloc(sizeof(MPI_Request) * size);
      for (int i = 1; i < size; i++) MPI_Irecv(&v[i], 1, MPI_INT, MPI_ANY_SOURCE, MPI_ANY_TAG, MPI_COMM_WORLD, &q[i - 1]);
      for (int i = 1; i < size; i++) {
        int idx; MPI_Status st;
        MPI_Waitany(size - 1, q, &idx, &st);
        CHECK(idx != MPI_UNDEFINED && v[idx + 1] == st.MPI_SOURCE * 10 && st.MPI_TAG == 100 + st.MPI_SOURCE);
        seen++;
      }
      int idx; MPI_Waitany(size - 1, q, &idx, MPI_STATUS_IGNORE);
      CHECK(idx == MPI_UNDEFINED && seen == size - 1);
      free(v); free(q);
    } else { int v = rank * 10; MPI_Send(&v, 1, MPI_INT, 0, 100 + rank, MPI_COMM_WORLD); }
    MPI_Barrier(MPI_COMM_WORLD);                          /* nobody's later messages may meet the wildcard receives */
  }
  /* persistent requests, restarted three times (VecScatterBegin/End pattern) */
  {
    double sb[4], rb[4];
    MPI_Request q[2];
    MPI_Recv_init(rb, 4, MPI_DOUBLE, left, 9, MPI_COMM_WORLD, &q[0]);
    MPI_Send_init(sb, 4, MPI_DOUBLE, right, 9, MPI_COMM_WORLD, &q[1]);
    for (int it = 0; it < 3; it++) {
      for (int i = 0; i < 4; i++) sb[i] = rank + 100.0 * it + i;
      MPI_Startall(2, q);
      int idx; MPI_Status st;
      MPI_Waitany(1, q, &idx, &st);
      CHECK(idx == 0 && st.MPI_SOURCE == left);
      MPI_Waitall(1, q + 1, MPI_STATUSES_IGNORE);
      for (int i = 0; i < 4; i++) CHECK(rb[i] == left + 100.0 * it + i);
      MPI_Waitany(2, q, &idx, &st); CHECK(idx == MPI_UNDEFINED);      /* both inactive now */
    }
    MPI_Request_free(&q[0]); MPI_Request_free(&q[1]);
  }
  /* probe */
  {
    int v = rank + 5, w = -1, f = 0; MPI_Status st;
    MPI_Send(&v, 1, MPI_INT, right, 21, MPI_COMM_WORLD);
    MPI_Probe(left, 21, MPI_COMM_WORLD, &st);
    int cnt; MPI_Get_count(&st, MPI_INT, &cnt); CHECK(cnt == 1 && st.MPI_SOURCE == left);
    MPI_Recv(&w, 1, MPI_INT, left, 21, MPI_COMM_WORLD, MPI_STATUS_IGNORE); CHECK(w == left + 5);
    MPI_Iprobe(MPI_ANY_SOURCE, 21, MPI_COMM_WORLD, &f, &st); CHECK(!f);
  }
  /* collectives */
  {
    double x = rank + 1.0, s = 0;
    MPI_Allreduce(&x, &s, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD); CHECK(s == size * (size + 1) / 2.0);
    int m = rank, mx = -1, mn = -1;
    MPI_Allreduce(&m, &mx, 1, MPI_INT, MPI_MAX, MPI_COMM_WORLD); CHECK(mx == size - 1);
    MPI_Reduce(&m, &mn, 1, MPI_INT, MPI_MIN, size - 1, MPI_COMM_WORLD); if (rank == size - 1) CHECK(mn == 0);
    int sc = 0; MPI_Scan(&m, &sc, 1, MPI_INT, MPI_SUM, MPI_COMM_WORLD); CHECK(sc == rank * (rank + 1) / 2);
    int *all = malloc(sizeof(int) * size * 2);
    MPI_Allgather(&m, 1, MPI_INT, all, 1, MPI_INT, MPI_COMM_WORLD);
    for (int i = 0; i < size; i++) CHECK(all[i] == i);
    int *cnt = malloc(sizeof(int) * size), *dis = malloc(sizeof(int) * size), tot = 0;
    for (int i = 0; i < size; i++) { cnt[i] = i % 3; dis[i] = tot; tot += cnt[i]; }
    int mine[3] = {rank, rank, rank}, *gv = malloc(sizeof(int) * (tot + 1));
    MPI_Allgatherv(mine, rank % 3, MPI_INT, gv, cnt, dis, MPI_INT, MPI_COMM_WORLD);
    for (int i = 0; i < size; i++) for (int k = 0; k < cnt[i]; k++) CHECK(gv[dis[i] + k] == i);
    memset(gv, 0xff, sizeof(int) * (tot + 1));
    MPI_Gatherv(mine, rank % 3, MPI_INT, gv, cnt, dis, MPI_INT, 0, MPI_COMM_WORLD);
    if (rank == 0) for (int i = 0; i < size; i++) for (int k = 0; k < cnt[i]; k++) CHECK(gv[dis[i] + k] == i);
    int got[3] = {-1, -1, -1};
    MPI_Scatterv(gv, cnt, dis, MPI_INT, got, rank % 3, MPI_INT, 0, MPI_COMM_WORLD);
    for (int k = 0; k < rank % 3; k++) CHECK(got[k] == rank);
    int b = rank == 1 % size ? 77 : 0; MPI_Bcast(&b, 1, MPI_INT, 1 % size, MPI_COMM_WORLD); CHECK(b == 77);
    /* alltoallv: rank r sends (r*size + d) to d */
    int *sv = malloc(sizeof(int) * size), *rv = malloc(sizeof(int) * size), *one = malloc(sizeof(int) * size), *off = malloc(sizeof(int) * size);
    for (int d = 0; d < size; d++) { sv[d] = rank * size + d; one[d] = 1; off[d] = d; }
    MPI_Alltoallv(sv, one, off, MPI_INT, rv, one, off, MPI_INT, MPI_COMM_WORLD);
    for (int s2 = 0; s2 < size; s2++) CHECK(rv[s2] == s2 * size + rank);
    /* user op on a contiguous derived type (PetscMaxSum_Op on MPIU_2INT) and MPI_MAXLOC on MPI_2INT */
    MPI_Datatype t2; MPI_Op op;
    MPI_Type_contiguous(2, MPI_INT, &t2); MPI_Type_commit(&t2);
    MPI_Op_create(maxsum, 1, &op);
    int *pin = malloc(sizeof(int) * 2 * size), *pout = malloc(sizeof(int) * 2 * size);
    for (int i = 0; i < size; i++) { pin[2 * i] = (rank + i) % size; pin[2 * i + 1] = 1; }
    MPI_Allreduce(pin, pout, size, t2, op, MPI_COMM_WORLD);
    for (int i = 0; i < size; i++) CHECK(pout[2 * i] == size - 1 && pout[2 * i + 1] == size);
    MPI_Op_free(&op); MPI_Type_free(&t2);
    int loc[2] = {(rank * 7) % size, rank}, lo[2];
    MPI_Allreduce(loc, lo, 1, MPI_2INT, MPI_MAXLOC, MPI_COMM_WORLD);
    CHECK(lo[0] == (lo[1] * 7) % size);
    for (int r2 = 0; r2 < size; r2++) CHECK((r2 * 7) % size <= lo[0]);
    MPI_Barrier(MPI_COMM_WORLD);
    free(all); free(cnt); free(dis); free(gv); free(sv); free(rv); free(one); free(off); free(pin); free(pout);
  }
  /* communicators: dup (own matching context), split, attributes with delete callbacks, MPI_TAG_UB */
  {
    MPI_Comm d, half;
    MPI_Comm_dup(MPI_COMM_WORLD, &d);
    int v = rank, w = -1, u = -1;
    MPI_Request q[2];
    MPI_Irecv(&w, 1, MPI_INT, left, 1, d, &q[0]);
    MPI_Irecv(&u, 1, MPI_INT, left, 1, MPI_COMM_WORLD, &q[1]);
    int v2 = rank + 1000;
    MPI_Send(&v2, 1, MPI_INT, right, 1, MPI_COMM_WORLD);      /* same tag, different communicator: must not cross */
    MPI_Send(&v, 1, MPI_INT, right, 1, d);
    MPI_Waitall(2, q, MPI_STATUSES_IGNORE);
    CHECK(w == left && u == left + 1000);
    int key; void *p; int *ub;
    MPI_Keyval_create(MPI_NULL_COPY_FN, del_fn, &key, NULL);
    MPI_Attr_get(d, key, &p, &flag); CHECK(!flag);
    MPI_Attr_put(d, key, (void *)0x1234);
    MPI_Attr_get(d, key, &p, &flag); CHECK(flag && p == (void *)0x1234);
    MPI_Attr_get(MPI_COMM_WORLD, MPI_TAG_UB, &ub, &flag); CHECK(flag && *ub >= 32767);
    MPI_Comm_free(&d); CHECK(d == MPI_COMM_NULL && ndel == 1);
    MPI_Comm_split(MPI_COMM_WORLD, rank % 2, -rank, &half);
    int hr, hs; MPI_Comm_rank(half, &hr); MPI_Comm_size(half, &hs);
    CHECK(hs == (size + 1 - rank % 2) / 2);
    int sum = 0, one = 1; MPI_Allreduce(&one, &sum, 1, MPI_INT, MPI_SUM, half); CHECK(sum == hs);
    int top = -1; MPI_Allreduce(&rank, &top, 1, MPI_INT, MPI_MAX, half);
    if (hr == 0) CHECK(rank == top);                       /* keys were -rank: the highest world rank comes first */
    MPI_Comm_free(&half);
    int s1; MPI_Comm_size(MPI_COMM_SELF, &s1); CHECK(s1 == 1);
    int z = 5, zz = 0; MPI_Allreduce(&z, &zz, 1, MPI_INT, MPI_SUM, MPI_COMM_SELF); CHECK(zz == 5);
    MPI_Send(&z, 1, MPI_INT, 0, 2, MPI_COMM_SELF); zz = 0; MPI_Recv(&zz, 1, MPI_INT, 0, 2, MPI_COMM_SELF, MPI_STATUS_IGNORE); CHECK(zz == 5);
  }
  int ok = 1, allok = 0;
  MPI_Allreduce(&ok, &allok, 1, MPI_INT, MPI_LAND, MPI_COMM_WORLD);
  if (rank == 0) printf("b200mpi selftest ok %d\n", size);
  MPI_Finalize();
  return 0;
}
