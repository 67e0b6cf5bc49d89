/* mpi.h — b200mpi: an N-rank MPI subset for ONE node, ranks = processes, transport = shared memory.
 *
 * Why it exists: the host side of the B200 Mat/Vec plugin stays C and runs one process per GPU inside an MPI program
 * (PETSc's MPIAIJ / VecScatter setup code talks MPI: src/mat/impls/aij/mpi/mmaij.c, src/vec/vec/utils/vpscat.c,
 * src/mat/utils/matstash.c, src/sys/utils/mpimesg.c, src/sys/objects/tagm.c).  The reference ships only a ONE-rank
 * stand-in (include/mpiuni/mpi.h + src/sys/mpiuni/mpi.c); this box has no MPI at all.  b200mpi covers the surface
 * SURVEY.md A.3 lists so that (i) the unmodified reference runs on P ranks (parity oracle for garray / colmap / scatter
 * lists, CPU baseline with a real MatMult_MPIAIJ), and (ii) the plugin runs one rank per GPU from plain C programs.
 *
 * Model: `b200mpiexec -n P prog args...` creates one shared segment (memfd) holding P x P single-producer/single-consumer
 * byte rings and starts P processes; every message is a 32-byte header + payload streamed through ring[src][dst];
 * matching (communicator context, source, tag; MPI_ANY_SOURCE / MPI_ANY_TAG; non-overtaking) is done by the receiver,
 * unmatched messages are parked in an unexpected queue, so a send completes as soon as the receiver enters ANY MPI call.
 * Collectives are built on the point-to-point layer with a private context.  A program started without the launcher
 * is a 1-rank job.  Handles are small integers.  Not thread safe (neither is PETSc).
 */
#ifndef B200MPI_MPI_H
#define B200MPI_MPI_H

#ifdef __cplusplus
extern "C" {
#endif

#define B200MPI_VERSION 1
#define MPI_VERSION     2
#define MPI_SUBVERSION  2

typedef int       MPI_Comm;
typedef int       MPI_Datatype;
typedef int       MPI_Op;
typedef int       MPI_Request;
typedef int       MPI_Group;
typedef int       MPI_Errhandler;
typedef int       MPI_Info;
typedef int       MPI_Win;
typedef int       MPI_Fint;
typedef long      MPI_Aint;
typedef long long MPI_Offset;

typedef struct {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
  int b200_bytes;               /* payload size, for MPI_Get_count */
  int b200_cancelled;
} MPI_Status;

/* ---- constants ---- */
#define MPI_SUCCESS        0
#define MPI_ERR_BUFFER     1
#define MPI_ERR_COUNT      2
#define MPI_ERR_TYPE       3
#define MPI_ERR_TAG        4
#define MPI_ERR_COMM       5
#define MPI_ERR_RANK       6
#define MPI_ERR_REQUEST    7
#define MPI_ERR_ROOT       8
#define MPI_ERR_GROUP      9
#define MPI_ERR_OP         10
#define MPI_ERR_ARG        12
#define MPI_ERR_UNKNOWN    13
#define MPI_ERR_TRUNCATE   14
#define MPI_ERR_OTHER      15
#define MPI_ERR_INTERN     16
#define MPI_ERR_LASTCODE   64
#define MPI_MAX_ERROR_STRING   256
#define MPI_MAX_PROCESSOR_NAME 256

#define MPI_COMM_NULL      0
#define MPI_COMM_WORLD     1
#define MPI_COMM_SELF      2
#define MPI_GROUP_NULL     0
#define MPI_GROUP_EMPTY    1
#define MPI_REQUEST_NULL   0
#define MPI_OP_NULL        0
#define MPI_DATATYPE_NULL  0
#define MPI_INFO_NULL      0
#define MPI_WIN_NULL       0
#define MPI_ERRHANDLER_NULL 0
#define MPI_ERRORS_ARE_FATAL 1
#define MPI_ERRORS_RETURN    2

#define MPI_ANY_SOURCE     (-1)
#define MPI_ANY_TAG        (-1)
#define MPI_PROC_NULL      (-2)
#define MPI_ROOT           (-3)
#define MPI_UNDEFINED      (-32766)
#define MPI_KEYVAL_INVALID (-1)
#define MPI_IN_PLACE       ((void *)1)
#define MPI_BOTTOM         ((void *)0)
#define MPI_STATUS_IGNORE   ((MPI_Status *)0)
#define MPI_STATUSES_IGNORE ((MPI_Status *)0)

#define MPI_IDENT     0
#define MPI_CONGRUENT 1
#define MPI_SIMILAR   2
#define MPI_UNEQUAL   3

/* predefined attribute keys (MPI_Attr_get on MPI_COMM_WORLD) */
#define MPI_TAG_UB          0
#define MPI_HOST            1
#define MPI_IO              2
#define MPI_WTIME_IS_GLOBAL 3

#define MPI_THREAD_SINGLE     0
#define MPI_THREAD_FUNNELED   1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE   3

/* datatypes (handles 1..; derived types get handles >= 64) */
#define MPI_CHAR               1
#define MPI_SIGNED_CHAR        2
#define MPI_UNSIGNED_CHAR      3
#define MPI_BYTE               4
#define MPI_SHORT              5
#define MPI_UNSIGNED_SHORT     6
#define MPI_INT                7
#define MPI_UNSIGNED           8
#define MPI_LONG               9
#define MPI_UNSIGNED_LONG      10
#define MPI_LONG_LONG_INT      11
#define MPI_LONG_LONG          MPI_LONG_LONG_INT
#define MPI_UNSIGNED_LONG_LONG 12
#define MPI_FLOAT              13
#define MPI_DOUBLE             14
#define MPI_LONG_DOUBLE        15
#define MPI_PACKED             16
#define MPI_2INT               17
#define MPI_FLOAT_INT          18
#define MPI_DOUBLE_INT         19
#define MPI_LONG_INT           20
#define MPI_SHORT_INT          21
#define MPI_2DOUBLE            22     /* not standard C MPI, harmless */
#define MPI_C_DOUBLE_COMPLEX   23
#define MPI_C_COMPLEX          24
#define MPI_INT64_T            MPI_LONG
#define MPI_INT32_T            MPI_INT

/* reduction operations (user operations get handles >= 32) */
#define MPI_SUM     1
#define MPI_MAX     2
#define MPI_MIN     3
#define MPI_PROD    4
#define MPI_LAND    5
#define MPI_LOR     6
#define MPI_LXOR    7
#define MPI_BAND    8
#define MPI_BOR     9
#define MPI_BXOR    10
#define MPI_MAXLOC  11
#define MPI_MINLOC  12
#define MPI_REPLACE 13

typedef void (MPI_User_function)(void *invec, void *inoutvec, int *len, MPI_Datatype *datatype);
typedef int  (MPI_Copy_function)(MPI_Comm oldcomm, int keyval, void *extra_state, void *attr_in, void *attr_out, int *flag);
typedef int  (MPI_Delete_function)(MPI_Comm comm, int keyval, void *attr, void *extra_state);
typedef void (MPI_Handler_function)(MPI_Comm *, int *, ...);
#define MPI_NULL_COPY_FN   ((MPI_Copy_function *)0)
#define MPI_NULL_DELETE_FN ((MPI_Delete_function *)0)
#define MPI_DUP_FN         ((MPI_Copy_function *)1)

/* ---- environment ---- */
int    MPI_Init(int *argc, char ***argv);
int    MPI_Init_thread(int *argc, char ***argv, int required, int *provided);
int    MPI_Initialized(int *flag);
int    MPI_Finalize(void);
int    MPI_Finalized(int *flag);
int    MPI_Abort(MPI_Comm comm, int errorcode);
double MPI_Wtime(void);
double MPI_Wtick(void);
int    MPI_Get_processor_name(char *name, int *resultlen);
int    MPI_Error_string(int errorcode, char *string, int *resultlen);
int    MPI_Error_class(int errorcode, int *errorclass);
int    MPI_Errhandler_create(MPI_Handler_function *fn, MPI_Errhandler *eh);
int    MPI_Errhandler_set(MPI_Comm comm, MPI_Errhandler eh);
int    MPI_Errhandler_get(MPI_Comm comm, MPI_Errhandler *eh);
int    MPI_Errhandler_free(MPI_Errhandler *eh);
int    MPI_Pcontrol(const int level, ...);

/* ---- communicators, groups, attributes ---- */
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_dup(MPI_Comm comm, MPI_Comm *newcomm);
int MPI_Comm_split(MPI_Comm comm, int color, int key, MPI_Comm *newcomm);
int MPI_Comm_create(MPI_Comm comm, MPI_Group group, MPI_Comm *newcomm);
int MPI_Comm_free(MPI_Comm *comm);
int MPI_Comm_compare(MPI_Comm a, MPI_Comm b, int *result);
int MPI_Comm_group(MPI_Comm comm, MPI_Group *group);
int MPI_Comm_test_inter(MPI_Comm comm, int *flag);
int MPI_Group_size(MPI_Group group, int *size);
int MPI_Group_rank(MPI_Group group, int *rank);
int MPI_Group_incl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup);
int MPI_Group_excl(MPI_Group group, int n, const int *ranks, MPI_Group *newgroup);
int MPI_Group_translate_ranks(MPI_Group g1, int n, const int *ranks1, MPI_Group g2, int *ranks2);
int MPI_Group_free(MPI_Group *group);
int MPI_Keyval_create(MPI_Copy_function *copy_fn, MPI_Delete_function *delete_fn, int *keyval, void *extra_state);
int MPI_Keyval_free(int *keyval);
int MPI_Attr_put(MPI_Comm comm, int keyval, void *attr);
int MPI_Attr_get(MPI_Comm comm, int keyval, void *attr_ptr, int *flag);
int MPI_Attr_delete(MPI_Comm comm, int keyval);
#define MPI_Comm_create_keyval(c, d, k, e) MPI_Keyval_create(c, d, k, e)
#define MPI_Comm_free_keyval(k)            MPI_Keyval_free(k)
#define MPI_Comm_set_attr(c, k, a)         MPI_Attr_put(c, k, a)
#define MPI_Comm_get_attr(c, k, a, f)      MPI_Attr_get(c, k, a, f)
#define MPI_Comm_delete_attr(c, k)         MPI_Attr_delete(c, k)
#define MPI_Comm_f2c(c) ((MPI_Comm)(c))
#define MPI_Comm_c2f(c) ((MPI_Fint)(c))

/* ---- datatypes, operations ---- */
int MPI_Type_size(MPI_Datatype dt, int *size);
int MPI_Type_extent(MPI_Datatype dt, MPI_Aint *extent);
int MPI_Type_get_extent(MPI_Datatype dt, MPI_Aint *lb, MPI_Aint *extent);
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_vector(int count, int blocklength, int stride, MPI_Datatype old, MPI_Datatype *newtype);   /* stride == blocklength only */
int MPI_Type_dup(MPI_Datatype old, MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *dt);
int MPI_Type_free(MPI_Datatype *dt);
int MPI_Op_create(MPI_User_function *fn, int commute, MPI_Op *op);
int MPI_Op_free(MPI_Op *op);
int MPI_Get_count(const MPI_Status *status, MPI_Datatype dt, int *count);
int MPI_Get_elements(const MPI_Status *status, MPI_Datatype dt, int *count);
int MPI_Address(void *location, MPI_Aint *address);

/* ---- point to point ---- */
int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Ssend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Rsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Bsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm);
int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Isend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Issend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irsend(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Irecv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Sendrecv(const void *sbuf, int scount, MPI_Datatype sdt, int dest, int stag,
                 void *rbuf, int rcount, MPI_Datatype rdt, int source, int rtag, MPI_Comm comm, MPI_Status *status);
int MPI_Send_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Ssend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Rsend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Bsend_init(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Recv_init(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm comm, MPI_Request *req);
int MPI_Start(MPI_Request *req);
int MPI_Startall(int count, MPI_Request *reqs);
int MPI_Wait(MPI_Request *req, MPI_Status *status);
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *statuses);
int MPI_Waitany(int count, MPI_Request *reqs, int *index, MPI_Status *status);
int MPI_Waitsome(int incount, MPI_Request *reqs, int *outcount, int *indices, MPI_Status *statuses);
int MPI_Test(MPI_Request *req, int *flag, MPI_Status *status);
int MPI_Testall(int count, MPI_Request *reqs, int *flag, MPI_Status *statuses);
int MPI_Testany(int count, MPI_Request *reqs, int *index, int *flag, MPI_Status *status);
int MPI_Request_free(MPI_Request *req);
int MPI_Cancel(MPI_Request *req);
int MPI_Test_cancelled(const MPI_Status *status, int *flag);
int MPI_Probe(int source, int tag, MPI_Comm comm, MPI_Status *status);
int MPI_Iprobe(int source, int tag, MPI_Comm comm, int *flag, MPI_Status *status);

/* ---- collectives ---- */
int MPI_Barrier(MPI_Comm comm);
int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm comm);
int MPI_Reduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, int root, MPI_Comm comm);
int MPI_Allreduce(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Scan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Exscan(const void *sbuf, void *rbuf, int count, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Reduce_scatter(const void *sbuf, void *rbuf, const int *rcounts, MPI_Datatype dt, MPI_Op op, MPI_Comm comm);
int MPI_Gather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Gatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Scatter(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Scatterv(const void *sbuf, const int *scounts, const int *displs, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, int root, MPI_Comm comm);
int MPI_Allgather(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm);
int MPI_Allgatherv(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, const int *rcounts, const int *displs, MPI_Datatype rdt, MPI_Comm comm);
int MPI_Alltoall(const void *sbuf, int scount, MPI_Datatype sdt, void *rbuf, int rcount, MPI_Datatype rdt, MPI_Comm comm);
int MPI_Alltoallv(const void *sbuf, const int *scounts, const int *sdispls, MPI_Datatype sdt,
                  void *rbuf, const int *rcounts, const int *rdispls, MPI_Datatype rdt, MPI_Comm comm);

/* ---- b200mpi extensions (used by tests / the launcher) ---- */
int b200mpi_world_rank(void);      /* valid before MPI_Init: read from the launcher's environment */
int b200mpi_world_size(void);

#ifdef __cplusplus
}
#endif
#endif /* B200MPI_MPI_H */
