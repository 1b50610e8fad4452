/* mpit_msg -- C ABI of the host message layer (libmpit_host.so, csrc/msg/hostcomm.cpp).
 *
 * A tagged, communicator-scoped, MPI-style message layer implemented from scratch over TCP with a
 * progress thread: the CPU control plane and the CPU-worker data plane of mpit_b200.  It stands
 * where the reference binds an MPI library to Lua (/root/reference/lua-mpi.h, mpifuncs.c; the
 * functions actually used by the reference are listed in SURVEY.md section 2.1).
 *
 * Conventions: every function returns MPIT_SUCCESS (0) or an MPIT_ERR_* code, results go through
 * out-pointers -- the same "return the error code, write outputs into caller storage" shape as the
 * reference wrappers (/root/reference/mpifuncs.c:1-2512).  Buffers are HOST pointers; `count` is in
 * elements of `dtype`.  Communicators, requests are small integer handles.
 */
#ifndef MPIT_MSG_H_
#define MPIT_MSG_H_

#ifdef __cplusplus
extern "C" {
#endif

/* error classes (subset of /root/reference/lua-mpi.h:210-230) */
#define MPIT_SUCCESS 0
#define MPIT_ERR_BUFFER 1
#define MPIT_ERR_COUNT 2
#define MPIT_ERR_TYPE 3
#define MPIT_ERR_TAG 4
#define MPIT_ERR_COMM 5
#define MPIT_ERR_RANK 6
#define MPIT_ERR_REQUEST 7
#define MPIT_ERR_ROOT 8
#define MPIT_ERR_OP 10
#define MPIT_ERR_ARG 13
#define MPIT_ERR_TRUNCATE 15
#define MPIT_ERR_OTHER 16
#define MPIT_ERR_INTERN 17

/* datatypes (/root/reference/lua-mpi.h:130-151) */
enum { MPIT_BYTE = 0, MPIT_CHAR = 1, MPIT_SHORT = 2, MPIT_INT = 3, MPIT_LONG = 4, MPIT_FLOAT = 5,
       MPIT_DOUBLE = 6, MPIT_UNSIGNED_CHAR = 7, MPIT_UNSIGNED = 8, MPIT_LONG_LONG = 9 };
/* reduce ops (/root/reference/lua-mpi.h:179-190) */
enum { MPIT_SUM = 0, MPIT_PROD = 1, MPIT_MAX = 2, MPIT_MIN = 3, MPIT_LAND = 4, MPIT_LOR = 5,
       MPIT_BAND = 6, MPIT_BOR = 7 };

#define MPIT_COMM_WORLD 0
#define MPIT_COMM_SELF 1
#define MPIT_COMM_NULL (-1)
#define MPIT_ANY_SOURCE (-1)
#define MPIT_ANY_TAG (-1)
#define MPIT_REQUEST_NULL (-1)
#define MPIT_UNDEFINED (-32766)

/* MPI_Status: {source, tag, error, count_bytes, cancelled} */
typedef struct { int source; int tag; int error; long long count_bytes; int cancelled; } mpit_status;

/* ---- environment (/root/reference/lua-mpi.h:80-121: Init, Init_MTF/MTS/MTM) ------------------- */
/* Rendezvous through master_addr:master_port (rank 0 listens). thread_level 0..3 mirrors
 * MPI_THREAD_SINGLE..MULTIPLE; this implementation always provides MULTIPLE. */
int mpit_msg_init(int rank, int size, const char* master_addr, int master_port, int thread_level,
                  int* provided);
int mpit_msg_initialized(int* flag);
int mpit_msg_finalize(void);
int mpit_msg_abort(int comm, int errorcode);
double mpit_msg_wtime(void);
double mpit_msg_wtick(void);
int mpit_msg_get_processor_name(char* name64, int* resultlen);
int mpit_msg_type_size(int dtype, int* size);

/* ---- communicators ------------------------------------------------------------------------------ */
int mpit_msg_comm_rank(int comm, int* rank);
int mpit_msg_comm_size(int comm, int* size);
int mpit_msg_comm_dup(int comm, int* newcomm);
int mpit_msg_comm_split(int comm, int color, int key, int* newcomm);
int mpit_msg_comm_free(int* comm);

/* ---- point to point ------------------------------------------------------------------------------ */
int mpit_msg_send(const void* buf, long long count, int dtype, int dest, int tag, int comm);
int mpit_msg_recv(void* buf, long long count, int dtype, int source, int tag, int comm, mpit_status* status);
int mpit_msg_isend(const void* buf, long long count, int dtype, int dest, int tag, int comm, int* request);
int mpit_msg_irecv(void* buf, long long count, int dtype, int source, int tag, int comm, int* request);
int mpit_msg_sendrecv(const void* sbuf, long long scount, int sdtype, int dest, int stag, void* rbuf,
                      long long rcount, int rdtype, int source, int rtag, int comm, mpit_status* status);
int mpit_msg_iprobe(int source, int tag, int comm, int* flag, mpit_status* status);
int mpit_msg_probe(int source, int tag, int comm, mpit_status* status);
int mpit_msg_test(int* request, int* flag, mpit_status* status);
int mpit_msg_wait(int* request, mpit_status* status);
int mpit_msg_waitall(int n, int* requests, mpit_status* statuses);
int mpit_msg_testall(int n, int* requests, int* flag, mpit_status* statuses);
int mpit_msg_waitany(int n, int* requests, int* index, mpit_status* status);
int mpit_msg_cancel(int* request);
int mpit_msg_test_cancelled(const mpit_status* status, int* flag);
int mpit_msg_request_free(int* request);
int mpit_msg_get_count(const mpit_status* status, int dtype, long long* count);

/* ---- collectives ------------------------------------------------------------------------------------ */
int mpit_msg_barrier(int comm);
int mpit_msg_bcast(void* buf, long long count, int dtype, int root, int comm);
int mpit_msg_reduce(const void* sbuf, void* rbuf, long long count, int dtype, int op, int root, int comm);
int mpit_msg_allreduce(const void* sbuf, void* rbuf, long long count, int dtype, int op, int comm);
int mpit_msg_iallreduce(const void* sbuf, void* rbuf, long long count, int dtype, int op, int comm,
                        int* request);
int mpit_msg_gather(const void* sbuf, long long scount, int dtype, void* rbuf, int root, int comm);
int mpit_msg_allgather(const void* sbuf, long long scount, int dtype, void* rbuf, int comm);
int mpit_msg_scatter(const void* sbuf, long long count, int dtype, void* rbuf, int root, int comm);
int mpit_msg_alltoall(const void* sbuf, long long count, int dtype, void* rbuf, int comm);
int mpit_msg_reduce_scatter_block(const void* sbuf, void* rbuf, long long count, int dtype, int op, int comm);
int mpit_msg_scan(const void* sbuf, void* rbuf, long long count, int dtype, int op, int comm);

/* ---- introspection (tests, tracing) ------------------------------------------------------------------- */
/* stats8: {msgs_sent, bytes_sent, msgs_recv, bytes_recv, unexpected_now, posted_now, requests_live, 0} */
int mpit_msg_stats(long long* stats8);

#ifdef __cplusplus
}
#endif
#endif /* MPIT_MSG_H_ */
