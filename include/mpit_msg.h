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
/* classes of the MPI-IO chapter (csrc/msg/hostio.cpp) */
#define MPIT_ERR_ACCESS 20
#define MPIT_ERR_AMODE 21
#define MPIT_ERR_FILE_EXISTS 25
#define MPIT_ERR_FILE 27
#define MPIT_ERR_IO 32
#define MPIT_ERR_NO_SPACE 36
#define MPIT_ERR_NO_SUCH_FILE 37
#define MPIT_ERR_READ_ONLY 40

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
/* Non-collective: a communicator over an explicit world-rank list; all members pass the same list and
 * seed.  Used for inter-communicator bridges (MPI_Intercomm_create, MPI_Comm_accept / connect). */
int mpit_msg_comm_from_ranks(const int* world_ranks, int n, int seed, int* newcomm);
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
/* MPI_Request_get_status: completion flag and status WITHOUT releasing the request
 * (/root/reference/mpifuncs.c Request_get_status). */
int mpit_msg_request_get_status(int request, int* flag, mpit_status* status);
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
/* stats8: {msgs_sent, bytes_sent, msgs_recv, bytes_recv, unexpected_now, posted_now, requests_live,
 *          rendezvous_peers (peers that may pull this process's large Isends with process_vm_readv)} */
int mpit_msg_stats(long long* stats8);

/* ---- derived-datatype engine (csrc/msg/hostio.cpp; /root/reference/mpifuncs.c Pack, Unpack, Type_*) --------
 * A committed type is `nblk` byte blocks {disp[i], len[i]} in type-map order, replicated every `extent`
 * bytes.  pack gathers `count` replicas from `base` into the contiguous stream `out`; unpack scatters.
 * `limit` bounds the stream bytes moved; the number moved is returned. */
long long mpit_dt_pack(const void* base, long long count, long long extent, int nblk, const long long* disp,
                       const long long* len, void* out, long long limit);
long long mpit_dt_unpack(void* base, long long count, long long extent, int nblk, const long long* disp,
                         const long long* len, const void* in, long long limit);

/* ---- file IO under MPI-IO semantics (csrc/msg/hostio.cpp; /root/reference/mpifuncs.c:557-1080) ------------
 * amode bits as in MPI (1 RDONLY, 2 RDWR, 4 WRONLY, 8 CREATE, 16 EXCL, 128 APPEND).  Offsets are bytes. */
int mpit_io_open(const char* path, int amode, int* fd);
int mpit_io_close(int fd);
int mpit_io_delete(const char* path);
int mpit_io_get_size(int fd, long long* size);
int mpit_io_set_size(int fd, long long size);
int mpit_io_preallocate(int fd, long long size);
int mpit_io_sync(int fd);
int mpit_io_lock(int fd, int mode);   /* 0 unlock, 1 shared, 2 exclusive (atomic mode) */
int mpit_io_rw(int fd, int write, void* buf, long long nbytes, long long offset, long long* moved);
/* transfer through a view: tiles of `fextent` bytes starting at `view_disp`, each exposing blocks
 * {fdisp[i], flen[i]} (`fsize` visible bytes per tile); `pos` counts visible bytes from the view start */
int mpit_io_view_rw(int fd, int write, void* buf, long long nbytes, long long view_disp, long long fextent,
                    int nblk, const long long* fdisp, const long long* flen, long long fsize, long long pos,
                    long long* moved);
int mpit_io_view_byte_offset(long long view_disp, long long fextent, int nblk, const long long* fdisp,
                             const long long* flen, long long fsize, long long pos, long long* byte_off);
/* shared file pointer kept in a side-car file: op 0 fetch-and-add, 1 set, 2 read */
int mpit_io_shared_pointer(int sidecar_fd, int op, long long value, long long* old);

/* ---- background mini-batch assembler (csrc/io/loader.cpp) ---------------------------------------------------
 * Stands in for the per-sample batch assembly of /root/reference/asyncsgd/goot.lua:164,172-184 (per-epoch
 * permutation, random crop position per sample), moved to a worker thread.  `images` = n_samples records of
 * `sample_bytes`, `labels` = n_samples int32; the ring buffers (caller-owned, ideally pinned) hold `nslots`
 * batches: ring_images [nslots][batch][sample_bytes], ring_labels [nslots][batch], ring_crop [nslots][batch][2]
 * (y, x in [0, crop_range)).  One consumer thread: next() blocks for a filled slot, release() recycles it. */
int mpit_loader_create(const unsigned char* images, const int* labels, long long n_samples, long long sample_bytes,
                       int batch, int nslots, unsigned char* ring_images, int* ring_labels, int* ring_crop,
                       int crop_range, unsigned long long seed, int shuffle, void** handle);
int mpit_loader_next(void* handle, int* slot, long long* epoch);
int mpit_loader_release(void* handle, int slot);
int mpit_loader_destroy(void* handle);

#ifdef __cplusplus
}
#endif
#endif /* MPIT_MSG_H_ */
