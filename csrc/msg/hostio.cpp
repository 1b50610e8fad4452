// Native back end of two MPI chapters that the reference only reaches through its MPI library:
// derived datatypes (/root/reference/mpifuncs.c: Type_vector, Type_indexed, Pack, Unpack ... -- "Datatypes
// 26/26" in SURVEY.md 2.1) and MPI-IO (/root/reference/mpifuncs.c:557-1080, File_* -- "MPI-IO 52/52").
//
// Both chapters reduce to the same primitive: move bytes between a contiguous stream and a strided
// layout described by a list of (displacement, length) blocks that tiles with a fixed extent.
//   * datatype engine: the strided side is user memory (mpit_dt_pack / mpit_dt_unpack);
//   * file views: the strided side is the file (mpit_io_view_rw walks the tiling with pread/pwrite).
// The Python front end (mpit_b200/msg/datatypes.py, fileio.py) flattens type maps to block lists
// once per committed type and calls in here for the data movement.
//
// The shared file pointer of MPI-IO lives in a small side-car file next to the data file and is
// advanced under an fcntl record lock, so it is shared by every process that opens the file, not only
// by threads of one process.
#include "mpit_msg.h"

#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include <algorithm>

namespace {

long long full_pread(int fd, char* buf, long long n, long long off) {
  long long got = 0;
  while (got < n) {
    ssize_t r = pread(fd, buf + got, (size_t)(n - got), (off_t)(off + got));
    if (r < 0) {
      if (errno == EINTR) continue;
      return -(long long)errno;
    }
    if (r == 0) break;   // end of file: a short read is a valid MPI-IO outcome
    got += r;
  }
  return got;
}

long long full_pwrite(int fd, const char* buf, long long n, long long off) {
  long long put = 0;
  while (put < n) {
    ssize_t r = pwrite(fd, buf + put, (size_t)(n - put), (off_t)(off + put));
    if (r < 0) {
      if (errno == EINTR) continue;
      return -(long long)errno;
    }
    put += r;
  }
  return put;
}

int lock_whole(int fd, short type) {
  struct flock fl;
  memset(&fl, 0, sizeof fl);
  fl.l_type = type;
  fl.l_whence = SEEK_SET;
  fl.l_start = 0;
  fl.l_len = 0;
  for (;;) {
    if (fcntl(fd, F_SETLKW, &fl) == 0) return 0;
    if (errno != EINTR) return -errno;
  }
}

}  // namespace

extern "C" {

// ---- datatype engine ---------------------------------------------------------------------------------
// `count` replicas of a type whose one instance is `nblk` blocks {disp[i], len[i]} (bytes, relative to
// `base`, in type-map order) and whose replicas are `extent` bytes apart.  `out` / `in` is the packed
// stream.  `limit` bounds the number of stream bytes moved (a partial trailing element is legal when
// unpacking a short message); the number of bytes moved is returned.
long long mpit_dt_pack(const void* base, long long count, long long extent, int nblk, const long long* disp,
                       const long long* len, void* out, long long limit) {
  const char* b = (const char*)base;
  char* o = (char*)out;
  long long done = 0;
  for (long long c = 0; c < count && done < limit; ++c) {
    const char* inst = b + c * extent;
    for (int i = 0; i < nblk && done < limit; ++i) {
      long long n = std::min(len[i], limit - done);
      memcpy(o + done, inst + disp[i], (size_t)n);
      done += n;
    }
  }
  return done;
}

long long mpit_dt_unpack(void* base, long long count, long long extent, int nblk, const long long* disp,
                         const long long* len, const void* in, long long limit) {
  char* b = (char*)base;
  const char* s = (const char*)in;
  long long done = 0;
  for (long long c = 0; c < count && done < limit; ++c) {
    char* inst = b + c * extent;
    for (int i = 0; i < nblk && done < limit; ++i) {
      long long n = std::min(len[i], limit - done);
      memcpy(inst + disp[i], s + done, (size_t)n);
      done += n;
    }
  }
  return done;
}

// ---- files --------------------------------------------------------------------------------------------
// amode bits follow MPI: 1 RDONLY, 2 RDWR, 4 WRONLY, 8 CREATE, 16 EXCL, 32 DELETE_ON_CLOSE (handled by the
// caller), 64 UNIQUE_OPEN, 128 APPEND, 256 SEQUENTIAL.
int mpit_io_open(const char* path, int amode, int* fd) {
  int flags = 0;
  if (amode & 2) flags |= O_RDWR;
  else if (amode & 4) flags |= O_WRONLY;
  else flags |= O_RDONLY;
  if (amode & 8) flags |= O_CREAT;
  if (amode & 16) flags |= O_EXCL;
  int f = open(path, flags | O_CLOEXEC, 0644);
  if (f < 0) return errno == ENOENT ? MPIT_ERR_NO_SUCH_FILE : errno == EEXIST ? MPIT_ERR_FILE_EXISTS
             : errno == EACCES ? MPIT_ERR_ACCESS : MPIT_ERR_IO;
  *fd = f;
  return MPIT_SUCCESS;
}

int mpit_io_close(int fd) { return close(fd) == 0 ? MPIT_SUCCESS : MPIT_ERR_IO; }

int mpit_io_delete(const char* path) {
  if (unlink(path) == 0) return MPIT_SUCCESS;
  return errno == ENOENT ? MPIT_ERR_NO_SUCH_FILE : MPIT_ERR_IO;
}

int mpit_io_get_size(int fd, long long* size) {
  struct stat st;
  if (fstat(fd, &st) != 0) return MPIT_ERR_IO;
  *size = (long long)st.st_size;
  return MPIT_SUCCESS;
}

int mpit_io_set_size(int fd, long long size) { return ftruncate(fd, (off_t)size) == 0 ? MPIT_SUCCESS : MPIT_ERR_IO; }

// Grow (never shrink) the file and make sure the blocks exist.
int mpit_io_preallocate(int fd, long long size) {
  long long cur = 0;
  int rc = mpit_io_get_size(fd, &cur);
  if (rc) return rc;
  if (cur >= size) return MPIT_SUCCESS;
  int e = posix_fallocate(fd, 0, (off_t)size);
  if (e == 0) return MPIT_SUCCESS;
  return ftruncate(fd, (off_t)size) == 0 ? MPIT_SUCCESS : MPIT_ERR_NO_SPACE;
}

int mpit_io_sync(int fd) { return fsync(fd) == 0 ? MPIT_SUCCESS : MPIT_ERR_IO; }

// Whole-file advisory record lock (atomic mode): mode 0 unlock, 1 shared, 2 exclusive.
int mpit_io_lock(int fd, int mode) {
  return lock_whole(fd, mode == 0 ? F_UNLCK : mode == 1 ? F_RDLCK : F_WRLCK) == 0 ? MPIT_SUCCESS : MPIT_ERR_IO;
}

// Contiguous transfer at an absolute byte offset; *moved = bytes transferred (short at end of file).
int mpit_io_rw(int fd, int write, void* buf, long long nbytes, long long offset, long long* moved) {
  long long r = write ? full_pwrite(fd, (const char*)buf, nbytes, offset) : full_pread(fd, (char*)buf, nbytes, offset);
  if (r < 0) { *moved = 0; return r == -ENOSPC ? MPIT_ERR_NO_SPACE : MPIT_ERR_IO; }
  *moved = r;
  return MPIT_SUCCESS;
}

// Transfer through a file view.  The view starts at absolute byte `view_disp` and is a tiling, every
// `fextent` bytes, of `nblk` visible blocks {fdisp[i], flen[i]} (sorted, non-overlapping, `fsize` visible
// bytes per tile).  `pos` is the position in the stream of VISIBLE bytes at which to start; `nbytes` are
// moved between that stream and the contiguous `buf`.
int mpit_io_view_rw(int fd, int write, void* buf, long long nbytes, long long view_disp, long long fextent,
                    int nblk, const long long* fdisp, const long long* flen, long long fsize, long long pos,
                    long long* moved) {
  *moved = 0;
  if (nbytes == 0) return MPIT_SUCCESS;
  if (fsize <= 0 || nblk <= 0) return MPIT_ERR_ARG;
  char* b = (char*)buf;
  long long tile = pos / fsize, within = pos % fsize;
  int i = 0;
  while (i < nblk && within >= flen[i]) { within -= flen[i]; ++i; }
  long long done = 0;
  while (done < nbytes) {
    if (i == nblk) { i = 0; ++tile; within = 0; }
    long long n = std::min(flen[i] - within, nbytes - done);
    long long off = view_disp + tile * fextent + fdisp[i] + within;
    long long r = write ? full_pwrite(fd, b + done, n, off) : full_pread(fd, b + done, n, off);
    if (r < 0) return r == -ENOSPC ? MPIT_ERR_NO_SPACE : MPIT_ERR_IO;
    done += r;
    if (r < n) break;   // end of file while reading
    within = 0;
    ++i;
  }
  *moved = done;
  return MPIT_SUCCESS;
}

// Absolute byte offset of visible-stream position `pos` under the same view description.
int mpit_io_view_byte_offset(long long view_disp, long long fextent, int nblk, const long long* fdisp,
                             const long long* flen, long long fsize, long long pos, long long* byte_off) {
  if (fsize <= 0 || nblk <= 0) return MPIT_ERR_ARG;
  long long tile = pos / fsize, within = pos % fsize;
  int i = 0;
  while (i < nblk && within >= flen[i]) { within -= flen[i]; ++i; }
  if (i == nblk) { i = 0; ++tile; within = 0; }
  *byte_off = view_disp + tile * fextent + fdisp[i] + within;
  return MPIT_SUCCESS;
}

// Shared file pointer: an 8-byte counter in a side-car file.  op 0: fetch-and-add `value`, op 1: set to
// `value`, op 2: read.  *old receives the value before the operation.
int mpit_io_shared_pointer(int sidecar_fd, int op, long long value, long long* old) {
  if (lock_whole(sidecar_fd, F_WRLCK) != 0) return MPIT_ERR_IO;
  long long cur = 0;
  long long r = full_pread(sidecar_fd, (char*)&cur, 8, 0);
  int rc = MPIT_SUCCESS;
  if (r < 0) rc = MPIT_ERR_IO;
  else {
    if (r < 8) cur = 0;
    *old = cur;
    long long next = op == 0 ? cur + value : op == 1 ? value : cur;
    if (op != 2 && full_pwrite(sidecar_fd, (const char*)&next, 8, 0) != 8) rc = MPIT_ERR_IO;
  }
  lock_whole(sidecar_fd, F_UNLCK);
  return rc;
}

}  // extern "C"
