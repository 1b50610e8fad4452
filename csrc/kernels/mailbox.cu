// Device transport of the message layer: tagged two-sided messages between GPUs (sm_100a + host helpers).
//
// Reference: any Torch Storage, a CudaStorage included, goes straight to the transport
// (/root/reference/lua-mpi.h:70-78: MPI_Send/Recv/Isend/Irecv on the raw device pointer), and the coroutine IO
// layer polls MPI_Iprobe / MPI_Test thousands of times per message (/root/reference/init.lua:41-108).
//
// Split of labour here:
//   * payloads never leave the GPUs: small ones are stored by the sender's kernel into a slot of the receiver's
//     symmetric memory over NVLink, large ones stay in the sender's tensor and the receiver's kernel pulls them
//     through a CUDA-IPC mapping of that allocation (rendezvous, zero copies);
//   * descriptors {seq, tag, bytes, kind, state, done} live in a POSIX shared-memory segment that every process
//     maps and registers with CUDA: the sender's GPU publishes `seq` with st.release.sys in stream order (after
//     the payload stores), while probing, matching (ANY_SOURCE / ANY_TAG, non-overtaking per pair), cancelling
//     (compare-and-swap on `state`) and acknowledging are plain CPU atomics on the same words -- an Iprobe or a
//     Test is a memory read, not a kernel launch and not a device synchronisation.
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cuda.h>

#include "mpit_b200.h"
#include "comm.cuh"

namespace mpit {

// One descriptor = 128 bytes.  Ring layout in the segment: [dst][src][K].
struct MbxDesc {
  uint32_t seq;        // 1-based message number of the (src, dst) pair; written LAST, by the sender's GPU
  int32_t tag;
  uint64_t nbytes;
  uint64_t rv_offset;  // rendezvous: offset of the payload inside the exported allocation
  uint32_t kind;       // 0 eager (payload in the receiver's slot), 1 rendezvous (payload in the sender's tensor)
  uint32_t state;      // 1 posted, 2 claimed by a receive, 3 cancelled by the sender (CPU compare-and-swap)
  uint32_t done;       // receiver: = seq once the payload has been copied out (slot / sender buffer reusable)
  uint32_t pad[7];
  unsigned char ipc[64];   // rendezvous: cudaIpcMemHandle_t of the sender's allocation
};
static_assert(sizeof(MbxDesc) == 128, "descriptor is 128 bytes");

// Byte copy with 16-byte accesses where both sides allow it; `src` or `dst` may be peer memory.
__global__ void __launch_bounds__(256) mbx_copy_kernel(unsigned char* dst, const unsigned char* src, long long n) {
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec = (((uintptr_t)dst | (uintptr_t)src) & 15) == 0;
  const long long n16 = vec ? n >> 4 : 0;
  for (long long q = t0; q < n16; q += stride) {
    float4 v = ld_sys4(reinterpret_cast<const float*>(src) + 4 * q);
    st_sys4(reinterpret_cast<float*>(dst) + 4 * q, v);
  }
  for (long long i = (n16 << 4) + t0; i < n; i += stride) dst[i] = src[i];
}

// Stream-ordered publication of a descriptor: everything this GPU stored before (the eager payload in the
// peer's slot, or the producer kernels of a rendezvous payload) is visible system-wide before `seq` appears.
__global__ void mbx_publish_kernel(uint32_t* seq_word, uint32_t seq) {
  __threadfence_system();
  st_release_sys_u32(seq_word, seq);
}

}  // namespace mpit

using namespace mpit;

extern "C" {

int mpit_mbx_desc_bytes(void) { return (int)sizeof(MbxDesc); }

/* Maps (and with create != 0 first creates and zeroes) the POSIX shared-memory segment `name` of `bytes` bytes and
 * registers it with CUDA; *host_ptr is the CPU address, *dev_ptr the address kernels of the current device use. */
int mpit_shm_map(const char* name, long long bytes, int create, void** host_ptr, void** dev_ptr) {
  if (!name || bytes <= 0 || !host_ptr || !dev_ptr) return MPIT_E_ARG;
  int fd = shm_open(name, create ? (O_CREAT | O_RDWR) : O_RDWR, 0600);
  if (fd < 0) return MPIT_E_UNSUPPORTED;
  if (create && ftruncate(fd, bytes) != 0) { close(fd); return MPIT_E_UNSUPPORTED; }
  void* p = mmap(nullptr, (size_t)bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) return MPIT_E_UNSUPPORTED;
  if (create) memset(p, 0, (size_t)bytes);
  cudaError_t e = cudaHostRegister(p, (size_t)bytes, cudaHostRegisterMapped | cudaHostRegisterPortable);
  if (e != cudaSuccess) { munmap(p, (size_t)bytes); return (int)e; }
  void* d = nullptr;
  e = cudaHostGetDevicePointer(&d, p, 0);
  if (e != cudaSuccess) { cudaHostUnregister(p); munmap(p, (size_t)bytes); return (int)e; }
  *host_ptr = p;
  *dev_ptr = d;
  return 0;
}

int mpit_shm_unmap(const char* name, void* host_ptr, long long bytes, int unlink_it) {
  if (host_ptr) {
    cudaHostUnregister(host_ptr);
    munmap(host_ptr, (size_t)bytes);
  }
  if (unlink_it && name) shm_unlink(name);
  return 0;
}

int mpit_mbx_copy(void* dst, const void* src, long long nbytes, void* stream) {
  if (nbytes <= 0) return 0;
  if (!dst || !src) return MPIT_E_ARG;
  int grid = ew_grid(nbytes >> 4, 256, 4, 4);
  mbx_copy_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(static_cast<unsigned char*>(dst),
                                                          static_cast<const unsigned char*>(src), nbytes);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_mbx_publish(void* desc_dev, unsigned seq, void* stream) {
  if (!desc_dev) return MPIT_E_ARG;
  mbx_publish_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(&static_cast<MbxDesc*>(desc_dev)->seq, seq);
  MPIT_LAUNCH_CHECK();
  return 0;
}

/* Sender, CPU: fill a free descriptor (everything but `seq`, which the GPU publishes). */
int mpit_mbx_fill(void* desc_host, int tag, long long nbytes, int kind, unsigned long long rv_offset,
                  const void* ipc64) {
  if (!desc_host) return MPIT_E_ARG;
  MbxDesc* d = static_cast<MbxDesc*>(desc_host);
  d->tag = tag;
  d->nbytes = (uint64_t)nbytes;
  d->rv_offset = rv_offset;
  d->kind = (uint32_t)kind;
  if (ipc64) memcpy(d->ipc, ipc64, 64);
  __atomic_store_n(&d->state, 1u, __ATOMIC_RELEASE);
  return 0;
}

/* Zero-byte messages need no GPU at all: the CPU publishes the descriptor itself. */
int mpit_mbx_publish_host(void* desc_host, unsigned seq) {
  if (!desc_host) return MPIT_E_ARG;
  __atomic_store_n(&static_cast<MbxDesc*>(desc_host)->seq, seq, __ATOMIC_RELEASE);
  return 0;
}

/* Descriptor words, CPU side.  which: 0 seq, 1 state, 2 done. */
int mpit_mbx_word(const void* desc_host, int which) {
  const MbxDesc* d = static_cast<const MbxDesc*>(desc_host);
  const uint32_t* w = which == 0 ? &d->seq : (which == 1 ? &d->state : &d->done);
  return (int)__atomic_load_n(w, __ATOMIC_ACQUIRE);
}
int mpit_mbx_set_done(void* desc_host, unsigned seq) {
  __atomic_store_n(&static_cast<MbxDesc*>(desc_host)->done, seq, __ATOMIC_RELEASE);
  return 0;
}
/* Compare-and-swap on `state`; returns the value found (== expect on success).  A receive claims a message with
 * 1 -> 2, MPI_Cancel of a send withdraws it with 1 -> 3: exactly one of the two wins. */
int mpit_mbx_cas_state(void* desc_host, unsigned expect, unsigned val) {
  uint32_t e = expect;
  __atomic_compare_exchange_n(&static_cast<MbxDesc*>(desc_host)->state, &e, val, false, __ATOMIC_ACQ_REL,
                              __ATOMIC_ACQUIRE);
  return (int)e;
}

/* Receiver, CPU: find the oldest matching message.  `ring` = my [nsrc][K] descriptors; base[s] = lowest message
 * number of source s not yet consumed; mask[s] bit j set = message base[s] + j already consumed (out-of-order
 * matches by tag).  Per source the scan follows message order and stops at the first message that has not arrived,
 * which keeps MPI's non-overtaking rule.  want_src < 0 = ANY_SOURCE (scan starts at start_src), want_tag < 0 =
 * ANY_TAG.  Returns 1 and fills out5 = {src, seq, tag, kind, state} / *out_bytes, or 0. */
int mpit_mbx_scan(const void* ring, int nsrc, int K, const unsigned* base, const unsigned long long* mask,
                  int want_src, int want_tag, int start_src, int* out5, long long* out_bytes) {
  const MbxDesc* r = static_cast<const MbxDesc*>(ring);
  for (int i = 0; i < nsrc; ++i) {
    int s = want_src >= 0 ? want_src : (start_src + i) % nsrc;
    if (want_src >= 0 && i > 0) break;
    for (int j = 0; j < K; ++j) {
      const uint32_t seq = base[s] + (uint32_t)j;
      const MbxDesc* d = r + (size_t)s * K + seq % (uint32_t)K;
      if (__atomic_load_n(&d->seq, __ATOMIC_ACQUIRE) != seq) break;      // not arrived yet: nothing behind it counts
      if ((mask[s] >> j) & 1ull) continue;                                // consumed earlier (other tag)
      const uint32_t st = __atomic_load_n(&d->state, __ATOMIC_ACQUIRE);
      if (st == 3u || (st == 1u && (want_tag < 0 || d->tag == want_tag))) {
        out5[0] = s; out5[1] = (int)seq; out5[2] = d->tag; out5[3] = (int)d->kind; out5[4] = (int)st;
        *out_bytes = (long long)d->nbytes;
        return 1;                                                         // cancelled ones are reported for clean-up
      }
    }
  }
  return 0;
}

int mpit_mbx_desc_info(const void* desc_host, unsigned long long* rv_offset, void* ipc64) {
  const MbxDesc* d = static_cast<const MbxDesc*>(desc_host);
  if (rv_offset) *rv_offset = d->rv_offset;
  if (ipc64) memcpy(ipc64, d->ipc, 64);
  return 0;
}

/* CUDA-IPC handle of the allocation that contains `ptr` and the offset of `ptr` inside it (rendezvous sends). */
int mpit_ipc_export_range(unsigned long long ptr, void* handle64, unsigned long long* offset) {
  typedef CUresult (*GetRange)(CUdeviceptr*, size_t*, CUdeviceptr);
  static GetRange get_range = nullptr;
  if (!get_range) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &p, cudaEnableDefault, &q) != cudaSuccess || !p)
      return MPIT_E_UNSUPPORTED;
    get_range = reinterpret_cast<GetRange>(p);
  }
  CUdeviceptr base = 0;
  size_t size = 0;
  CUresult r = get_range(&base, &size, (CUdeviceptr)ptr);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void*)base);
  if (e != cudaSuccess) { cudaGetLastError(); return (int)e; }
  memcpy(handle64, &h, 64);
  *offset = (unsigned long long)(ptr - base);
  return 0;
}

}  // extern "C"
