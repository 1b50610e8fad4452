// Symmetric heap: same-size device buffers on every rank, each mapped into every other rank's
// address space, optionally with an NVLS multicast mapping over all of them.
//
// This is the data plane that replaces the reference's "hand the device pointer to a CUDA-aware
// MPI" bridge (/root/reference/lua-mpi.h:70-78).  Two sharing mechanisms:
//   * legacy CUDA IPC on cudaMalloc memory (cudaIpcGetMemHandle / cudaIpcOpenMemHandle), and
//   * the VMM API (cuMemCreate + POSIX file-descriptor export) which is what multicast needs;
//     descriptors travel between ranks with pidfd_getfd(2), so no SCM_RIGHTS socket dance.
// The driver API is reached through cudaGetDriverEntryPoint: the library never links libcuda
// and therefore still loads on CPU-only boxes.
#include "mpit_b200.h"
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <unistd.h>
#include <sys/syscall.h>
#include <mutex>
#include <unordered_map>

#ifndef SYS_pidfd_open
#define SYS_pidfd_open 434
#endif
#ifndef SYS_pidfd_getfd
#define SYS_pidfd_getfd 438
#endif

namespace {

struct Driver {
  bool ok = false;
  decltype(&cuMemCreate) MemCreate = nullptr;
  decltype(&cuMemRelease) MemRelease = nullptr;
  decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
  decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
  decltype(&cuMemAddressReserve) AddrReserve = nullptr;
  decltype(&cuMemAddressFree) AddrFree = nullptr;
  decltype(&cuMemMap) MemMap = nullptr;
  decltype(&cuMemUnmap) MemUnmap = nullptr;
  decltype(&cuMemSetAccess) SetAccess = nullptr;
  decltype(&cuMemGetAllocationGranularity) Granularity = nullptr;
  decltype(&cuMulticastCreate) McCreate = nullptr;
  decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
  decltype(&cuMulticastBindMem) McBindMem = nullptr;
  decltype(&cuMulticastGetGranularity) McGranularity = nullptr;
  decltype(&cuDeviceGetAttribute) DevAttr = nullptr;
  decltype(&cuDeviceGet) DevGet = nullptr;
};

template <typename T>
static bool entry(const char* name, T& fn) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || !p) return false;
  fn = reinterpret_cast<T>(p);
  return true;
}

static Driver& drv() {
  static Driver d;
  static std::once_flag once;
  std::call_once(once, [] {
    bool ok = true;
    ok &= entry("cuMemCreate", d.MemCreate);
    ok &= entry("cuMemRelease", d.MemRelease);
    ok &= entry("cuMemExportToShareableHandle", d.MemExport);
    ok &= entry("cuMemImportFromShareableHandle", d.MemImport);
    ok &= entry("cuMemAddressReserve", d.AddrReserve);
    ok &= entry("cuMemAddressFree", d.AddrFree);
    ok &= entry("cuMemMap", d.MemMap);
    ok &= entry("cuMemUnmap", d.MemUnmap);
    ok &= entry("cuMemSetAccess", d.SetAccess);
    ok &= entry("cuMemGetAllocationGranularity", d.Granularity);
    ok &= entry("cuDeviceGetAttribute", d.DevAttr);
    ok &= entry("cuDeviceGet", d.DevGet);
    // multicast is optional
    entry("cuMulticastCreate", d.McCreate);
    entry("cuMulticastAddDevice", d.McAddDevice);
    entry("cuMulticastBindMem", d.McBindMem);
    entry("cuMulticastGetGranularity", d.McGranularity);
    d.ok = ok;
  });
  return d;
}

struct VmmRec {
  CUmemGenericAllocationHandle handle;
  size_t size;
  int fd;       // exported descriptor (-1 for imported mappings)
  bool is_mc;
};
static std::mutex g_mu;
static std::unordered_map<unsigned long long, VmmRec> g_vmm;

static int map_handle(CUmemGenericAllocationHandle h, size_t size, int dev, unsigned long long* out_ptr) {
  Driver& d = drv();
  CUdeviceptr va = 0;
  CUresult r = d.AddrReserve(&va, size, 0, 0, 0);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  r = d.MemMap(va, size, 0, h, 0);
  if (r != CUDA_SUCCESS) { d.AddrFree(va, size); return 1000 + (int)r; }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = d.SetAccess(va, size, &acc, 1);
  if (r != CUDA_SUCCESS) { d.MemUnmap(va, size); d.AddrFree(va, size); return 1000 + (int)r; }
  *out_ptr = (unsigned long long)va;
  return 0;
}

}  // namespace

extern "C" {

// ---- legacy IPC path --------------------------------------------------------------------
int mpit_symm_malloc(unsigned long long bytes, unsigned long long* out_ptr) {
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  e = cudaMemset(p, 0, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  *out_ptr = (unsigned long long)p;
  return 0;
}

int mpit_symm_free(unsigned long long ptr) { return (int)cudaFree((void*)ptr); }

int mpit_symm_ipc_export(unsigned long long ptr, void* handle64) {
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, (void*)ptr);
  if (e != cudaSuccess) return (int)e;
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return 0;
}

int mpit_symm_ipc_import(const void* handle64, unsigned long long* out_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return (int)e;
  *out_ptr = (unsigned long long)p;
  return 0;
}

int mpit_symm_ipc_close(unsigned long long ptr) { return (int)cudaIpcCloseMemHandle((void*)ptr); }

// Single-process multi-GPU (tests, one driver process steering several GPUs).
int mpit_enable_peer_access(int dev, int peer) {
  int can = 0;
  cudaError_t e = cudaDeviceCanAccessPeer(&can, dev, peer);
  if (e != cudaSuccess) return (int)e;
  if (!can) return MPIT_E_UNSUPPORTED;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(dev);
  e = cudaDeviceEnablePeerAccess(peer, 0);
  cudaSetDevice(cur);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  return (int)e;
}

// ---- VMM path ------------------------------------------------------------------------------
// Returns the allocation granularity to which sizes must be rounded (multicast-aware when
// with_multicast != 0), or a negative error.
long long mpit_vmm_granularity(int dev, int nranks, int with_multicast) {
  Driver& d = drv();
  if (!d.ok) return MPIT_E_UNSUPPORTED;
  cudaFree(0);  // make sure a context exists
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t g = 0;
  if (d.Granularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS) return MPIT_E_UNSUPPORTED;
  if (with_multicast && d.McGranularity) {
    CUmulticastObjectProp mp{};
    mp.numDevices = nranks;
    mp.size = g;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (d.McGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  return (long long)g;
}

int mpit_vmm_multicast_supported(int dev) {
  Driver& d = drv();
  if (!d.ok || !d.McCreate) return 0;
  cudaFree(0);
  CUdevice cd;
  if (d.DevGet(&cd, dev) != CUDA_SUCCESS) return 0;
  int v = 0;
  if (d.DevAttr(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cd) != CUDA_SUCCESS) return 0;
  return v;
}

// Allocate `bytes` (already rounded to the granularity) of shareable device memory, map it
// locally, zero it, and export a POSIX fd for peers.
int mpit_vmm_alloc(int dev, unsigned long long bytes, unsigned long long* out_ptr, int* out_fd) {
  Driver& d = drv();
  if (!d.ok) return MPIT_E_UNSUPPORTED;
  cudaFree(0);
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemCreate(&h, (size_t)bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  int rc = map_handle(h, (size_t)bytes, dev, out_ptr);
  if (rc) { d.MemRelease(h); return rc; }
  int fd = -1;
  r = d.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  cudaError_t e = cudaMemset((void*)*out_ptr, 0, (size_t)bytes);
  if (e != cudaSuccess) return (int)e;
  *out_fd = fd;
  std::lock_guard<std::mutex> lk(g_mu);
  g_vmm[*out_ptr] = VmmRec{h, (size_t)bytes, fd, false};
  return 0;
}

// Duplicate descriptor `remote_fd` of process `pid` into this process (pidfd_getfd).
int mpit_fd_from_pid(int pid, int remote_fd) {
  if (pid == (int)getpid()) return dup(remote_fd);
  int pidfd = (int)syscall(SYS_pidfd_open, pid, 0);
  if (pidfd < 0) return -1;
  int fd = (int)syscall(SYS_pidfd_getfd, pidfd, remote_fd, 0);
  close(pidfd);
  return fd;
}

// Map a peer's allocation (given a local duplicate of its fd) on device `dev`.
int mpit_vmm_import(int dev, int fd, unsigned long long bytes, unsigned long long* out_ptr) {
  Driver& d = drv();
  if (!d.ok) return MPIT_E_UNSUPPORTED;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  int rc = map_handle(h, (size_t)bytes, dev, out_ptr);
  if (rc) { d.MemRelease(h); return rc; }
  std::lock_guard<std::mutex> lk(g_mu);
  g_vmm[*out_ptr] = VmmRec{h, (size_t)bytes, -1, false};
  return 0;
}

// Multicast object life cycle: create (rank 0) -> fd -> import (others) -> add_device (all) ->
// [barrier] -> bind (all, own local allocation) -> [barrier] -> map (all).
int mpit_mc_create(int nranks, unsigned long long bytes, unsigned long long* out_handle, int* out_fd) {
  Driver& d = drv();
  if (!d.ok || !d.McCreate) return MPIT_E_UNSUPPORTED;
  CUmulticastObjectProp mp{};
  mp.numDevices = nranks;
  mp.size = (size_t)bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUmemGenericAllocationHandle h;
  CUresult r = d.McCreate(&h, &mp);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  int fd = -1;
  r = d.MemExport(&fd, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  *out_handle = (unsigned long long)h;
  *out_fd = fd;
  return 0;
}

int mpit_mc_import(int fd, unsigned long long* out_handle) {
  Driver& d = drv();
  if (!d.ok || !d.McCreate) return MPIT_E_UNSUPPORTED;
  CUmemGenericAllocationHandle h;
  CUresult r = d.MemImport(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  *out_handle = (unsigned long long)h;
  return 0;
}

int mpit_mc_add_device(unsigned long long mc_handle, int dev) {
  Driver& d = drv();
  CUdevice cd;
  if (d.DevGet(&cd, dev) != CUDA_SUCCESS) return MPIT_E_ARG;
  CUresult r = d.McAddDevice((CUmemGenericAllocationHandle)mc_handle, cd);
  return r == CUDA_SUCCESS ? 0 : 1000 + (int)r;
}

int mpit_mc_bind(unsigned long long mc_handle, unsigned long long local_ptr, unsigned long long bytes) {
  Driver& d = drv();
  CUmemGenericAllocationHandle mem;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_vmm.find(local_ptr);
    if (it == g_vmm.end()) return MPIT_E_ARG;
    mem = it->second.handle;
  }
  CUresult r = d.McBindMem((CUmemGenericAllocationHandle)mc_handle, 0, mem, 0, (size_t)bytes, 0);
  return r == CUDA_SUCCESS ? 0 : 1000 + (int)r;
}

int mpit_mc_map(unsigned long long mc_handle, int dev, unsigned long long bytes, unsigned long long* out_ptr) {
  int rc = map_handle((CUmemGenericAllocationHandle)mc_handle, (size_t)bytes, dev, out_ptr);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  g_vmm[*out_ptr] = VmmRec{(CUmemGenericAllocationHandle)mc_handle, (size_t)bytes, -1, true};
  return 0;
}

int mpit_vmm_free(unsigned long long ptr) {
  Driver& d = drv();
  VmmRec rec;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_vmm.find(ptr);
    if (it == g_vmm.end()) return MPIT_E_ARG;
    rec = it->second;
    g_vmm.erase(it);
  }
  d.MemUnmap((CUdeviceptr)ptr, rec.size);
  d.AddrFree((CUdeviceptr)ptr, rec.size);
  d.MemRelease(rec.handle);
  if (rec.fd >= 0) close(rec.fd);
  return 0;
}

int mpit_close_fd(int fd) { return close(fd); }

}  // extern "C"
