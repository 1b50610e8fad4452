// Derived-datatype pack / unpack on the device: the strided side of an MPI datatype (a block list that
// tiles with a fixed extent, see csrc/msg/hostio.cpp for the host engine) gathered into, or scattered
// from, a contiguous stream in device memory.
//
// The reference hands device pointers of CudaTensors straight to a CUDA-aware MPI
// (/root/reference/lua-mpi.h:70-78), where a derived datatype on a device buffer is packed by the MPI
// library; here the message layer stages device buffers through pinned memory, so the strided gather
// happens on the GPU first and one contiguous copy crosses PCIe.
//
// Work decomposition: the stream is cut into `unit`-byte words (unit = the largest power of two <= 16 that
// divides every displacement, length, the extent and both base addresses; chosen by the caller).  One
// thread moves one word per grid-stride step: stream word -> (replica, offset inside the replica) ->
// binary search of the block prefix table -> strided address.  The block table is read through L1/L2
// (`__ldg`); with a few dozen blocks it stays cached.  Bandwidth-bound; no tensor-core work here.
#include <cuda_runtime.h>
#include <stdint.h>

#include "mpit_b200.h"

namespace {

template <typename W>
__global__ void __launch_bounds__(256) dt_move_kernel(char* __restrict__ strided, char* __restrict__ stream,
                                                      long long total_words, long long words_per_replica,
                                                      long long extent, int nblk,
                                                      const long long* __restrict__ disp,   // [nblk] bytes
                                                      const long long* __restrict__ pre,    // [nblk+1] bytes
                                                      int to_stream) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x; w < total_words; w += stride) {
    const long long rep = w / words_per_replica;
    const long long within = (w - rep * words_per_replica) * (long long)sizeof(W);   // byte offset in the replica
    // largest i with pre[i] <= within
    int lo = 0, hi = nblk - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (__ldg(pre + mid) <= within) lo = mid; else hi = mid - 1;
    }
    char* s = strided + rep * extent + __ldg(disp + lo) + (within - __ldg(pre + lo));
    char* t = stream + w * (long long)sizeof(W);
    if (to_stream) *reinterpret_cast<W*>(t) = *reinterpret_cast<const W*>(s);
    else *reinterpret_cast<W*>(s) = *reinterpret_cast<const W*>(t);
  }
}

template <typename W>
int launch(char* strided, char* stream, long long nbytes, long long size, long long extent, int nblk,
           const long long* disp, const long long* pre, int to_stream, cudaStream_t st) {
  const long long total = nbytes / (long long)sizeof(W);
  const long long per = size / (long long)sizeof(W);
  long long blocks = (total + 255) / 256;
  if (blocks > 148LL * 8) blocks = 148LL * 8;       // grid-stride beyond eight CTAs per SM
  if (blocks < 1) blocks = 1;
  dt_move_kernel<W><<<(unsigned)blocks, 256, 0, st>>>(strided, stream, total, per, extent, nblk, disp, pre, to_stream);
  return (int)cudaGetLastError();
}

}  // namespace

extern "C" int mpit_dt_move_device(void* strided_base, void* stream_buf, long long nbytes, long long size,
                                   long long extent, int nblk, const long long* d_disp, const long long* d_pre,
                                   int unit, int to_stream, void* stream) {
  if (nbytes == 0) return 0;
  if (nbytes < 0 || size <= 0 || nblk <= 0 || !strided_base || !stream_buf || !d_disp || !d_pre) return MPIT_E_ARG;
  if (nbytes % unit || size % unit) return MPIT_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  char* a = (char*)strided_base;
  char* b = (char*)stream_buf;
  switch (unit) {
    case 16: return launch<uint4>(a, b, nbytes, size, extent, nblk, d_disp, d_pre, to_stream, st);
    case 8: return launch<uint2>(a, b, nbytes, size, extent, nblk, d_disp, d_pre, to_stream, st);
    case 4: return launch<uint32_t>(a, b, nbytes, size, extent, nblk, d_disp, d_pre, to_stream, st);
    case 2: return launch<uint16_t>(a, b, nbytes, size, extent, nblk, d_disp, d_pre, to_stream, st);
    case 1: return launch<uint8_t>(a, b, nbytes, size, extent, nblk, d_disp, d_pre, to_stream, st);
    default: return MPIT_E_ARG;
  }
}
