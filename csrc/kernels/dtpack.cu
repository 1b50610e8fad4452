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

// Regular block lists (every block `blen` bytes, block i at d0 + i*bstride: MPI_Type_vector, 2-d sub-arrays,
// contiguous runs) need no table: the block index is a division.  Four independent words per thread and step keep
// enough loads in flight to cover HBM latency (the table-search kernel above is a chain of dependent loads and
// reaches about 30 % of the copy bandwidth on a 64 MiB gather; this form is what the large cases use).
template <typename W, bool SMALL>   // SMALL: every word index fits 32 bits (64-bit division is emulated and slow)
__global__ void __launch_bounds__(256) dt_move_regular_kernel(char* __restrict__ strided, char* __restrict__ stream,
                                                              long long total_words, long long words_per_replica,
                                                              long long extent, long long d0, long long blen_words,
                                                              long long bstride, int to_stream) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  long long w = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  auto addr = [&](long long word) -> char* {
    if (SMALL) {
      const unsigned wd = (unsigned)word, wpr = (unsigned)words_per_replica, bl = (unsigned)blen_words;
      const unsigned rep = wd / wpr;
      const unsigned within = wd - rep * wpr;
      const unsigned blk = within / bl;
      return strided + (long long)rep * extent + d0 + (long long)blk * bstride +
             (long long)(within - blk * bl) * (long long)sizeof(W);
    }
    const long long rep = word / words_per_replica;
    const long long within = word - rep * words_per_replica;
    const long long blk = within / blen_words;
    return strided + rep * extent + d0 + blk * bstride + (within - blk * blen_words) * (long long)sizeof(W);
  };
  for (; w + 3 * stride < total_words; w += 4 * stride) {
    char* a0 = addr(w); char* a1 = addr(w + stride); char* a2 = addr(w + 2 * stride); char* a3 = addr(w + 3 * stride);
    W* t = reinterpret_cast<W*>(stream);
    if (to_stream) {
      const W v0 = *reinterpret_cast<const W*>(a0), v1 = *reinterpret_cast<const W*>(a1),
              v2 = *reinterpret_cast<const W*>(a2), v3 = *reinterpret_cast<const W*>(a3);
      t[w] = v0; t[w + stride] = v1; t[w + 2 * stride] = v2; t[w + 3 * stride] = v3;
    } else {
      const W v0 = t[w], v1 = t[w + stride], v2 = t[w + 2 * stride], v3 = t[w + 3 * stride];
      *reinterpret_cast<W*>(a0) = v0; *reinterpret_cast<W*>(a1) = v1;
      *reinterpret_cast<W*>(a2) = v2; *reinterpret_cast<W*>(a3) = v3;
    }
  }
  for (; w < total_words; w += stride) {
    char* a = addr(w);
    W* t = reinterpret_cast<W*>(stream) + w;
    if (to_stream) *t = *reinterpret_cast<const W*>(a); else *reinterpret_cast<W*>(a) = *t;
  }
}

template <typename W>
int launch_regular(char* strided, char* stream, long long nbytes, long long size, long long extent, long long d0,
                   long long blen, long long bstride, int to_stream, cudaStream_t st) {
  const long long total = nbytes / (long long)sizeof(W);
  long long blocks = (total + 4 * 256 - 1) / (4 * 256);
  if (blocks > 148LL * 8) blocks = 148LL * 8;
  if (blocks < 1) blocks = 1;
  const long long per = size / (long long)sizeof(W), bw = blen / (long long)sizeof(W);
  if (total + 4 * blocks * 256 < (1LL << 32) && per < (1LL << 32))
    dt_move_regular_kernel<W, true><<<(unsigned)blocks, 256, 0, st>>>(strided, stream, total, per, extent, d0, bw,
                                                                      bstride, to_stream);
  else
    dt_move_regular_kernel<W, false><<<(unsigned)blocks, 256, 0, st>>>(strided, stream, total, per, extent, d0, bw,
                                                                       bstride, to_stream);
  return (int)cudaGetLastError();
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

// Same transfer for a regular block list: nblk blocks of `blen` bytes, block i at byte d0 + i*bstride of a replica
// (so size = nblk*blen).  No device tables.
extern "C" int mpit_dt_move_device_regular(void* strided_base, void* stream_buf, long long nbytes, long long size,
                                           long long extent, long long d0, long long blen, long long bstride,
                                           int unit, int to_stream, void* stream) {
  if (nbytes == 0) return 0;
  if (nbytes < 0 || size <= 0 || blen <= 0 || size % blen || !strided_base || !stream_buf) return MPIT_E_ARG;
  if (nbytes % unit || blen % unit) return MPIT_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  char* a = (char*)strided_base;
  char* b = (char*)stream_buf;
  switch (unit) {
    case 16: return launch_regular<uint4>(a, b, nbytes, size, extent, d0, blen, bstride, to_stream, st);
    case 8: return launch_regular<uint2>(a, b, nbytes, size, extent, d0, blen, bstride, to_stream, st);
    case 4: return launch_regular<uint32_t>(a, b, nbytes, size, extent, d0, blen, bstride, to_stream, st);
    case 2: return launch_regular<uint16_t>(a, b, nbytes, size, extent, d0, blen, bstride, to_stream, st);
    case 1: return launch_regular<uint8_t>(a, b, nbytes, size, extent, d0, blen, bstride, to_stream, st);
    default: return MPIT_E_ARG;
  }
}
