// All-reduce family over the symmetric heap (sm_100a): SUM / AVG, one-shot and two-shot, NVLS
// multimem.ld_reduce / multimem.st when the buffers have a multicast mapping, peer ld/st otherwise,
// and the variant fused with the elastic update  x -= alpha (x - mean)  (synchronous EASGD).
//
// Reference call sites: MPI_Allreduce (/root/reference/test/testreduceall.lua:28,
// /root/reference/asyncsgd/testreduceall.lua:72-77), MPI_Iallreduce (/root/reference/test/testireduceall.lua:32).
//
// Synchronisation is built from monotonically increasing epoch words instead of self-resetting
// CAS pads:
//   * every launch reads this rank's epoch word E and works under e = E + 1;
//   * START: CTA 0 stores e into slot [me] of every rank's START row (one system-scope store per
//     peer); every CTA polls its LOCAL START row until all slots are >= e.  A rank's inputs are
//     written by earlier work on its stream, so "rank r has started" == "rank r's input is ready";
//   * plain all-reduce END: the last CTA of a rank to finish (device-scope ticket) stores e into
//     the DONE row of every rank and waits for every peer's DONE -- the kernel cannot complete
//     before every slice has landed in this rank's output and every peer has stopped reading this
//     rank's input;
//   * elastic variant: the owner of a 64 KiB chunk publishes flag[chunk] = e on every rank after
//     its stores; dedicated tail CTAs of each rank consume chunks as their flags arrive, so the elastic
//     update runs concurrently with the reduction and no global barrier or serial second pass is
//     needed.  Seeing every flag also implies every owner has finished reading this
//     rank's input, so no END barrier is needed either.
// All spins are bounded and set the pad's error word on timeout.
#include "mpit_b200.h"
#include "comm.cuh"
#include "optim_math.cuh"

namespace mpit {

constexpr int kArThreads = 512;
constexpr int kArUnroll = 8;                                 // 16-byte accesses in flight per thread
constexpr long long kArChunkQuads = kArThreads * kArUnroll;  // 4096 quads = 64 KiB per chunk

struct FlagTable {
  uint32_t* peer[MPIT_MAX_RANKS];
};

struct Ar2Args {
  PeerTable in;        // symmetric input buffers (+ multicast address)
  PeerTable out;       // symmetric output buffers (two-shot gathers into every rank's copy)
  SignalTable sig;
  FlagTable flags;     // per-chunk arrival flags (elastic two-shot only)
  float* local_out;    // one-shot: private output (may be null)
  float* x;            // elastic variant: local parameters pulled toward the mean (may be null)
  long long n;         // elements
  float scale;         // 1 for SUM, 1/nranks for AVG
  float alpha;         // elastic moving rate
  // synchronous-EASGD step: after the elastic update also the Nesterov look-ahead of the next step
  // (v *= mom; x += v) and the refresh of the tensor-core operand copies, all while x is in registers
  float* v;            // momentum (null: no look-ahead)
  float mom;
  float* wpk;          // packed operand copy (null: none)
  const int* wmap;     // flat index -> packed slot, nmaps tables of n entries (null: identity)
  int nmaps;
};

__device__ __forceinline__ bool epoch_reached(uint32_t v, uint32_t e) { return (int32_t)(v - e) >= 0; }

// Bounded wait on a local word written by a peer with st.release.sys.
__device__ __forceinline__ void wait_epoch(const uint32_t* p, uint32_t e, uint32_t* err_word) {
  long long t0 = clock64();
  while (!epoch_reached(ld_acquire_sys_u32(p), e)) {
    if (clock64() - t0 > MPIT_SPIN_TIMEOUT_CYCLES) { atomicExch(err_word, 1u); break; }
  }
}

// Returns this launch's epoch after every rank has started.
__device__ __forceinline__ uint32_t ar_begin(const SignalTable& sig) {
  __shared__ uint32_t s_epoch;
  uint32_t* mine = sig.pad[sig.rank];
  if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile uint32_t*>(mine + MPIT_SIG_EPOCH) + 1u;
  __syncthreads();
  const uint32_t e = s_epoch;
  const int t = threadIdx.x;
  if (t < sig.nranks) {
    if (blockIdx.x == 0) {
      __threadfence_system();
      st_release_sys_u32(sig.pad[t] + MPIT_SIG_START + sig.rank, e);
    }
    wait_epoch(mine + MPIT_SIG_START + t, e, mine + MPIT_SIGNAL_ERR_WORD);
  }
  __syncthreads();
  return e;
}

// Device-scope ticket; the last CTA optionally runs the cross-rank DONE exchange, then publishes the
// epoch for the next launch.
__device__ __forceinline__ void ar_end(const SignalTable& sig, uint32_t e, bool cross_rank) {
  __shared__ uint32_t s_last;
  uint32_t* mine = sig.pad[sig.rank];
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();                      // this CTA's stores before the ticket
    uint32_t ticket = atomicAdd(mine + MPIT_SIG_TICKET, 1u);
    __threadfence();
    s_last = ticket == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;
  const int t = threadIdx.x;
  if (cross_rank && t < sig.nranks) {
    __threadfence_system();
    st_release_sys_u32(sig.pad[t] + MPIT_SIG_DONE + sig.rank, e);
    wait_epoch(mine + MPIT_SIG_DONE + t, e, mine + MPIT_SIGNAL_ERR_WORD);
  }
  __syncthreads();
  if (t == 0) {
    mine[MPIT_SIG_TICKET] = 0u;
    mine[MPIT_SIG_EPOCH] = e;
    __threadfence();
  }
}

__device__ __forceinline__ float4 scale4(float4 v, float s) {
  v.x *= s; v.y *= s; v.z *= s; v.w *= s;
  return v;
}
__device__ __forceinline__ float elastic1(float x, float m, float alpha) {
  return sub_rn(x, mul_rn(alpha, sub_rn(x, m)));
}

// Sum of quad q over all ranks.
__device__ __forceinline__ float4 reduce_quad(const PeerTable& in, long long q) {
  if (in.mc) return multimem_ld_reduce_add4(in.mc + 4 * q);   // reduced inside the NVSwitch
  float4 r[MPIT_MAX_RANKS];
#pragma unroll
  for (int p = 0; p < MPIT_MAX_RANKS; ++p)
    if (p < in.nranks) {
      int src = in.rank + p;
      if (src >= in.nranks) src -= in.nranks;                  // stagger peers across ranks
      r[p] = ld_sys4(in.peer[src] + 4 * q);
    }
  float4 acc = r[0];
#pragma unroll
  for (int p = 1; p < MPIT_MAX_RANKS; ++p)
    if (p < in.nranks) { acc.x += r[p].x; acc.y += r[p].y; acc.z += r[p].z; acc.w += r[p].w; }
  return acc;
}

// x[4q .. 4q+cnt) -= alpha (x - m), with a scalar path for the ragged end / unaligned x.
__device__ __forceinline__ void elastic_quad(const Ar2Args& a, long long q, const float4& m, bool vec_ok) {
  float* x = a.x;
  const long long n = a.n;
  const float alpha = a.alpha;
  const long long i0 = 4 * q;
  if (a.v || a.wpk) {
    // synchronous-EASGD step boundary: elastic update, look-ahead, operand refresh
    const float mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long long i = i0 + k;
      if (i >= n) break;
      float xv = elastic1(x[i], mm[k], alpha);
      if (a.v) {
        float vv = a.v[i];
        nesterov_lookahead_elem(xv, vv, a.mom);
        a.v[i] = vv;
      }
      x[i] = xv;
      if (a.wpk) pack_store(a.wpk, a.wmap, a.nmaps, n, i, xv);
    }
    return;
  }
  if (vec_ok && i0 + 4 <= n) {
    float4 v = ld4(x + i0);
    v.x = elastic1(v.x, m.x, alpha); v.y = elastic1(v.y, m.y, alpha);
    v.z = elastic1(v.z, m.z, alpha); v.w = elastic1(v.w, m.w, alpha);
    st4(x + i0, v);
  } else {
    const float mm[4] = {m.x, m.y, m.z, m.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i0 + k < n) x[i0 + k] = elastic1(x[i0 + k], mm[k], alpha);
  }
}
__device__ __forceinline__ void store_quad(float* dst, long long q, long long n, const float4& v, bool vec_ok) {
  const long long i0 = 4 * q;
  if (vec_ok && i0 + 4 <= n) st4(dst + i0, v);
  else {
    const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) if (i0 + k < n) dst[i0 + k] = vv[k];
  }
}

// One-shot: every rank reduces the whole vector for itself (latency path).  Symmetric buffers are
// padded to 16 bytes, so the ragged last quad is read whole and only the private destinations are
// written element-wise.
__global__ void __launch_bounds__(kArThreads) allreduce_oneshot_kernel(Ar2Args a) {
  const uint32_t e = ar_begin(a.sig);
  const long long nq = (a.n + 3) >> 2;
  const bool x_vec = a.x && ((reinterpret_cast<uintptr_t>(a.x) & 15) == 0);
  const bool o_vec = a.local_out && ((reinterpret_cast<uintptr_t>(a.local_out) & 15) == 0);
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += stride) {
    float4 acc = reduce_quad(a.in, q);
    if (a.scale != 1.f) acc = scale4(acc, a.scale);
    if (a.x) elastic_quad(a, q, acc, x_vec);
    if (a.local_out) store_quad(a.local_out, q, a.n, acc, o_vec);
  }
  ar_end(a.sig, e, true);    // nobody overwrites an input that is still being read
}

// Two-shot: chunk c (64 KiB) is reduced by rank c % R and written into every rank's output; each
// byte crosses NVLink once in and once out.  With x != null the grid is split by role: the first `nred` CTAs
// reduce (NVLink-bound; with NVLS a few dozen CTAs already saturate the switch), the others apply the elastic
// update chunk by chunk as the owners' flags arrive (HBM-bound), so the two run concurrently and the tail costs
// almost nothing beyond the reduction.  All CTAs must be resident (the host caps the grid at 2 per SM).
template <bool MC, int U>
__global__ void __launch_bounds__(kArThreads, 2) allreduce_twoshot_kernel(Ar2Args a, int nred) {
  // U = quads batched per thread: one multimem instruction per quad with NVLS; without it a quad is already R
  // peer loads, so fewer quads are batched (registers)
  const uint32_t e = ar_begin(a.sig);
  const int R = a.in.nranks, me = a.in.rank, t = threadIdx.x;
  const long long nq = (a.n + 3) >> 2;
  const long long nchunks = (nq + kArChunkQuads - 1) / kArChunkQuads;
  const bool elastic = a.x != nullptr;

  if ((int)blockIdx.x < nred) {
    // ---- reduce-scatter + all-gather of the chunks this rank owns -----------------------------
    for (long long c = me + (long long)R * blockIdx.x; c < nchunks; c += (long long)R * nred) {
      for (int u0 = 0; u0 < kArUnroll; u0 += U) {
        const long long q0 = c * kArChunkQuads + (long long)u0 * kArThreads + t;
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long q = q0 + (long long)u * kArThreads;
          if (q < nq) v[u] = MC ? multimem_ld_reduce_add4(a.in.mc + 4 * q) : reduce_quad(a.in, q);   // U loads in flight
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long long q = q0 + (long long)u * kArThreads;
          if (q < nq) {
            float4 r = a.scale != 1.f ? scale4(v[u], a.scale) : v[u];
            if (MC) {
              multimem_st4(a.out.mc + 4 * q, r);             // one store, the switch replicates it
            } else {
#pragma unroll
              for (int p = 0; p < MPIT_MAX_RANKS; ++p)
                if (p < R) {
                  int dst = me + p;
                  if (dst >= R) dst -= R;
                  st_sys4(a.out.peer[dst] + 4 * q, r);
                }
            }
          }
        }
      }
      if (elastic) {
        __syncthreads();                                     // every thread's stores are issued ...
        if (t < R) {
          __threadfence_system();                            // ... and ordered before the flag
          st_release_sys_u32(a.flags.peer[t] + c, e);
        }
      }
    }
  } else {
    // ---- elastic tail: consume chunks in the order the owners finish them ---------------------------
    const bool x_vec = (reinterpret_cast<uintptr_t>(a.x) & 15) == 0;
    const float* mean = a.out.peer[me];
    uint32_t* err = a.sig.pad[me] + MPIT_SIGNAL_ERR_WORD;
    const int ntail = (int)gridDim.x - nred;
    for (long long c = (long long)blockIdx.x - nred; c < nchunks; c += ntail) {
      if (t == 0) wait_epoch(a.flags.peer[me] + c, e, err);
      __syncthreads();
      const long long q0 = c * kArChunkQuads + t;
      float4 m[kArUnroll];
#pragma unroll
      for (int u = 0; u < kArUnroll; ++u) {
        const long long q = q0 + (long long)u * kArThreads;
        if (q < nq) m[u] = ld_sys4(mean + 4 * q);            // written by a peer: bypass L1
      }
#pragma unroll
      for (int u = 0; u < kArUnroll; ++u) {
        const long long q = q0 + (long long)u * kArThreads;
        if (q < nq) elastic_quad(a, q, m[u], x_vec);
      }
    }
  }
  ar_end(a.sig, e, !elastic);
}

static PeerTable ar_table(const unsigned long long* ptrs, unsigned long long mc, int nranks, int rank) {
  PeerTable t{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) t.peer[i] = (ptrs && i < nranks) ? reinterpret_cast<float*>(ptrs[i]) : nullptr;
  t.mc = reinterpret_cast<float*>(mc);
  t.nranks = nranks;
  t.rank = rank;
  return t;
}

}  // namespace mpit

using namespace mpit;

extern "C" {

long long mpit_allreduce_flag_words(long long capacity_floats) {
  long long nq = (capacity_floats + 3) / 4;
  return (nq + kArChunkQuads - 1) / kArChunkQuads + 1;
}

static int allreduce_impl(const unsigned long long* in_ptrs, unsigned long long in_mc,
                          const unsigned long long* out_ptrs, unsigned long long out_mc,
                          const unsigned long long* sig_ptrs, const unsigned long long* flag_ptrs, int nranks, int rank,
                          float* local_out, float* x_elastic, long long n, float scale, float alpha, int algo,
                          int max_blocks, float* v, float mom, float* wpk, const int* wmap, int nmaps, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || !in_ptrs || !sig_ptrs) return MPIT_E_ARG;
  Ar2Args a{};
  a.v = v; a.mom = mom; a.wpk = wpk; a.wmap = wmap; a.nmaps = nmaps;
  a.in = ar_table(in_ptrs, in_mc, nranks, rank);
  a.sig.nranks = nranks; a.sig.rank = rank;
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) a.sig.pad[i] = i < nranks ? reinterpret_cast<uint32_t*>(sig_ptrs[i]) : nullptr;
  a.local_out = local_out; a.x = x_elastic; a.n = n; a.scale = scale; a.alpha = alpha;
  const long long nq = (n + 3) >> 2;
  cudaStream_t st = (cudaStream_t)stream;
  if (algo == 0) {
    long long need = (nq + kArThreads - 1) / kArThreads;
    int cap = max_blocks > 0 ? max_blocks : MPIT_NUM_SMS;
    int grid = (int)(need < cap ? need : cap);
    allreduce_oneshot_kernel<<<grid < 1 ? 1 : grid, kArThreads, 0, st>>>(a);
  } else {
    if (!out_ptrs) return MPIT_E_ARG;
    if (x_elastic && !flag_ptrs) return MPIT_E_ARG;
    a.out = ar_table(out_ptrs, out_mc, nranks, rank);
    for (int i = 0; i < MPIT_MAX_RANKS; ++i)
      a.flags.peer[i] = (flag_ptrs && i < nranks) ? reinterpret_cast<uint32_t*>(flag_ptrs[i]) : nullptr;
    // Grid.  Reducers: one CTA per owned chunk, at most 32 with NVLS (16-64 CTAs saturate the switch; more only
    // add contention: 256 MiB on 8 GPUs 624 us with 16-32 CTAs, 678 us with 296) and at most 2 per SM with peer
    // loads (which want every SM).  The elastic variant adds tail CTAs up to 2 per SM in total; every CTA must be
    // resident because tail CTAs wait on flags that the reducers of the other ranks produce.
    const bool mc = in_mc && out_mc;
    const long long nchunks = (nq + kArChunkQuads - 1) / kArChunkQuads;
    const long long owned = (nchunks + nranks - 1) / nranks;
    const int full = 2 * MPIT_NUM_SMS;
    int cap = max_blocks > 0 ? (max_blocks < full ? max_blocks : full) : (mc ? 32 : (x_elastic ? full * 5 / 8 : full));
    if (x_elastic && cap > full * 3 / 4) cap = full * 3 / 4;      // the tail needs CTAs of its own
    int nred = (int)(owned < cap ? owned : cap);
    if (nred < 1) nred = 1;
    int grid = nred;
    if (x_elastic) {
      long long ntail = nchunks < full - nred ? nchunks : full - nred;
      if (ntail < 1) ntail = 1;
      grid = nred + (int)ntail;
    }
    if (mc) allreduce_twoshot_kernel<true, kArUnroll><<<grid, kArThreads, 0, st>>>(a, nred);
    else {
      a.in.mc = nullptr; a.out.mc = nullptr;
      if (nranks <= 2) allreduce_twoshot_kernel<false, 4><<<grid, kArThreads, 0, st>>>(a, nred);
      else if (nranks <= 4) allreduce_twoshot_kernel<false, 2><<<grid, kArThreads, 0, st>>>(a, nred);
      else allreduce_twoshot_kernel<false, 1><<<grid, kArThreads, 0, st>>>(a, nred);
    }
  }
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_allreduce(const unsigned long long* in_ptrs, unsigned long long in_mc,
                   const unsigned long long* out_ptrs, unsigned long long out_mc,
                   const unsigned long long* sig_ptrs, const unsigned long long* flag_ptrs, int nranks, int rank,
                   float* local_out, float* x_elastic, long long n, float scale, float alpha, int algo,
                   int max_blocks, void* stream) {
  return allreduce_impl(in_ptrs, in_mc, out_ptrs, out_mc, sig_ptrs, flag_ptrs, nranks, rank, local_out, x_elastic, n,
                        scale, alpha, algo, max_blocks, nullptr, 0.f, nullptr, nullptr, 0, stream);
}

int mpit_sync_easgd_step(const unsigned long long* in_ptrs, unsigned long long in_mc,
                         const unsigned long long* out_ptrs, unsigned long long out_mc,
                         const unsigned long long* sig_ptrs, const unsigned long long* flag_ptrs, int nranks, int rank,
                         float* x, float* v, float* wpk, const int* wmap, int nmaps, long long n, float alpha, float mom,
                         int algo, int max_blocks, void* stream) {
  if (!x) return MPIT_E_ARG;
  return allreduce_impl(in_ptrs, in_mc, out_ptrs, out_mc, sig_ptrs, flag_ptrs, nranks, rank, nullptr, x, n,
                        1.f / (float)nranks, alpha, algo, max_blocks, v, mom, wpk, wmap, nmaps, stream);
}

}  // extern "C"
