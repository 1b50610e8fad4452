// Fused exchange kernels over NVLink peer memory (sm_100a).
//
// Every GPU is a worker *and* the owner of one shard of the centre variable; the reference's
// parameter-server round trips collapse into one-sided accesses issued from inside the kernel
// that also does the math:
//
//   pull   (pclient.lua:74-84  <-> pserver.lua:65-79)   -> ld.relaxed.sys on the owner's shard
//   push   (pclient.lua:49-59  <-> pserver.lua:83-98)   -> red.relaxed.sys.add (p += g on owner)
//   seed   (pclient.lua:62-72  <-> pserver.lua:101-111) -> st.relaxed.sys
//   elastic step (optim-eamsgd.lua:58-70 + :26-47)      -> one pass: pull, d = a(x - c), push d,
//                                                          Nesterov apply/lookahead, operand pack
//   allreduce (test/testreduceall.lua:28)               -> one-shot / two-shot over peer loads or
//                                                          NVLS multimem.ld_reduce / multimem.st
//
// (paths relative to /root/reference/asyncsgd unless noted).  No NCCL or MPI call is made here.
#include "mpit_b200.h"
#include "comm.cuh"
#include "optim_math.cuh"

namespace mpit {

// Walks the flat vector in units of 4 consecutive elements ("quads") that never straddle a shard.
// The walk is organised in chunks of kChunkQuads quads that are dealt round-robin over the shards,
// starting at the shard after the caller's own: concurrently running CTAs therefore talk to all
// owners at once and the ranks never gang up on one owner (every rank walking shard 0, then shard 1,
// ... in lockstep would serialise the whole exchange on one GPU's links at a time).
constexpr long long kChunkQuads = 1024;   // 16 KiB per chunk
struct QuadWalk {
  Partition part;
  long long cps;    // chunks per shard (enough for the largest shard)
  long long total;  // virtual quads to visit: S * cps * kChunkQuads (some fall outside a shard: cnt = 0)
  int rank;
  __host__ __device__ static QuadWalk make(long long n, int S, int rank = 0) {
    QuadWalk w;
    w.part.n = n; w.part.S = S; w.part.base = n / S;
    long long max_quads = (w.part.size(S - 1) + 3) / 4;
    w.cps = (max_quads + kChunkQuads - 1) / kChunkQuads;
    if (w.cps < 1) w.cps = 1;
    w.total = w.cps * S * kChunkQuads;
    w.rank = rank;
    return w;
  }
  __device__ __forceinline__ void at(long long v, int& s, long long& j0, long long& i0, int& cnt) const {
    // chunk index fits 32 bits for any n < 2^40; 64-bit division by a run-time value is ~200 instructions,
    // which on the 350k-parameter step boundary cost more than the memory traffic
    const long long c = v / kChunkQuads, within = v - c * kChunkQuads;
    const unsigned c32 = (unsigned)c, S32 = (unsigned)part.S;
    const unsigned k = S32 == 1 ? c32 : c32 / S32;
    const int lane_shard = (int)(c32 - k * S32);
    s = lane_shard + rank + 1;
    if (s >= part.S) s -= part.S;
    if (s >= part.S) s -= part.S;
    j0 = ((long long)k * kChunkQuads + within) * 4;
    i0 = part.begin(s) + j0;
    long long rem = part.size(s) - j0;
    cnt = rem >= 4 ? 4 : (rem > 0 ? (int)rem : 0);
  }
};

__device__ __forceinline__ void load4(const float* p, long long i0, int cnt, float (&r)[4]) {
  if (cnt == 4 && ((i0 & 3) == 0)) {
    float4 t = ld4(p + i0);
    r[0] = t.x; r[1] = t.y; r[2] = t.z; r[3] = t.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = k < cnt ? p[i0 + k] : 0.f;
  }
}
__device__ __forceinline__ void store4(float* p, long long i0, int cnt, const float (&r)[4]) {
  if (cnt == 4 && ((i0 & 3) == 0)) {
    st4(p + i0, make_float4(r[0], r[1], r[2], r[3]));
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) if (k < cnt) p[i0 + k] = r[k];
  }
}
// Owner-side accesses: j0 is always a multiple of 4 and shard buffers are 16-byte aligned and
// padded to a multiple of 4 floats, so full quads are always vector accesses over NVLink.
__device__ __forceinline__ void peer_load4(const float* p, int cnt, float (&r)[4]) {
  if (cnt == 4) {
    float4 t = ld_sys4(p);
    r[0] = t.x; r[1] = t.y; r[2] = t.z; r[3] = t.w;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = k < cnt ? ld_sys(p + k) : 0.f;
  }
}
__device__ __forceinline__ void peer_store4(float* p, int cnt, const float (&r)[4]) {
  if (cnt == 4) st_sys4(p, make_float4(r[0], r[1], r[2], r[3]));
  else {
#pragma unroll
    for (int k = 0; k < 4; ++k) if (k < cnt) st_sys(p + k, r[k]);
  }
}
__device__ __forceinline__ void peer_red4(float* p, int cnt, const float (&r)[4]) {
  if (cnt == 4) red_add_sys4(p, make_float4(r[0], r[1], r[2], r[3]));
  else {
#pragma unroll
    for (int k = 0; k < 4; ++k) if (k < cnt) red_add_sys(p + k, r[k]);
  }
}

enum ExchMode : int { EX_PULL = 0, EX_PUT = 1, EX_PUSH = 2, EX_ELASTIC = 3, EX_FUSED = 4, EX_DOWNPOUR = 5 };

struct ExchArgs {
  PeerTable tab;
  QuadWalk walk;
  float* x;            // local parameters (flat)
  float* v;            // momentum
  const float* g;      // gradient source (flat, or packed with gmap)
  const int* gmap;
  float* d;            // elastic difference buffer (read by apply, written by sync)
  float* src;          // push source / pull destination when not x
  float* wpk;
  const int* wmap;
  int nmaps;
  long long n;
  long long map_base;  // pull: elements below map_base have no operand copy; the tables cover [map_base, n)
  float alpha, clr, l2wd, mom_next, scale, l1, l2, clip;
  int do_apply, use_mom, do_lookahead, do_sync, has_d_in, zero_src;
  int la_in;           // x holds the parameters BEFORE the pending look-ahead and v is already scaled: start from x + v
};

template <int MODE>
__global__ void __launch_bounds__(256) exchange_kernel(ExchArgs a) {
  pdl_trigger();   // the step kernel that follows may start its set-up; it waits for this grid's completion
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < a.walk.total; q += stride) {
    int s, cnt;
    long long j0, i0;
    a.walk.at(q, s, j0, i0, cnt);
    if (cnt == 0) continue;
    float* cptr = a.tab.peer[s] + j0;
    if (MODE == EX_PULL) {
      // x <- centre (MPI_Get).  Issue the remote load first; everything else hides under it.
      float c[4];
      peer_load4(cptr, cnt, c);
      store4(a.src, i0, cnt, c);
      if (a.wpk) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (k < cnt && i0 + k >= a.map_base)
            pack_store(a.wpk, a.wmap, a.nmaps, a.n - a.map_base, i0 + k - a.map_base, c[k]);
      }
    } else if (MODE == EX_PUT) {
      float r[4];
      load4(a.src, i0, cnt, r);
      peer_store4(cptr, cnt, r);
    } else if (MODE == EX_PUSH) {
      // centre += scale * clip(reg(g))   (MPI_Accumulate); optionally clears the accumulator.
      float r[4];
      load4(a.src, i0, cnt, r);
      if (a.l1 != 0.f || a.l2 != 0.f || a.clip > 0.f) {
        float p[4];
        load4(a.x, i0, cnt, p);
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = grad_reg_clip(r[k], p[k], a.l1, a.l2, a.clip);
      }
      if (a.scale != 1.f) {
#pragma unroll
        for (int k = 0; k < 4; ++k) r[k] = mul_rn(r[k], a.scale);
      }
      peer_red4(cptr, cnt, r);
      if (a.zero_src) {
        float z[4] = {0.f, 0.f, 0.f, 0.f};
        store4(a.src, i0, cnt, z);
      }
    } else if (MODE == EX_DOWNPOUR) {
      // Downpour step boundary in one pass (optim-downpour.lua:24-54): g' = -lr (g + l2wd x); acc += g';
      // sync step: centre += acc on the owner and x = the centre right after that addition (the push followed
      // by the pull, one NVLink round trip: the accumulate returns the value it found), acc = 0;
      // otherwise the worker moves on its own, x += g'.  Operand copies refreshed in the same pass.
      float x[4], acc[4], g[4];
      load4(a.x, i0, cnt, x);
      load4(a.src, i0, cnt, acc);
      if (a.gmap) {
#pragma unroll
        for (int k = 0; k < 4; ++k) g[k] = k < cnt ? a.g[a.gmap[i0 + k]] : 0.f;
      } else {
        load4(a.g, i0, cnt, g);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float gk = g[k];
        if (a.l2wd != 0.f) gk = add_rn(gk, mul_rn(a.l2wd, x[k]));
        g[k] = mul_rn(gk, a.clr);                 // clr = -lr
        acc[k] = add_rn(acc[k], g[k]);
      }
      if (a.do_sync) {
        float old[4] = {0.f, 0.f, 0.f, 0.f};
        if (cnt == 4) {
          float4 o = atom_add_sys4(cptr, make_float4(acc[0], acc[1], acc[2], acc[3]));
          old[0] = o.x; old[1] = o.y; old[2] = o.z; old[3] = o.w;
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) if (k < cnt) old[k] = atom_add_sys(cptr + k, acc[k]);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) { x[k] = add_rn(old[k], acc[k]); acc[k] = 0.f; }
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) x[k] = add_rn(x[k], g[k]);
      }
      store4(a.x, i0, cnt, x);
      store4(a.src, i0, cnt, acc);
      if (a.wpk) {
#pragma unroll
        for (int k = 0; k < 4; ++k) if (k < cnt) pack_store(a.wpk, a.wmap, a.nmaps, a.n, i0 + k, x[k]);
      }
    } else {
      // EX_ELASTIC / EX_FUSED
      float c[4] = {0.f, 0.f, 0.f, 0.f};
      if (a.do_sync) peer_load4(cptr, cnt, c);  // remote load in flight while we do the local step
      // operand-copy slots of these four parameters: fetched now, used after the update (the dependent
      // index -> store chain otherwise starts only once x is final)
      int wslot[2][4];
      const bool repack = MODE == EX_FUSED && a.wpk && a.wmap;
      if (repack) {
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
          for (int k = 0; k < 4; ++k) wslot[m][k] = (m < a.nmaps && k < cnt) ? a.wmap[(long long)m * a.n + i0 + k] : -1;
      }
      float x[4], v[4] = {0.f, 0.f, 0.f, 0.f};
      load4(a.x, i0, cnt, x);
      if (a.use_mom) load4(a.v, i0, cnt, v);
      if (MODE == EX_FUSED && a.do_apply) {
        if (a.la_in && a.use_mom) {
          // the previous step finished inside the weight-gradient reduction (umma_wgrad.cu), which keeps x at the
          // parameters proper and folds the look-ahead into the operand copy only: the pending gradient was taken
          // at x + v (v already scaled)
#pragma unroll
          for (int k = 0; k < 4; ++k) x[k] = add_rn(x[k], v[k]);
        }
        float g[4], din[4] = {0.f, 0.f, 0.f, 0.f};
        if (a.gmap) {
#pragma unroll
          for (int k = 0; k < 4; ++k) g[k] = k < cnt ? a.g[a.gmap[i0 + k]] : 0.f;
        } else {
          load4(a.g, i0, cnt, g);
        }
        if (a.has_d_in) load4(a.d, i0, cnt, din);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          msgd_apply_elem(x[k], v[k], g[k], din[k], a.clr, a.l2wd, a.use_mom != 0, a.has_d_in != 0);
      }
      if (a.do_sync) {
        float dd[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) dd[k] = mul_rn(a.alpha, sub_rn(x[k], c[k]));
        peer_red4(cptr, cnt, dd);   // centre += d, fire-and-forget: completes under fwd/bwd
        store4(a.d, i0, cnt, dd);
      }
      if (MODE == EX_FUSED) {
        if (a.do_lookahead && a.use_mom) {
          if (a.la_in && !a.do_apply) {
            // v was scaled when the previous step was finished by the reduction: only re-create the look-ahead point
#pragma unroll
            for (int k = 0; k < 4; ++k) x[k] = add_rn(x[k], v[k]);
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) nesterov_lookahead_elem(x[k], v[k], a.mom_next);
          }
        }
        store4(a.x, i0, cnt, x);
        if (a.use_mom) store4(a.v, i0, cnt, v);
        if (repack && a.nmaps <= 2) {
#pragma unroll
          for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int k = 0; k < 4; ++k) if (wslot[m][k] >= 0) a.wpk[wslot[m][k]] = x[k];
        } else if (a.wpk) {
#pragma unroll
          for (int k = 0; k < 4; ++k) if (k < cnt) pack_store(a.wpk, a.wmap, a.nmaps, a.n, i0 + k, x[k]);
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
// Broadcast / barrier over symmetric buffers.  (The all-reduce family lives in allreduce.cu.)
// ---------------------------------------------------------------------------------------
struct ArArgs {
  PeerTable out;      // symmetric buffers
  SignalTable sig;
  long long n;
};

// Broadcast (MPI_Bcast, wrapped at /root/reference/mpifuncs.c:145): the root pushes.
__global__ void __launch_bounds__(512) bcast_kernel(ArArgs a, int root) {
  int R = a.out.nranks, me = a.out.rank;
  long long stride = (long long)gridDim.x * blockDim.x;
  if (me == root) {
    const float* src = a.out.peer[me];
    long long n4 = a.n >> 2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
      float4 v = ld4(src + 4 * i);
      if (a.out.mc) multimem_st4(a.out.mc + 4 * i, v);
      else {
#pragma unroll
        for (int p = 0; p < MPIT_MAX_RANKS; ++p)
          if (p < R && p != me) st_sys4(a.out.peer[p] + 4 * i, v);
      }
    }
    if (blockIdx.x == 0)
      for (long long i = (n4 << 2) + threadIdx.x; i < a.n; i += blockDim.x)
        for (int p = 0; p < R; ++p) if (p != me) st_sys(a.out.peer[p] + i, src[i]);
  }
  block_barrier_sys(a.sig, blockIdx.x);
}

__global__ void barrier_kernel(SignalTable sig) { block_barrier_sys(sig, blockIdx.x); }

static PeerTable make_table(const unsigned long long* ptrs, unsigned long long mc, int nranks, int rank) {
  PeerTable t{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) t.peer[i] = i < nranks ? reinterpret_cast<float*>(ptrs[i]) : nullptr;
  t.mc = reinterpret_cast<float*>(mc);
  t.nranks = nranks;
  t.rank = rank;
  return t;
}
static SignalTable make_sig(const unsigned long long* ptrs, int nranks, int rank) {
  SignalTable t{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) t.pad[i] = i < nranks ? reinterpret_cast<uint32_t*>(ptrs[i]) : nullptr;
  t.nranks = nranks;
  t.rank = rank;
  return t;
}

// ---- one-sided window operations (MPI_Put / MPI_Get / MPI_Accumulate, mpifuncs.c:1656,1131,9) ----
// `remote` is an address inside a peer's symmetric buffer mapped into this process.  Vector (16-byte) accesses
// when both sides allow it, scalar otherwise (arbitrary displacements are legal in a window).
enum RmaOp : int { RMA_PUT = 0, RMA_GET = 1, RMA_ACC = 2 };

__global__ void __launch_bounds__(256) peer_rma_kernel(int op, float* remote, float* local, long long n, float scale) {
  const long long stride = (long long)gridDim.x * blockDim.x, t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool vec = (((uintptr_t)remote | (uintptr_t)local) & 15) == 0;
  const long long n4 = vec ? n >> 2 : 0;
  for (long long q = t0; q < n4; q += stride) {
    if (op == RMA_GET) {
      st4(local + 4 * q, ld_sys4(remote + 4 * q));
    } else {
      float4 v = ld4(local + 4 * q);
      if (scale != 1.f) { v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale; }
      if (op == RMA_PUT) st_sys4(remote + 4 * q, v);
      else red_add_sys4(remote + 4 * q, v);
    }
  }
  for (long long i = 4 * n4 + t0; i < n; i += stride) {
    if (op == RMA_GET) local[i] = ld_sys(remote + i);
    else if (op == RMA_PUT) st_sys(remote + i, local[i] * scale);
    else red_add_sys(remote + i, local[i] * scale);
  }
}

// Gradient push for a Linear layer whose weight W [N][K] is row-sharded over the owners (ops/pulled_linear.py):
// W_owner += scale * g on the owner's rows (vector red.add along k) AND on the owner's transposed copy
// [K][rows_per_owner] that the data-gradient GEMM reads (scalar red.add), so both layouts stay equal without a
// transpose pass on the owner.  Reference: pclient.lua:49-59 + pserver.lua:91 (p += g on the server).
__global__ void __launch_bounds__(256)
pulled_push_kernel(PeerTable wf, PeerTable wd, const float* __restrict__ g, int N, int K, int rpo, float scale) {
  const long long nq = (long long)N * (K >> 2);
  const long long stride = (long long)gridDim.x * blockDim.x;
  // rows are dealt starting at the owner after this rank, like the other exchanges
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += stride) {
    const int kq = K >> 2;
    int n = (int)(q / kq);
    const int k4 = (int)(q - (long long)n * kq) * 4;
    n += (wf.rank + 1) * rpo;
    if (n >= N) n -= N;
    const int owner = n / rpo, nl = n - owner * rpo;
    float4 v = ld4_stream(g + (size_t)n * K + k4);
    v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
    red_add_sys4(wf.peer[owner] + (size_t)nl * K + k4, v);
    float* t = wd.peer[owner] + (size_t)k4 * rpo + nl;
    red_add_sys(t, v.x);
    red_add_sys(t + rpo, v.y);
    red_add_sys(t + 2 * rpo, v.z);
    red_add_sys(t + 3 * rpo, v.w);
  }
}

static int exch_grid(const QuadWalk& w) { return ew_grid(w.total, 256, 1, 8); }

}  // namespace mpit

using namespace mpit;

extern "C" {

int mpit_signal_words(void) { return MPIT_SIGNAL_WORDS; }
int mpit_max_ranks(void) { return MPIT_MAX_RANKS; }

int mpit_param_pull(const unsigned long long* shard_ptrs, int nranks, int rank, float* dst, long long n,
                    float* wpk, const int* wmap, int nmaps, long long map_base, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || map_base < 0 || map_base > n) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.src = dst; a.wpk = wpk; a.wmap = wmap; a.nmaps = nmaps; a.n = n; a.map_base = map_base;
  exchange_kernel<EX_PULL><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_param_put(const unsigned long long* shard_ptrs, int nranks, int rank, const float* src, long long n,
                   void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.src = const_cast<float*>(src);
  exchange_kernel<EX_PUT><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_push_add(const unsigned long long* shard_ptrs, int nranks, int rank, float* src, const float* params,
                  long long n, float scale, float l1, float l2, float clip, int zero_src, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.src = src; a.x = const_cast<float*>(params);
  a.scale = scale; a.l1 = l1; a.l2 = l2; a.clip = clip; a.zero_src = zero_src;
  if ((l1 != 0.f || l2 != 0.f) && !params) return MPIT_E_ARG;
  exchange_kernel<EX_PUSH><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_elastic_exchange(const unsigned long long* shard_ptrs, int nranks, int rank, const float* x, float* d,
                          long long n, float alpha, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.x = const_cast<float*>(x); a.d = d; a.alpha = alpha; a.do_sync = 1;
  exchange_kernel<EX_ELASTIC><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_eamsgd_fused_step(const unsigned long long* shard_ptrs, int nranks, int rank, float* x, float* v,
                           const float* g, const int* gmap, float* d, float* wpk, const int* wmap,
                           int nmaps, long long n, float alpha, float clr, float l2wd, float mom_next, int do_apply,
                           int use_mom, int do_lookahead, int do_sync, int has_d_in, int la_in, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.x = x; a.v = v; a.g = g; a.gmap = gmap; a.d = d; a.wpk = wpk; a.wmap = wmap; a.nmaps = nmaps; a.n = n;
  a.alpha = alpha; a.clr = clr; a.l2wd = l2wd; a.mom_next = mom_next;
  a.do_apply = do_apply; a.use_mom = use_mom; a.do_lookahead = do_lookahead; a.do_sync = do_sync;
  a.has_d_in = has_d_in; a.la_in = la_in;
  exchange_kernel<EX_FUSED><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

/* Downpour step boundary (optim-downpour.lua:24-54) in one kernel: see EX_DOWNPOUR. */
int mpit_downpour_fused_step(const unsigned long long* shard_ptrs, int nranks, int rank, float* x, float* acc,
                             const float* g, const int* gmap, float* wpk, const int* wmap, int nmaps, long long n,
                             float neg_lr, float l2wd, int do_sync, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || !x || !acc || !g) return MPIT_E_ARG;
  ExchArgs a{};
  a.tab = make_table(shard_ptrs, 0, nranks, rank);
  a.walk = QuadWalk::make(n, nranks, rank);
  a.x = x; a.src = acc; a.g = g; a.gmap = gmap; a.wpk = wpk; a.wmap = wmap; a.nmaps = nmaps; a.n = n;
  a.clr = neg_lr; a.l2wd = l2wd; a.do_sync = do_sync;
  exchange_kernel<EX_DOWNPOUR><<<exch_grid(a.walk), 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_pulled_linear_push(const unsigned long long* wf_ptrs, const unsigned long long* wd_ptrs, int nranks, int rank,
                            const float* g, int N, int K, float scale, void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || !g || N % nranks || K % 4) return MPIT_E_ARG;
  PeerTable wf = make_table(wf_ptrs, 0, nranks, rank), wd = make_table(wd_ptrs, 0, nranks, rank);
  pulled_push_kernel<<<ew_grid((long long)N * (K / 4), 256, 2, 4), 256, 0, (cudaStream_t)stream>>>(wf, wd, g, N, K,
                                                                                                N / nranks, scale);
  MPIT_LAUNCH_CHECK();
  return 0;
}

static int ar_grid(long long n, int max_blocks) {
  long long need = ((n >> 2) + 511) / 512;
  if (need < 1) need = 1;
  if (need > max_blocks) need = max_blocks;
  return (int)need;
}

int mpit_bcast(const unsigned long long* buf_ptrs, unsigned long long buf_mc, const unsigned long long* sig_ptrs,
               int nranks, int rank, int root, long long n, int max_blocks, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || root < 0 || root >= nranks) return MPIT_E_ARG;
  if (max_blocks <= 0 || max_blocks > MPIT_BARRIER_BLOCKS) max_blocks = MPIT_BARRIER_BLOCKS;
  ArArgs a{};
  a.out = make_table(buf_ptrs, buf_mc, nranks, rank);
  a.sig = make_sig(sig_ptrs, nranks, rank);
  a.n = n;
  bcast_kernel<<<ar_grid(n, max_blocks), 512, 0, (cudaStream_t)stream>>>(a, root);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_device_barrier(const unsigned long long* sig_ptrs, int nranks, int rank, void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(make_sig(sig_ptrs, nranks, rank));
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_peer_rma(int op, float* remote, float* local, long long n, float scale, void* stream) {
  if (op < RMA_PUT || op > RMA_ACC || !remote || !local) return MPIT_E_ARG;
  if (n <= 0) return 0;
  peer_rma_kernel<<<ew_grid(n, 256, 8), 256, 0, (cudaStream_t)stream>>>(op, remote, local, n, scale);
  MPIT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
