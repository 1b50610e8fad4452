// Bandwidth-path exchange kernels: the same pull / push / elastic semantics as comm.cu, but moved by
// the bulk-copy engine (TMA) in 16 KiB pieces instead of per-thread 16-byte accesses.
//
// Why: over NVLink a per-thread `ld`/`red` holds a load-store tracking entry for ~2 us, and an SM
// runs out of entries long before the links are full -- the fused elastic pass (one remote load and
// one remote red per element) ended up at the *sum* of the pull and push times.  Here each CTA keeps a
// few 16 KiB transactions in flight per direction:
//     cp.async.bulk.shared::cluster.global        peer centre shard -> shared memory   (mbarrier)
//     cp.reduce.async.bulk.global.shared::cta.add.f32   shared memory -> += peer centre shard
// and the SM only does the elementwise math on shared memory.  Used for large, 16-byte-aligned
// problems (n % (4*S) == 0); the per-thread kernels in comm.cu remain the latency path and the
// general fallback.
//
// Semantics: pull pclient.lua:74-84, push pserver.lua:91, elastic optim-eamsgd.lua:58-65
// (paths under /root/reference/asyncsgd).
#include "mpit_b200.h"
#include "comm.cuh"
#include "umma.cuh"

namespace mpit {
using umma::smem_u32;
using umma::mbar_init;
using umma::mbar_expect_tx;
using umma::mbar_try_wait;
using umma::fence_barrier_init;

constexpr int kBulkFloats = 4096;            // 16 KiB per transaction
constexpr int kBulkBytes = kBulkFloats * 4;
constexpr int kBulkStages = 4;

enum BulkMode : int { BULK_PULL = 0, BULK_PUSH = 1, BULK_ELASTIC = 2 };

struct BulkArgs {
  PeerTable tab;
  Partition part;
  long long cps;        // chunks per shard
  float* x;             // pull: destination; elastic: local parameters (read)
  float* d;             // elastic: difference out; push: source
  float alpha, scale;
  int zero_src;
};

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_reduce_add_f32(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// chunk v -> (owner shard, element offset inside the shard, length); owners are visited round-robin
// starting after the caller's own rank so that concurrent CTAs and ranks spread over all links.
__device__ __forceinline__ bool bulk_chunk(const BulkArgs& a, long long v, int& s, long long& j0, int& len) {
  int S = a.part.S;
  s = (int)((v % S + a.tab.rank + 1) % S);
  j0 = (v / S) * kBulkFloats;
  long long rem = a.part.size(s) - j0;
  if (rem <= 0) return false;
  len = rem < kBulkFloats ? (int)rem : kBulkFloats;
  return true;
}

template <int MODE>
__global__ void __launch_bounds__(256) bulk_exchange_kernel(BulkArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  float* cbuf = reinterpret_cast<float*>(smem);                               // [stages][4096] incoming
  float* obuf = cbuf + kBulkStages * kBulkFloats;                             // [stages][4096] outgoing
  float* xbuf = obuf + kBulkStages * kBulkFloats;                             // [stages][4096] local x (elastic)
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + (MODE == BULK_ELASTIC ? kBulkStages * kBulkFloats : 0));
  const long long total = a.cps * a.part.S;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int i = 0; i < kBulkStages; ++i) mbar_init(&full[i], 1);
    fence_barrier_init();
  }
  __syncthreads();

  // what a stage's "incoming" transaction is: pull/elastic = the peer's centre chunk, push = nothing
  auto issue = [&](long long v, int stage) {
    int s, len;
    long long j0;
    if (!bulk_chunk(a, v, s, j0, len)) return;
    const long long i0 = a.part.begin(s) + j0;
    const uint32_t bytes = (uint32_t)len * 4;
    if (MODE == BULK_PULL) {
      mbar_expect_tx(&full[stage], bytes);
      bulk_load(cbuf + stage * kBulkFloats, a.tab.peer[s] + j0, bytes, &full[stage]);
    } else if (MODE == BULK_PUSH) {          // incoming = the local source chunk
      mbar_expect_tx(&full[stage], bytes);
      bulk_load(cbuf + stage * kBulkFloats, a.d + i0, bytes, &full[stage]);
    } else {                                 // elastic: the peer's centre chunk and the local x chunk
      mbar_expect_tx(&full[stage], 2 * bytes);
      bulk_load(cbuf + stage * kBulkFloats, a.tab.peer[s] + j0, bytes, &full[stage]);
      bulk_load(xbuf + stage * kBulkFloats, a.x + i0, bytes, &full[stage]);
    }
  };

  long long v0 = blockIdx.x;
  if (tid == 0)
    for (int i = 0; i < kBulkStages; ++i)
      if (v0 + (long long)i * gridDim.x < total) issue(v0 + (long long)i * gridDim.x, i);

  int stage = 0;
  uint32_t phase = 0;
  for (long long v = v0; v < total; v += gridDim.x) {
    int s, len;
    long long j0;
    const bool live = bulk_chunk(a, v, s, j0, len);
    if (live) {
      const long long i0 = a.part.begin(s) + j0;
      float* c = cbuf + stage * kBulkFloats;
      float* o = obuf + stage * kBulkFloats;
      // the outgoing buffer of this stage was handed to the copy engine kBulkStages iterations ago
      if (tid == 0) bulk_wait_read<kBulkStages - 1>();
      while (!mbar_try_wait(&full[stage], phase)) {}
      __syncthreads();
      if (MODE == BULK_PULL) {
        // centre chunk is in shared memory: hand it straight back to the copy engine, local store
        if (tid == 0) {
          fence_proxy_async();
          bulk_store(a.x + i0, c, (uint32_t)len * 4);
          bulk_commit();
        }
      } else if (MODE == BULK_PUSH) {
        for (int e = tid * 4; e < len; e += 256 * 4) {
          float4 g = *reinterpret_cast<const float4*>(c + e);
          if (a.scale != 1.f) { g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale; }
          *reinterpret_cast<float4*>(o + e) = g;
          if (a.zero_src) st4(a.d + i0 + e, make_float4(0.f, 0.f, 0.f, 0.f));
        }
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
          bulk_reduce_add_f32(a.tab.peer[s] + j0, o, (uint32_t)len * 4);   // centre += g on the owner
          bulk_commit();
        }
      } else {  // BULK_ELASTIC: d = alpha (x - centre); centre += d
        const float* xs = xbuf + stage * kBulkFloats;
        for (int e = tid * 4; e < len; e += 256 * 4) {
          float4 xv = *reinterpret_cast<const float4*>(xs + e);
          float4 cv = *reinterpret_cast<const float4*>(c + e);
          float4 dv;
          dv.x = __fmul_rn(a.alpha, __fsub_rn(xv.x, cv.x));
          dv.y = __fmul_rn(a.alpha, __fsub_rn(xv.y, cv.y));
          dv.z = __fmul_rn(a.alpha, __fsub_rn(xv.z, cv.z));
          dv.w = __fmul_rn(a.alpha, __fsub_rn(xv.w, cv.w));
          *reinterpret_cast<float4*>(o + e) = dv;
        }
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
          bulk_reduce_add_f32(a.tab.peer[s] + j0, o, (uint32_t)len * 4);   // centre += d on the owner
          bulk_store(a.d + i0, o, (uint32_t)len * 4);                      // and keep d for the apply step
          bulk_commit();
        }
      }
    }
    // refill this stage with the chunk kBulkStages iterations ahead
    long long nxt = v + (long long)kBulkStages * gridDim.x;
    if (MODE == BULK_PULL) {
      if (tid == 0 && nxt < total) {
        bulk_wait_read<0>();          // the store just issued must have drained this buffer
        issue(nxt, stage);
      }
    } else {
      // every thread finished reading this stage's incoming buffers before the barrier above
      if (tid == 0 && nxt < total) issue(nxt, stage);
    }
    if (++stage == kBulkStages) { stage = 0; phase ^= 1; }
  }
  if (tid == 0) bulk_wait_all();
}

static BulkArgs make_bulk(const unsigned long long* ptrs, int nranks, int rank, long long n) {
  BulkArgs a{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) a.tab.peer[i] = i < nranks ? reinterpret_cast<float*>(ptrs[i]) : nullptr;
  a.tab.nranks = nranks; a.tab.rank = rank;
  a.part.n = n; a.part.S = nranks; a.part.base = n / nranks;
  long long maxsz = a.part.size(nranks - 1);
  a.cps = (maxsz + kBulkFloats - 1) / kBulkFloats;
  return a;
}

template <int MODE>
static int launch_bulk(const BulkArgs& a, cudaStream_t st) {
  constexpr int smem = (MODE == BULK_ELASTIC ? 3 : 2) * kBulkStages * kBulkBytes + 64;
  static unsigned long long done = 0;   // bit per device
  MPIT_CUDA_TRY(set_smem_limit_per_device(bulk_exchange_kernel<MODE>, smem, done));
  long long total = a.cps * a.part.S;
  int grid = (int)(total < MPIT_NUM_SMS ? total : MPIT_NUM_SMS);
  bulk_exchange_kernel<MODE><<<grid, 256, smem, st>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

static bool bulk_ok(long long n, int nranks, const void* p0, const void* p1) {
  return n >= (long long)nranks * kBulkFloats && n % (4ll * nranks) == 0 &&
         (reinterpret_cast<uintptr_t>(p0) & 15) == 0 && (reinterpret_cast<uintptr_t>(p1) & 15) == 0;
}

}  // namespace mpit

using namespace mpit;

extern "C" {

// 1 when the bulk path can serve this problem (large enough, every shard a multiple of 16 bytes).
int mpit_bulk_supported(long long n, int nranks) { return bulk_ok(n, nranks, nullptr, nullptr) ? 1 : 0; }

int mpit_param_pull_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, float* dst, long long n,
                         void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  if (!bulk_ok(n, nranks, dst, nullptr)) return MPIT_E_UNSUPPORTED;
  BulkArgs a = make_bulk(shard_ptrs, nranks, rank, n);
  a.x = dst;
  return launch_bulk<BULK_PULL>(a, (cudaStream_t)stream);
}

int mpit_push_add_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, float* src, long long n,
                       float scale, int zero_src, void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  if (!bulk_ok(n, nranks, src, nullptr)) return MPIT_E_UNSUPPORTED;
  BulkArgs a = make_bulk(shard_ptrs, nranks, rank, n);
  a.d = src; a.scale = scale; a.zero_src = zero_src;
  return launch_bulk<BULK_PUSH>(a, (cudaStream_t)stream);
}

int mpit_elastic_exchange_bulk(const unsigned long long* shard_ptrs, int nranks, int rank, const float* x, float* d,
                               long long n, float alpha, void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  if (!bulk_ok(n, nranks, x, d)) return MPIT_E_UNSUPPORTED;
  BulkArgs a = make_bulk(shard_ptrs, nranks, rank, n);
  a.x = const_cast<float*>(x); a.d = d; a.alpha = alpha;
  return launch_bulk<BULK_ELASTIC>(a, (cudaStream_t)stream);
}

}  // extern "C"
