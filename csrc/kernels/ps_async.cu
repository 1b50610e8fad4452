// Asynchronous parameter server with server-side update rules on the device path (sm_100a).
//
// Reference: the BiCNN servers apply rmsprop / adam / adamax / adagrad / adadelta to their shard once per
// incoming client gradient, in the order the gradients arrive, without ever waiting for a client that has
// not pushed (/root/reference/BiCNN/pserver.lua:115-205; per-client receive buffers storage.g[crank] :49-58);
// the clients push a raw (optionally su-step accumulated) gradient and pull the parameters
// (/root/reference/BiCNN/optim-adam.lua:24-57 and its adamax/adagrad/adadelta/rmsprop siblings).
//
// Here every GPU is a client and the owner of shard `rank` (parameters + rule state).  Three kernels:
//
//   ps_push   (client)  gradient prologue (packed->flat gather, L1/L2/clamp, su-step accumulation), then
//                       st.relaxed.sys.v4 of the shard pieces into THIS client's staging slot on every owner;
//                       the last CTA draws an arrival ticket on each owner (atom.add.sys) and publishes it with
//                       st.release.sys.  Flow control: a one-warp kernel in front of it waits (bounded, on words in
//                       the client's OWN memory) until every owner has acknowledged the previous push.
//   ps_poll   (owner)   one warp: snapshot of the arrival words, ready slots sorted by arrival ticket, the
//                       update count and the per-update step sizes (adam bias correction etc.) computed on the
//                       device -- nothing about a drain depends on host state, so it can sit in a CUDA graph.
//   ps_apply  (owner)   one pass over the shard: state in registers, the ready gradients applied one client at
//                       a time in arrival order (the rules are elementwise, so the per-element chain is the
//                       server's sequential order), slots acknowledged by the last CTA.
//
// The pull is the ordinary one-sided parameter pull (comm.cu); it never waits.  An owner drains either inline
// (its own stream, once per step) or from a native server thread (mpit_ps_server_start) that relaunches
// poll+apply on a high-priority stream, which keeps a slow worker from delaying the gradients of the others.
#include <atomic>
#include <chrono>
#include <thread>

#include "mpit_b200.h"
#include "comm.cuh"
#include "optim_math.cuh"

namespace mpit {

// ---- control block: one per rank, in symmetric memory (uint32 words) -----------------------------
enum PsWord : int {
  PS_ARR = 0,          // [8]  owner side: arrival ticket + 1 of client c's newest gradient (written by c)
  PS_ARRIVALS = 8,     //      owner side: arrival counter (clients atom.add)
  PS_DONE = 16,        // [8]  owner side: PS_ARR value of client c already applied (local)
  PS_ACKED = 24,       // [8]  owner side: gradients of client c applied so far (local)
  PS_APPLIED = 32,     //      owner side: total updates applied to this shard (adam_t / pversion)
  PS_APPLY_TICKET = 50,//      CTA tickets of the running apply kernel
  PS_NTODO = 56,       //      snapshot of the running drain
  PS_TBASE = 57,
  PS_TODO_CLIENT = 64, // [8]
  PS_TODO_ARR = 72,    // [8]
  PS_TODO_STEP = 80,   // [8] float bits
  // client side, in a 128-byte line of its own (the flow-control wait polls it while the owner side is busy)
  PS_ACK = 96,         // [8]  pushes of mine that owner s has applied (written by s)
  PS_NPUSH = 104,      //      pushes issued
  PS_PUSH_TICKET = 105,//      CTA tickets of the running push kernel
  PS_ERR = 106,        //      a bounded wait gave up (1 + owner)
  PS_MAXWAIT = 107,    //      longest flow-control wait of a push so far, in units of 1024 clocks
  PS_WORDS = 128,
};

struct PsCtlTable {
  uint32_t* peer[MPIT_MAX_RANKS];
  int nranks, rank;
};

struct PsRule {
  int rule;                 // RuleId
  float a, b, c, d, eps;    // constant part of RuleHyper
  double lr, beta1, beta2, lrd;
  int step_div;             // adam: bias-correction exponent floor(t/step_div)+1 (pserver.lua:151-152); 0 = t
};

// Step size of update number t (1-based), in double and narrowed once like the host restatement
// (mpit_b200/ops/optim_ops.py rule_hyper).
__device__ float ps_step_size(const PsRule& r, unsigned t) {
  switch (r.rule) {
    case RULE_ADAM: {
      double k = r.step_div > 0 ? (double)(t / (unsigned)r.step_div + 1u) : (double)t;
      return (float)(-(r.lr * sqrt(1.0 - pow(r.beta2, k)) / (1.0 - pow(r.beta1, k))));
    }
    case RULE_ADAMAX:
      return (float)(-(r.lr / (1.0 - pow(r.beta1, (double)t))));
    case RULE_ADAGRAD:
      return (float)(-(r.lr / (1.0 + (double)(t - 1u) * r.lrd)));
    default:
      return (float)(-r.lr);
  }
}

__device__ __forceinline__ uint32_t atom_add_sys_u32(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.relaxed.sys.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

// ---- client: push ----------------------------------------------------------------------------------
struct PsPushArgs {
  PeerTable slot;       // peer[s] = base of MY staging slot inside owner s
  PsCtlTable ctl;
  Partition part;
  long long cps;        // chunks per shard (walk, see below)
  const float* gflat;   // gradient in flat order (elements below map_base, or all of them when gmap == null)
  const float* gpk;     // packed gradient of the elements from map_base on, read through gmap
  const int* gmap;
  long long map_base;
  const float* x;       // parameters (L1 / L2 terms)
  float* acc;           // su-step accumulator (null: none)
  const int* active;    // device flag: 0 = this step produced no gradient (no regularisation either); null = 1
  float* reg_loss;      // [2] += l1*|x|_1, l2*|x|^2/2 of this step (null: not wanted)
  long long n;
  float l1, l2, clip, gscale;
  int do_push;          // 0: accumulate only
};

constexpr long long kPsChunkQuads = 1024;   // 16 KiB runs, dealt round-robin over the owners starting after me

// Flow control, as a kernel of its own in front of the push: one warp waits (bounded, with back-off, on words in
// this client's OWN memory) until every owner has applied the previous push, i.e. until the staging slots are free.
// While it waits the rest of the GPU stays available to the owner-side kernels of this very GPU.
__global__ void __launch_bounds__(32) ps_wait_kernel(uint32_t* mine, int nranks) {
  const int t = threadIdx.x;
  if (t < nranks) {
    const uint32_t want = *reinterpret_cast<volatile uint32_t*>(mine + PS_NPUSH);
    const long long t0 = clock64();
    unsigned backoff = 32;
    while ((int32_t)(ld_acquire_sys_u32(mine + PS_ACK + t) - want) < 0) {
      if (clock64() - t0 > MPIT_SPIN_TIMEOUT_CYCLES) { atomicExch(mine + PS_ERR, 1u + (uint32_t)t); break; }
      __nanosleep(backoff);
      if (backoff < 1024) backoff <<= 1;
    }
    atomicMax(mine + PS_MAXWAIT, (uint32_t)((clock64() - t0) >> 10));
  }
}

__global__ void __launch_bounds__(256) ps_push_kernel(PsPushArgs a) {
  const int R = a.ctl.nranks, me = a.ctl.rank, t = threadIdx.x;
  uint32_t* mine = a.ctl.peer[me];
  const bool act = a.active ? (*a.active != 0) : true;
  float rl1 = 0.f, rl2 = 0.f;
  const long long total = a.cps * R * kPsChunkQuads;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + t; q < total; q += stride) {
    const long long c = q / kPsChunkQuads, within = q - c * kPsChunkQuads;
    const unsigned c32 = (unsigned)c, k = c32 / (unsigned)R;
    int s = (int)(c32 - k * R) + me + 1;
    if (s >= R) s -= R;
    if (s >= R) s -= R;
    const long long j0 = ((long long)k * kPsChunkQuads + within) * 4;
    const long long rem = a.part.size(s) - j0;
    if (rem <= 0) continue;
    const int cnt = rem >= 4 ? 4 : (int)rem;
    const long long i0 = a.part.begin(s) + j0;
    float g[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (e >= cnt) break;
      const long long i = i0 + e;
      float v = (a.gmap && i >= a.map_base) ? a.gpk[a.gmap[i - a.map_base]] : a.gflat[i];
      if (act && (a.l1 != 0.f || a.l2 != 0.f || a.clip > 0.f)) {
        const float p = (a.l1 != 0.f || a.l2 != 0.f) ? a.x[i] : 0.f;
        v = grad_reg_clip(v, p, a.l1, a.l2, a.clip);
        if (a.reg_loss) { rl1 += fabsf(p); rl2 += p * p; }
      }
      if (a.gscale != 1.f) v = mul_rn(v, a.gscale);     // e.g. -lr for Downpour (optim-downpour.lua:24)
      if (a.acc) {
        v = add_rn(a.acc[i], v);
        a.acc[i] = a.do_push ? 0.f : v;
      }
      g[e] = v;
    }
    if (a.do_push) {
      float* dst = a.slot.peer[s] + j0;
      if (cnt == 4) st_sys4(dst, make_float4(g[0], g[1], g[2], g[3]));
      else
        for (int e = 0; e < cnt; ++e) st_sys(dst + e, g[e]);
    }
  }
  if (a.reg_loss && act && (a.l1 != 0.f || a.l2 != 0.f)) {
    rl1 = warp_sum(rl1); rl2 = warp_sum(rl2);
    if ((t & 31) == 0) {
      if (a.l1 != 0.f) atomicAdd(a.reg_loss, a.l1 * rl1);
      if (a.l2 != 0.f) atomicAdd(a.reg_loss + 1, 0.5f * a.l2 * rl2);
    }
  }
  if (!a.do_push) return;
  // publish: payload of every CTA before the arrival words
  __shared__ uint32_t s_last;
  __syncthreads();
  if (t == 0) {
    __threadfence_system();
    uint32_t ticket = atomicAdd(mine + PS_PUSH_TICKET, 1u);
    __threadfence();
    s_last = ticket == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;
  if (t < R) {
    __threadfence_system();
    int s = t + me + 1;                      // start with the neighbour, like the payload walk
    if (s >= R) s -= R;
    if (s >= R) s -= R;
    uint32_t* owner = a.ctl.peer[s];
    const uint32_t tk = atom_add_sys_u32(owner + PS_ARRIVALS, 1u);
    st_release_sys_u32(owner + PS_ARR + me, tk + 1u);
  }
  __syncthreads();
  if (t == 0) {
    mine[PS_PUSH_TICKET] = 0u;
    mine[PS_NPUSH] = mine[PS_NPUSH] + 1u;
    __threadfence();
  }
}

// ---- owner: poll -----------------------------------------------------------------------------------
__global__ void __launch_bounds__(32) ps_poll_kernel(uint32_t* ctl, int nranks, PsRule rule, int max_take) {
  pdl_trigger();
  pdl_wait();
  const int c = threadIdx.x;
  uint32_t a = 0, d = 0;
  bool ready = false;
  if (c < nranks) {
    a = ld_acquire_sys_u32(ctl + PS_ARR + c);
    d = ctl[PS_DONE + c];
    ready = a != d;
  }
  const unsigned mask = __ballot_sync(0xffffffffu, ready);
  // arrival order: rank of my ticket among the ready ones (tickets are unique per owner)
  int pos = 0;
  for (int o = 0; o < nranks; ++o) {
    const uint32_t ao = __shfl_sync(0xffffffffu, a, o);
    if (((mask >> o) & 1u) && o != c && (int32_t)(ao - a) < 0) ++pos;
  }
  int ntodo = __popc(mask);
  if (max_take > 0 && ntodo > max_take) ntodo = max_take;
  const uint32_t tbase = ctl[PS_APPLIED];
  if (ready && pos < ntodo) {
    ctl[PS_TODO_CLIENT + pos] = (uint32_t)c;
    ctl[PS_TODO_ARR + pos] = a;
  }
  if (c < ntodo) ctl[PS_TODO_STEP + c] = __float_as_uint(ps_step_size(rule, tbase + (uint32_t)c + 1u));
  if (c == 0) {
    ctl[PS_NTODO] = (uint32_t)ntodo;
    ctl[PS_TBASE] = tbase;
    ctl[PS_APPLIED] = tbase + (uint32_t)ntodo;
  }
}

// ---- owner: apply ----------------------------------------------------------------------------------
struct PsApplyArgs {
  PsCtlTable ctl;
  float* p;             // my shard of the centre
  float* s0;
  float* s1;
  float* s2;
  const float* slots;   // [nranks][cap] staging slots in my memory
  long long cap;        // floats per slot
  long long size;       // elements of my shard
  PsRule rule;
};

template <int RULE>
__global__ void __launch_bounds__(256) ps_apply_kernel(PsApplyArgs a) {
  pdl_wait();
  const int me = a.ctl.rank, t = threadIdx.x;
  uint32_t* ctl = a.ctl.peer[me];
  __shared__ uint32_t s_n, s_client[MPIT_MAX_RANKS], s_arr[MPIT_MAX_RANKS];
  __shared__ float s_step[MPIT_MAX_RANKS];
  __shared__ uint32_t s_last;
  if (t == 0) s_n = ctl[PS_NTODO];
  if (t < MPIT_MAX_RANKS) {
    s_client[t] = ctl[PS_TODO_CLIENT + t];
    s_arr[t] = ctl[PS_TODO_ARR + t];
    s_step[t] = __uint_as_float(ctl[PS_TODO_STEP + t]);
  }
  __syncthreads();
  const int ntodo = (int)s_n;
  if (ntodo == 0) return;
  constexpr bool kS0 = RULE != RULE_ADD;
  constexpr bool kS1 = RULE == RULE_RMSPROP || RULE == RULE_ADAM || RULE == RULE_ADAMAX || RULE == RULE_ADADELTA;
  constexpr bool kS2 = RULE == RULE_RMSPROP;
  RuleHyper h{a.rule.a, a.rule.b, a.rule.c, a.rule.d, a.rule.eps, 0.f};
  const long long nq = (a.size + 3) >> 2;       // shard buffers are padded to whole quads
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long q = (long long)blockIdx.x * blockDim.x + t; q < nq; q += stride) {
    const long long i = 4 * q;
    float4 g[MPIT_MAX_RANKS];
#pragma unroll
    for (int j = 0; j < MPIT_MAX_RANKS; ++j)
      if (j < ntodo) g[j] = ld_sys4(a.slots + (long long)s_client[j] * a.cap + i);     // all gradients in flight
    float4 p = ld4(a.p + i);
    float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 s0 = kS0 ? ld4(a.s0 + i) : z, s1 = kS1 ? ld4(a.s1 + i) : z, s2 = kS2 ? ld4(a.s2 + i) : z;
#pragma unroll
    for (int j = 0; j < MPIT_MAX_RANKS; ++j) {
      if (j < ntodo) {
        h.step = s_step[j];
        rule_apply<RULE>(p.x, g[j].x, s0.x, s1.x, s2.x, h);
        rule_apply<RULE>(p.y, g[j].y, s0.y, s1.y, s2.y, h);
        rule_apply<RULE>(p.z, g[j].z, s0.z, s1.z, s2.z, h);
        rule_apply<RULE>(p.w, g[j].w, s0.w, s1.w, s2.w, h);
      }
    }
    st4(a.p + i, p);
    if (kS0) st4(a.s0 + i, s0);
    if (kS1) st4(a.s1 + i, s1);
    if (kS2) st4(a.s2 + i, s2);
  }
  __syncthreads();
  if (t == 0) {
    __threadfence_system();
    uint32_t ticket = atomicAdd(ctl + PS_APPLY_TICKET, 1u);
    __threadfence();
    s_last = ticket == gridDim.x - 1 ? 1u : 0u;
  }
  __syncthreads();
  if (!s_last) return;
  if (t < ntodo) {
    const uint32_t c = s_client[t];
    ctl[PS_DONE + c] = s_arr[t];
    const uint32_t cnt = ctl[PS_ACKED + c] + 1u;
    ctl[PS_ACKED + c] = cnt;
    __threadfence_system();
    st_release_sys_u32(a.ctl.peer[c] + PS_ACK + me, cnt);    // the client may overwrite its slot
  }
  __syncthreads();
  if (t == 0) {
    ctl[PS_APPLY_TICKET] = 0u;
    ctl[PS_NTODO] = 0u;
    __threadfence();
  }
}

static PsCtlTable make_ctl(const unsigned long long* ptrs, int nranks, int rank) {
  PsCtlTable t{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) t.peer[i] = i < nranks ? reinterpret_cast<uint32_t*>(ptrs[i]) : nullptr;
  t.nranks = nranks;
  t.rank = rank;
  return t;
}

static int fill_rule(PsRule& r, int rule, const double* hyper10) {
  if (rule < RULE_ADD || rule > RULE_ADADELTA || !hyper10) return MPIT_E_ARG;
  r.rule = rule;
  r.a = (float)hyper10[0]; r.b = (float)hyper10[1]; r.c = (float)hyper10[2]; r.d = (float)hyper10[3];
  r.eps = (float)hyper10[4];
  r.lr = hyper10[5]; r.beta1 = hyper10[6]; r.beta2 = hyper10[7]; r.lrd = hyper10[8];
  r.step_div = (int)hyper10[9];
  return 0;
}

static int launch_drain(const PsCtlTable& ctl, const PsRule& rule, float* p, float* s0, float* s1, float* s2,
                        const float* slots, long long cap, long long size, int max_take, cudaStream_t st) {
  MPIT_CUDA_TRY(launch_k(ps_poll_kernel, dim3(1), dim3(32), 0, st, ctl.peer[ctl.rank], ctl.nranks, rule, max_take));
  PsApplyArgs a{};
  a.ctl = ctl; a.p = p; a.s0 = s0; a.s1 = s1; a.s2 = s2; a.slots = slots; a.cap = cap; a.size = size; a.rule = rule;
  const int grid = ew_grid((size + 3) / 4, 256, 2, 4);
  switch (rule.rule) {
    case RULE_ADD: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_ADD>, dim3(grid), dim3(256), 0, st, a)); break;
    case RULE_RMSPROP: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_RMSPROP>, dim3(grid), dim3(256), 0, st, a)); break;
    case RULE_ADAM: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_ADAM>, dim3(grid), dim3(256), 0, st, a)); break;
    case RULE_ADAMAX: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_ADAMAX>, dim3(grid), dim3(256), 0, st, a)); break;
    case RULE_ADAGRAD: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_ADAGRAD>, dim3(grid), dim3(256), 0, st, a)); break;
    case RULE_ADADELTA: MPIT_CUDA_TRY(launch_k(ps_apply_kernel<RULE_ADADELTA>, dim3(grid), dim3(256), 0, st, a)); break;
    default: return MPIT_E_ARG;
  }
  return 0;
}

// ---- native server thread -----------------------------------------------------------------------------
struct PsServer {
  std::thread th;
  std::atomic<int> stop{0};
  std::atomic<long long> rounds{0};
  std::atomic<int> error{0};
};

}  // namespace mpit

using namespace mpit;

extern "C" {

int mpit_ps_ctl_words(void) { return PS_WORDS; }

/* counters for tests / monitoring: which = 0 applied updates on my shard, 1 pushes issued by me, 2 error word,
 * 3 arrivals seen on my shard, 4 longest flow-control wait of my pushes (units of 1024 clocks).  `ctl_local` is this rank's control block (DEVICE); the value is copied to *out
 * (HOST) with a synchronous cudaMemcpy. */
int mpit_ps_counter(const void* ctl_local, int which, unsigned* out) {
  static const int word[5] = {PS_APPLIED, PS_NPUSH, PS_ERR, PS_ARRIVALS, PS_MAXWAIT};
  if (which < 0 || which > 4 || !ctl_local || !out) return MPIT_E_ARG;
  MPIT_CUDA_TRY(cudaMemcpy(out, static_cast<const uint32_t*>(ctl_local) + word[which], 4, cudaMemcpyDeviceToHost));
  return 0;
}

int mpit_ps_push(const unsigned long long* slot_ptrs, const unsigned long long* ctl_ptrs, int nranks, int rank,
                 const float* gflat, const float* gpk, const int* gmap, long long map_base, const float* x, float* acc,
                 const int* active, float* reg_loss, long long n, float l1, float l2, float clip, float gscale,
                 int do_push, void* stream) {
  if (n <= 0) return 0;
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || !slot_ptrs || !ctl_ptrs) return MPIT_E_ARG;
  if ((l1 != 0.f || l2 != 0.f) && !x) return MPIT_E_ARG;
  if (!gflat && !(gmap && map_base == 0)) return MPIT_E_ARG;
  PsPushArgs a{};
  for (int i = 0; i < MPIT_MAX_RANKS; ++i) a.slot.peer[i] = i < nranks ? reinterpret_cast<float*>(slot_ptrs[i]) : nullptr;
  a.slot.nranks = nranks; a.slot.rank = rank;
  a.ctl = make_ctl(ctl_ptrs, nranks, rank);
  a.part.n = n; a.part.S = nranks; a.part.base = n / nranks;
  const long long max_quads = (a.part.size(nranks - 1) + 3) / 4;
  a.cps = (max_quads + kPsChunkQuads - 1) / kPsChunkQuads;
  if (a.cps < 1) a.cps = 1;
  a.gflat = gflat; a.gpk = gpk; a.gmap = gmap; a.map_base = map_base; a.x = x; a.acc = acc; a.active = active;
  a.reg_loss = reg_loss; a.n = n; a.l1 = l1; a.l2 = l2; a.clip = clip; a.gscale = gscale; a.do_push = do_push;
  const int grid = ew_grid(a.cps * nranks * kPsChunkQuads, 256, 1, 4);
  if (do_push) {
    ps_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a.ctl.peer[rank], nranks);
    MPIT_LAUNCH_CHECK();
  }
  ps_push_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_ps_drain(const unsigned long long* ctl_ptrs, int nranks, int rank, int rule, const double* hyper10, float* p,
                  float* s0, float* s1, float* s2, const float* slots, long long cap, long long size, int max_take,
                  void* stream) {
  if (nranks < 1 || nranks > MPIT_MAX_RANKS || !ctl_ptrs || !p || !slots || size <= 0) return MPIT_E_ARG;
  PsRule r{};
  int rc = fill_rule(r, rule, hyper10);
  if (rc) return rc;
  return launch_drain(make_ctl(ctl_ptrs, nranks, rank), r, p, s0, s1, s2, slots, cap, size, max_take,
                      (cudaStream_t)stream);
}

/* Starts a host thread that keeps draining this rank's staging slots on a high-priority stream of `dev`
 * until mpit_ps_server_stop; `interval_us` is the pause between two drains.  Returns an opaque handle in
 * *out_handle (HOST). */
int mpit_ps_server_start(int dev, const unsigned long long* ctl_ptrs, int nranks, int rank, int rule,
                         const double* hyper10, float* p, float* s0, float* s1, float* s2, const float* slots,
                         long long cap, long long size, int interval_us, void** out_handle) {
  if (!out_handle || nranks < 1 || nranks > MPIT_MAX_RANKS) return MPIT_E_ARG;
  PsRule r{};
  int rc = fill_rule(r, rule, hyper10);
  if (rc) return rc;
  PsCtlTable ctl = make_ctl(ctl_ptrs, nranks, rank);
  PsServer* srv = new PsServer();
  srv->th = std::thread([=]() {
    if (cudaSetDevice(dev) != cudaSuccess) { srv->error = 1; return; }
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStream_t st;
    if (cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, hi) != cudaSuccess) { srv->error = 2; return; }
    while (!srv->stop.load(std::memory_order_acquire)) {
      int e = launch_drain(ctl, r, p, s0, s1, s2, slots, cap, size, 0, st);
      if (e != 0 || cudaStreamSynchronize(st) != cudaSuccess) { srv->error = e ? e : 3; break; }
      srv->rounds.fetch_add(1, std::memory_order_relaxed);
      // pauses below the timer slack of sleep_for (~60 us) are spent polling the clock, like the reference's
      // servers busy-poll their coroutine queue (co_wait with usec = 0, /root/reference/init.lua:185-192)
      if (interval_us >= 100) {
        std::this_thread::sleep_for(std::chrono::microseconds(interval_us));
      } else if (interval_us > 0) {
        const auto until = std::chrono::steady_clock::now() + std::chrono::microseconds(interval_us);
        while (std::chrono::steady_clock::now() < until && !srv->stop.load(std::memory_order_relaxed)) std::this_thread::yield();
      }
    }
    // a last drain so that nothing pushed before the stop request is left behind
    if (!srv->error.load()) { launch_drain(ctl, r, p, s0, s1, s2, slots, cap, size, 0, st); cudaStreamSynchronize(st); }
    cudaStreamDestroy(st);
  });
  *out_handle = srv;
  return 0;
}

/* Stops and joins the server thread; *rounds_out (HOST, may be NULL) receives the number of drains it ran.
 * Returns the thread's error code (0 = none). */
int mpit_ps_server_stop(void* handle, long long* rounds_out) {
  if (!handle) return MPIT_E_ARG;
  PsServer* srv = static_cast<PsServer*>(handle);
  srv->stop.store(1, std::memory_order_release);
  if (srv->th.joinable()) srv->th.join();
  if (rounds_out) *rounds_out = srv->rounds.load();
  int e = srv->error.load();
  delete srv;
  return e;
}

}  // extern "C"
