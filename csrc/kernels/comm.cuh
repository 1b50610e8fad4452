// Device-side primitives of the NVLink transport: peer tables, system-scope loads/stores/
// reductions on peer-mapped pointers, multimem (NVLS) accessors and the cross-GPU block barrier.
//
// The reference moves every one of these bytes with MPI_Isend/Iprobe/Irecv/Test polled from Lua
// (/root/reference/init.lua:41-108); here a kernel issues ld/st/red directly on the owner's
// memory over NVLink 5 (one-sided Get / Put / Accumulate, cf. the wrapped-but-unused
// MPI_Get/MPI_Put/MPI_Accumulate at /root/reference/mpifuncs.c:1131,1656,9).
#pragma once
#include "common.cuh"

#define MPIT_MAX_RANKS 8
#define MPIT_BARRIER_BLOCKS 256            // signal-pad rows (max CTAs taking part in a barrier)
#define MPIT_SPIN_TIMEOUT_CYCLES (4000000000ll)  // ~2 s at 1.9 GHz: bounded spin, never hang a box

namespace mpit {

// Same-layout buffer on every rank: peer[r] is rank r's copy mapped into this process.
struct PeerTable {
  float* peer[MPIT_MAX_RANKS];
  float* mc;        // NVLS multicast mapping of the same buffer (null when unavailable)
  int nranks;
  int rank;
};

// Signal pads: pad[r] is rank r's pad, uint32 [MPIT_BARRIER_BLOCKS][MPIT_MAX_RANKS], plus an
// error word at the end that a timed-out spin sets.
struct SignalTable {
  uint32_t* pad[MPIT_MAX_RANKS];
  int nranks;
  int rank;
};
#define MPIT_SIGNAL_WORDS (MPIT_BARRIER_BLOCKS * MPIT_MAX_RANKS + 64)
#define MPIT_SIGNAL_ERR_WORD (MPIT_BARRIER_BLOCKS * MPIT_MAX_RANKS)
// Epoch-based synchronisation of the all-reduce family (allreduce.cu), in the spare words after the error word:
#define MPIT_SIG_START (MPIT_SIGNAL_ERR_WORD + 8)    // [MPIT_MAX_RANKS] "rank r has started launch e"
#define MPIT_SIG_DONE (MPIT_SIGNAL_ERR_WORD + 16)    // [MPIT_MAX_RANKS] "rank r has finished launch e"
#define MPIT_SIG_EPOCH (MPIT_SIGNAL_ERR_WORD + 24)   // launches completed on this pad (local only)
#define MPIT_SIG_TICKET (MPIT_SIGNAL_ERR_WORD + 25)  // CTAs of the running launch that have finished (local only)

// Reference partition (asyncsgd/pclient.lua:118-128): S contiguous shards of floor(N/S)
// elements, the last one extended to N.  0-based here (the reference's wire format is 1-based).
struct Partition {
  long long n;
  long long base;  // floor(n / S)
  int S;
  __host__ __device__ long long begin(int s) const { return base * s; }
  __host__ __device__ long long size(int s) const { return s == S - 1 ? n - base * (S - 1) : base; }
};

// ---- system-scope memory ops on (possibly peer) global addresses ---------------------------
__device__ __forceinline__ float ld_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_sys4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_sys(float* p, float v) {
  asm volatile("st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void st_sys4(float* p, const float4& v) {
  asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}
// Fire-and-forget accumulate into the owner's memory (MPI_Accumulate SUM).
__device__ __forceinline__ void red_add_sys(float* p, float v) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void red_add_sys4(float* p, const float4& v) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w) : "memory");
}

// Accumulate that also returns what was there before (MPI_Get_accumulate / Fetch_and_op SUM): one NVLink round trip
// gives the owner's value including nothing but what arrived before this addition.
__device__ __forceinline__ float atom_add_sys(float* p, float v) {
  float old;
  asm volatile("atom.relaxed.sys.global.add.f32 %0, [%1], %2;" : "=f"(old) : "l"(p), "f"(v) : "memory");
  return old;
}
__device__ __forceinline__ float4 atom_add_sys4(float* p, const float4& v) {
  float4 old;
  asm volatile("atom.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4], {%5,%6,%7,%8};"
               : "=f"(old.x), "=f"(old.y), "=f"(old.z), "=f"(old.w)
               : "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
  return old;
}

// ---- NVLS multimem accessors (valid only on a multicast mapping) ---------------------------
__device__ __forceinline__ float4 multimem_ld_reduce_add4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_red_add4(float* mc, const float4& v) {
  asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---- flags ----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cas_release_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.release.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t cas_acquire_sys(uint32_t* p, uint32_t cmp, uint32_t val) {
  uint32_t old;
  asm volatile("atom.acquire.sys.global.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(p), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Cross-GPU barrier between the CTAs with the same blockIdx on every rank (the analogue of
// MPI_Win_fence, /root/reference/mpifuncs.c:2367).  Thread r < nranks of the CTA signals rank r
// (0 -> 1 on r's pad[block][me]) and then consumes rank r's signal on its own pad (1 -> 0), so
// the pads reset themselves.  All ranks must launch the same grid.  Spins are bounded: on
// timeout the error word is set and the kernel carries on (the host checks it).
__device__ __forceinline__ void block_barrier_sys(const SignalTable& sig, int row) {
  __syncthreads();
  int t = threadIdx.x;
  if (t < sig.nranks) {
    __threadfence_system();
    uint32_t* remote = sig.pad[t] + row * MPIT_MAX_RANKS + sig.rank;
    uint32_t* local = sig.pad[sig.rank] + row * MPIT_MAX_RANKS + t;
    long long t0 = clock64();
    bool ok = true;
    while (cas_release_sys(remote, 0u, 1u) != 0u) {
      if (clock64() - t0 > MPIT_SPIN_TIMEOUT_CYCLES) { ok = false; break; }
    }
    while (ok && cas_acquire_sys(local, 1u, 0u) != 1u) {
      if (clock64() - t0 > MPIT_SPIN_TIMEOUT_CYCLES) { ok = false; break; }
    }
    if (!ok) atomicExch(sig.pad[sig.rank] + MPIT_SIGNAL_ERR_WORD, 1u);
    __threadfence_system();
  }
  __syncthreads();
}

}  // namespace mpit
