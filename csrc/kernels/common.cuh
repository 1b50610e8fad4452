// Shared device/host helpers for the mpit_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define MPIT_NUM_SMS 148  // B200: 2 dies x 74 SMs

#define MPIT_LAUNCH_CHECK()                         \
  do {                                              \
    cudaError_t e__ = cudaGetLastError();           \
    if (e__ != cudaSuccess) return (int)e__;        \
  } while (0)

#define MPIT_CUDA_TRY(expr)                         \
  do {                                              \
    cudaError_t e__ = (expr);                       \
    if (e__ != cudaSuccess) return (int)e__;        \
  } while (0)

namespace mpit {

// cudaFuncSetAttribute is per device, not per process: a process that drives several GPUs has to raise the
// dynamic shared-memory limit on each of them.  One 64-bit mask per kernel instantiation remembers the devices
// already done (the call itself is idempotent, so a lost race costs one redundant call).
template <typename K>
static inline cudaError_t set_smem_limit_per_device(K kernel, int bytes, unsigned long long& done_mask) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const unsigned long long bit = 1ull << (dev & 63);
  if (__atomic_load_n(&done_mask, __ATOMIC_ACQUIRE) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) __atomic_fetch_or(&done_mask, bit, __ATOMIC_RELEASE);
  return e;
}

// Grid sizing for streaming elementwise kernels: enough CTAs for >= 2 waves on
// 148 SMs but never more than the work needs.
static inline int ew_grid(long long n_items, int threads, int items_per_thread = 4, int max_ctas_per_sm = 8) {
  long long per_cta = (long long)threads * items_per_thread;
  long long need = (n_items + per_cta - 1) / per_cta;
  long long cap = (long long)MPIT_NUM_SMS * max_ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// ---- programmatic dependent launch (PDL) -------------------------------------------------------
// A training step is ~20 short kernels; each spends its first 1.5-2 us on set-up (barrier init, TMEM
// allocation, descriptor fetch) before it touches data.  Launched with the programmatic-serialization
// attribute a kernel may start while its predecessor is still running: it does its set-up, then
// `pdl_wait()` blocks until the predecessor grid has completed and its writes are visible.
// `pdl_trigger()` (first statement of every participating kernel) lets the *next* kernel start early.
// Rule: a kernel launched through launch_k() must execute pdl_wait() in every thread before its first
// global-memory access that could alias anything an earlier kernel reads or writes.
extern int g_pdl_enabled;   // defined in optim.cu; mpit_set_pdl()
extern int g_x3_mode;       // defined in optim.cu; mpit_set_tf32x3(): 0 plain TF32, 1 / 2 3xTF32 (umma.cuh)

__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                   Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl_enabled ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, const float4& v) { *reinterpret_cast<float4*>(p) = v; }
// 16-byte float add into this GPU's memory, no return value (one L2 atomic unit operation per quad)
__device__ __forceinline__ void red_add4(float* p, const float4& v) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// Streaming (read-once) 128-bit load that does not allocate in L1.
__device__ __forceinline__ float4 ld4_stream(const float* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}

// Non-fused fp32 arithmetic: the optimizer kernels promise the same rounding as a
// plain CPU fp32 restatement (one rounding per op, no FMA contraction), so that a
// single-worker trajectory is bit-reproducible against the NumPy/torch oracle.
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }

// Philox4x32-10 counter RNG (stateless; used for dropout masks and sampling).
struct Philox {
  __device__ __forceinline__ static uint4 round(uint4 c, uint2 k) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    return make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
  }
  __device__ __forceinline__ static uint4 gen(uint64_t seed, uint64_t ctr, uint32_t stream) {
    uint2 k = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
    uint4 c = make_uint4((uint32_t)ctr, (uint32_t)(ctr >> 32), stream, 0u);
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      c = round(c, k);
      k.x += 0x9E3779B9u;
      k.y += 0xBB67AE85u;
    }
    return c;
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace mpit
