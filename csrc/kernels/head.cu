// Fused classifier head (sm_100a, CUDA cores): Linear(K -> ncls) + LogSoftMax + ClassNLLCriterion forward,
// and the backward of all three down to the gradient w.r.t. the PRE-activation of the layer below
// (asyncsgd/models/7-layers-cnn.lua:22-25, asyncsgd/goot.lua:40,200-206):
//
//   z      = x W^T + b                     logits                        [B][ld]   (ncls valid columns)
//   loss  += sum_r (logsumexp(z_r) - z_r[y_r]),   confusion[y][argmax] += 1
//   dz     = (softmax(z) - onehot(y)) * gscale                           [B][ld]   (feeds the weight gradient)
//   db_cls = column sums of dz                                           [ld]
//   dx     = (dz W) * (x > 0 ? in_scale : 0)                             [B][K]    gradient through the ReLU(+dropout)
//                                                                                  that produced x
//   db_in  = column sums of dx                                           [K]       bias gradient of the layer below
//
// With ten classes this is 2 MFLOP: as three launches (tensor-core GEMM, loss kernel, tensor-core data
// gradient) it cost 3 x ~6 us of launch / pipeline-fill latency on the critical path of a 160 us step.  Here
// every CTA keeps W in shared memory and a warp owns a row end to end (x stays in registers between the forward
// dot products and the backward mask).  Rows are spread over up to 64 CTAs (a single CTA needed ~15 us for 128
// rows); each CTA leaves its partial sums in a scratch row and the last one to finish (ticket counter) adds the
// rows up in a fixed order -- no atomics on floats, bit-reproducible results.
#include "mpit_b200.h"
#include "common.cuh"

namespace mpit {

constexpr int kHeadRows = 2;   // rows per warp in flight (their loads are issued together)
constexpr int kHeadWarps = 8;
constexpr int kHeadMaxCtas = 64;

struct HeadArgs {
  const float* x;        // [B][K]
  const float* w;        // [>= ncls][K]
  const float* bias;     // [>= ncls] or null
  const int* labels;     // [B]
  float* logits;         // [B][ld]
  float* dlogits;        // [B][ld]  (null: forward only)
  float* dx;             // [B][K]   (null: forward only)
  float* loss_sum;
  int* confusion;        // [ncls][ncls] or null
  float* db_cls;         // [ld] or null
  float* db_in;          // [K] or null
  unsigned long long* rng_inc;
  float* scratch;        // [kHeadMaxCtas][K + 64] partial rows, then one int ticket (zero before the first launch)
  int B, K, ncls, ld, accumulate;
  float gscale, in_scale;
};

template <int KPL, int NC>   // K / 32: x values per lane; NC: class count padded to 16 or 32 (unrolled)
__global__ void __launch_bounds__(kHeadWarps * 32) classifier_head_kernel(HeadArgs a) {
  extern __shared__ float hsm[];
  float* ws = hsm;                        // [NC][K], rows >= ncls are zero
  float* cs = hsm + NC * a.K;             // [warps][K]  column-sum partials of dx
  __shared__ float lpart[kHeadWarps];
  __shared__ float bpart[kHeadWarps][33];
  __shared__ int last_flag;
  pdl_trigger();
  pdl_wait();
  if (a.rng_inc && blockIdx.x == 0 && threadIdx.x == 0) a.rng_inc[1] += 1ull;   // dropout step counter lives on the device
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const bool train = a.dlogits != nullptr;
  // W is fetched into registers now (16-byte loads; K % 32 == 0 and w is 16-byte aligned) and only parked in shared
  // memory after the first rows' x loads have been issued, so both global-load latencies overlap
  constexpr int WV = NC * KPL * 32 / (kHeadWarps * 32 * 4) > 0 ? NC * KPL * 32 / (kHeadWarps * 32 * 4) : 1;
  float4 wreg[WV];
#pragma unroll
  for (int u = 0; u < WV; ++u) {
    const int i = (u * kHeadWarps * 32 + threadIdx.x) * 4;
    wreg[u] = (i < a.ncls * a.K) ? ld4(a.w + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  auto park_w = [&]() {
#pragma unroll
    for (int u = 0; u < WV; ++u) {
      const int i = (u * kHeadWarps * 32 + threadIdx.x) * 4;
      if (i < NC * a.K) *reinterpret_cast<float4*>(ws + i) = wreg[u];
    }
    __syncthreads();
  };
  const float bz = (a.bias && lane < a.ncls) ? a.bias[lane] : 0.f;
  bool w_ready = false;

  float lsum = 0.f, bsum = 0.f;
  float csum[KPL];
#pragma unroll
  for (int j = 0; j < KPL; ++j) csum[j] = 0.f;

  const int rstride = gridDim.x * nw;
  for (int r0 = blockIdx.x * nw + wib; r0 < a.B; r0 += kHeadRows * rstride) {
    float xv[kHeadRows][KPL];
    int y[kHeadRows];
#pragma unroll
    for (int q = 0; q < kHeadRows; ++q) {
      const int r = r0 + q * rstride;
      y[q] = r < a.B ? a.labels[r] : 0;
#pragma unroll
      for (int j = 0; j < KPL; ++j) xv[q][j] = r < a.B ? a.x[(size_t)r * a.K + lane + 32 * j] : 0.f;
    }
    if (!w_ready) { park_w(); w_ready = true; }            // warp-uniform: r0 depends on the warp only
#pragma unroll
    for (int q = 0; q < kHeadRows; ++q) {
      const int r = r0 + q * rstride;
      if (r >= a.B) break;
      // ---- logits: NC independent dot products (ILP), then NC interleaved butterflies; lane c keeps z_c ----
      float p[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const float* wr = ws + c * a.K + lane;
        float t = 0.f;
#pragma unroll
        for (int j = 0; j < KPL; ++j) t = fmaf(xv[q][j], wr[32 * j], t);
        p[c] = t;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int c = 0; c < NC; ++c) p[c] += __shfl_xor_sync(0xffffffffu, p[c], o);
      }
      float z = -INFINITY;
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (lane == c && c < a.ncls) z = p[c] + bz;
      // ---- log-softmax, loss, prediction ----
      const float m = warp_max(z);
      const float e = lane < a.ncls ? __expf(z - m) : 0.f;
      const float s = warp_sum(e);
      const float lse = m + __logf(s);
      const unsigned ball = __ballot_sync(0xffffffffu, lane < a.ncls && z == m);
      const int pred = __ffs(ball) - 1;              // first max wins, like torch.max
      const float zy = __shfl_sync(0xffffffffu, z, y[q] & 31);
      if (lane < a.ld) a.logits[(size_t)r * a.ld + lane] = lane < a.ncls ? z : 0.f;
      if (lane == 0) {
        lsum += lse - zy;
        if (a.confusion) atomicAdd(a.confusion + y[q] * a.ncls + pred, 1);
      }
      if (!train) continue;
      // ---- backward: dz on lane c, then dx = (dz W) masked by the activation that fed this layer ----
      const float dz = lane < a.ncls ? (e / s - (lane == y[q] ? 1.f : 0.f)) * a.gscale : 0.f;
      if (lane < a.ld) a.dlogits[(size_t)r * a.ld + lane] = dz;
      bsum += dz;
      float acc[KPL];
#pragma unroll
      for (int j = 0; j < KPL; ++j) acc[j] = 0.f;
#pragma unroll
      for (int c = 0; c < NC; ++c) {       // dz is zero on lanes >= ncls, and so are those rows of ws
        const float dzc = __shfl_sync(0xffffffffu, dz, c);
        const float* wr = ws + c * a.K + lane;
#pragma unroll
        for (int j = 0; j < KPL; ++j) acc[j] = fmaf(dzc, wr[32 * j], acc[j]);
      }
#pragma unroll
      for (int j = 0; j < KPL; ++j) {
        const float g = xv[q][j] > 0.f ? acc[j] * a.in_scale : 0.f;
        if (a.dx) a.dx[(size_t)r * a.K + lane + 32 * j] = g;
        csum[j] += g;
      }
    }
  }
  if (!w_ready) park_w();                                  // warps that had no row still owe their part of W
  // ---- fixed-order reduction: warps of this CTA -> one partial row; last CTA (ticket) -> results ----
  if (lane == 0) lpart[wib] = lsum;
  bpart[wib][lane] = bsum;
#pragma unroll
  for (int j = 0; j < KPL; ++j) cs[wib * a.K + lane + 32 * j] = csum[j];
  __syncthreads();
  const int pitch = a.K + 64;
  float* mine = a.scratch + (size_t)blockIdx.x * pitch;      // [0] loss, [32..63] db_cls, [64..64+K) db_in
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += lpart[w];
    mine[0] = t;
  }
  if (wib == 1) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += bpart[w][lane];
    mine[32 + lane] = t;
  }
  for (int k = threadIdx.x; k < a.K; k += blockDim.x) {
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += cs[w * a.K + k];
    mine[64 + k] = t;
  }
  __threadfence();
  __syncthreads();
  int* ticket = reinterpret_cast<int*>(a.scratch + (size_t)kHeadMaxCtas * pitch);
  if (threadIdx.x == 0) last_flag = atomicAdd(ticket, 1) == (int)gridDim.x - 1;
  __syncthreads();
  if (!last_flag) return;
  __threadfence();
  const int G = gridDim.x;
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int g = 0; g < G; ++g) t += __ldcg(a.scratch + (size_t)g * pitch);
    *a.loss_sum = a.accumulate ? *a.loss_sum + t : t;
    *ticket = 0;                                   // ready for the next launch / graph replay
  }
  if (train && a.db_cls && wib == 1 && lane < a.ld) {
    float t = 0.f;
    for (int g = 0; g < G; ++g) t += __ldcg(a.scratch + (size_t)g * pitch + 32 + lane);
    a.db_cls[lane] = t;
  }
  if (train && a.db_in) {
    for (int k = threadIdx.x; k < a.K; k += blockDim.x) {
      float t = 0.f;
      for (int g = 0; g < G; ++g) t += __ldcg(a.scratch + (size_t)g * pitch + 64 + k);
      a.db_in[k] = t;
    }
  }
}

}  // namespace mpit

using namespace mpit;

extern "C" int mpit_classifier_head(const float* x, const float* w, const float* bias, const int* labels, float* logits,
                                    float* dlogits, float* dx, float* loss_sum, int* confusion, float* db_cls,
                                    float* db_in, unsigned long long* rng_inc, float* scratch, int B, int K, int ncls,
                                    int ld, float gscale, float in_scale, int accumulate, void* stream) {
  if (!x || !w || !labels || !logits || !loss_sum || !scratch) return MPIT_E_ARG;
  if ((uintptr_t)w & 15) return MPIT_E_ARG;
  if (B < 1 || ncls < 1 || ncls > 32 || ld < ncls || ld > 32 || K % 32 != 0) return MPIT_E_ARG;
  if ((dx || db_in || db_cls) && !dlogits) return MPIT_E_ARG;
  const int kpl = K / 32;
  const int nc = ncls <= 16 ? 16 : 32;
  const size_t smem = sizeof(float) * ((size_t)nc * K + kHeadWarps * (size_t)K);
  if (smem > 160 * 1024 || (kpl != 2 && kpl != 4 && kpl != 8 && kpl != 16)) return MPIT_E_UNSUPPORTED;
  HeadArgs a{x, w, bias, labels, logits, dlogits, dx, loss_sum, confusion, db_cls, db_in, rng_inc, scratch,
             B, K, ncls, ld, accumulate, gscale, in_scale};
  int grid = (B + kHeadWarps - 1) / kHeadWarps;      // one row per warp while the CTAs last
  if (grid > kHeadMaxCtas) grid = kHeadMaxCtas;
  cudaStream_t st = (cudaStream_t)stream;
  auto go = [&](auto kern) -> int {
    // set on every launch: a static flag inside a generic lambda turned out to be shared by all of its
    // instantiations, which left every kernel variant but the first at the default 48 KB limit
    MPIT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    MPIT_CUDA_TRY(launch_k(kern, dim3(grid), dim3(kHeadWarps * 32), smem, st, a));
    return 0;
  };
#define MPIT_HEAD(KPL) case KPL: return nc == 16 ? go(classifier_head_kernel<KPL, 16>) : go(classifier_head_kernel<KPL, 32>);
  switch (kpl) {
    MPIT_HEAD(2)
    MPIT_HEAD(4)
    MPIT_HEAD(8)
    MPIT_HEAD(16)
  }
#undef MPIT_HEAD
  return MPIT_E_UNSUPPORTED;
}

// Floats the caller must provide (zero-filled once) as `scratch` of mpit_classifier_head for a given K.
extern "C" int mpit_classifier_head_scratch_floats(int K) { return kHeadMaxCtas * (K + 64) + 16; }
