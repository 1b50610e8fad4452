// BiCNN sentence-CNN kernels that are not GEMMs (sm_100a): embedding gather / tied scatter-add, the
// finish of the fused TemporalConvolution->Max->ReLU->Normalize tower, its backward, and the fused
// GESD similarity + margin-ranking loss.  The GEMMs themselves (Linear+Tanh, TemporalConvolution with
// the max-over-time epilogue, and their gradients) run on umma_igemm.cu / umma_wgrad.cu.
//
// Reference: /root/reference/BiCNN/bicnn.lua:30-121 (model), :87-105,121 (GESD, MarginRankingCriterion),
// BiCNN/Normalize.lua (forward :11-38; backward restated in closed form instead of the d x d Jacobian
// of :53-69), nn.LookupTable with shared gradWeight (:30-40).
#include "mpit_b200.h"
#include "common.cuh"

namespace mpit {

// out[row][0..ld) = E[tokens[row]][0..dim) zero-padded to ld (ld multiple of 4; dim multiple of 4).
__global__ void __launch_bounds__(256)
embedding_gather_kernel(const float* __restrict__ E, const int* __restrict__ tokens, float* __restrict__ out,
                        long long rows, int dim, int ldE, int ld) {
  const int per_row = ld >> 2;
  long long total = rows * per_row;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long r = i / per_row;
    int c = (int)(i - r * per_row) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < dim) v = ld4(E + (size_t)tokens[r] * ldE + c);
    st4(out + (size_t)r * ld + c, v);
  }
}

// dE[tokens[row]][c] += g[row][c] for c < dim.  The four towers share one table and one gradient
// buffer (bicnn.lua:35-40), so every occurrence of a word simply accumulates: red.add.
__global__ void __launch_bounds__(256)
embedding_scatter_add_kernel(float* __restrict__ dE, const int* __restrict__ tokens, const float* __restrict__ g,
                             const int* __restrict__ row_valid, long long rows, int dim, int ldE, int ld) {
  long long total = rows * dim;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long r = i / dim;
    int c = (int)(i - r * dim);
    if (row_valid && !row_valid[r]) continue;
    float v = g[(size_t)r * ld + c];
    if (v != 0.f) atomicAdd(dE + (size_t)tokens[r] * ldE + c, v);
  }
}

// One block per sentence: unpack the (value, argmax) keys of the max-over-time epilogue, L2-normalise
// (Normalize.lua:24-31: normp = sum x^2 + eps, norm = sqrt(normp), y = x / norm).
__global__ void __launch_bounds__(256)
encoder_finish_kernel(const unsigned long long* __restrict__ mx, float* __restrict__ r, unsigned char* __restrict__ arg,
                      float* __restrict__ y, float* __restrict__ normp, int F, int ld, float eps) {
  const int s = blockIdx.x;
  __shared__ float red[8];
  __shared__ float s_norm;
  float acc = 0.f;
  for (int c = threadIdx.x; c < F; c += blockDim.x) {
    unsigned long long k = mx[(size_t)s * ld + c];
    float v = __uint_as_float((unsigned)(k >> 32));
    r[(size_t)s * ld + c] = v;
    arg[(size_t)s * ld + c] = (unsigned char)(k & 0xffu);
    acc += v * v;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = eps;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
    normp[s] = t;
    s_norm = sqrtf(t);
  }
  __syncthreads();
  const float inv = 1.f / s_norm;
  for (int c = threadIdx.x; c < F; c += blockDim.x) y[(size_t)s * ld + c] = r[(size_t)s * ld + c] * inv;
}

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.f;
  for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
  __syncthreads();
  return t;
}

// GESD of P (query, answer) pairs of normalised embeddings, one block per pair (bicnn.lua:98-105,442):
//   sim = 1/(1 + ||q - a||) * 1/(1 + exp(-(q.a + 1))).   stats[p] = {dot, l2}.
__global__ void __launch_bounds__(256)
gesd_pairs_kernel(const float* __restrict__ y, const int* __restrict__ qi, const int* __restrict__ ai,
                  float* __restrict__ sim, float* __restrict__ stats, int F, int ld) {
  const int p = blockIdx.x;
  const float* q = y + (size_t)qi[p] * ld;
  const float* a = y + (size_t)ai[p] * ld;
  __shared__ float red[8];
  float dot = 0.f, d2 = 0.f;
  for (int c = threadIdx.x; c < F; c += blockDim.x) {
    float qv = q[c], av = a[c];
    dot += qv * av;
    d2 += (qv - av) * (qv - av);
  }
  dot = block_sum(dot, red);
  d2 = block_sum(d2, red);
  if (threadIdx.x == 0) {
    float l2 = sqrtf(d2);
    sim[p] = 1.f / (1.f + l2) * (1.f / (1.f + __expf(-(dot + 1.f))));
    if (stats) { stats[2 * p] = dot; stats[2 * p + 1] = l2; }
  }
}

// MarginRankingCriterion with y=1 over (s+, s-) pairs and the gradient w.r.t. the normalised
// embeddings, fused (bicnn.lua:121,380-393).  triple t: pairs (2t, 2t+1) share the query.
//   loss_t = max(0, margin - (s+ - s-)); dL/ds+ = -1, dL/ds- = +1 when active.
//   ds/dq = -A^2 B (q-a)/l2 + A B (1-B) a ,  ds/da = +A^2 B (q-a)/l2 + A B (1-B) q   (A=1/(1+l2), B=sigmoid(dot+1))
// dy is accumulated with atomics (a sentence can take part in several pairs); must be zeroed by the caller.
__global__ void __launch_bounds__(256)
gesd_margin_bwd_kernel(const float* __restrict__ y, const int* __restrict__ qi, const int* __restrict__ ai,
                       const float* __restrict__ sim, const float* __restrict__ stats, float* __restrict__ dy,
                       float* __restrict__ loss_sum, int F, int ld, float margin) {
  const int t = blockIdx.x;
  const float sp = sim[2 * t], sn = sim[2 * t + 1];
  const float l = margin - (sp - sn);
  if (l <= 0.f) return;
  if (threadIdx.x == 0) atomicAdd(loss_sum, l);
  for (int k = 0; k < 2; ++k) {
    const int p = 2 * t + k;
    const float gs = k == 0 ? -1.f : 1.f;
    const float dot = stats[2 * p], l2 = stats[2 * p + 1];
    const float A = 1.f / (1.f + l2), B = 1.f / (1.f + __expf(-(dot + 1.f)));
    const float cd = (l2 > 0.f) ? (-A * A * B / l2) : 0.f;   // coefficient of (q - a)
    const float cb = A * B * (1.f - B);
    const float* q = y + (size_t)qi[p] * ld;
    const float* a = y + (size_t)ai[p] * ld;
    float* dq = dy + (size_t)qi[p] * ld;
    float* da = dy + (size_t)ai[p] * ld;
    for (int c = threadIdx.x; c < F; c += blockDim.x) {
      float qv = q[c], av = a[c];
      atomicAdd(dq + c, gs * (cd * (qv - av) + cb * av));
      atomicAdd(da + c, gs * (-cd * (qv - av) + cb * qv));
    }
  }
}

// Backward of Normalize (closed form) -> ReLU -> Max over time.  The gradient of sentence s, column c lands in
// frame arg[s][c] of the dense dC [S][T][ld]; this kernel writes EVERY frame of its sentence (zeros elsewhere), so
// the caller does not clear the 100 MB tensor first, and it also leaves the per-sentence gradient in gmax [S][ld]
// (the temporal convolution's bias gradient is its column sum -- 3 MB instead of a pass over dC).
//   dx = g/norm - x (x.g)/(normp*norm)  with x = r = relu(max), then masked by r > 0.
__global__ void __launch_bounds__(256)
encoder_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ r, const unsigned char* __restrict__ arg,
                   const float* __restrict__ normp, float* __restrict__ dC, float* __restrict__ gmax, int F, int ld, int T) {
  const int s = blockIdx.x;
  __shared__ float red[8];
  float xg = 0.f;
  for (int c = threadIdx.x; c < F; c += blockDim.x) xg += r[(size_t)s * ld + c] * dy[(size_t)s * ld + c];
  xg = block_sum(xg, red);
  const float np = normp[s], nrm = sqrtf(np);
  const float k1 = 1.f / nrm, k2 = xg / (np * nrm);
  for (int c4 = threadIdx.x * 4; c4 < ld; c4 += blockDim.x * 4) {
    float g[4];
    int at[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = c4 + j;
      const float x = c < F ? r[(size_t)s * ld + c] : 0.f;
      g[j] = x > 0.f ? dy[(size_t)s * ld + c] * k1 - x * k2 : 0.f;
      at[j] = x > 0.f ? (int)arg[(size_t)s * ld + c] : -1;
    }
    if (gmax) st4(gmax + (size_t)s * ld + c4, make_float4(g[0], g[1], g[2], g[3]));
    for (int t = 0; t < T; ++t)
      st4(dC + ((size_t)s * T + t) * ld + c4, make_float4(at[0] == t ? g[0] : 0.f, at[1] == t ? g[1] : 0.f,
                                                         at[2] == t ? g[2] : 0.f, at[3] == t ? g[3] : 0.f));
  }
}

}  // namespace mpit

using namespace mpit;

extern "C" {

int mpit_embedding_gather(const float* E, const int* tokens, float* out, long long rows, int dim, int ldE, int ld,
                          void* stream) {
  if (rows <= 0) return 0;
  if ((dim & 3) || (ld & 3) || (ldE & 3)) return MPIT_E_ARG;
  embedding_gather_kernel<<<ew_grid(rows * (ld >> 2), 256, 2), 256, 0, (cudaStream_t)stream>>>(E, tokens, out, rows,
                                                                                             dim, ldE, ld);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_embedding_scatter_add(float* dE, const int* tokens, const float* g, const int* row_valid, long long rows,
                               int dim, int ldE, int ld, void* stream) {
  if (rows <= 0) return 0;
  embedding_scatter_add_kernel<<<ew_grid(rows * dim, 256, 2), 256, 0, (cudaStream_t)stream>>>(dE, tokens, g, row_valid,
                                                                                           rows, dim, ldE, ld);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_encoder_finish(const unsigned long long* mx, float* r, unsigned char* arg, float* y, float* normp, int S,
                        int F, int ld, float eps, void* stream) {
  if (S <= 0) return 0;
  encoder_finish_kernel<<<S, 256, 0, (cudaStream_t)stream>>>(mx, r, arg, y, normp, F, ld, eps);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_gesd_pairs(const float* y, const int* qi, const int* ai, float* sim, float* stats, int P, int F, int ld,
                    void* stream) {
  if (P <= 0) return 0;
  gesd_pairs_kernel<<<P, 256, 0, (cudaStream_t)stream>>>(y, qi, ai, sim, stats, F, ld);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_gesd_margin_bwd(const float* y, const int* qi, const int* ai, const float* sim, const float* stats, float* dy,
                         float* loss_sum, int triples, int F, int ld, float margin, void* stream) {
  if (triples <= 0) return 0;
  gesd_margin_bwd_kernel<<<triples, 256, 0, (cudaStream_t)stream>>>(y, qi, ai, sim, stats, dy, loss_sum, F, ld, margin);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_encoder_bwd(const float* dy, const float* r, const unsigned char* arg, const float* normp, float* dC,
                     float* gmax, int S, int F, int ld, int T, void* stream) {
  if (S <= 0) return 0;
  if (ld & 3) return MPIT_E_ARG;
  encoder_bwd_kernel<<<S, 256, 0, (cudaStream_t)stream>>>(dy, r, arg, normp, dC, gmax, F, ld, T);
  MPIT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
