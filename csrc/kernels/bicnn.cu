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

// ---- device-side mini-batch assembly and negative sampling (bicnn.lua:295-359) ------------------------------
// The training set, the answer table and the per-sample target lists live on the device; a step only receives
// the indices of its samples.  Sentence s of the step's [S][Lk] token matrix is: question of sample s (s < B),
// its positive answer (B <= s < 2B), or candidate negative k of sample b (s = 2B + b*K + k) -- a uniformly drawn
// answer that is not one of the sample's targets (rejection sampling, bicnn.lua:327-334), keyed by
// (seed, step, s) through Philox so that graph replays draw fresh candidates.
struct BatchArgs {
  const int* q_tok; const int* q_len; const int* a_tok; const int* a_len;   // training set [N][Lk], [N]
  const int* targets; int tmax;                                           // [N][tmax], -1 padded
  const int* ans_tok; const int* ans_len; int A;                          // answer table [A][Lk], [A]
  int Lk, kw;
  const int* batch; int B, K, S;                                          // sample indices of this step (-1 = none)
  unsigned long long seed; const unsigned long long* step;
  int* tokens; int* lens; int* frames; int* row_valid; int* neg_idx;
};

__global__ void __launch_bounds__(64) bicnn_build_batch_kernel(BatchArgs a) {
  const int s = blockIdx.x;
  __shared__ int s_src, s_len;
  if (threadIdx.x == 0) {
    int src = -1, len = 0;      // src: row offset into one of the three token tables, encoded below
    if (s < a.B) {
      const int idx = a.batch[s];
      if (idx >= 0) { src = idx; len = a.q_len[idx]; }
    } else if (s < 2 * a.B) {
      const int idx = a.batch[s - a.B];
      if (idx >= 0) { src = idx; len = a.a_len[idx]; }
    } else if (s < 2 * a.B + a.B * a.K) {
      const int b = (s - 2 * a.B) / a.K;
      const int idx = a.batch[b];
      if (idx >= 0) {
        const int* tg = a.targets + (size_t)idx * a.tmax;
        int j = 0;
        bool found = false;
        for (unsigned attempt = 0; attempt < 16 && !found; ++attempt) {
          const uint4 r = Philox::gen(a.seed, *a.step * (unsigned long long)a.S + (unsigned long long)s, attempt);
          const unsigned rv[4] = {r.x, r.y, r.z, r.w};
          for (int u = 0; u < 4 && !found; ++u) {
            j = (int)(rv[u] % (unsigned)a.A);
            found = true;
            for (int t = 0; t < a.tmax; ++t) if (tg[t] == j) { found = false; break; }
          }
        }
        a.neg_idx[s - 2 * a.B] = j;
        src = j;
        len = a.ans_len[j];
      }
    }
    s_src = src;
    s_len = len;
    a.lens[s] = src >= 0 ? len : a.kw;
    const int fr = len - a.kw + 1;
    a.frames[s] = (src >= 0 && fr > 0) ? fr : 0;
  }
  __syncthreads();
  const int src = s_src, len = s_len;
  const int* table = s < a.B ? a.q_tok : (s < 2 * a.B ? a.a_tok : a.ans_tok);
  for (int l = threadIdx.x; l < a.Lk; l += blockDim.x) {
    const bool in = src >= 0 && l < len;
    a.tokens[(size_t)s * a.Lk + l] = in ? table[(size_t)src * a.Lk + l] : 0;
    a.row_valid[(size_t)s * a.Lk + l] = in ? 1 : 0;
  }
}

// First margin violator per sample (bicnn.lua:321-359): sim holds s+ of the B samples followed by the B*K
// candidate similarities.  Emits the (q,a+),(q,a-) pair list of the backward pass; a sample without a violator
// (or an empty batch slot) gets the similarity pair (1, 0), which the margin kernel treats as inactive.
__global__ void __launch_bounds__(128)
bicnn_select_kernel(const float* __restrict__ sim, const float* __restrict__ stats, const int* __restrict__ batch, int B,
                    int K, float margin, int* __restrict__ ai2, float* __restrict__ sim2, float* __restrict__ stats2,
                    int* __restrict__ active, float* __restrict__ counters, unsigned long long* step) {
  __shared__ int s_any;
  if (threadIdx.x == 0) s_any = 0;
  __syncthreads();
  int mine = 0;
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    int first = -1;
    if (batch[b] >= 0) {
      const float sp = sim[b];
      for (int k = 0; k < K; ++k)
        if (sp - sim[B + b * K + k] < margin) { first = k; break; }
    }
    ai2[2 * b] = B + b;
    if (first >= 0) {
      const int p = B + b * K + first;
      ai2[2 * b + 1] = 2 * B + b * K + first;
      sim2[2 * b] = sim[b]; sim2[2 * b + 1] = sim[p];
      stats2[4 * b] = stats[2 * b]; stats2[4 * b + 1] = stats[2 * b + 1];
      stats2[4 * b + 2] = stats[2 * p]; stats2[4 * b + 3] = stats[2 * p + 1];
      ++mine;
    } else {
      ai2[2 * b + 1] = 2 * B + b * K;
      sim2[2 * b] = 1.f; sim2[2 * b + 1] = 0.f;
      stats2[4 * b] = stats2[4 * b + 1] = stats2[4 * b + 2] = stats2[4 * b + 3] = 0.f;
    }
  }
  if (mine) atomicAdd(&s_any, mine);
  __syncthreads();
  if (threadIdx.x == 0) {
    *active = s_any > 0 ? 1 : 0;
    if (counters) { counters[0] += (float)s_any; counters[1] += 1.f; }   // triples trained on, steps
    *step += 1ull;
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

/* Device-side mini-batch assembly with negative sampling; ptrs18 is a HOST array of the 18 device pointers
 * {q_tok, q_len, a_tok, a_len, targets, ans_tok, ans_len, batch, step, tokens, lens, frames, row_valid, neg_idx}
 * (14 used, 4 reserved) and dims8 = {tmax, A, Lk, kw, B, K, S, 0}. */
int mpit_bicnn_build_batch(const unsigned long long* ptrs18, const int* dims8, unsigned long long seed, void* stream) {
  if (!ptrs18 || !dims8) return MPIT_E_ARG;
  BatchArgs a{};
  a.q_tok = (const int*)ptrs18[0]; a.q_len = (const int*)ptrs18[1]; a.a_tok = (const int*)ptrs18[2];
  a.a_len = (const int*)ptrs18[3]; a.targets = (const int*)ptrs18[4]; a.ans_tok = (const int*)ptrs18[5];
  a.ans_len = (const int*)ptrs18[6]; a.batch = (const int*)ptrs18[7]; a.step = (const unsigned long long*)ptrs18[8];
  a.tokens = (int*)ptrs18[9]; a.lens = (int*)ptrs18[10]; a.frames = (int*)ptrs18[11]; a.row_valid = (int*)ptrs18[12];
  a.neg_idx = (int*)ptrs18[13];
  a.tmax = dims8[0]; a.A = dims8[1]; a.Lk = dims8[2]; a.kw = dims8[3]; a.B = dims8[4]; a.K = dims8[5]; a.S = dims8[6];
  a.seed = seed;
  if (a.A < 1 || a.B < 1 || a.K < 0 || a.S < 2 * a.B + a.B * a.K || a.Lk < 1) return MPIT_E_ARG;
  bicnn_build_batch_kernel<<<a.S, 64, 0, (cudaStream_t)stream>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_bicnn_select(const float* sim, const float* stats, const int* batch, int B, int K, float margin, int* ai2,
                      float* sim2, float* stats2, int* active, float* counters, unsigned long long* step, void* stream) {
  if (B < 1 || K < 1) return MPIT_E_ARG;
  bicnn_select_kernel<<<1, 128, 0, (cudaStream_t)stream>>>(sim, stats, batch, B, K, margin, ai2, sim2, stats2, active,
                                                          counters, step);
  MPIT_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
