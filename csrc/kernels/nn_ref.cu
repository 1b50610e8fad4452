// CUDA-core fp32 reference kernels for the conv/linear layers (NHWC activations, packed
// [CO][KH][KWP][C] weights).  These are the *exact-fp32* path: they exist (a) as the on-device
// oracle the tcgen05 kernels in umma_conv.cu are checked against, and (b) as the fallback for
// shapes the tensor-core tiles do not cover.  The hot path of the CIFAR workload is umma_conv.cu.
//
// Layer semantics follow asyncsgd/models/7-layers-cnn.lua:8-24 (SpatialConvolution, ReLU,
// SpatialMaxPooling(2,2,2,2), Linear, Dropout(0.5)); crop semantics follow
// asyncsgd/goot.lua:177-178 (random 28x28 window) and :283-285 (centre window).
#include "mpit_b200.h"
#include "common.cuh"

namespace mpit {

struct ConvGeom {
  int B;          // images
  int SH, SW;     // source image height/width (before crop)
  int H, W;       // logical input height/width (after crop; == SH,SW when no crop)
  int C;          // physical input channels (padded)
  int KH, KW;     // logical kernel
  int KWP;        // physical kernel width (padded)
  int OH, OW;     // conv output (before pooling)
  int CO;         // physical output channels
};

__device__ __forceinline__ float conv_point(const ConvGeom& g, const float* __restrict__ x,
                                            const float* __restrict__ w, int n, int y0, int x0, int oh,
                                            int ow, int co) {
  float acc = 0.f;
  const float* wrow = w + (size_t)co * g.KH * g.KWP * g.C;
  for (int kh = 0; kh < g.KH; ++kh) {
    const float* xr = x + (((size_t)n * g.SH + (y0 + oh + kh)) * g.SW + (x0 + ow)) * g.C;
    const float* wr = wrow + (size_t)kh * g.KWP * g.C;
    for (int kc = 0; kc < g.KW * g.C; ++kc) acc = fmaf(xr[kc], wr[kc], acc);
  }
  return acc;
}

// One thread per (pooled) output element; co is the fastest index -> coalesced stores and
// warp-uniform activation reads.
__global__ void __launch_bounds__(256)
conv_fwd_ref_kernel(ConvGeom g, const float* __restrict__ x, const float* __restrict__ w,
                    const float* __restrict__ bias, const int* __restrict__ crop, float* __restrict__ out,
                    unsigned char* __restrict__ pool_idx, int relu, int pool, int dropout,
                    const unsigned long long* __restrict__ rng) {
  unsigned long long seed = dropout ? rng[0] : 0ull, step = dropout ? rng[1] : 0ull;
  int PH = pool ? g.OH / 2 : g.OH, PW = pool ? g.OW / 2 : g.OW;
  long long total = (long long)g.B * PH * PW * g.CO;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int co = (int)(i % g.CO);
    long long r = i / g.CO;
    int pw = (int)(r % PW); r /= PW;
    int ph = (int)(r % PH);
    int n = (int)(r / PH);
    int y0 = crop ? crop[2 * n] : 0, x0 = crop ? crop[2 * n + 1] : 0;
    float b = bias ? bias[co] : 0.f;
    float v;
    if (pool) {
      float best = -INFINITY;
      int bi = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float z = conv_point(g, x, w, n, y0, x0, 2 * ph + (q >> 1), 2 * pw + (q & 1), co) + b;
        if (z > best) { best = z; bi = q; }
      }
      v = best;
      pool_idx[i] = (unsigned char)bi;
    } else {
      v = conv_point(g, x, w, n, y0, x0, ph, pw, co) + b;
    }
    if (relu) v = fmaxf(v, 0.f);
    if (dropout) {
      uint4 rnd = Philox::gen(seed, (unsigned long long)(i >> 2), (uint32_t)step);
      uint32_t bits = (i & 3) == 0 ? rnd.x : (i & 3) == 1 ? rnd.y : (i & 3) == 2 ? rnd.z : rnd.w;
      v = (bits & 1u) ? v * 2.f : 0.f;
    }
    out[i] = v;
  }
}

// dX[n,ih,iw,c] = sum_{kh,kw,co} dZ[n,ih-kh,iw-kw,co] * W[co][kh][kw][c]
__global__ void __launch_bounds__(256)
conv_dgrad_ref_kernel(ConvGeom g, const float* __restrict__ dz, const float* __restrict__ w,
                      float* __restrict__ dx) {
  long long total = (long long)g.B * g.H * g.W * g.C;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int c = (int)(i % g.C);
    long long r = i / g.C;
    int iw = (int)(r % g.W); r /= g.W;
    int ih = (int)(r % g.H);
    int n = (int)(r / g.H);
    float acc = 0.f;
    for (int kh = 0; kh < g.KH; ++kh) {
      int oh = ih - kh;
      if (oh < 0 || oh >= g.OH) continue;
      for (int kw = 0; kw < g.KW; ++kw) {
        int ow = iw - kw;
        if (ow < 0 || ow >= g.OW) continue;
        const float* dzr = dz + (((size_t)n * g.OH + oh) * g.OW + ow) * g.CO;
        const float* wr = w + ((size_t)kh * g.KWP + kw) * g.C + c;
        size_t wstride = (size_t)g.KH * g.KWP * g.C;
        for (int co = 0; co < g.CO; ++co) acc = fmaf(dzr[co], wr[(size_t)co * wstride], acc);
      }
    }
    dx[i] = acc;
  }
}

// dW[co][kh][kw][c] = sum_{n,oh,ow} dZ[n,oh,ow,co] * X[n,y0+oh+kh,x0+ow+kw,c].
// Block = 16(co) x 16(c) tile for one (kh,kw); gridDim.z splits the pixel range and the
// partial sums are combined with red.add (dW must be zeroed by the caller).
__global__ void __launch_bounds__(256)
conv_wgrad_ref_kernel(ConvGeom g, const float* __restrict__ dz, const float* __restrict__ x,
                      const int* __restrict__ crop, float* __restrict__ dw) {
  int tc = threadIdx.x & 15, tco = threadIdx.x >> 4;
  int c = blockIdx.x * 16 + tc;
  int co = blockIdx.y % ((g.CO + 15) / 16) * 16 + tco;
  int khw = blockIdx.y / ((g.CO + 15) / 16);
  int kh = khw / g.KW, kw = khw % g.KW;
  long long npix = (long long)g.B * g.OH * g.OW;
  long long chunk = (npix + gridDim.z - 1) / gridDim.z;
  long long p0 = (long long)blockIdx.z * chunk, p1 = p0 + chunk < npix ? p0 + chunk : npix;
  if (c >= g.C || co >= g.CO) return;
  float acc = 0.f;
  for (long long p = p0; p < p1; ++p) {
    int ow = (int)(p % g.OW);
    long long r = p / g.OW;
    int oh = (int)(r % g.OH);
    int n = (int)(r / g.OH);
    int y0 = crop ? crop[2 * n] : 0, x0 = crop ? crop[2 * n + 1] : 0;
    float a = dz[(size_t)p * g.CO + co];
    float b = x[(((size_t)n * g.SH + (y0 + oh + kh)) * g.SW + (x0 + ow + kw)) * g.C + c];
    acc = fmaf(a, b, acc);
  }
  atomicAdd(dw + (((size_t)co * g.KH + kh) * g.KWP + kw) * g.C + c, acc);
}

// dZ[n,oh,ow,c] = dP[n,oh/2,ow/2,c] if this position won the max and the pooled ReLU output is
// positive, else 0.  Without pooling it is the plain ReLU(+dropout) mask:
// dZ = dOut * scale if out > 0.
__global__ void __launch_bounds__(256)
act_bwd_kernel(const float* __restrict__ dout, const float* __restrict__ out,
               const unsigned char* __restrict__ pool_idx, float* __restrict__ dz, int B, int OH, int OW,
               int C, int pool, float scale) {
  long long total = (long long)B * OH * OW * C;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    if (!pool) {
      dz[i] = out[i] > 0.f ? dout[i] * scale : 0.f;
      continue;
    }
    int c = (int)(i % C);
    long long r = i / C;
    int ow = (int)(r % OW); r /= OW;
    int oh = (int)(r % OH);
    int n = (int)(r / OH);
    size_t pi = (((size_t)n * (OH / 2) + oh / 2) * (OW / 2) + ow / 2) * C + c;
    int q = (oh & 1) * 2 + (ow & 1);
    dz[i] = (pool_idx[pi] == q && out[pi] > 0.f) ? dout[pi] * scale : 0.f;
  }
}

// db[c] = sum_rows dZ[row][c]; dZ is [rows][C] with C contiguous.
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ dz, float* __restrict__ db, long long rows, int C) {
  // blockDim = (32 cols, 8 row-lanes)
  int c = blockIdx.x * 32 + threadIdx.x;
  float acc = 0.f;
  if (c < C) {
    for (long long r = (long long)blockIdx.y * blockDim.y + threadIdx.y; r < rows; r += (long long)gridDim.y * blockDim.y)
      acc += dz[(size_t)r * C + c];
  }
  __shared__ float sm[8][33];
  sm[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += sm[k][threadIdx.x];
    atomicAdd(db + c, s);
  }
}

// uint8 CHW [B,3,SH,SW] -> fp32 "pair-expanded" NHWC8 [B,SH,SW,8], scaled by 1/255
// (asyncsgd/goot.lua:87-88): pixel x holds (r,g,b,0) of pixel x followed by (r,g,b,0) of pixel x+1.
// 32 consecutive floats starting at pixel x are then pixels x,x+1 | x+1,x+2 | x+2,x+3 | x+3,x+4, i.e.
// one 128-byte TMA row carries all five kernel columns of the first 5x5 convolution, and every row
// starts on a 32-byte boundary (needed by the 32-byte-atom swizzle of the weight-gradient kernel).
__global__ void __launch_bounds__(256)
prep_u8_kernel(const unsigned char* __restrict__ src, float* __restrict__ dst, int B, int SH, int SW, int cin,
               const int* __restrict__ labels_src, int* __restrict__ labels_dst, const int* __restrict__ crop_src,
               int* __restrict__ crop_dst, const float* __restrict__ fwd_src, float* fwd_dst, int fwd_n) {
  pdl_trigger();
  pdl_wait();
  // a few floats forwarded from the previous step (its loss) to wherever the caller points -- in the end-to-end
  // step that is host-mapped pinned memory, which spares the stream a copy-engine operation per step
  if (fwd_src && blockIdx.x == 0 && threadIdx.x < fwd_n) fwd_dst[threadIdx.x] = fwd_src[threadIdx.x];
  // the batch's labels and crop offsets ride along (one launch stages a whole mini-batch: three separate
  // device-to-device copies in front of every step cost ~10 us of a 140 us step)
  {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (labels_src && t < B) labels_dst[t] = labels_src[t];
    if (crop_src && t < 2 * B) crop_dst[t] = crop_src[t];
  }
  long long total = (long long)B * SH * SW;
  long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    long long n = i / (SH * SW);
    int p = (int)(i % (SH * SW));
    int x = p % SW;
    const unsigned char* s = src + (size_t)n * cin * SH * SW + p;
    size_t plane = (size_t)SH * SW;
    float4 a, b;
    a.x = (float)s[0] / 255.f;
    a.y = cin > 1 ? (float)s[plane] / 255.f : 0.f;
    a.z = cin > 2 ? (float)s[2 * plane] / 255.f : 0.f;
    a.w = cin > 3 ? (float)s[3 * plane] / 255.f : 0.f;
    if (x + 1 < SW) {
      b.x = (float)s[1] / 255.f;
      b.y = cin > 1 ? (float)s[plane + 1] / 255.f : 0.f;
      b.z = cin > 2 ? (float)s[2 * plane + 1] / 255.f : 0.f;
      b.w = cin > 3 ? (float)s[3 * plane + 1] / 255.f : 0.f;
    } else {
      b.x = b.y = b.z = b.w = 0.f;
    }
    st4(dst + 8 * i, a);
    st4(dst + 8 * i + 4, b);
  }
}

// LogSoftMax + ClassNLLCriterion fused forward/backward (asyncsgd/models/7-layers-cnn.lua:25,
// asyncsgd/goot.lua:40,200-206), one warp per sample.  logits are [B][ld] with ncls valid
// columns.  dlogits = (softmax - onehot) * gscale.  Also accumulates the confusion matrix
// that the reference updates per sample (goot.lua:209).
__global__ void __launch_bounds__(1024)
nll_fwd_bwd_kernel(const float* __restrict__ logits, const int* __restrict__ labels,
                   float* __restrict__ dlogits, float* __restrict__ loss_sum, int* __restrict__ confusion,
                   int B, int ncls, int ld, float gscale, unsigned long long* __restrict__ rng_inc, int accumulate,
                   float* __restrict__ dbias) {
  __shared__ float part[32];
  __shared__ float bpart[32][33];
  pdl_trigger();
  pdl_wait();
  // The step counter that keys the dropout Philox stream lives on the device so that a
  // captured CUDA graph advances it on every replay.
  if (rng_inc && blockIdx.x == 0 && threadIdx.x == 0) rng_inc[1] += 1ull;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nw = blockDim.x >> 5;
  float mine = 0.f;
  float bsum = 0.f;     // this lane's class: gradient summed over the rows of this warp
  const int rstride = gridDim.x * nw;
  // four rows per warp in flight: their loads are issued together (one CTA serves a whole 128-sample batch,
  // so per-row load latency would otherwise be paid four times in sequence)
  for (int r0 = blockIdx.x * nw + wib; r0 < B; r0 += 4 * rstride) {
    float z[4];
    int y[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = r0 + k * rstride;
      z[k] = (r < B && lane < ncls) ? logits[(size_t)r * ld + lane] : -INFINITY;
      y[k] = r < B ? labels[r] : 0;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int r = r0 + k * rstride;
      if (r >= B) break;
      float m = warp_max(z[k]);
      float e = lane < ncls ? __expf(z[k] - m) : 0.f;
      float s = warp_sum(e);
      float lse = m + __logf(s);
      if (dlogits) {
        float gz = 0.f;
        if (lane < ncls) gz = (e / s - (lane == y[k] ? 1.f : 0.f)) * gscale;
        if (lane < ld) dlogits[(size_t)r * ld + lane] = gz;
        bsum += gz;
      }
      // argmax (first max wins, like torch.max)
      unsigned ball = __ballot_sync(0xffffffffu, lane < ncls && z[k] == m);
      int pred = __ffs(ball) - 1;
      float zy = __shfl_sync(0xffffffffu, z[k], y[k] & 31);
      if (lane == 0) {
        mine += lse - zy;
        if (confusion) atomicAdd(confusion + y[k] * ncls + pred, 1);
      }
    }
  }
  // fixed-order block reduction: with one CTA (the training configuration) the loss is a plain store,
  // bit-reproducible and with no zeroing launch in front of it
  if (lane == 0) part[wib] = mine;
  if (dbias) bpart[wib][lane] = bsum;
  __syncthreads();
  if (dbias && wib == 1 && lane < ld) {   // single-CTA launches only (host checks): fixed-order sum over warps
    float t = 0.f;
    for (int w = 0; w < nw; ++w) t += bpart[w][lane];
    dbias[lane] = t;
  }
  if (wib == 0) {
    float t = lane < nw ? part[lane] : 0.f;
    t = warp_sum(t);
    if (lane == 0) {
      if (gridDim.x == 1 && !accumulate) *loss_sum = t;
      else atomicAdd(loss_sum, t);
    }
  }
}

}  // namespace mpit

using namespace mpit;

static ConvGeom geom_from(const int* g) {
  ConvGeom c;
  c.B = g[0]; c.SH = g[1]; c.SW = g[2]; c.H = g[3]; c.W = g[4]; c.C = g[5];
  c.KH = g[6]; c.KW = g[7]; c.KWP = g[8]; c.OH = g[9]; c.OW = g[10]; c.CO = g[11];
  return c;
}

extern "C" {

int mpit_conv_fwd_ref(const int* geom12, const float* x, const float* w, const float* bias, const int* crop,
                      float* out, unsigned char* pool_idx, int relu, int pool, int dropout,
                      const unsigned long long* rng, void* stream) {
  ConvGeom g = geom_from(geom12);
  long long total = (long long)g.B * (pool ? g.OH / 2 : g.OH) * (pool ? g.OW / 2 : g.OW) * g.CO;
  conv_fwd_ref_kernel<<<ew_grid(total, 256, 1, 16), 256, 0, (cudaStream_t)stream>>>(
      g, x, w, bias, crop, out, pool_idx, relu, pool, dropout, rng);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_conv_dgrad_ref(const int* geom12, const float* dz, const float* w, float* dx, void* stream) {
  ConvGeom g = geom_from(geom12);
  long long total = (long long)g.B * g.H * g.W * g.C;
  conv_dgrad_ref_kernel<<<ew_grid(total, 256, 1, 16), 256, 0, (cudaStream_t)stream>>>(g, dz, w, dx);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_conv_wgrad_ref(const int* geom12, const float* dz, const float* x, const int* crop, float* dw,
                        void* stream) {
  ConvGeom g = geom_from(geom12);
  cudaStream_t st = (cudaStream_t)stream;
  MPIT_CUDA_TRY(cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)g.CO * g.KH * g.KWP * g.C, st));
  long long npix = (long long)g.B * g.OH * g.OW;
  int split = (int)((npix + 511) / 512);
  if (split > 64) split = 64;
  if (split < 1) split = 1;
  dim3 grid((g.C + 15) / 16, ((g.CO + 15) / 16) * g.KH * g.KW, split);
  conv_wgrad_ref_kernel<<<grid, 256, 0, st>>>(g, dz, x, crop, dw);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_act_bwd(const float* dout, const float* out, const unsigned char* pool_idx, float* dz, int B, int OH,
                 int OW, int C, int pool, float scale, void* stream) {
  long long total = (long long)B * OH * OW * C;
  act_bwd_kernel<<<ew_grid(total, 256, 4), 256, 0, (cudaStream_t)stream>>>(dout, out, pool_idx, dz, B, OH, OW,
                                                                            C, pool, scale);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_colsum(const float* dz, float* db, long long rows, int C, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  MPIT_CUDA_TRY(cudaMemsetAsync(db, 0, sizeof(float) * (size_t)C, st));
  int gy = (int)((rows + 63) / 64);
  if (gy > 128) gy = 128;
  if (gy < 1) gy = 1;
  colsum_kernel<<<dim3((C + 31) / 32, gy), dim3(32, 8), 0, st>>>(dz, db, rows, C);
  MPIT_LAUNCH_CHECK();
  return 0;
}

int mpit_prep_u8(const unsigned char* src, float* dst, int B, int SH, int SW, int cin, const int* labels_src,
                 int* labels_dst, const int* crop_src, int* crop_dst, void* stream) {
  if (cin < 1 || cin > 4) return MPIT_E_ARG;
  if ((labels_src && !labels_dst) || (crop_src && !crop_dst)) return MPIT_E_ARG;
  long long total = (long long)B * SH * SW;
  if (total < 2 * B) return MPIT_E_ARG;      // the label / crop copy uses the first 2B threads
  MPIT_CUDA_TRY(launch_k(prep_u8_kernel, dim3(ew_grid(total, 256, 1)), dim3(256), 0, (cudaStream_t)stream, src, dst,
                         B, SH, SW, cin, labels_src, labels_dst, crop_src, crop_dst, (const float*)nullptr,
                         (float*)nullptr, 0));
  return 0;
}

int mpit_prep_u8_fwd(const unsigned char* src, float* dst, int B, int SH, int SW, int cin, const int* labels_src,
                     int* labels_dst, const int* crop_src, int* crop_dst, const float* fwd_src, float* fwd_dst,
                     int fwd_n, void* stream) {
  if (cin < 1 || cin > 4) return MPIT_E_ARG;
  if ((labels_src && !labels_dst) || (crop_src && !crop_dst) || (fwd_src && !fwd_dst) || fwd_n < 0 || fwd_n > 256)
    return MPIT_E_ARG;
  long long total = (long long)B * SH * SW;
  if (total < 2 * B) return MPIT_E_ARG;
  MPIT_CUDA_TRY(launch_k(prep_u8_kernel, dim3(ew_grid(total, 256, 1)), dim3(256), 0, (cudaStream_t)stream, src, dst,
                         B, SH, SW, cin, labels_src, labels_dst, crop_src, crop_dst, fwd_src, fwd_dst, fwd_n));
  return 0;
}

int mpit_nll_fwd_bwd(const float* logits, const int* labels, float* dlogits, float* loss_sum, int* confusion,
                     int B, int ncls, int ld, float gscale, unsigned long long* rng_inc, int accumulate,
                     float* dbias, void* stream) {
  if (ncls > 32 || ld > 32) return (int)cudaErrorInvalidValue;
  if (dbias && (!dlogits || B > 4096)) return MPIT_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  // up to 4096 rows: one 1024-thread CTA (32 warps x <=128 rows each) stores the sum directly
  int grid = 1;
  if (B > 4096) {
    grid = (B + 1023) / 1024;
    if (grid > 4 * MPIT_NUM_SMS) grid = 4 * MPIT_NUM_SMS;
    if (!accumulate) MPIT_CUDA_TRY(cudaMemsetAsync(loss_sum, 0, sizeof(float), st));
    accumulate = 1;
  }
  MPIT_CUDA_TRY(launch_k(nll_fwd_bwd_kernel, dim3(grid), dim3(1024), 0, st, logits, labels, dlogits, loss_sum,
                         confusion, B, ncls, ld, gscale, rng_inc, accumulate, dbias));
  return 0;
}

}  // extern "C"
