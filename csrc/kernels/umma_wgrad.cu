// Weight gradients of the conv / linear layers on tcgen05 (sm_100a).
//
//   dW[m][n] = sum over pixels p of  Xcol[p][m] * dZ[p][n]
//
// m runs over the packed weight's K index (kernel row, kernel column, input channel), n over output
// channels and the reduction runs over output pixels (n_img, oh, ow).  Both operands are therefore
// "MN-major": a shared-memory tile is [64 pixels][32 floats], exactly what a TMA box over an NHWC
// tensor delivers -- for X the box is shifted by the kernel offset (and the crop offset), so im2col
// is again only a coordinate.  tcgen05.mma (kind::tf32, M=128, N=BLOCK_N, K=8) is issued with both
// major bits set; the accumulator lives in TMEM.  32-bit MN-major operands require the
// SWIZZLE_128B_BASE32B shared-memory layout (TMA mode SWIZZLE_128B_ATOM_32B).
//
// The pixel dimension is split over gridDim.z CTAs.  Each CTA writes its partial [128 x BLOCK_N]
// tile to a scratch workspace with plain 128-byte stores and mpit_wgrad_reduce sums the splits into
// the packed gradient (no atomics: bit-reproducible gradients, and the stores are 4x cheaper than
// red.add at this size).
//
// Reference counterpart: nn.SpatialConvolution / nn.Linear accGradParameters called once per sample
// (asyncsgd/goot.lua:206); here once per mini-batch.
#include "mpit_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "optim_math.cuh"
#include <string.h>
#include <stdio.h>
#include <mutex>

namespace mpit {
using namespace umma;

struct WgradParams {
  int tw, th, tn;            // pixel box of one k-block: tw*th*tn == 64
  int tiles_w, tiles_h;      // pixel sub-tiles per image
  int kw_it, cb;             // M-chunk -> (kh, kw, channel block)
  int total_chunks;          // K/32 of the packed weight
  int num_kb;                // total pixel k-blocks
  int B;
  int N;                     // output channels (physical)
  int m_pad;                 // m_tiles * 128
  float* part;               // scratch [split][m_pad][N]
  const int* crop;
  int* err;
  int x3;                    // 0 plain TF32; 1 / 2: 3xTF32, operands split in shared memory (umma.cuh)
};

constexpr int kWStages = 3;
constexpr int kChunkBytes = 64 * 128;   // 64 pixels x 32 floats
constexpr int kWABytes = 4 * kChunkBytes;

template <int BLOCK_N>
struct WSmem {
  static constexpr int kBBytes = (BLOCK_N / 32) * kChunkBytes;
  static constexpr int kStageBytes = kWABytes + kBBytes;
  static constexpr int kBarOff = kWStages * kStageBytes;
  static constexpr int kTotal = kBarOff + 256 + 1024;
  // 3xTF32: a stage is [A | B | A_lo | B_lo]; two stages of BLOCK_N <= 64 fit, BLOCK_N = 128 has room for one only
  static constexpr int kStages3 = (kBarOff / (2 * kStageBytes)) < 2 ? (kBarOff / (2 * kStageBytes)) : 2;
};

template <int BLOCK_N, bool X3 = false>
__global__ void __launch_bounds__(192)
umma_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_dz,
                  const WgradParams p) {
  using L = WSmem<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::kBarOff);
  uint64_t* empty_bar = full_bar + kWStages;
  uint64_t* split_bar = empty_bar + kWStages;
  uint64_t* tmem_full = split_bar + kWStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  const int x3 = X3 ? p.x3 : 0;
  constexpr int nstages = X3 ? L::kStages3 : kWStages;
  constexpr int stage_bytes = X3 ? 2 * L::kStageBytes : L::kStageBytes;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = blockIdx.x;
  const int n_tile0 = blockIdx.y * BLOCK_N;
  const int split = blockIdx.z, nsplit = gridDim.z;
  const int kb_per = (p.num_kb + nsplit - 1) / nsplit;
  const int kb_lo = split * kb_per;
  const int kb_hi = min(p.num_kb, kb_lo + kb_per);
  const int nkb = max(0, kb_hi - kb_lo);
  int nchunks = p.total_chunks - m_tile * 4;
  if (nchunks > 4) nchunks = 4;

  pdl_trigger();
  // parameter-only producer set-up (integer divisions) goes BEFORE the wait on the previous kernel
  const int pj = warp == 0 ? 0 : warp - 1;     // issuing threads: warps 0,2,3,4 -> 0,1,2,3 (3xTF32: warp 0 alone)
  const int per_kh = p.kw_it * p.cb;
  const int ch = m_tile * 4 + pj;
  const int ckh = ch / per_kh, rch = ch - ckh * per_kh;
  const int ckw = rch / p.cb, cc0 = (rch - ckw * p.cb) * 32;
  const int per_img = p.tiles_w * p.tiles_h;
  const int grp0 = kb_lo / per_img, r_img = kb_lo - grp0 * per_img;
  const int th0 = r_img / p.tiles_w, tw0 = r_img - th0 * p.tiles_w;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_x);
    prefetch_tmap(&tmap_dz);
    for (int i = 0; i < kWStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
      if (X3) mbar_init(&split_bar[i], 128);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, BLOCK_N);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();

  // ===== TMA producers =====
  // A k-block is up to 4 A-chunk loads + BLOCK_N/32 B-chunk loads; issued by one thread that is a > 1000-clock
  // dependent chain per k-block (the feed was issue-bound, not bandwidth-bound: tools/dev/tma_feed_probe.cu).
  // The loads are dealt to four issuing threads -- the elected lanes of warp 0 and of epilogue warps 2,3,4:
  // producer j issues A chunk j and B chunk j of EVERY k-block, producer 0 also posts the expect_tx.  Each
  // producer walks the k-blocks in order, so the empty-barrier parity test stays unambiguous.
  const bool has_a = pj < nchunks, has_b = pj < BLOCK_N / 32;
  if (X3) {
    // 3xTF32: warps 2..5 split the landed tiles, so warp 0 issues every chunk of a k-block itself (the MMA side
    // needs three times as long per k-block, the feed is not the bound)
    if (warp == 0 && elect_one()) {
      const uint32_t bytes = (uint32_t)(nchunks * kChunkBytes + L::kBBytes);
      int grp = grp0, th_i = th0, tw_i = tw0;
      int stage = 0, phase = 0;
      for (int kb = kb_lo; kb < kb_hi; ++kb) {
        const int n0 = grp * p.tn, oh0 = th_i * p.th, ow0 = tw_i * p.tw;
        int y0 = 0, x0 = 0;
        if (p.crop && n0 < p.B) { y0 = p.crop[2 * n0]; x0 = p.crop[2 * n0 + 1]; }
        if (kb - kb_lo >= nstages && !mbar_wait(&empty_bar[stage], phase ^ 1)) { if (p.err) *p.err = 11; break; }
        uint8_t* a_dst = smem + stage * stage_bytes;
        mbar_expect_tx(&full_bar[stage], bytes);
        for (int j = 0; j < nchunks; ++j) {
          const int chj = m_tile * 4 + j;
          const int kh_j = chj / per_kh, rj = chj - kh_j * per_kh;
          const int kw_j = rj / p.cb, cc_j = (rj - kw_j * p.cb) * 32;
          tma_load_4d(a_dst + j * kChunkBytes, &tmap_x, &full_bar[stage], cc_j, ow0 + kw_j + x0, oh0 + kh_j + y0, n0);
        }
        for (int j = 0; j < BLOCK_N / 32; ++j)
          tma_load_4d(a_dst + kWABytes + j * kChunkBytes, &tmap_dz, &full_bar[stage], n_tile0 + j * 32, ow0, oh0, n0);
        if (++stage == nstages) { stage = 0; phase ^= 1; }
        if (++tw_i == p.tiles_w) { tw_i = 0; if (++th_i == p.tiles_h) { th_i = 0; ++grp; } }
      }
    }
  } else if (warp != 1 && pj < 4 && (has_a || has_b || pj == 0) && elect_one()) {
    const uint32_t bytes = (uint32_t)(nchunks * kChunkBytes + L::kBBytes);
    // k-block -> (image group, tile row, tile column), kept incrementally
    int grp = grp0, th_i = th0, tw_i = tw0;
    int y0 = 0, x0 = 0;
    if (p.crop && has_a && grp * p.tn < p.B) { y0 = p.crop[2 * grp * p.tn]; x0 = p.crop[2 * grp * p.tn + 1]; }
    int stage = 0, phase = 0;
    for (int kb = kb_lo; kb < kb_hi; ++kb) {
      const int n0 = grp * p.tn, oh0 = th_i * p.th, ow0 = tw_i * p.tw;
      // next k-block's position (and its crop offsets, fetched before the wait so the load latency hides)
      int ngrp = grp, nth = th_i, ntw = tw_i + 1;
      if (ntw == p.tiles_w) { ntw = 0; if (++nth == p.tiles_h) { nth = 0; ++ngrp; } }
      int ny0 = y0, nx0 = x0;
      if (p.crop && has_a && ngrp != grp && kb + 1 < kb_hi && ngrp * p.tn < p.B) {
        ny0 = p.crop[2 * ngrp * p.tn]; nx0 = p.crop[2 * ngrp * p.tn + 1];
      }
      if (kb - kb_lo >= kWStages && !mbar_wait(&empty_bar[stage], phase ^ 1)) { if (p.err) *p.err = 11; break; }
      uint8_t* a_dst = smem + stage * L::kStageBytes;
      if (pj == 0) mbar_expect_tx(&full_bar[stage], bytes);
      if (has_a)
        tma_load_4d(a_dst + pj * kChunkBytes, &tmap_x, &full_bar[stage], cc0, ow0 + ckw + x0, oh0 + ckh + y0, n0);
      if (has_b)
        tma_load_4d(a_dst + kWABytes + pj * kChunkBytes, &tmap_dz, &full_bar[stage], n_tile0 + pj * 32, ow0, oh0, n0);
      if (++stage == kWStages) { stage = 0; phase ^= 1; }
      grp = ngrp; th_i = nth; tw_i = ntw; y0 = ny0; x0 = nx0;
    }
  }
  __syncwarp();
  if (X3 && warp >= 2) {
    // ===== 3xTF32: operand split, warps 2..5 (rows of chunks this tile does not own are never loaded: whatever
    // they hold only reaches accumulator rows that the epilogue zeroes) =====
    const int tid = threadIdx.x - 64;
    int st = 0, ph = 0;
    for (int i = 0; i < nkb; ++i) {
      if (!mbar_wait(&full_bar[st], ph)) { if (p.err) *p.err = 14; break; }
      const uint32_t base = smem_u32(smem) + st * stage_bytes;
      split_tile(base, base + L::kStageBytes, L::kStageBytes, tid, 128, x3);
      fence_proxy_async_smem();
      mbar_arrive(&split_bar[st]);
      if (++st == nstages) { st = 0; ph ^= 1; }
    }
  }
  if (warp == 0) {
    // warp 0 only issues loads
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_tf32(128, BLOCK_N, 1, 1);   // both operands MN-major
      // descriptor of stage 0; stages / k-steps only add to the 14-bit address field
      const uint64_t desc0 = make_smem_desc(smem_u32(smem), kChunkBytes, 512, 1);
      uint64_t ad = desc0;
      int stage = 0, phase = 0;
      for (int i = 0; i < nkb; ++i) {
        if (!mbar_wait(X3 ? &split_bar[stage] : &full_bar[stage], phase)) { if (p.err) *p.err = 12; break; }
        tc_fence_after();
        const uint64_t bd = ad + (kWABytes >> 4);
        if (X3) {
          const uint64_t lo = (uint64_t)(L::kStageBytes >> 4);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t o = ks * (1024 >> 4);
            mma_tf32(tmem_base, ad + lo + o, bd + o, idesc, (i | ks) ? 1u : 0u);     // A_lo * B_hi
            mma_tf32(tmem_base, ad + o, bd + lo + o, idesc, 1u);                     // A_hi * B_lo
            mma_tf32(tmem_base, ad + o, bd + o, idesc, 1u);                          // A_hi * B_hi
          }
        } else {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)     // 8 pixels per MMA
            mma_tf32(tmem_base, ad + ks * (1024 >> 4), bd + ks * (1024 >> 4), idesc, (i | ks) ? 1u : 0u);
        }
        mma_commit(&empty_bar[stage]);
        ad += (stage_bytes >> 4);
        if (++stage == nstages) { stage = 0; phase ^= 1; ad = desc0; }
      }
      mma_commit(tmem_full);
    }
  } else {
    const int quad = warp & 3;
    if (nkb > 0) {
      if (!mbar_wait(tmem_full, 0)) { if (p.err) *p.err = 13; }
    }
    tc_fence_after();
    // The partial tile is stored n-major, part[split][n][m]: a TMEM lane is one m row, so for each of its
    // 32 columns the warp's 32 lanes write 32 consecutive m -- one full 128-byte line per store, with no
    // transposition -- and the reduce kernel reads AND writes (gw[n][m]) along m, coalesced on both sides.
    // rows of chunks this tile does not own hold garbage: write zeros there (one chunk = one warp's 32 rows)
    const bool rows_ok = nkb > 0 && quad < nchunks;
    float* dst = p.part + ((size_t)split * p.N + n_tile0) * p.m_pad + m_tile * 128 + quad * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
      float v[32];
      if (rows_ok) {
        tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
      } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) v[c] = 0.f;
      }
#pragma unroll
      for (int c = 0; c < 32; ++c) dst[(size_t)(c0 + c) * p.m_pad] = v[c];
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
}

// dW[n][m] (packed forward layout, row pitch ldw) = sum_s part[s][n][m]; threads run along m.
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ part, float* __restrict__ gw, int nsplit, int m_pad, int K, int N,
                    int ldw) {
  const unsigned e = blockIdx.x * 256u + threadIdx.x;
  if (e >= (unsigned)K * (unsigned)N) return;
  const unsigned n = e / (unsigned)K, m = e - n * (unsigned)K;
  const float* p = part + (size_t)n * m_pad + m;
  const size_t stride = (size_t)N * m_pad;
  float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
  int s = 0;
  for (; s + 4 <= nsplit; s += 4) {
    acc0 += p[(size_t)s * stride];
    acc1 += p[(size_t)(s + 1) * stride];
    acc2 += p[(size_t)(s + 2) * stride];
    acc3 += p[(size_t)(s + 3) * stride];
  }
  for (; s < nsplit; ++s) acc0 += p[(size_t)s * stride];
  gw[(size_t)n * ldw + m] = (acc0 + acc1) + (acc2 + acc3);
}

// Sum the split-K partials of up to 16 segments (layers' weight gradients, and bias-gradient partial rows
// written by the data-gradient epilogues) in one launch.  An output element is owned by SL = 1..8 threads
// ("split lanes", chosen per segment from its split count) that each add every SL-th partial with four loads
// in flight, then combine through shared memory in a fixed order -- a single thread walking 64-144 partials
// with dependent L2 latencies made this kernel 10 us.  Partials are n-major (part[s][n][m]) and consecutive
// threads run along m, so reads and the gw[n][m] writes are both coalesced.
struct ReduceSeg { const float* part; float* gw; int nsplit, m_pad, K, N, first_block, sl, vec; };
// Optional optimizer epilogue: the reduction is the last kernel of a step and touches every gradient element exactly
// once, so on steps without an exchange it can finish the step itself -- Nesterov/EAMSGD apply of THIS step's
// gradient, the look-ahead of the next step and the refresh of the tensor-core operand copies
// (asyncsgd/optim-msgd.lua:28-37, optim-eamsgd.lua:37-44,70) -- and the separate boundary kernel of the next step
// is not launched at all.  The float operations and their order per parameter are those of the boundary kernel
// (comm.cu EX_FUSED / optim.cu local_step), so trajectories stay bit-identical.  State convention: x is left at the
// parameters PROPER (after the apply), v already multiplied by the next momentum, and only the operand copy holds the
// look-ahead point x + v; `la_in` says that this is also the state the step started from.
struct OptEpi {
  float* x; float* v; const float* d;
  const int* pk2flat;        // packed (forward layout) index -> flat parameter index, -1 = padding
  float* wpk; const int* wmap2;   // operand copies: forward layout has the gradient's index, wmap2[i] = data-gradient slot
  const float* gpk;          // base of the packed gradient buffer (segment outputs point into it)
  float clr, l2wd, mom_next;
  int use_mom, has_d, la_in, enabled;
};
struct ReduceArgs { ReduceSeg seg[16]; int nseg; OptEpi opt; };

__device__ __forceinline__ void opt_epilogue_elem(const OptEpi& o, const float* gw_elem, float g) {
  const int i = o.pk2flat[gw_elem - o.gpk];
  if (i < 0) return;
  float x = o.x[i], v = o.use_mom ? o.v[i] : 0.f;
  if (o.la_in && o.use_mom) x = add_rn(x, v);                        // the point this step's gradient was taken at
  msgd_apply_elem(x, v, g, o.has_d ? o.d[i] : 0.f, o.clr, o.l2wd, o.use_mom != 0, o.has_d != 0);
  o.x[i] = x;                                                        // parameters proper
  float la = x;
  if (o.use_mom) {
    v = mul_rn(v, o.mom_next);
    o.v[i] = v;
    la = add_rn(x, v);                                               // next step's look-ahead point: operands only
  }
  o.wpk[gw_elem - o.gpk] = la;
  const int s2 = o.wmap2[i];
  if (s2 >= 0) o.wpk[s2] = la;
}

// V = 4: an output "element" is four consecutive m (16-byte loads and stores; K % 4 == 0, buffers 16-byte
// aligned -- every weight-gradient segment); V = 1: scalar (bias rows).
template <int V>
__device__ __forceinline__ void reduce_segment(const ReduceSeg& sg, float* red, const int block, const OptEpi& opt) {
  // sl is a power of two, so outputs-per-block is a shift; all index math in 32 bits (a 64-bit division per
  // thread cost more than the loads: host checks K*N < 2^31)
  const int opb_shift = 8 - (31 - __clz(sg.sl));
  const int opb = 1 << opb_shift;                      // outputs per block
  const int o = threadIdx.x & (opb - 1), sl = threadIdx.x >> opb_shift;
  const unsigned e = (unsigned)block * (unsigned)opb + (unsigned)o;
  const unsigned kq = (unsigned)sg.K / V;              // output elements per n row
  const bool live = e < kq * (unsigned)sg.N;
  float acc[4][V];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int v = 0; v < V; ++v) acc[i][v] = 0.f;
  if (live) {
    const unsigned n = e / kq, m = (e - n * kq) * V;   // m fastest: e*V is also the gw index
    const float* p = sg.part + (size_t)n * sg.m_pad + m;
    const size_t stride = (size_t)sg.N * sg.m_pad;
    auto add = [&](int i, int sp) {
      if (V == 4) {
        const float4 t = ld4(p + (size_t)sp * stride);
        acc[i][0] += t.x; acc[i][V > 1 ? 1 : 0] += t.y; acc[i][V > 2 ? 2 : 0] += t.z; acc[i][V > 3 ? 3 : 0] += t.w;
      } else {
        acc[i][0] += p[(size_t)sp * stride];
      }
    };
    int sp = sl;
    for (; sp + 3 * sg.sl < sg.nsplit; sp += 4 * sg.sl) {
      add(0, sp); add(1, sp + sg.sl); add(2, sp + 2 * sg.sl); add(3, sp + 3 * sg.sl);
    }
    for (; sp < sg.nsplit; sp += sg.sl) add(0, sp);
  }
  float t[V];
#pragma unroll
  for (int v = 0; v < V; ++v) t[v] = (acc[0][v] + acc[1][v]) + (acc[2][v] + acc[3][v]);
  if (sg.sl > 1) {
#pragma unroll
    for (int v = 0; v < V; ++v) red[v * 256 + threadIdx.x] = t[v];
    __syncthreads();
    if (sl == 0)
      for (int k = 1; k < sg.sl; ++k)
#pragma unroll
        for (int v = 0; v < V; ++v) t[v] += red[v * 256 + k * opb + o];
  }
  if (live && sl == 0) {
    if (V == 4) st4(sg.gw + (size_t)e * 4, make_float4(t[0], t[V > 1 ? 1 : 0], t[V > 2 ? 2 : 0], t[V > 3 ? 3 : 0]));
    else sg.gw[e] = t[0];
    if (opt.enabled) {
#pragma unroll
      for (int v = 0; v < V; ++v) opt_epilogue_elem(opt, sg.gw + (size_t)e * V + v, t[v]);
    }
  }
}

__global__ void __launch_bounds__(256) multi_wgrad_reduce_kernel(ReduceArgs a) {
  __shared__ float red[4 * 256];
  pdl_trigger();
  pdl_wait();
  int s = 0;
  while (s + 1 < a.nseg && (int)blockIdx.x >= a.seg[s + 1].first_block) ++s;
  const ReduceSeg sg = a.seg[s];
  if (sg.vec) reduce_segment<4>(sg, red, blockIdx.x - sg.first_block, a.opt);
  else reduce_segment<1>(sg, red, blockIdx.x - sg.first_block, a.opt);
}

// Column sums of up to 8 [rows][C] matrices in one launch (bias gradients of every layer).
// A warp covers 128 consecutive floats with float4 loads; C must be a multiple of 4 and <= 256.
struct ColsumSeg { const float* src; float* dst; long long rows; int C; int first_block; int nblk; };
struct ColsumArgs { ColsumSeg seg[8]; int nseg; };

__global__ void __launch_bounds__(256) multi_colsum_kernel(ColsumArgs a) {
  int s = 0;
  while (s + 1 < a.nseg && (int)blockIdx.x >= a.seg[s + 1].first_block) ++s;
  const ColsumSeg sg = a.seg[s];
  const int b = blockIdx.x - sg.first_block;
  const int c4 = sg.C >> 2;                 // float4 columns per row
  const long long total4 = sg.rows * c4;    // float4 elements
  // thread t always visits the same column group: stride is a multiple of c4
  long long stride = (long long)sg.nblk * 256;
  stride = (stride + c4 - 1) / c4 * c4;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  long long i0 = (long long)b * 256 + threadIdx.x;
  for (long long i = i0; i < total4; i += stride) {
    float4 v = ld4(sg.src + 4 * i);
    acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
  }
  // combine threads of the block that share a column group through shared memory
  __shared__ float4 sm[256];
  sm[threadIdx.x] = acc;
  __syncthreads();
  if ((int)threadIdx.x < c4) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = threadIdx.x; k < 256; k += c4) { t.x += sm[k].x; t.y += sm[k].y; t.z += sm[k].z; t.w += sm[k].w; }
    const int col = threadIdx.x;   // (b*256 + t) % c4 == t because 256 % c4 == 0
    atomicAdd(sg.dst + 4 * col + 0, t.x);
    atomicAdd(sg.dst + 4 * col + 1, t.y);
    atomicAdd(sg.dst + 4 * col + 2, t.z);
    atomicAdd(sg.dst + 4 * col + 3, t.w);
  }
}

template <int BLOCK_N>
static int launch_wgrad(const CUtensorMap& tx, const CUtensorMap& tz, const WgradParams& p, int m_tiles,
                        int nsplit, cudaStream_t st) {
  using L = WSmem<BLOCK_N>;
  dim3 grid(m_tiles, p.N / BLOCK_N, nsplit);
  if (p.x3) {
    static unsigned long long attr_done3 = 0;   // bit per device
    MPIT_CUDA_TRY(set_smem_limit_per_device(umma_wgrad_kernel<BLOCK_N, true>, L::kTotal, attr_done3));
    MPIT_CUDA_TRY(launch_k(umma_wgrad_kernel<BLOCK_N, true>, grid, dim3(192), (size_t)L::kTotal, st, tx, tz, p));
    return 0;
  }
  static unsigned long long attr_done = 0;   // bit per device
  MPIT_CUDA_TRY(set_smem_limit_per_device(umma_wgrad_kernel<BLOCK_N, false>, L::kTotal, attr_done));
  MPIT_CUDA_TRY(launch_k(umma_wgrad_kernel<BLOCK_N, false>, grid, dim3(192), (size_t)L::kTotal, st, tx, tz, p));
  return 0;
}

}  // namespace mpit

using namespace mpit;

extern "C" {

// geom11 (HOST): {tw,th,tn,tiles_w,tiles_h,kw_it,cb,total_chunks,num_kb,B,N}.  `part` must hold
// nsplit * m_pad * N floats, m_pad = ceil(total_chunks/4)*128.
int mpit_umma_wgrad(const void* tmap_x128, const void* tmap_dz128, const int* geom11, float* part,
                    const int* crop, int* err, int block_n, int nsplit, void* stream) {
  WgradParams p{};
  p.tw = geom11[0]; p.th = geom11[1]; p.tn = geom11[2]; p.tiles_w = geom11[3]; p.tiles_h = geom11[4];
  p.kw_it = geom11[5]; p.cb = geom11[6]; p.total_chunks = geom11[7]; p.num_kb = geom11[8];
  p.B = geom11[9]; p.N = geom11[10];
  if (p.tw * p.th * p.tn != 64 || p.N % block_n != 0 || nsplit < 1) return MPIT_E_ARG;
  int m_tiles = (p.total_chunks + 3) / 4;
  p.m_pad = m_tiles * 128;
  p.part = part; p.crop = crop; p.err = err;
  p.x3 = g_x3_mode;
  CUtensorMap tx, tz;
  memcpy(&tx, tmap_x128, sizeof(tx));
  memcpy(&tz, tmap_dz128, sizeof(tz));
  cudaStream_t st = (cudaStream_t)stream;
  switch (block_n) {
    case 32: return launch_wgrad<32>(tx, tz, p, m_tiles, nsplit, st);
    case 64: return launch_wgrad<64>(tx, tz, p, m_tiles, nsplit, st);
    case 128: return launch_wgrad<128>(tx, tz, p, m_tiles, nsplit, st);
    default: return MPIT_E_ARG;
  }
}

// gw[n*ldw + m] = sum_s part[s][n][m] for m < K, n < N.
int mpit_wgrad_reduce(const float* part, float* gw, int nsplit, int m_pad, int K, int N, int ldw, void* stream) {
  const long long total = (long long)K * N;
  if (total <= 0) return 0;
  if (total >= (1ll << 31)) return MPIT_E_ARG;
  wgrad_reduce_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(part, gw, nsplit, m_pad, K, N,
                                                                                         ldw);
  MPIT_LAUNCH_CHECK();
  return 0;
}

// Bias gradients of up to 8 layers in one launch. srcs/dsts are HOST arrays of device pointers;
// dst vectors are zeroed here (memset nodes) and accumulated with atomics.
int mpit_multi_colsum(int nseg, const unsigned long long* srcs, const unsigned long long* dsts,
                      const long long* rows, const int* cols, void* stream) {
  if (nseg < 1 || nseg > 8) return MPIT_E_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  ColsumArgs a{};
  a.nseg = nseg;
  int blocks = 0;
  for (int i = 0; i < nseg; ++i) {
    if (cols[i] % 4 != 0 || 256 % (cols[i] / 4) != 0) return MPIT_E_ARG;
    a.seg[i].src = reinterpret_cast<const float*>(srcs[i]);
    a.seg[i].dst = reinterpret_cast<float*>(dsts[i]);
    a.seg[i].rows = rows[i];
    a.seg[i].C = cols[i];
    a.seg[i].first_block = blocks;
    long long total4 = rows[i] * (cols[i] / 4);
    long long nb = (total4 + 256 * 8 - 1) / (256 * 8);   // ~8 float4 per thread
    if (nb > 296) nb = 296;
    if (nb < 1) nb = 1;
    a.seg[i].nblk = (int)nb;
    blocks += (int)nb;
    MPIT_CUDA_TRY(cudaMemsetAsync(a.seg[i].dst, 0, sizeof(float) * (size_t)cols[i], st));
  }
  multi_colsum_kernel<<<blocks, 256, 0, st>>>(a);
  MPIT_LAUNCH_CHECK();
  return 0;
}

// Split-K reduction of up to 16 segments in one launch; every array is a HOST array of nseg entries.
// Segment i: gw_i[n*K_i + m] = sum_{s < nsplit_i} part_i[(s*N_i + n)*m_pad_i + m], m < K_i, n < N_i.
// (A bias gradient summed from T tile rows of width G*C is the segment {nsplit=T*G, m_pad=1, K=1, N=C}.)
static int multi_reduce_impl(int nseg, const unsigned long long* parts, const unsigned long long* gws,
                             const int* nsplit, const int* m_pad, const int* K, const int* N, const OptEpi* opt,
                             void* stream) {
  if (nseg < 1 || nseg > 16) return MPIT_E_ARG;
  ReduceArgs a{};
  a.nseg = nseg;
  if (opt) a.opt = *opt;
  int blocks = 0;
  for (int i = 0; i < nseg; ++i) {
    a.seg[i].part = reinterpret_cast<const float*>(parts[i]);
    a.seg[i].gw = reinterpret_cast<float*>(gws[i]);
    a.seg[i].nsplit = nsplit[i]; a.seg[i].m_pad = m_pad[i]; a.seg[i].K = K[i]; a.seg[i].N = N[i];
    if ((long long)K[i] * N[i] >= (1ll << 31)) return MPIT_E_ARG;
    a.seg[i].first_block = blocks;
    int sl = 1;
    while (sl < 8 && sl * 4 < nsplit[i]) sl *= 2;      // ~4+ partials per split lane
    a.seg[i].sl = sl;
    const int vec = (K[i] % 4 == 0 && m_pad[i] % 4 == 0 && ((parts[i] | gws[i]) & 15) == 0) ? 1 : 0;
    a.seg[i].vec = vec;
    const long long opb = 256 / sl;
    blocks += (int)(((long long)K[i] * N[i] / (vec ? 4 : 1) + opb - 1) / opb);
  }
  MPIT_CUDA_TRY(launch_k(multi_wgrad_reduce_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, a));
  return 0;
}

int mpit_multi_wgrad_reduce(int nseg, const unsigned long long* parts, const unsigned long long* gws,
                            const int* nsplit, const int* m_pad, const int* K, const int* N, void* stream) {
  return multi_reduce_impl(nseg, parts, gws, nsplit, m_pad, K, N, nullptr, stream);
}

/* The same reduction that also FINISHES the training step (see OptEpi): every gradient element it produces is applied
 * to its parameter (x, v; d subtracted when has_d), the next step's look-ahead is folded into the operand copies
 * wpk (forward slot = the gradient's own packed index, data-gradient slot = wmap2[flat index]).  Every output segment
 * must lie inside the packed gradient buffer `gpk`; pk2flat maps a packed index to its flat parameter (-1: none). */
int mpit_multi_wgrad_reduce_opt(int nseg, const unsigned long long* parts, const unsigned long long* gws,
                                const int* nsplit, const int* m_pad, const int* K, const int* N, float* x, float* v,
                                const float* d, const int* pk2flat, float* wpk, const int* wmap2, const float* gpk,
                                float clr, float l2wd, float mom_next, int use_mom, int has_d, int la_in, void* stream) {
  if (!x || !pk2flat || !wpk || !wmap2 || !gpk || (use_mom && !v) || (has_d && !d)) return MPIT_E_ARG;
  OptEpi o{};
  o.x = x; o.v = v; o.d = d; o.pk2flat = pk2flat; o.wpk = wpk; o.wmap2 = wmap2; o.gpk = gpk;
  o.clr = clr; o.l2wd = l2wd; o.mom_next = mom_next; o.use_mom = use_mom; o.has_d = has_d; o.la_in = la_in;
  o.enabled = 1;
  return multi_reduce_impl(nseg, parts, gws, nsplit, m_pad, K, N, &o, stream);
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// Weight gradient for callers that cannot allocate device memory.
//
// mpit_umma_wgrad needs a split-K workspace and normally the caller provides it.  Some callers have
// no way to do that: a host program or foreign-language binding that only holds this C ABI and owns
// no device allocator; code running next to a framework whose caching allocator has already claimed
// the whole device heap, where a raw cudaMalloc would fail or evict; and latency-sensitive loops that
// must not execute cudaMalloc/cudaFree because those calls synchronise the whole device.  For them
// the workspace lives in the library image itself: a fixed array in device global memory that every
// GPU loading this library carries until the process exits (MPIT_NOALLOC_SCRATCH_FLOATS floats,
// 8 MiB).  Requests that would need more are refused, nothing is ever allocated or freed here.
//
// `g_noalloc_scratch` below is, in host code, only a handle for the symbol: its device address is
// looked up once per device with cudaGetSymbolAddress (with that device current) and it is that
// address which the kernels receive.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_noalloc_scratch[MPIT_NOALLOC_SCRATCH_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};

// One-time set-up for the CURRENT device: resolve the scratch symbol's device address and raise the
// dynamic shared-memory limit of the kernel instantiations on this device.
int noalloc_setup_current_device(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return MPIT_E_ARG;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return MPIT_E_UNSUPPORTED;
    e = cudaFuncSetAttribute(mpit::umma_wgrad_kernel<32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             mpit::WSmem<32>::kTotal);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(mpit::umma_wgrad_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             mpit::WSmem<64>::kTotal);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(mpit::umma_wgrad_kernel<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             mpit::WSmem<128>::kTotal);
    if (e != cudaSuccess) return (int)e;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int mpit_umma_wgrad_noalloc(const void* tmap_x128, const void* tmap_dz128, const int* geom11,
                                       float* gw, int ldw, const int* crop, int* err, int block_n,
                                       int nsplit, void* stream) {
  if (!geom11 || !gw || nsplit < 1) return MPIT_E_ARG;
  const int total_chunks = geom11[7], N = geom11[10];
  if (total_chunks < 1 || N < 1) return MPIT_E_ARG;
  const int m_pad = ((total_chunks + 3) / 4) * 128;
  const unsigned long long need = (unsigned long long)nsplit * (unsigned long long)m_pad * (unsigned long long)N;
  if (need > (unsigned long long)MPIT_NOALLOC_SCRATCH_FLOATS) {
    fprintf(stderr,
            "mpit_umma_wgrad_noalloc: this problem needs %llu scratch floats (nsplit=%d x m_pad=%d x N=%d) but the "
            "built-in scratch holds %u; lower nsplit or call mpit_umma_wgrad with your own workspace\n",
            need, nsplit, m_pad, N, (unsigned)MPIT_NOALLOC_SCRATCH_FLOATS);
    return MPIT_E_SCRATCH;
  }
  float* scratch = nullptr;
  int rc = noalloc_setup_current_device(&scratch);
  if (rc) return rc;
  // Every partial tile that the reduction reads is written in full by the gradient kernel of this
  // same call, so whatever an earlier call left in the scratch is never observed: nothing to clear.
  rc = mpit_umma_wgrad(tmap_x128, tmap_dz128, geom11, scratch, crop, err, block_n, nsplit, stream);
  if (rc) return rc;
  return mpit_wgrad_reduce(scratch, gw, nsplit, m_pad, total_chunks * 32, N, ldw, stream);
}
