// Implicit-GEMM convolution / linear layers on the 5th-gen tensor cores (sm_100a):
//
//   D[128 x BLOCK_N] (TMEM, fp32) = sum_kb  A_kb[128 x 32] (smem, tf32)  *  B_kb[BLOCK_N x 32]^T
//
// * A rows are output pixels (n, oh, ow) of an NHWC activation tensor.  One k-block is one
//   (kernel-row, kernel-col, 32-channel) slice, fetched by TMA as a 4-D box whose (w, h) coordinates
//   are shifted by the kernel offset -- im2col never exists in memory.  Forward uses +offsets,
//   the data-gradient uses -offsets on the output gradient with TMA zero fill standing in for the
//   "full" padding.  The random/centre crop of the first layer is just another coordinate offset
//   (asyncsgd/goot.lua:177-178,283-285).  A Linear layer is the 1x1 case.
// * B rows are output channels of the packed weight matrix, fetched as a 2-D TMA box.
// * Both operands land in 128-byte-swizzled shared memory and are consumed by tcgen05.mma
//   (kind::tf32, M=128, N=BLOCK_N, K=8 per instruction) issued by one elected thread.
// * Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM allocator), warps 2..5 =
//   epilogue (tcgen05.ld -> bias/ReLU/2x2-maxpool/dropout or gradient masking -> 128-byte stores).
//
// Layers: asyncsgd/models/7-layers-cnn.lua:8-24 (SpatialConvolution+ReLU+SpatialMaxPooling,
// Linear+ReLU+Dropout, Linear).
#include <stdio.h>
#include "mpit_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include <string.h>
#include <stdlib.h>

namespace mpit {
using namespace umma;

enum EpiMode : int { EPI_STORE = 0, EPI_POOL = 1, EPI_MASK = 2, EPI_EXPAND = 3, EPI_MAXTIME = 4 };
// `relu` doubles as the activation selector: 0 none, 1 ReLU, 2 Tanh (forward: tanh(x); EPI_MASK: the
// gradient is multiplied by 1 - aux^2, aux being the stored tanh output).

struct IgemmParams {
  int nsub, tw, th, tn;      // sub-box: tw*th*tn rows, nsub sub-boxes per 128-row tile
  int tiles_w, tiles_h;      // sub-tiles per image
  int kw_it, cb;             // k-block -> (kh, kw, channel block)
  int num_kb;
  int sgn;                   // +1 forward, -1 data gradient
  int B, OH, OW;             // extent of the GEMM M side: rows are (n, oh, ow)
  int N;                     // total output columns
  int mode, relu, dropout;
  float scale;
  const float* bias;
  float* out;
  unsigned char* idx;
  const float* aux;
  const unsigned char* aux_idx;
  const int* crop;
  const unsigned long long* rng;
  int* err;                  // set to 1 when a bounded pipeline wait expires
  float* colpart;            // optional [m_tiles][N]: per-tile column sums of the values written (MASK / EXPAND)
  int stages;                // smem pipeline depth (2..kMaxStages), chosen per launch
  int x3;                    // 0: plain TF32; 1 / 2: 3xTF32 with the operand split done in shared memory (umma.cuh)
  int sb, kb_split, accum;   // pulled GEMM only: depth of the weight ring, k-blocks per split-K slice, red.add epilogue
  long long* dbg;            // optional per-CTA timeline (16 clock64 stamps per CTA), null in production
};

constexpr int kMaxStages = 12;
constexpr int kABytes = 128 * 128;  // 128 rows x 32 tf32
constexpr int kSmemBudget = 200 * 1024;

template <int BLOCK_N>
struct SmemLayout {
  static constexpr int kBBytes = BLOCK_N * 128;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kMaxFit = (kSmemBudget / kStageBytes) < kMaxStages ? (kSmemBudget / kStageBytes) : kMaxStages;
  // + barriers + align slack; `mul` = 2 in 3xTF32 mode (every stage also holds the lo tiles)
  static constexpr int total(int stages, int mul = 1) { return stages * kStageBytes * mul + 512 + 1024; }
};

// Out of line on purpose.  An epilogue runs once per CTA, so its code is fetched cold; with tanh and Philox
// inlined into each of the 8 unrolled store steps the taken path hopped over ~16 KB of skipped code per
// step and the instruction misses cost > 1 us per tile.  As functions their lines are fetched once.
__device__ __noinline__ float4 tanh4_ool(float4 x) { return make_float4(tanhf(x.x), tanhf(x.y), tanhf(x.z), tanhf(x.w)); }
// Keep-bits of four float4 groups at once: four independent Philox chains interleave, so the call costs
// about one chain's latency.  Nibble j of the result belongs to idx_j (bit k = element k of the group).
__device__ __noinline__ uint32_t keep4_ool(unsigned long long seed, uint32_t step, unsigned long long i0,
                                           unsigned long long i1, unsigned long long i2, unsigned long long i3) {
  const uint4 r0 = Philox::gen(seed, i0, step), r1 = Philox::gen(seed, i1, step);
  const uint4 r2 = Philox::gen(seed, i2, step), r3 = Philox::gen(seed, i3, step);
  auto nib = [](const uint4& r) { return (r.x & 1u) | ((r.y & 1u) << 1) | ((r.z & 1u) << 2) | ((r.w & 1u) << 3); };
  return nib(r0) | (nib(r1) << 4) | (nib(r2) << 8) | (nib(r3) << 12);
}

__device__ __forceinline__ void subtile_origin(const IgemmParams& p, int s, int& n0, int& oh0, int& ow0) {
  // integer division is ~100 clocks of dependent latency on the issue path: skip it in the common cases
  // (one sub-tile per image group, or a single column of sub-tiles)
  const int per_img = p.tiles_w * p.tiles_h;
  int grp = s, r = 0;
  if (per_img != 1) { grp = s / per_img; r = s - grp * per_img; }
  int th_i = r, tw_i = 0;
  if (p.tiles_w != 1) { th_i = r / p.tiles_w; tw_i = r - th_i * p.tiles_w; }
  n0 = grp * p.tn;
  oh0 = th_i * p.th;
  ow0 = tw_i * p.tw;
}

// Epilogue of one 128-row accumulator tile for the modes that store row-major tiles (STORE / MASK / EXPAND); shared
// by the streaming kernel below and the resident-input kernel (umma_resident.cu).  Called by the four epilogue
// warps.  (n, oh, ow, valid) describe the output row of this THREAD (TMEM lane); `tile` indexes the colpart row.
// `wait_acc`: wait for `tmem_full` (phase 0) after the side inputs have been requested.
template <int BLOCK_N, int MODE, bool ACCUM = false>
__device__ __forceinline__ void coalesced_epilogue(const IgemmParams& p, uint8_t* smem, const uint32_t tmem_base,
                                                   uint64_t* tmem_full, const bool wait_acc, const int quad, const int lane,
                                                   const int row, const int n, const int oh, const int ow, const bool valid,
                                                   const int n_tile0, const int tile, const unsigned long long seed,
                                                   const unsigned long long step, const float* bias_s, long long* dbg) {
  // ---- coalesced epilogue (STORE / MASK / EXPAND) ----
  // TMEM hands every lane one output ROW (32 columns = 128 B).  Storing (or loading the side inputs) in
  // that arrangement touches 32 different 128-byte lines per warp instruction -- measured 5.4 us for the
  // un-pooling epilogue of one 128x64 tile.  So each 32x32 block is transposed through a warp-private,
  // XOR-swizzled 4 KB tile in (by now idle) pipeline memory: afterwards 8 adjacent lanes cover one
  // 128-byte row segment and a warp instruction touches 4 full lines.
  __shared__ long long rowoff_s[128];   // element offset of each tile row's output row, -1 = out of range
  __shared__ long long rowdst_s[128];   // EXPAND: offset of the 2x2 window's top-left pixel
  __shared__ __align__(16) float colsum_s[2 * 128];
  const size_t row_base = (((size_t)n * p.OH + oh) * p.OW + ow) * p.N + n_tile0;
  rowoff_s[row] = valid ? (long long)row_base : -1;
  if (MODE == EPI_EXPAND)
    rowdst_s[row] = (long long)((((size_t)n * (2 * p.OH) + 2 * oh) * (2 * p.OW) + 2 * ow) * p.N + n_tile0);
  __syncwarp();
  const int sub = lane >> 3, ch = lane & 7;   // row within a group of 4, 16-byte chunk within the row
  long long off[8], dst[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    off[i] = rowoff_s[quad * 32 + i * 4 + sub];
    dst[i] = MODE == EPI_EXPAND ? rowdst_s[quad * 32 + i * 4 + sub] : 0;
  }
  // side inputs (stored activation, pool index) are fetched BEFORE the accumulator is waited for
  float4 ax[8];
  uint32_t ai[8];
  auto fetch_side = [&](int c0) {
    if (MODE == EPI_STORE) return;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (off[i] < 0) continue;
      ax[i] = ld4(p.aux + off[i] + c0 + 4 * ch);
      if (MODE == EPI_EXPAND) ai[i] = *reinterpret_cast<const uint32_t*>(p.aux_idx + off[i] + c0 + 4 * ch);
    }
  };
  fetch_side(0);
  // touch the output rows now: the first access of an SM to a page costs ~1 us (translation miss), which
  // otherwise lands on the critical path between "accumulator ready" and the first store
  if (off[0] >= 0) {
    const float* o0 = MODE == EPI_EXPAND ? p.out + (size_t)dst[0] : p.out + (size_t)off[0];
    asm volatile("prefetch.global.L2 [%0];" ::"l"(o0 + 4 * ch));
  }
  if (wait_acc) {
    if (!mbar_wait(tmem_full, 0)) { if (p.err) *p.err = 3; }
    if (dbg && threadIdx.x == 64) dbg[5] = clock64();   // accumulator ready
    tc_fence_after();
    asm volatile("bar.sync 1, 128;" ::: "memory");      // bias_s written by all four epilogue warps
  }
  const uint32_t xpose = smem_u32(smem) + quad * 4096;   // stage 0 is idle: every MMA has completed
  const size_t up = (size_t)2 * p.OW * p.N;                      // EXPAND: one image row down
#pragma unroll 1
  for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
    float v[32];
    if (c0) fetch_side(c0);
    tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
    if (dbg && threadIdx.x == 64 && c0 == 0) dbg[11] = clock64();   // first TMEM read done
#pragma unroll
    for (int c4 = 0; c4 < 8; ++c4)
      sts4(xpose + (lane * 8 + (c4 ^ (lane & 7))) * 16, v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]);
    __syncwarp();
    float4 xs[8];     // all eight reads first: explicit shared-space loads, nothing for them to alias with
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = i * 4 + sub;
      xs[i] = lds4(xpose + (r * 8 + (ch ^ (r & 7))) * 16);
    }
    if (dbg && threadIdx.x == 64 && c0 == 0) dbg[12] = clock64();   // first block transposed
    const float4 b4 = *reinterpret_cast<const float4*>(&bias_s[c0 + 4 * ch]);
    float4 csum = make_float4(0.f, 0.f, 0.f, 0.f);   // this lane's 4 columns, summed over its 8 rows
    uint32_t keep = 0;
    if (MODE == EPI_STORE && p.dropout) {
      auto gi = [&](int i) { return (unsigned long long)(((size_t)off[i] + c0 + 4 * ch) >> 2); };
      keep = keep4_ool(seed, (uint32_t)step, gi(0), gi(1), gi(2), gi(3)) |
             (keep4_ool(seed, (uint32_t)step, gi(4), gi(5), gi(6), gi(7)) << 16);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float4 x = xs[i];
      if (off[i] < 0) continue;
      const size_t e0 = (size_t)off[i] + c0 + 4 * ch;
      if (MODE == EPI_STORE) {
        x.x += b4.x; x.y += b4.y; x.z += b4.z; x.w += b4.w;
        if (p.relu == 1) {
          x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
        } else if (p.relu == 2) {
          x = tanh4_ool(x);
        }
        if (p.dropout) {
          const uint32_t kb = keep >> (4 * i);
          x.x = (kb & 1u) ? x.x * 2.f : 0.f;
          x.y = (kb & 2u) ? x.y * 2.f : 0.f;
          x.z = (kb & 4u) ? x.z * 2.f : 0.f;
          x.w = (kb & 8u) ? x.w * 2.f : 0.f;
        }
        if (ACCUM && p.accum) red_add4(p.out + e0, x);   // split-K slice: partial tile into the zeroed output
        else st4(p.out + e0, x);
      } else if (MODE == EPI_MASK) {
        const float4 h = ax[i];
        float4 o;
        if (p.relu == 2) {   // d tanh: g * (1 - h^2)
          o.x = x.x * (1.f - h.x * h.x) * p.scale;
          o.y = x.y * (1.f - h.y * h.y) * p.scale;
          o.z = x.z * (1.f - h.z * h.z) * p.scale;
          o.w = x.w * (1.f - h.w * h.w) * p.scale;
        } else {
          o.x = h.x > 0.f ? x.x * p.scale : 0.f;
          o.y = h.y > 0.f ? x.y * p.scale : 0.f;
          o.z = h.z > 0.f ? x.z * p.scale : 0.f;
          o.w = h.w > 0.f ? x.w * p.scale : 0.f;
        }
        st4(p.out + e0, o);
        csum.x += o.x; csum.y += o.y; csum.z += o.z; csum.w += o.w;
      } else {  // EPI_EXPAND: un-pool + ReLU mask; tile rows are pooled positions
        const float4 h = ax[i];
        const uint32_t bsel = ai[i];   // four arg-max codes (0..3), one byte per column
        float* d0 = p.out + (size_t)dst[i] + c0 + 4 * ch;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float4 o;
          o.x = ((bsel & 0xffu) == (uint32_t)q && h.x > 0.f) ? x.x * p.scale : 0.f;
          o.y = (((bsel >> 8) & 0xffu) == (uint32_t)q && h.y > 0.f) ? x.y * p.scale : 0.f;
          o.z = (((bsel >> 16) & 0xffu) == (uint32_t)q && h.z > 0.f) ? x.z * p.scale : 0.f;
          o.w = ((bsel >> 24) == (uint32_t)q && h.w > 0.f) ? x.w * p.scale : 0.f;
          st4(d0 + (q >> 1) * up + (q & 1) * p.N, o);
        }
        // exactly one of the four window positions receives the gradient (when the unit was active)
        csum.x += h.x > 0.f ? x.x * p.scale : 0.f;
        csum.y += h.y > 0.f ? x.y * p.scale : 0.f;
        csum.z += h.z > 0.f ? x.z * p.scale : 0.f;
        csum.w += h.w > 0.f ? x.w * p.scale : 0.f;
      }
    }
    if ((MODE == EPI_MASK || MODE == EPI_EXPAND) && p.colpart) {
      // Bias gradient of the layer below = column sums of what was just written.  Fixed-order reduction:
      // 8 rows in registers, 4 row groups by shuffle, 4 warps through shared memory -> one partial row per
      // tile, summed over tiles by the split-K reduce kernel.  No atomics, no zeroing, and the 19 MB
      // gradient tensor is not read again by a separate column-sum kernel.
      csum.x += __shfl_xor_sync(0xffffffffu, csum.x, 8);  csum.y += __shfl_xor_sync(0xffffffffu, csum.y, 8);
      csum.z += __shfl_xor_sync(0xffffffffu, csum.z, 8);  csum.w += __shfl_xor_sync(0xffffffffu, csum.w, 8);
      csum.x += __shfl_xor_sync(0xffffffffu, csum.x, 16); csum.y += __shfl_xor_sync(0xffffffffu, csum.y, 16);
      csum.z += __shfl_xor_sync(0xffffffffu, csum.z, 16); csum.w += __shfl_xor_sync(0xffffffffu, csum.w, 16);
      float* cs = colsum_s + ((c0 >> 5) & 1) * 128;      // double-buffered by iteration parity
      if (sub == 0) *reinterpret_cast<float4*>(cs + quad * 32 + 4 * ch) = csum;
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (quad == 0)
        p.colpart[(size_t)tile * p.N + n_tile0 + c0 + lane] = (cs[lane] + cs[32 + lane]) + (cs[64 + lane] + cs[96 + lane]);
    }
    if (dbg && threadIdx.x == 64 && c0 == 0) dbg[13] = clock64();   // first iteration's stores issued
    __syncwarp();
  }
}

// Body shared by the plain kernel and the pull-fused kernel.  `tmap_a_p`/`tmap_b_p` point into
// kernel-parameter space; `b_row0` is the row coordinate of this CTA's B tile inside *its* tensor map
// (equal to the output column offset unless B lives in a peer's shard).
// MODE (the epilogue) is a template parameter: with all five epilogues, 32-column unrolling, tanh and Philox
// inlined in one body the kernel was ~100 KB of straight-line code that every CTA walks once, and the cold
// instruction fetches cost more than the math (the "math" phase of a 32-column epilogue measured 1.2 us).
// X3 (compile time, so that the TF32 kernels carry none of its code): 3xTF32 mode, see umma.cuh.
template <int BLOCK_N, int MODE, bool X3 = false>
__device__ __forceinline__ void igemm_body(const CUtensorMap* tmap_a_p, const CUtensorMap* tmap_b_p, const int b_row0,
                                           const IgemmParams& p) {
  const CUtensorMap& tmap_a = *tmap_a_p;
  const CUtensorMap& tmap_b = *tmap_b_p;
  using L = SmemLayout<BLOCK_N>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int kStages = p.stages;
  const int x3 = X3 ? p.x3 : 0;
  constexpr int stage_bytes = X3 ? 2 * L::kStageBytes : L::kStageBytes;  // [A | B] or [A | B | A_lo | B_lo]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kStages * stage_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* split_bar = empty_bar + kMaxStages;     // 3xTF32: lo tiles of the stage are written (128 arrivals)
  uint64_t* tmem_full = split_bar + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int n_tile0 = blockIdx.y * BLOCK_N;
  const int rows_per_sub = 128 / p.nsub;

  long long* dbg = p.dbg ? p.dbg + 16 * (blockIdx.y * gridDim.x + blockIdx.x) : nullptr;
  pdl_trigger();   // let the next kernel of the step begin its own set-up now
  // Everything that depends on the kernel parameters only -- tile origins (integer divisions), each producer's
  // first pipeline slot and tap -- is computed here, BEFORE the wait on the previous kernel, so that under
  // programmatic dependent launch it costs nothing (it used to be ~0.4 us between the wait and the first load).
  int org_n[2], org_h[2], org_w[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    org_n[j] = org_h[j] = org_w[j] = 0;
    if (j < p.nsub) subtile_origin(p, tile * p.nsub + j, org_n[j], org_h[j], org_w[j]);
  }
  // 3xTF32: the epilogue warps split the landed tiles instead of issuing loads; the MMA side takes three times as
  // long per k-block, so one issuing thread keeps up
  const int kProducers = X3 ? 1 : (kStages < 5 ? kStages : 5);
  const int pi = warp == 0 ? 0 : warp - 1;          // warps 0,2,3,4,5 -> 0,1,2,3,4 (warp 1 issues the MMAs)
  int stage = 0, phase = 0, kh = 0, kw = 0, cblk = 0;
  auto advance = [&](int by) {
    stage += by;
    while (stage >= kStages) { stage -= kStages; phase ^= 1; }
    cblk += by;
    while (cblk >= p.cb) { cblk -= p.cb; if (++kw == p.kw_it) { kw = 0; ++kh; } }
  };
  if (warp != 1) advance(pi);
  if (dbg && threadIdx.x == 0) dbg[0] = clock64();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    for (int i = 0; i < kStages; ++i) {
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
  pdl_wait();      // everything above overlapped the previous kernel; its results are visible from here on
  if (dbg && threadIdx.x == 0) dbg[1] = clock64();   // prologue done
  // The epilogue's bias row goes to shared memory now, so its global-load latency hides under the main loop
  // (fetched after the accumulator was ready, 32 scalar loads cost the fc layers > 2 us each).
  __shared__ __align__(16) float bias_s[256];
  if (warp >= 2)
    for (int c = threadIdx.x - 64; c < BLOCK_N; c += 128) bias_s[c] = p.bias ? p.bias[n_tile0 + c] : 0.f;

  // ===== TMA producers =====
  // Issuing a k-block (wait for the slot, expect_tx, 2-3 tensor loads with their coordinate arithmetic) is a
  // ~400-clock dependent chain in ONE thread -- measured (tools/dev/tma_feed_probe.cu) to be what bounded the
  // operand feed at ~35 B/clk/SM while the TMA unit itself sustains > 120 B/clk/SM.  So k-blocks are dealt
  // round-robin to kProducers issuing threads: the elected lane of warp 0 and of the four epilogue warps,
  // which have nothing else to do until the accumulator is complete.  No integer division on the issue path.
  // Never more producers than stages: a producer that has passed its wait for k-block kb knows k-block
  // kb - kStages is consumed; its next one (kb + NP) needs kb + NP - 2*kStages consumed for the parity test
  // to be unambiguous, which in-order consumption guarantees iff NP <= kStages.
  if (warp != 1 && pi < kProducers && elect_one()) {
    int n0[2], oh0[2], ow0[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      n0[j] = org_n[j]; oh0[j] = org_h[j]; ow0[j] = org_w[j];
      // (in MAXTIME mode `crop` carries per-sentence frame counts for the epilogue, not offsets)
      if (j < p.nsub && p.crop && MODE != EPI_MAXTIME && n0[j] < p.B) {
        oh0[j] += p.crop[2 * n0[j]];
        ow0[j] += p.crop[2 * n0[j] + 1];
      }
    }
    const int sub_bytes = rows_per_sub * 128;
    if (dbg && warp == 0) dbg[8] = clock64();   // producer set-up done
    for (int kb = pi; kb < p.num_kb; kb += kProducers) {
      // the first pass over the ring needs no wait: those slots have never been used
      if (kb >= kStages && !mbar_wait(&empty_bar[stage], phase ^ 1)) { if (p.err) *p.err = 1; break; }
      if (dbg && kb == 0) dbg[9] = clock64();   // first slot free
      uint8_t* a_dst = smem + stage * stage_bytes;
      mbar_expect_tx(&full_bar[stage], L::kStageBytes);
      tma_load_4d(a_dst, &tmap_a, &full_bar[stage], cblk * 32, ow0[0] + p.sgn * kw, oh0[0] + p.sgn * kh, n0[0]);
      if (p.nsub == 2)
        tma_load_4d(a_dst + sub_bytes, &tmap_a, &full_bar[stage], cblk * 32, ow0[1] + p.sgn * kw, oh0[1] + p.sgn * kh,
                    n0[1]);
      tma_load_2d(a_dst + kABytes, &tmap_b, &full_bar[stage], kb * 32, b_row0);
      if (dbg && kb == 0) dbg[10] = clock64();  // first k-block's loads issued
      advance(kProducers);
    }
    if (dbg && warp == 0) dbg[2] = clock64();   // warp 0's last TMA issued
  }
  __syncwarp();
  if (X3 && warp >= 2) {
    // ===== 3xTF32: operand split, warps 2..5 =====
    const int tid = threadIdx.x - 64;
    int st = 0, ph = 0;
    for (int kb = 0; kb < p.num_kb; ++kb) {
      if (!mbar_wait(&full_bar[st], ph)) { if (p.err) *p.err = 4; break; }
      const uint32_t base = smem_u32(smem) + st * stage_bytes;
      split_tile(base, base + L::kStageBytes, L::kStageBytes, tid, 128, x3);
      fence_proxy_async_smem();            // generic-proxy writes -> visible to the tensor core's async-proxy reads
      mbar_arrive(&split_bar[st]);
      if (++st == kStages) { st = 0; ph ^= 1; }
    }
  }
  if (warp == 0) {
    // nothing further: warp 0 only issues loads
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_tf32(128, BLOCK_N, 0, 0);
      // descriptor of stage 0; later stages / k-steps only add to the 14-bit address field (smem < 256 KB)
      const uint64_t desc0 = make_smem_desc(smem_u32(smem), 16, 1024);
      uint64_t ad = desc0;
      int stage = 0, phase = 0;
      for (int kb = 0; kb < p.num_kb; ++kb) {
        if (!mbar_wait(X3 ? &split_bar[stage] : &full_bar[stage], phase)) { if (p.err) *p.err = 2; break; }
        if (dbg && kb == 0) dbg[3] = clock64();   // first operands landed
        tc_fence_after();
        const uint64_t bd = ad + (kABytes >> 4);
        if (X3) {
          const uint64_t lo = (uint64_t)(L::kStageBytes >> 4);     // lo tiles sit one [A | B] block further
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            mma_tf32(tmem_base, ad + lo + 2 * k, bd + 2 * k, idesc, (kb | k) ? 1u : 0u);     // A_lo * B_hi
            mma_tf32(tmem_base, ad + 2 * k, bd + lo + 2 * k, idesc, 1u);                     // A_hi * B_lo
            mma_tf32(tmem_base, ad + 2 * k, bd + 2 * k, idesc, 1u);                          // A_hi * B_hi
          }
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_tf32(tmem_base, ad + 2 * k, bd + 2 * k, idesc, (kb | k) ? 1u : 0u);
        }
        mma_commit(&empty_bar[stage]);   // frees the smem slot once these MMAs have read it
        ad += (stage_bytes >> 4);
        if (++stage == kStages) { stage = 0; phase ^= 1; ad = desc0; }
      }
      mma_commit(tmem_full);             // accumulator complete
      if (dbg) dbg[4] = clock64();       // last MMA issued
    }
  } else {
    // ===== epilogue: warps 2..5, TMEM lane quadrant = warp % 4 =====
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    const int j = row / rows_per_sub, rr = row - j * rows_per_sub;
    const int n0 = org_n[j], oh0 = org_h[j], ow0 = org_w[j];
    const int w_ = rr % p.tw, h_ = (rr / p.tw) % p.th, nn = rr / (p.tw * p.th);
    const int n = n0 + nn, oh = oh0 + h_, ow = ow0 + w_;
    const bool valid = n < p.B && oh < p.OH && ow < p.OW;
    unsigned long long seed = 0, step = 0;
    if (p.dropout) { seed = p.rng[0]; step = p.rng[1]; }
    if constexpr (MODE == EPI_POOL || MODE == EPI_MAXTIME) {
      // ---- lane == output row (the reductions run across rows with warp shuffles) ----
      if (!mbar_wait(tmem_full, 0)) { if (p.err) *p.err = 3; }
      if (dbg && threadIdx.x == 64) dbg[5] = clock64();   // accumulator ready
      tc_fence_after();
      asm volatile("bar.sync 1, 128;" ::: "memory");      // bias_s written by all four epilogue warps
      __align__(16) unsigned char bi[32];
#pragma unroll 1
      for (int c0 = 0; c0 < BLOCK_N; c0 += 32) {
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)c0, v);
        if (dbg && threadIdx.x == 64 && c0 == 0) dbg[11] = clock64();   // first TMEM read done
        const int col0 = n_tile0 + c0;
        if (MODE == EPI_POOL) {
        // 2x2 max over (oh^1, ow^1): partners are lane+1, lane+tw, lane+tw+1 of the even/even lane
        const bool writer = ((h_ | w_) & 1) == 0;
        const int PH = p.OH >> 1, PW = p.OW >> 1;
        size_t obase = ((((size_t)n * PH + (oh >> 1)) * PW + (ow >> 1)) * p.N) + col0;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          float x = v[c] + bias_s[c0 + c];
          float x01 = __shfl_down_sync(0xffffffffu, x, 1);
          float x10 = __shfl_down_sync(0xffffffffu, x, p.tw);
          float x11 = __shfl_down_sync(0xffffffffu, x, p.tw + 1);
          float best = x;
          int b = 0;
          if (x01 > best) { best = x01; b = 1; }
          if (x10 > best) { best = x10; b = 2; }
          if (x11 > best) { best = x11; b = 3; }
          if (p.relu) best = fmaxf(best, 0.f);
          v[c] = best;
          bi[c] = (unsigned char)b;
        }
        if (valid && writer) {
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4)
            st4(p.out + obase + 4 * c4, make_float4(v[4 * c4], v[4 * c4 + 1], v[4 * c4 + 2], v[4 * c4 + 3]));
          uint4* ip = reinterpret_cast<uint4*>(p.idx + obase);
          const uint4* bp = reinterpret_cast<const uint4*>(bi);
          ip[0] = bp[0];
          ip[1] = bp[1];
        }
        } else {
        // Temporal convolution -> Max over time -> ReLU (BiCNN/bicnn.lua:59,76-84).  A warp holds the
        // 32 time steps of one sentence (tile = tn sentences x 32 frames): bias, mask frames past the
        // sentence end, butterfly max with the arg-max, then lane c publishes column c with one
        // 64-bit atomicMax of (bits(max(v,0)) << 32 | t): relu(max) = max over t of max(v,0), and
        // non-negative floats order like their bit patterns, so several w-tiles of a long sentence
        // combine correctly.  `crop` carries the per-sentence frame counts here.
        const int frames = (n < p.B) ? p.crop[n] : 0;
        const bool live = valid && ow < frames;
        unsigned long long mine = 0ull;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          float x = live ? fmaxf(v[c] + bias_s[c0 + c], 0.f) : 0.f;
          unsigned long long key = ((unsigned long long)__float_as_uint(x) << 32) | (unsigned)(live ? ow : 0);
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
            key = other > key ? other : key;
          }
          if (lane == c) mine = key;
        }
        if (n < p.B && (col0 + lane) < p.N)
          atomicMax(reinterpret_cast<unsigned long long*>(p.out) + (size_t)n * p.N + col0 + lane, mine);
        }
      }
    } else {
      coalesced_epilogue<BLOCK_N, MODE>(p, smem, tmem_base, tmem_full, true, quad, lane, row, n, oh, ow, valid, n_tile0,
                                        tile, seed, step, bias_s, dbg);
    }
    tc_fence_before();
    if (dbg && threadIdx.x == 64) dbg[6] = clock64();   // epilogue done
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BLOCK_N);
  }
  if (dbg && threadIdx.x == 0) dbg[7] = clock64();
}

template <int BLOCK_N, int MODE, bool X3>
__global__ void __launch_bounds__(192)
umma_igemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const IgemmParams p) {
  igemm_body<BLOCK_N, MODE, X3>(&tmap_a, &tmap_b, blockIdx.y * BLOCK_N, p);
}

// ---------------------------------------------------------------------------------------------------------
// Resident-input convolution (forward or data gradient) for layers whose per-image input fits in shared memory.
//
// The streaming kernel above fetches, for every kernel tap, the shifted 128-row window of the input again: for the
// 5x5 data gradient of the second convolution that is 25x the same 32 KB image through L2 -> SM (2.4 MB per CTA),
// and the kernel ran at the chip's L2 feed limit.  Here one CTA owns one image: its (zero-padded) input is loaded
// ONCE into shared memory as `cb` channel-chunk planes of RH x RW pixels x 128 B, and the tap shift is applied to
// the *descriptor*: with q = y*RW + x enumerating output pixels in the padded row pitch, tap (kh,kw) reads plane
// rows q + shift(kh,kw), shift = (sgn*kh - y_org)*RW + (sgn*kw - x_org), i.e. the A operand of every tap is the same
// plane at a different start row.  A K-major SWIZZLE_128B operand may start at any 128-byte row of a TMA-written
// tile (the swizzle is a function of absolute shared-memory address bits; tools/dev/desc_shift_probe.cu).  Only the
// weights are streamed (BLOCK_N x 32 floats per k-block), and each weight tile feeds all `ntiles` accumulators.
// Columns x >= OW of the padded pitch are computed and discarded (25 % for RW = 16, OW = 12).
struct ResidentParams {
  int RW, RH, x_org, y_org, cb, KH, KW, sgn, ntiles, stages;
};

template <int BLOCK_N, int MODE>
__global__ void __launch_bounds__(192)
umma_resident_kernel(const __grid_constant__ CUtensorMap tmap_in, const __grid_constant__ CUtensorMap tmap_w,
                     const IgemmParams p, const ResidentParams r) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int plane_bytes = r.RW * r.RH * 128;
  constexpr int kBBytes = BLOCK_N * 128;
  uint8_t* bstage = smem + r.cb * plane_bytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bstage + r.stages * kBBytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* res_bar = empty_bar + kMaxStages;        // one per channel-chunk plane (<= 8)
  uint64_t* tmem_full = res_bar + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  __shared__ __align__(16) float bias_s[256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x;                        // image
  const int n_tile0 = blockIdx.y * BLOCK_N;
  const int num_kb = r.KH * r.KW * r.cb;
  const int kStages = r.stages;
  const uint32_t tmem_cols = (uint32_t)(r.ntiles * BLOCK_N) <= 32 ? 32 : (r.ntiles * BLOCK_N <= 64 ? 64 :
                             (r.ntiles * BLOCK_N <= 128 ? 128 : (r.ntiles * BLOCK_N <= 256 ? 256 : 512)));
  pdl_trigger();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_in);
    prefetch_tmap(&tmap_w);
    for (int i = 0; i < kStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < r.cb; ++i) mbar_init(&res_bar[i], 1);
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, tmem_cols);
  const int kProducers = kStages < 5 ? kStages : 5;
  const int pi = warp == 0 ? 0 : warp - 1;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  if (warp >= 2)
    for (int c = threadIdx.x - 64; c < BLOCK_N; c += 128) bias_s[c] = p.bias ? p.bias[n_tile0 + c] : 0.f;

  if (warp != 1 && pi < kProducers && elect_one()) {
    if (warp == 0) {       // the resident planes: one box per channel chunk, zero-filled outside the image
      for (int c = 0; c < r.cb; ++c) {
        mbar_expect_tx(&res_bar[c], plane_bytes);
        tma_load_4d(smem + c * plane_bytes, &tmap_in, &res_bar[c], c * 32, r.x_org, r.y_org, n);
      }
    }
    int stage = pi, phase = 0;
    while (stage >= kStages) { stage -= kStages; phase ^= 1; }
    for (int kb = pi; kb < num_kb; kb += kProducers) {
      if (kb >= kStages && !mbar_wait(&empty_bar[stage], phase ^ 1)) { if (p.err) *p.err = 21; break; }
      mbar_expect_tx(&full_bar[stage], kBBytes);
      tma_load_2d(bstage + stage * kBBytes, &tmap_w, &full_bar[stage], kb * 32, n_tile0);
      stage += kProducers;
      while (stage >= kStages) { stage -= kStages; phase ^= 1; }
    }
  }
  __syncwarp();
  if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_tf32(128, BLOCK_N, 0, 0);
      const uint64_t desc0 = make_smem_desc(smem_u32(smem), 16, 1024);
      const uint64_t bdesc0 = make_smem_desc(smem_u32(bstage), 16, 1024);
      int stage = 0, phase = 0, kh = 0, kw = 0, c = 0;
      for (int kb = 0; kb < num_kb; ++kb) {
        if (kb < r.cb) {     // first touch of plane c
          if (!mbar_wait(&res_bar[c], 0)) { if (p.err) *p.err = 22; break; }
        }
        if (!mbar_wait(&full_bar[stage], phase)) { if (p.err) *p.err = 23; break; }
        tc_fence_after();
        const int shift = (r.sgn * kh - r.y_org) * r.RW + (r.sgn * kw - r.x_org);
        const uint64_t ad = desc0 + (uint64_t)((c * plane_bytes + shift * 128) >> 4);
        const uint64_t bd = bdesc0 + (uint64_t)((stage * kBBytes) >> 4);
        for (int t = 0; t < r.ntiles; ++t) {
          const uint64_t at = ad + (uint64_t)((t * 128 * 128) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_tf32(tmem_base + (uint32_t)(t * BLOCK_N), at + 2 * k, bd + 2 * k, idesc, (kb | k) ? 1u : 0u);
        }
        mma_commit(&empty_bar[stage]);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
        if (++c == r.cb) { c = 0; if (++kw == r.KW) { kw = 0; ++kh; } }
      }
      mma_commit(tmem_full);
    }
  } else if (warp >= 2) {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    for (int t = 0; t < r.ntiles; ++t) {
      const int q = t * 128 + row;
      const int y = q / r.RW, x = q - y * r.RW;
      const bool valid = n < p.B && y < p.OH && x < p.OW;
      // plane 0 doubles as the transposition buffer: every MMA has completed once tmem_full fired
      coalesced_epilogue<BLOCK_N, MODE>(p, smem, tmem_base + (uint32_t)(t * BLOCK_N), tmem_full, t == 0, quad, lane, row, n,
                                        y, x, valid, n_tile0, n * r.ntiles + t, 0ull, 0ull, bias_s, nullptr);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// Parameter pull fused with the GEMM that consumes the weights: the B operand (rows of a Linear /
// convolution weight) is NOT first copied out of the sharded centre; each CTA's TMA loads fetch its
// weight tile straight from the owner's shard over NVLink (a tensor map per owner over the peer-mapped
// shard) while the MMAs of earlier tiles run.  Reference: pclient.lua:74-84 (pull) followed by the
// first nn.Linear forward of the model.
struct PeerMaps {
  CUtensorMap map[8];
  int rows_per_owner;
};

template <int BLOCK_N>
__global__ void __launch_bounds__(192)
umma_igemm_pull_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ PeerMaps pm,
                       const IgemmParams p) {
  const int n_tile0 = blockIdx.y * BLOCK_N;
  const int owner = n_tile0 / pm.rows_per_owner;
  igemm_body<BLOCK_N, EPI_STORE>(&tmap_a, &pm.map[owner], n_tile0 - owner * pm.rows_per_owner, p);
}

// The same idea as a plain GEMM kernel with reuse inside the CTA: one CTA owns up to MT 128-row M-tiles of the
// activations (MT accumulators of BLOCK_N TMEM columns each), so every weight tile that crosses NVLink is fetched
// ONCE per CTA and feeds MT * 4 MMAs -- the single-tile kernel above re-fetches it for every M-tile, which is what
// made it lose to gather + GEMM from 512 rows up.  (TMA multicast loads here read only this GPU's memory, so sharing
// a remote tile between CTAs is not an option; reuse across accumulators is.)
//   BY_K = false (forward, y = act(x W^T + b)): W row-sharded, owner chosen by the output-column tile;
//   BY_K = true  (data gradient, dx = dy W):   the reduction runs over W's rows, i.e. ACROSS owners; each owner also
//                keeps its rows transposed ([in][rows_per_owner]), and the owner is chosen by the k-block.
// Two rings instead of one: the local activation tiles (MT x 16 KB per k-block, L2 latency) and the remote weight
// tiles (4 / 8 KB per k-block, NVLink latency) have very different latency x size products.  With one ring of
// [A | B] stages only two stages fit at MT = 4, i.e. two weight tiles in flight per CTA, and every k-block then
// waited a whole NVLink round trip.  The B ring is `p.sb` slots deep (up to 16) and runs `sb - stages` k-blocks
// ahead of the A ring, issued by the same thread: slot B(j + sb - stages) and slot A(j) are both released by the
// MMAs of k-block j - stages, so running ahead costs no extra waiting.
// gridDim.z > 1 (data gradient only): split-K -- slice z reduces k-blocks [z * kb_split, (z+1) * kb_split) and the
// epilogue adds its partial tile into the (zeroed) output with red.add.v4.f32, which gives the few-output-tile,
// long-reduction shape of dx = dy W enough CTAs.
constexpr int kMaxBStages = 16;

template <int BLOCK_N, int MT, bool BY_K>
__global__ void __launch_bounds__(192)
umma_pulled_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ PeerMaps pm,
                        const IgemmParams p) {
  constexpr int kBBytes = BLOCK_N * 128;
  constexpr int kAStage = MT * kABytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int SA = p.stages, SB = p.sb;
  uint8_t* smem_b = smem + SA * kAStage;
  uint64_t* full_a = reinterpret_cast<uint64_t*>(smem_b + SB * kBBytes);
  uint64_t* empty_a = full_a + kMaxStages;
  uint64_t* full_b = empty_a + kMaxStages;
  uint64_t* empty_b = full_b + kMaxBStages;
  uint64_t* tmem_full = empty_b + kMaxBStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
  __shared__ __align__(16) float bias_s[256];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tile0 = blockIdx.y * BLOCK_N;
  const int m_tiles = (p.B + 127) / 128;
  const int mt0 = blockIdx.x * MT;
  const int mt_here = m_tiles - mt0 < MT ? m_tiles - mt0 : MT;
  const int kb0 = BY_K ? (int)blockIdx.z * p.kb_split : 0;          // this CTA's slice of the reduction
  const int nkb = BY_K ? (p.num_kb - kb0 < p.kb_split ? p.num_kb - kb0 : p.kb_split) : p.num_kb;
  pdl_trigger();
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    for (int i = 0; i < SA; ++i) { mbar_init(&full_a[i], 1); mbar_init(&empty_a[i], 1); }
    for (int i = 0; i < SB; ++i) { mbar_init(&full_b[i], 1); mbar_init(&empty_b[i], 1); }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, MT * BLOCK_N);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();
  if (warp >= 2)
    for (int c = threadIdx.x - 64; c < BLOCK_N; c += 128)
      bias_s[c] = (p.bias && (!BY_K || blockIdx.z == 0)) ? p.bias[n_tile0 + c] : 0.f;

  if (warp == 0) {
    if (elect_one()) {
      const int owner_n = BY_K ? 0 : n_tile0 / pm.rows_per_owner;
      int as = 0, aph = 0, bs = 0, bph = 0, jb = 0;
      bool ok = true;
      auto issue_b = [&]() {   // weight tile of local k-block jb, from its owner's shard
        if (jb >= SB && !mbar_wait(&empty_b[bs], bph ^ 1)) { ok = false; return; }
        mbar_expect_tx(&full_b[bs], (uint32_t)kBBytes);
        const int kb = kb0 + jb;
        if (BY_K) {
          const int owner = (kb * 32) / pm.rows_per_owner;
          tma_load_2d(smem_b + bs * kBBytes, &pm.map[owner], &full_b[bs], kb * 32 - owner * pm.rows_per_owner, n_tile0);
        } else {
          tma_load_2d(smem_b + bs * kBBytes, &pm.map[owner_n], &full_b[bs], kb * 32, n_tile0 - owner_n * pm.rows_per_owner);
        }
        ++jb;
        if (++bs == SB) { bs = 0; bph ^= 1; }
      };
      const int ahead = SB - SA;
      while (jb < ahead && jb < nkb) issue_b();          // never waits: jb < SB
      for (int j = 0; j < nkb && ok; ++j) {
        if (jb < nkb) issue_b();
        if (!ok) break;
        if (j >= SA && !mbar_wait(&empty_a[as], aph ^ 1)) { ok = false; break; }
        uint8_t* dst = smem + as * kAStage;
        mbar_expect_tx(&full_a[as], (uint32_t)(mt_here * kABytes));
        for (int mt = 0; mt < mt_here; ++mt)
          tma_load_4d(dst + mt * kABytes, &tmap_a, &full_a[as], (kb0 + j) * 32, 0, 0, (mt0 + mt) * 128);
        if (++as == SA) { as = 0; aph ^= 1; }
      }
      if (!ok && p.err) *p.err = 1;
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_tf32(128, BLOCK_N, 0, 0);
      const uint64_t desc_a0 = make_smem_desc(smem_u32(smem), 16, 1024);
      const uint64_t desc_b0 = make_smem_desc(smem_u32(smem_b), 16, 1024);
      int as = 0, aph = 0, bs = 0, bph = 0;
      for (int j = 0; j < nkb; ++j) {
        if (!mbar_wait(&full_b[bs], bph) || !mbar_wait(&full_a[as], aph)) { if (p.err) *p.err = 2; break; }
        tc_fence_after();
        const uint64_t ad = desc_a0 + (uint64_t)((as * kAStage) >> 4);
        const uint64_t bd = desc_b0 + (uint64_t)((bs * kBBytes) >> 4);
        for (int mt = 0; mt < mt_here; ++mt) {
          const uint64_t am = ad + (uint64_t)((mt * kABytes) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            mma_tf32(tmem_base + (uint32_t)(mt * BLOCK_N), am + 2 * k, bd + 2 * k, idesc, (j | k) ? 1u : 0u);
        }
        mma_commit(&empty_a[as]);
        mma_commit(&empty_b[bs]);
        if (++as == SA) { as = 0; aph ^= 1; }
        if (++bs == SB) { bs = 0; bph ^= 1; }
      }
      mma_commit(tmem_full);
    }
  } else {
    const int quad = warp & 3;
    const int row = quad * 32 + lane;
    for (int mt = 0; mt < mt_here; ++mt) {
      const int n = (mt0 + mt) * 128 + row;
      coalesced_epilogue<BLOCK_N, EPI_STORE, BY_K>(p, smem, tmem_base + (uint32_t)(mt * BLOCK_N), tmem_full, mt == 0, quad,
                                                   lane, row, n, 0, 0, n < p.B, n_tile0, mt0 + mt, 0ull, 0ull, bias_s,
                                                   nullptr);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, MT * BLOCK_N);
  }
}

// ---- host side -----------------------------------------------------------------------------------
static decltype(&cuTensorMapEncodeTiled) get_encode() {
  static decltype(&cuTensorMapEncodeTiled) fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && p)
      fn = reinterpret_cast<decltype(&cuTensorMapEncodeTiled)>(p);
  }
  return fn;
}

template <int BLOCK_N, int MODE, bool X3>
static int launch_igemm_x(const CUtensorMap* ta, const CUtensorMap* tb, IgemmParams p, int m_tiles,
                          cudaStream_t st) {
  using L = SmemLayout<BLOCK_N>;
  static unsigned long long attr_done = 0;   // bit per device
  MPIT_CUDA_TRY(set_smem_limit_per_device(umma_igemm_kernel<BLOCK_N, MODE, X3>, L::total(L::kMaxFit), attr_done));
  dim3 grid(m_tiles, p.N / BLOCK_N);
  // Pipeline depth.  Grids of up to one CTA per SM get the deepest pipeline that fits.  Larger grids are better
  // served by residency: make the stages small enough that ceil(ctas/148) CTAs (up to 4) share an SM, so the whole
  // grid is one wave and one CTA's epilogue overlaps another's loads (conv1: 576 short CTAs, 2 stages each).
  int ctas = m_tiles * (p.N / BLOCK_N);
  constexpr int mul = X3 ? 2 : 1;               // 3xTF32: every stage also holds the lo tiles
  const int max_fit = L::kMaxFit / mul;
  int stages = max_fit;
  if (ctas > MPIT_NUM_SMS) {
    int per_sm = (ctas + MPIT_NUM_SMS - 1) / MPIT_NUM_SMS;
    if (per_sm > 4) per_sm = 4;
    int fit = ((226 * 1024) / per_sm - 2560) / (L::kStageBytes * mul);
    if (fit < stages) stages = fit;
  }
  if (stages > p.num_kb) stages = p.num_kb;
  if (stages < 2) stages = 2;
  if (stages > max_fit) stages = max_fit;
  p.stages = stages;
  MPIT_CUDA_TRY(launch_k(umma_igemm_kernel<BLOCK_N, MODE, X3>, grid, dim3(192), (size_t)L::total(stages, mul), st, *ta, *tb, p));
  return 0;
}

template <int BLOCK_N, int MODE>
static int launch_igemm(const CUtensorMap* ta, const CUtensorMap* tb, const IgemmParams& p, int m_tiles, cudaStream_t st) {
  if (p.x3) {
    if (MODE == EPI_MAXTIME) return MPIT_E_UNSUPPORTED;      // the sentence-CNN epilogue has no 3xTF32 instantiation
    return launch_igemm_x<BLOCK_N, MODE == EPI_MAXTIME ? EPI_STORE : MODE, true>(ta, tb, p, m_tiles, st);
  }
  return launch_igemm_x<BLOCK_N, MODE, false>(ta, tb, p, m_tiles, st);
}

template <int BLOCK_N, int MT, bool BY_K>
static int launch_pulled(dim3 grid, size_t smem, cudaStream_t st, const CUtensorMap& ta, const PeerMaps& pm,
                         const IgemmParams& p) {
  static unsigned long long done = 0;     // bit per device
  MPIT_CUDA_TRY(set_smem_limit_per_device(umma_pulled_gemm_kernel<BLOCK_N, MT, BY_K>, 220 * 1024, done));
  MPIT_CUDA_TRY(launch_k(umma_pulled_gemm_kernel<BLOCK_N, MT, BY_K>, grid, dim3(192), smem, st, ta, pm, p));
  return 0;
}

}  // namespace mpit

using namespace mpit;

extern "C" {

// Encode a rank-`rank` fp32 tiled tensor map; swizzle128: 0 none, 1 SWIZZLE_128B (K-major operands),
// 2 SWIZZLE_128B_ATOM_32B (tf32 MN-major operands).  dims/box are innermost-first;
// strides_bytes has rank-1 entries (stride of dim 1..rank-1).  out128 is a HOST buffer of 128 bytes.
int mpit_tmap_encode_f32(void* out128, const void* gptr, int rank, const unsigned long long* dims,
                         const unsigned long long* strides_bytes, const unsigned* box, int swizzle128) {
  auto enc = get_encode();
  if (!enc) return MPIT_E_UNSUPPORTED;
  if (rank < 1 || rank > 5) return MPIT_E_ARG;
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) gstr[i] = strides_bytes[i];
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(gptr), gdim, gstr, bx,
                   es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 == 2 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                   : (swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE),
                   CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return 1000 + (int)r;
  memcpy(out128, &m, sizeof(m));
  static_assert(sizeof(CUtensorMap) == 128, "CUtensorMap is 128 bytes");
  return 0;
}

// geom17 (HOST): {nsub,tw,th,tn,tiles_w,tiles_h,kw_it,cb,num_kb,sgn,B,OH,OW,N,mode,relu,dropout}
int mpit_umma_igemm(const void* tmap_a128, const void* tmap_b128, const int* geom17, float scale,
                    const float* bias, float* out, unsigned char* idx, const float* aux,
                    const unsigned char* aux_idx, const int* crop, const unsigned long long* rng, int* err,
                    float* colpart, int block_n, void* stream) {
  IgemmParams p{};
  {
    // MPIT_IGEMM_DBG=<device address of a long long[16 * ctas] buffer>: per-CTA timeline for tuning
    static long long* dbg_buf = []() -> long long* {
      const char* e = getenv("MPIT_IGEMM_DBG");
      return e ? reinterpret_cast<long long*>(strtoull(e, nullptr, 0)) : nullptr;
    }();
    p.dbg = dbg_buf;
  }
  p.nsub = geom17[0]; p.tw = geom17[1]; p.th = geom17[2]; p.tn = geom17[3];
  p.tiles_w = geom17[4]; p.tiles_h = geom17[5]; p.kw_it = geom17[6]; p.cb = geom17[7];
  p.num_kb = geom17[8]; p.sgn = geom17[9]; p.B = geom17[10]; p.OH = geom17[11]; p.OW = geom17[12];
  p.N = geom17[13]; p.mode = geom17[14]; p.relu = geom17[15]; p.dropout = geom17[16];
  p.scale = scale; p.bias = bias; p.out = out; p.idx = idx; p.aux = aux; p.aux_idx = aux_idx;
  p.crop = crop; p.rng = rng; p.err = err; p.colpart = colpart;
  p.x3 = g_x3_mode;
  if (p.nsub < 1 || p.nsub > 2 || p.nsub * p.tw * p.th * p.tn != 128 || p.num_kb < 1 || p.N % block_n != 0)
    return MPIT_E_ARG;
  if (p.mode == EPI_POOL && (p.tw > 16 || (p.tw & 1) || (p.th & 1))) return MPIT_E_ARG;
  int subtiles = ((p.B + p.tn - 1) / p.tn) * p.tiles_w * p.tiles_h;
  int m_tiles = (subtiles + p.nsub - 1) / p.nsub;
  CUtensorMap ta_, tb_;   // the caller's buffers need not be 64-byte aligned
  memcpy(&ta_, tmap_a128, sizeof(CUtensorMap));
  memcpy(&tb_, tmap_b128, sizeof(CUtensorMap));
  const CUtensorMap* ta = &ta_;
  const CUtensorMap* tb = &tb_;
  cudaStream_t st = (cudaStream_t)stream;
  switch (block_n) {
#define MPIT_IGEMM_MODES(BN)                                                         \
  switch (p.mode) {                                                                 \
    case EPI_STORE: return launch_igemm<BN, EPI_STORE>(ta, tb, p, m_tiles, st);     \
    case EPI_POOL: return launch_igemm<BN, EPI_POOL>(ta, tb, p, m_tiles, st);       \
    case EPI_MASK: return launch_igemm<BN, EPI_MASK>(ta, tb, p, m_tiles, st);       \
    case EPI_EXPAND: return launch_igemm<BN, EPI_EXPAND>(ta, tb, p, m_tiles, st);   \
    case EPI_MAXTIME: return launch_igemm<BN, EPI_MAXTIME>(ta, tb, p, m_tiles, st); \
    default: return MPIT_E_ARG;                                                     \
  }
    case 32: MPIT_IGEMM_MODES(32)
    case 64: MPIT_IGEMM_MODES(64)
    case 128: MPIT_IGEMM_MODES(128)
    case 256: MPIT_IGEMM_MODES(256)
#undef MPIT_IGEMM_MODES
    default: return MPIT_E_ARG;
  }
}

// Resident-input convolution / data gradient: one CTA per image, input planes loaded once, weights streamed.
// geom15 (HOST) = {RW, RH, x_org, y_org, cb, KH, KW, sgn, ntiles, B, OH, OW, N, mode, relu}; OH/OW are the extents of
// the GEMM M side (rows q = y*RW + x with y < OH are computed, x >= OW discarded).  tmap_in: 4-D NHWC map with box
// (32, RW, RH, 1) and zero OOB fill; tmap_w: 2-D [N][K] map with box (32, block_n), k-blocks ordered (kh, kw, chunk).
// Modes 0 (store), 2 (mask), 3 (un-pool + mask).  Returns MPIT_E_UNSUPPORTED when the planes do not fit.
int mpit_umma_conv_resident(const void* tmap_in128, const void* tmap_w128, const int* geom15, float scale,
                            const float* bias, float* out, const float* aux, const unsigned char* aux_idx, int* err,
                            float* colpart, int block_n, void* stream) {
  ResidentParams r{};
  IgemmParams p{};
  r.RW = geom15[0]; r.RH = geom15[1]; r.x_org = geom15[2]; r.y_org = geom15[3]; r.cb = geom15[4];
  r.KH = geom15[5]; r.KW = geom15[6]; r.sgn = geom15[7]; r.ntiles = geom15[8];
  p.B = geom15[9]; p.OH = geom15[10]; p.OW = geom15[11]; p.N = geom15[12]; p.mode = geom15[13]; p.relu = geom15[14];
  p.scale = scale; p.bias = bias; p.out = out; p.aux = aux; p.aux_idx = aux_idx; p.err = err; p.colpart = colpart;
  if (r.cb < 1 || r.cb > 8 || r.ntiles < 1 || r.KH < 1 || r.KW < 1 || (r.sgn != 1 && r.sgn != -1) || p.N % block_n != 0)
    return MPIT_E_ARG;
  if ((r.RW * r.RH) % 8 != 0 || r.RW > 256 || r.RH > 256) return MPIT_E_ARG;
  if (r.ntiles * block_n > 512) return MPIT_E_UNSUPPORTED;
  // every row a tile can touch must lie inside the plane
  const int max_shift = (r.sgn > 0 ? (r.KH - 1) - r.y_org : -r.y_org) * r.RW + (r.sgn > 0 ? (r.KW - 1) - r.x_org : -r.x_org);
  const int min_shift = (r.sgn > 0 ? -r.y_org : -(r.KH - 1) - r.y_org) * r.RW + (r.sgn > 0 ? -r.x_org : -(r.KW - 1) - r.x_org);
  if (min_shift < 0 || r.ntiles * 128 + max_shift > r.RW * r.RH) return MPIT_E_ARG;
  const int plane_bytes = r.RW * r.RH * 128;
  const int stage_bytes = block_n * 128;
  int stages = (220 * 1024 - r.cb * plane_bytes - 1536) / stage_bytes;
  if (stages > 8) stages = 8;
  const int num_kb = r.KH * r.KW * r.cb;
  if (stages > num_kb) stages = num_kb;
  if (stages < 2) return MPIT_E_UNSUPPORTED;
  r.stages = stages;
  const size_t smem = (size_t)r.cb * plane_bytes + (size_t)stages * stage_bytes + 512 + 1024;   // + barriers + align
  CUtensorMap ti, tw;
  memcpy(&ti, tmap_in128, sizeof(ti));
  memcpy(&tw, tmap_w128, sizeof(tw));
  dim3 grid(p.B, p.N / block_n);
  cudaStream_t st = (cudaStream_t)stream;
  auto go = [&](auto kern) -> int {
    // the opt-in limit is raised on every launch to what it needs (the kernel's static shared memory counts against the
    // same 227 KB); a static cache in this generic lambda would be shared by all of its instantiations
    MPIT_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    MPIT_CUDA_TRY(launch_k(kern, grid, dim3(192), smem, st, ti, tw, p, r));
    return 0;
  };
  if (block_n == 64) {
    if (p.mode == EPI_EXPAND) return go(umma_resident_kernel<64, EPI_EXPAND>);
    if (p.mode == EPI_MASK) return go(umma_resident_kernel<64, EPI_MASK>);
    if (p.mode == EPI_STORE) return go(umma_resident_kernel<64, EPI_STORE>);
  } else if (block_n == 32) {
    if (p.mode == EPI_EXPAND) return go(umma_resident_kernel<32, EPI_EXPAND>);
    if (p.mode == EPI_MASK) return go(umma_resident_kernel<32, EPI_MASK>);
    if (p.mode == EPI_STORE) return go(umma_resident_kernel<32, EPI_STORE>);
  }
  return MPIT_E_UNSUPPORTED;
}

// Linear layer whose weight rows live in the NVLink-sharded centre: y = act(x W^T + b), W fetched by
// the GEMM's own TMA loads from each owner's shard (tmaps_b: HOST array of nranks x 128 bytes, one map
// per owner over its [rows_per_owner][K] block).  N = nranks * rows_per_owner; BLOCK_N 32 or 64 must
// divide rows_per_owner.
int mpit_umma_linear_pull(const void* tmap_a128, const void* tmaps_b, int nranks, int rows_per_owner,
                          const int* geom17, const float* bias, float* out, int* err, int block_n, void* stream) {
  if (nranks < 1 || nranks > 8 || rows_per_owner % block_n != 0) return MPIT_E_ARG;
  IgemmParams p{};
  p.nsub = geom17[0]; p.tw = geom17[1]; p.th = geom17[2]; p.tn = geom17[3];
  p.tiles_w = geom17[4]; p.tiles_h = geom17[5]; p.kw_it = geom17[6]; p.cb = geom17[7];
  p.num_kb = geom17[8]; p.sgn = geom17[9]; p.B = geom17[10]; p.OH = geom17[11]; p.OW = geom17[12];
  p.N = geom17[13]; p.mode = geom17[14]; p.relu = geom17[15]; p.dropout = 0;
  p.scale = 1.f; p.bias = bias; p.out = out; p.err = err;
  if (p.nsub < 1 || p.nsub > 2 || p.nsub * p.tw * p.th * p.tn != 128 || p.num_kb < 1 ||
      p.N != nranks * rows_per_owner || p.mode != EPI_STORE)
    return MPIT_E_ARG;
  int subtiles = ((p.B + p.tn - 1) / p.tn) * p.tiles_w * p.tiles_h;
  int m_tiles = (subtiles + p.nsub - 1) / p.nsub;
  CUtensorMap ta;
  PeerMaps pm;
  memset(&pm, 0, sizeof(pm));
  memcpy(&ta, tmap_a128, sizeof(CUtensorMap));
  for (int i = 0; i < nranks; ++i) memcpy(&pm.map[i], static_cast<const char*>(tmaps_b) + 128 * i, sizeof(CUtensorMap));
  pm.rows_per_owner = rows_per_owner;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(m_tiles, p.N / block_n);
  if (block_n == 32) {
    using L = SmemLayout<32>;
    p.stages = p.num_kb < 4 ? (p.num_kb < 2 ? 2 : p.num_kb) : 4;
    MPIT_CUDA_TRY(cudaFuncSetAttribute(umma_igemm_pull_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       L::total(4)));
    umma_igemm_pull_kernel<32><<<grid, 192, L::total(p.stages), st>>>(ta, pm, p);
  } else if (block_n == 64) {
    using L = SmemLayout<64>;
    p.stages = p.num_kb < 4 ? (p.num_kb < 2 ? 2 : p.num_kb) : 4;
    MPIT_CUDA_TRY(cudaFuncSetAttribute(umma_igemm_pull_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       L::total(4)));
    umma_igemm_pull_kernel<64><<<grid, 192, L::total(p.stages), st>>>(ta, pm, p);
  } else {
    return MPIT_E_ARG;
  }
  MPIT_LAUNCH_CHECK();
  return 0;
}

// y[M][N] = act(x[M][K] W^T + b) (by_k = 0) or dx[M][N] = dy[M][K] W (by_k = 1) with W living row-sharded in the owners'
// symmetric memory; see umma_pulled_gemm_kernel.  `tmap_a128`: 4-D map [K,1,1,Mpad] box [32,1,1,128] over the local
// activations; `tmaps_b`: HOST array of nranks x 128 bytes, one map per owner (by_k = 0: [rows_per_owner][K] box
// [32, block_n]; by_k = 1: the transposed shard [N][rows_per_owner] box [32, block_n]).  K is the reduction length,
// N the number of output columns; mt = M-tiles per CTA (1, 2 or 4; block_n * mt <= 256).  ksplit > 1 (by_k = 1, no
// bias, no activation, K / 32 divisible by ksplit): split-K over ksplit CTAs per output tile; `out` is zeroed on the
// stream first and the slices add into it.
int mpit_umma_pulled_gemm(const void* tmap_a128, const void* tmaps_b, int nranks, int rows_per_owner, int by_k, int M,
                          int K, int N, int relu, const float* bias, float* out, int* err, int block_n, int mt,
                          int ksplit, void* stream) {
  if (nranks < 1 || nranks > 8 || M < 1 || K % 32 || N % block_n || (block_n != 32 && block_n != 64) ||
      (mt != 1 && mt != 2 && mt != 4) || ksplit < 1)
    return MPIT_E_ARG;
  if (by_k ? (rows_per_owner % 32 != 0 || K != nranks * rows_per_owner)
           : (rows_per_owner % block_n != 0 || N != nranks * rows_per_owner))
    return MPIT_E_ARG;
  if (ksplit > 1 && (!by_k || relu || bias || (K / 32) % ksplit)) return MPIT_E_ARG;
  IgemmParams p{};
  p.nsub = 1; p.tw = 1; p.th = 1; p.tn = 128; p.tiles_w = 1; p.tiles_h = 1; p.kw_it = 1; p.cb = K / 32;
  p.num_kb = K / 32; p.sgn = 1; p.B = M; p.OH = 1; p.OW = 1; p.N = N; p.mode = EPI_STORE; p.relu = relu;
  p.scale = 1.f; p.bias = bias; p.out = out; p.err = err;
  p.kb_split = p.num_kb / ksplit;
  p.accum = ksplit > 1;
  CUtensorMap ta;
  PeerMaps pm;
  memset(&pm, 0, sizeof(pm));
  memcpy(&ta, tmap_a128, sizeof(CUtensorMap));
  for (int i = 0; i < nranks; ++i) memcpy(&pm.map[i], static_cast<const char*>(tmaps_b) + 128 * i, sizeof(CUtensorMap));
  pm.rows_per_owner = rows_per_owner;
  const int m_tiles = (M + 127) / 128;
  dim3 grid((m_tiles + mt - 1) / mt, N / block_n, ksplit);
  cudaStream_t st = (cudaStream_t)stream;
  // ring depths: the activation ring as shallow as keeps the tensor core fed from L2 (4 / 3 / 2 stages of
  // mt x 16 KB), everything else of ~216 KB for weight tiles in flight over NVLink
  const int a_stage = mt * kABytes, b_tile = block_n * 128;
  int sa = mt == 1 ? 4 : (mt == 2 ? 3 : 2);
  if (sa > p.kb_split) sa = p.kb_split;
  int sb = (216 * 1024 - sa * a_stage) / b_tile;
  if (sb > kMaxBStages) sb = kMaxBStages;
  if (sb > p.kb_split) sb = p.kb_split;
  if (sb < sa) sb = sa;
  p.stages = sa;
  p.sb = sb;
  const size_t smem = (size_t)sa * a_stage + (size_t)sb * b_tile + 512 + 1024;
  if (ksplit > 1) MPIT_CUDA_TRY(cudaMemsetAsync(out, 0, (size_t)M * N * sizeof(float), st));
#define MPIT_PULLED(BN, MTV)                                                                              \
  if (block_n == BN && mt == MTV)                                                                         \
    return by_k ? launch_pulled<BN, MTV, true>(grid, smem, st, ta, pm, p) : launch_pulled<BN, MTV, false>(grid, smem, st, ta, pm, p);
  MPIT_PULLED(32, 1) MPIT_PULLED(32, 2) MPIT_PULLED(32, 4)
  MPIT_PULLED(64, 1) MPIT_PULLED(64, 2) MPIT_PULLED(64, 4)
#undef MPIT_PULLED
  return MPIT_E_ARG;
}

}  // extern "C"
