// Micro-benchmark: how fast can one SM (and the whole chip) be fed by TMA box loads of the shapes the
// implicit-GEMM kernels use?  Producer only -- no MMA -- so the numbers are the ceiling of the operand
// feed: bytes/clk/SM as a function of (box shape, loads in flight, CTAs).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc/kernels -o build/tma_feed_probe \
//        tools/dev/tma_feed_probe.cu && build/tma_feed_probe
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "umma.cuh"

using namespace mpit::umma;

struct Args {
  int kind;      // 0: 2-D box (32 x rows), 1: 4-D NHWC box (32, w, h, n)
  int depth;     // loads in flight
  int iters;
  int bytes;     // per load
  int w, h, n;   // 4-D box
  int rows;      // 2-D box rows
  int taps_w, taps_h, cblocks;
  long long* out;   // per-CTA clocks
};

__global__ void __launch_bounds__(128) probe_kernel(const __grid_constant__ CUtensorMap tm, Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar[16];
  if (threadIdx.x == 0) {
    prefetch_tmap(&tm);
    for (int i = 0; i < a.depth; ++i) mbar_init(&bar[i], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    int stage = 0, phase = 0;
    // incremental coordinates: no integer division on the issue path
    int cb = 0, kw = 0, kh = 0, r0 = ((blockIdx.x * 7) & 7) * a.rows;
    const int n0 = (blockIdx.x * a.n) % 128;
    for (int it = 0; it < a.iters + a.depth; ++it) {
      if (it >= a.depth) {            // wait for the load issued `depth` iterations ago in this stage
        while (!mbar_try_wait(&bar[stage], phase)) {}
      }
      if (it < a.iters) {
        mbar_expect_tx(&bar[stage], a.bytes);
        uint8_t* dst = smem + (size_t)stage * a.bytes;
        if (a.kind == 0) {
          tma_load_2d(dst, &tm, &bar[stage], cb * 32, r0);
          if (++cb == a.cblocks) { cb = 0; r0 += a.rows; if (r0 >= 8 * a.rows) r0 = 0; }
        } else {
          tma_load_4d(dst, &tm, &bar[stage], cb * 32, kw, kh, n0);
          if (++cb == a.cblocks) { cb = 0; if (++kw == a.taps_w) { kw = 0; if (++kh == a.taps_h) kh = 0; } }
        }
      }
      if (++stage == a.depth) { stage = 0; if (it >= a.depth) phase ^= 1; }
    }
    a.out[blockIdx.x] = clock64() - t0;
  }
}

// First-load latency: where does the ~1.5 us between "prologue done" and "first operands landed" in the real
// kernels go?  Variant 0: tensor map passed by value (__grid_constant__, lives in the launch's parameter
// buffer); variant 1: tensor map in ordinary global memory that stays put across launches.
__global__ void __launch_bounds__(128) first_load_kernel(const __grid_constant__ CUtensorMap tm_param,
                                                         const CUtensorMap* tm_global, int use_global, int do_prefetch,
                                                         long long* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar[2];
  const CUtensorMap* tm = use_global ? tm_global : &tm_param;
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    if (do_prefetch) prefetch_tmap(tm);
    mbar_init(&bar[0], 1);
    mbar_init(&bar[1], 1);
    fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long t1 = clock64();
    mbar_expect_tx(&bar[0], 16384);
    tma_load_2d(smem, tm, &bar[0], 0, (blockIdx.x & 7) * 128);
    while (!mbar_try_wait(&bar[0], 0)) {}
    long long t2 = clock64();
    mbar_expect_tx(&bar[1], 16384);
    tma_load_2d(smem + 16384, tm, &bar[1], 32, (blockIdx.x & 7) * 128);
    while (!mbar_try_wait(&bar[1], 0)) {}
    long long t3 = clock64();
    out[blockIdx.x * 4 + 0] = t1 - t0;
    out[blockIdx.x * 4 + 1] = t2 - t1;
    out[blockIdx.x * 4 + 2] = t3 - t2;
  }
}

static decltype(&cuTensorMapEncodeTiled) get_encode() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  return reinterpret_cast<decltype(&cuTensorMapEncodeTiled)>(p);
}

int main() {
  auto enc = get_encode();
  if (!enc) { printf("no encode\n"); return 1; }
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 210 * 1024);
  // sources: NHWC [128][12][12][64] (conv2 input, 4.7 MB) and a 2-D [1024][1600] weight-like matrix (6.5 MB)
  float *x4, *x2;
  cudaMalloc(&x4, sizeof(float) * 128 * 12 * 12 * 64);
  cudaMalloc(&x2, sizeof(float) * 1024 * 1600);
  cudaMemset(x4, 0, sizeof(float) * 128 * 12 * 12 * 64);
  cudaMemset(x2, 0, sizeof(float) * 1024 * 1600);
  long long* out;
  cudaMalloc(&out, sizeof(long long) * 1024);
  long long hout[1024];

  {
    CUtensorMap tm;
    cuuint64_t dims[2] = {1600, 1024};
    cuuint64_t strides[1] = {1600 * 4};
    cuuint32_t box[2] = {32, 128};
    cuuint32_t es[2] = {1, 1};
    enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, x2, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUtensorMap* dtm;
    cudaMalloc(&dtm, sizeof(tm));
    cudaMemcpy(dtm, &tm, sizeof(tm), cudaMemcpyHostToDevice);
    cudaFuncSetAttribute(first_load_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int use_global = 0; use_global < 2; ++use_global)
      for (int pf = 0; pf < 2; ++pf)
        for (int g : {1, 148})
          for (int rep = 0; rep < 3; ++rep) {
            first_load_kernel<<<g, 128, 40 * 1024>>>(tm, dtm, use_global, pf, out);
            cudaDeviceSynchronize();
            cudaMemcpy(hout, out, sizeof(long long) * 4 * g, cudaMemcpyDeviceToHost);
            double a = 0, b = 0, c = 0;
            for (int i = 0; i < g; ++i) { a += hout[4 * i]; b += hout[4 * i + 1]; c += hout[4 * i + 2]; }
            printf("first-load tmap=%s prefetch=%d ctas %3d rep %d : setup %6.0f  first %6.0f  second %6.0f clk\n",
                   use_global ? "global" : "param ", pf, g, rep, a / g, b / g, c / g);
          }
  }
  struct Case { const char* name; int kind, rows, w, h, n; };
  Case cases[] = {
      {"2D 32x128 (16KB)", 0, 128, 0, 0, 0},
      {"2D 32x64  (8KB)", 0, 64, 0, 0, 0},
      {"2D 32x256 (32KB)", 0, 256, 0, 0, 0},
      {"4D 32x8x8x2 (16KB)", 1, 0, 8, 8, 2},
      {"4D 32x4x4x8 (16KB)", 1, 0, 4, 4, 8},
      {"4D 32x2x2x32 (16KB)", 1, 0, 2, 2, 32},
      {"4D 32x12x8x1 (12KB)", 1, 0, 12, 8, 1},
  };
  int depths[] = {1, 2, 4, 8, 12};
  int ctas[] = {1, 8, 64, 148};
  for (auto& c : cases) {
    CUtensorMap tm;
    int bytes;
    if (c.kind == 0) {
      cuuint64_t dims[2] = {1600, 1024};
      cuuint64_t strides[1] = {1600 * 4};
      cuuint32_t box[2] = {32, (cuuint32_t)c.rows};
      cuuint32_t es[2] = {1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, x2, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r) { printf("encode2 %d\n", (int)r); return 1; }
      bytes = 128 * c.rows;
    } else {
      cuuint64_t dims[4] = {64, 12, 12, 128};
      cuuint64_t strides[3] = {64 * 4, 64 * 4 * 12, 64 * 4 * 144};
      cuuint32_t box[4] = {32, (cuuint32_t)c.w, (cuuint32_t)c.h, (cuuint32_t)c.n};
      cuuint32_t es[4] = {1, 1, 1, 1};
      CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, x4, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r) { printf("encode4 %d\n", (int)r); return 1; }
      bytes = 128 * c.w * c.h * c.n;
    }
    for (int d : depths) {
      if ((size_t)d * bytes > 200 * 1024) continue;
      for (int g : ctas) {
        Args a{};
        a.kind = c.kind; a.depth = d; a.iters = 400; a.bytes = bytes; a.w = c.w; a.h = c.h; a.n = c.n; a.rows = c.rows;
        a.taps_w = 12 - c.w + 1 > 5 ? 5 : (12 - c.w + 1 < 1 ? 1 : 12 - c.w + 1);
        a.taps_h = 12 - c.h + 1 > 5 ? 5 : (12 - c.h + 1 < 1 ? 1 : 12 - c.h + 1);
        a.cblocks = c.kind == 0 ? 50 : 2;
        a.out = out;
        for (int rep = 0; rep < 2; ++rep) {   // second run is the measurement (L2 warm)
          probe_kernel<<<g, 128, (size_t)d * bytes + 2048>>>(tm, a);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("launch error %s\n", cudaGetErrorString(e)); return 1; }
        }
        cudaMemcpy(hout, out, sizeof(long long) * g, cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int i = 0; i < g; ++i) mx = hout[i] > mx ? hout[i] : mx;
        double clk_per_load = (double)mx / a.iters;
        printf("%-22s depth %2d ctas %3d : %7.1f clk/load  %6.1f B/clk/SM  chip %7.0f B/clk\n", c.name, d, g, clk_per_load,
               bytes / clk_per_load, g * bytes / clk_per_load);
      }
    }
  }
  return 0;
}
