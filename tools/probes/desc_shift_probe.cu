// Probe: can a K-major SWIZZLE_128B operand start at an arbitrary 128-byte ROW of a TMA-written tile?
// (needed to feed tcgen05.mma from a resident, padded image with a per-tap row shift instead of re-fetching the
// shifted window for every tap).  A = rows [s, s+128) of a 256-row tile, B = 32x32 identity, so D[r][n] must equal
// src[s+r][n].  Tried with the descriptor's base-offset field (bits 49..51) = 0 and = (start_address >> 7) & 7.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc/kernels -o build/desc_shift_probe tools/dev/desc_shift_probe.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

#include "umma.cuh"

using namespace mpit::umma;

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
                                             int shift, int use_base_offset, float* out) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_s = smem;                 // 256 rows x 128 B = 32 KB
  uint8_t* b_s = smem + 32768;         // 32 rows x 128 B
  __shared__ uint64_t bar, done;
  __shared__ uint32_t tslot;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&done, 1); fence_barrier_init(); }
  if (warp == 0) tmem_alloc(&tslot, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tslot;
  if (threadIdx.x == 0) {
    mbar_expect_tx(&bar, 32768 + 4096);
    tma_load_2d(a_s, &tm_a, &bar, 0, 0);
    tma_load_2d(b_s, &tm_b, &bar, 0, 0);
    while (!mbar_try_wait(&bar, 0)) {}
    tc_fence_after();
    constexpr uint32_t idesc = make_idesc_tf32(128, 32, 0, 0);
    const uint32_t a_addr = smem_u32(a_s) + shift * 128;
    for (int k = 0; k < 4; ++k) {
      uint64_t ad = make_smem_desc(a_addr + k * 32, 16, 1024);
      if (use_base_offset) ad |= (uint64_t)((a_addr >> 7) & 7) << 49;
      uint64_t bd = make_smem_desc(smem_u32(b_s) + k * 32, 16, 1024);
      mma_tf32(tmem, ad, bd, idesc, k ? 1u : 0u);
    }
    mma_commit(&done);
  }
  __syncthreads();
  while (!mbar_try_wait(&done, 0)) {}
  tc_fence_after();
  float v[32];
  tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16), v);
  for (int c = 0; c < 32; ++c) out[threadIdx.x * 32 + c] = v[c];
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc(tmem, 32); }
}

int main() {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  auto enc = reinterpret_cast<decltype(&cuTensorMapEncodeTiled)>(fp);
  std::vector<float> ha(256 * 32), hb(32 * 32, 0.f);
  for (int r = 0; r < 256; ++r) for (int c = 0; c < 32; ++c) ha[r * 32 + c] = (float)((r * 37 + c * 5) % 1021);   // integers < 2^11 are exact in tf32
  for (int i = 0; i < 32; ++i) hb[i * 32 + i] = 1.f;     // B[n][k] = delta
  float *da, *db, *dout;
  cudaMalloc(&da, ha.size() * 4); cudaMalloc(&db, hb.size() * 4); cudaMalloc(&dout, 128 * 32 * 4);
  cudaMemcpy(da, ha.data(), ha.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(db, hb.data(), hb.size() * 4, cudaMemcpyHostToDevice);
  CUtensorMap ta, tb;
  cuuint64_t dims[2] = {32, 256}; cuuint64_t st[1] = {128}; cuuint32_t box[2] = {32, 256}; cuuint32_t es[2] = {1, 1};
  enc(&ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, da, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  cuuint64_t dimsb[2] = {32, 32}; cuuint32_t boxb[2] = {32, 32};
  enc(&tb, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, db, dimsb, st, boxb, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
      CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024);
  std::vector<float> ho(128 * 32);
  for (int ubo = 0; ubo < 2; ++ubo)
    for (int shift : {0, 1, 2, 3, 4, 5, 7, 8, 9, 13, 16, 21, 68, 100}) {
      cudaMemset(dout, 0, 128 * 32 * 4);
      probe<<<1, 128, 40 * 1024>>>(ta, tb, shift, ubo, dout);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("base_offset=%d shift=%d: CUDA error %s\n", ubo, shift, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost);
      int bad = 0, first = -1;
      for (int r = 0; r < 128; ++r) for (int c = 0; c < 32; ++c)
        if (ho[r * 32 + c] != ha[(r + shift) * 32 + c]) { if (first < 0) first = r * 32 + c; ++bad; }
      printf("base_offset=%d shift=%3d: %s (%d wrong", ubo, shift, bad ? "MISMATCH" : "ok", bad);
      if (bad) printf(", first at row %d col %d: got %.0f want %.0f", first / 32, first % 32, ho[first], ha[first + shift * 32]);
      printf(")\n");
    }
  return 0;
}
