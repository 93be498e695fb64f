// Experiment: does a tcgen05 shared-memory descriptor (K-major, SWIZZLE_128B) whose start address is shifted by
// r rows (r*128 bytes, not a multiple of the 1024-byte swizzle atom) read rows r..r+127 of a tile that was
// written with the absolute-address 128B swizzle?  Tried with base_offset = 0 and base_offset = (start>>7)&7.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I mlx_audio_b200/csrc experiments/umma_rowshift.cu -o /tmp/umma_rowshift
#include <cstdio>
#include "ptx_sm100.cuh"
#include <cuda_bf16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
using namespace kk;

constexpr int ROWS = 256, N = 64, K = 64;

__global__ void __launch_bounds__(128) probe(const __nv_bfloat16* A, const __nv_bfloat16* Bm, float* D, int shift, int use_bo) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                 // ROWS x 128 B
  uint8_t* sB = smem + ROWS * 128;    // N x 128 B
  uint64_t* bar = (uint64_t*)(sB + N * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  for (int i = threadIdx.x; i < ROWS * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    *(__nv_bfloat16*)(sA + r * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2) = A[i];
  }
  for (int i = threadIdx.x; i < N * K; i += blockDim.x) {
    int r = i / K, k = i % K;
    *(__nv_bfloat16*)(sB + r * 128 + (((k >> 3) ^ (r & 7)) << 4) + (k & 7) * 2) = Bm[i];
  }
  ptx::fence_proxy_async();
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) ptx::tmem_alloc<64>(slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, N);
    uint32_t a0 = ptx::smem_u32(sA) + shift * 128;
    uint32_t b0 = ptx::smem_u32(sB);
    uint32_t bo = use_bo ? ((a0 >> 7) & 7) : 0;
    for (int k = 0; k < K / 16; ++k) {
      uint64_t ad = ptx::umma_desc(a0 + k * 32, 16, 1024, ptx::UMMA_LAYOUT_SW128, bo);
      uint64_t bd = ptx::umma_desc(b0 + k * 32, 16, 1024, ptx::UMMA_LAYOUT_SW128);
      ptx::umma_bf16(tmem, ad, bd, idesc, k != 0);
    }
    ptx::umma_commit(bar);
  }
  ptx::mbar_wait(bar, 0);
  ptx::tc_fence_after();
  int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int ch = 0; ch < N / 32; ++ch) {
    float v[32];
    ptx::tmem_ld_32x32(tmem + ((uint32_t)(warp * 32) << 16) + ch * 32, v);
    ptx::tmem_ld_wait();
    for (int i = 0; i < 32; ++i) D[(warp * 32 + lane) * N + ch * 32 + i] = v[i];
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc<64>(tmem);
}

int main() {
  std::vector<__nv_bfloat16> hA(ROWS * K), hB(N * K);
  std::vector<float> fA(ROWS * K), fB(N * K);
  srand(1);
  for (int i = 0; i < ROWS * K; ++i) { fA[i] = (float)(rand() % 17 - 8); hA[i] = __float2bfloat16(fA[i]); }
  for (int i = 0; i < N * K; ++i) { fB[i] = (float)(rand() % 9 - 4); hB[i] = __float2bfloat16(fB[i]); }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * N * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  int smem = ROWS * 128 + N * 128 + 64 + 1024;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> hD(128 * N);
  for (int bo = 0; bo < 2; ++bo)
    for (int shift = 0; shift <= 17; ++shift) {
      probe<<<1, 128, smem>>>(dA, dB, dD, shift, bo);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("shift %d bo %d: CUDA error %s\n", shift, bo, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
      int bad = 0;
      for (int i = 0; i < 128; ++i)
        for (int n = 0; n < N; ++n) {
          float ref = 0;
          for (int k = 0; k < K; ++k) ref += fA[(shift + i) * K + k] * fB[n * K + k];
          if (ref != hD[i * N + n]) ++bad;
        }
      printf("shift %2d base_offset_mode %d: %s (%d mismatches)\n", shift, bo, bad ? "WRONG" : "ok", bad);
    }
  return 0;
}
