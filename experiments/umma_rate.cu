// Probe: cycles per tcgen05.mma (kind::f16, bf16, M = 128, K = 16, operands K-major SWIZZLE_128B in shared memory) as a
// function of N, issued back to back by one thread - into ONE accumulator (the dependent chain of a K loop) or into TWO
// alternating accumulators - on one CTA and on 148 CTAs at once.  Answers: is there a fixed per-instruction cost that a
// narrow N cannot get under (conv_post: ~107 clk per 128 x 32 x 16 MMA), and what does a 128 x 256 x 16 MMA cost when
// nothing else uses shared memory (the fused conv kernel sees ~216 clk per MMA at k = 11)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I mlx_audio_b200/csrc experiments/umma_rate.cu -o /tmp/umma_rate
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "ptx_sm100.cuh"
using namespace kk;

constexpr int NMMA = 1024;

template <int N>
__global__ void __launch_bounds__(128) rate(long long* out, int alt, int shift_rows) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = smem;                       // 320 rows x 128 B (zeros)
  uint8_t* sB = smem + 320 * 128;           // 256 rows x 128 B
  uint64_t* bar = (uint64_t*)(sB + 256 * 128);
  uint32_t* slot = (uint32_t*)(bar + 1);
  for (int i = threadIdx.x; i < (320 + 256) * 128 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
  ptx::fence_proxy_async();
  if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) ptx::tmem_alloc<512>(slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, N);
    const uint64_t ad0 = ptx::umma_desc(ptx::smem_u32(sA), 16, 1024, ptx::UMMA_LAYOUT_SW128);
    const uint64_t bd0 = ptx::umma_desc(ptx::smem_u32(sB), 16, 1024, ptx::UMMA_LAYOUT_SW128);
    const long long t0 = clock64();
    // groups of 8 MMAs, fully unrolled with constant descriptor offsets: the issue loop itself costs a few instructions
    // per MMA.  shift_rows: the second four of a group read A through a view shifted by 3 rows (the conv kernel's taps)
    const uint32_t acc1 = alt ? 256u : 0u;
    const uint64_t sh = (uint64_t)(shift_rows * 3 * 8);
    for (int i = 0; i < NMMA; i += 8) {
#pragma unroll
      for (int u = 0; u < 8; ++u)
        ptx::umma_bf16(tmem + ((u & 4) ? acc1 : 0u), ad0 + (uint64_t)((u & 3) * 2) + ((u & 4) ? sh : 0), bd0 + (uint64_t)((u & 3) * 2), idesc,
                       (i | u) != 0);
    }
    const long long t1 = clock64();
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0);
    const long long t2 = clock64();
    out[blockIdx.x * 2] = t1 - t0;        // issue time
    out[blockIdx.x * 2 + 1] = t2 - t0;    // issue + drain
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) ptx::tmem_dealloc<512>(tmem);
}

template <int N>
static void run(long long* d_out, int grid) {
  const int smem = (320 + 256) * 128 + 64 + 1024;
  cudaFuncSetAttribute(rate<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<long long> h(grid * 2);
  for (int shift = 0; shift <= 1; ++shift)
    for (int alt = 0; alt <= 1; ++alt) {
      rate<N><<<grid, 128, smem>>>(d_out, alt, shift);   // warm-up
      rate<N><<<grid, 128, smem>>>(d_out, alt, shift);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); exit(1); }
      cudaMemcpy(h.data(), d_out, h.size() * 8, cudaMemcpyDeviceToHost);
      double issue = 0, total = 0;
      for (int i = 0; i < grid; ++i) { issue += h[2 * i]; total += h[2 * i + 1]; }
      printf("M=128 N=%3d grid=%3d %s %s: issue %6.1f clk/MMA, issue+drain %6.1f clk/MMA (ideal %d)\n", N, grid,
             alt ? "two accumulators" : "one accumulator ", shift ? "row-shifted A views" : "fixed operands     ",
             issue / grid / NMMA, total / grid / NMMA, 128 * N / 256);
    }
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 148 * 2 * 8);
  for (int grid : {1, 148}) {
    run<32>(d_out, grid);
    run<64>(d_out, grid);
    run<128>(d_out, grid);
    run<256>(d_out, grid);
  }
  return 0;
}
