// Experiment: HBM read throughput of 148 persistent CTAs as a function of bytes in flight per SM.
//  mode 0: register loads - each of T threads keeps U x 16 B loads in flight (issue U, consume, repeat)
//  mode 1: cp.async.bulk (TMA 1-D bulk) global->shared ring of NST stages x CH bytes, one issuing thread
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 experiments/inflight_bw.cu -o experiments/inflight_bw.bin
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>

template <int U>
__global__ void reg_loads(const float4* __restrict__ x, size_t n_per_cta, float* sink) {
  const float4* p = x + (size_t)blockIdx.x * n_per_cta;
  float acc = 0.f;
  for (size_t i = threadIdx.x; i + (size_t)(U - 1) * blockDim.x < n_per_cta; i += (size_t)U * blockDim.x) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) v[u] = __ldg(p + i + (size_t)u * blockDim.x);
#pragma unroll
    for (int u = 0; u < U; ++u) acc += v[u].x + v[u].y + v[u].z + v[u].w;
  }
  if (acc == 123.456f) sink[0] = acc;
}

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void wait_par(uint64_t* bar, uint32_t ph) {
  uint32_t ok = 0;
  while (!ok)
    asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.b32 %0,1,0,p;}"
                 : "=r"(ok) : "r"(s32(bar)), "r"(ph) : "memory");
}

__global__ void bulk_loads(const char* __restrict__ x, size_t bytes_per_cta, int nst, int ch, float* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);
  uint64_t* empty = full + 16;
  uint8_t* buf = smem + 256;
  if (threadIdx.x == 0) {
    for (int s = 0; s < nst; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&full[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 128;" ::"r"(s32(&empty[s])));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const char* p = x + (size_t)blockIdx.x * bytes_per_cta;
  const int nchunks = (int)(bytes_per_cta / ch);
  if (threadIdx.x == 0) {
    for (int i = 0; i < nchunks; ++i) {
      int s = i % nst; uint32_t ph = (i / nst) & 1;
      wait_par(&empty[s], ph ^ 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(&full[s])), "r"(ch) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(s32(buf + (size_t)s * ch)), "l"(p + (size_t)i * ch), "r"(ch), "r"(s32(&full[s])) : "memory");
    }
  } else if (threadIdx.x >= 32 && threadIdx.x < 160) {
    float acc = 0.f;
    const int t = threadIdx.x - 32;
    for (int i = 0; i < nchunks; ++i) {
      int s = i % nst; uint32_t ph = (i / nst) & 1;
      wait_par(&full[s], ph);
      const float4* b4 = reinterpret_cast<const float4*>(buf + (size_t)s * ch);
      for (int j = t; j < ch / 16; j += 128) { float4 v = b4[j]; acc += v.x + v.y + v.z + v.w; }
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(&empty[s])) : "memory");
    }
    if (acc == 123.456f) sink[0] = acc;
  }
}

int main() {
  const size_t total = (size_t)148 * 8 * 1024 * 1024;  // 1.24 GB
  char* x; float* sink;
  cudaMalloc(&x, total); cudaMalloc(&sink, 4); cudaMemset(x, 0, total);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const size_t per = total / 148;
  auto report = [&](const char* name, int kb) {
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    printf("%-28s in-flight/SM %4d KB : %.3f ms  %.2f TB/s  (%s)\n", name, kb, ms, total / ms / 1e9,
           cudaGetErrorString(cudaGetLastError()));
  };
  for (int rep = 0; rep < 2; ++rep) {
    for (int T : {128, 192, 256, 512}) {
      cudaEventRecord(e0); reg_loads<4><<<148, T>>>((const float4*)x, per / 16, sink); report("reg U=4", T * 4 * 16 / 1024);
      cudaEventRecord(e0); reg_loads<8><<<148, T>>>((const float4*)x, per / 16, sink); report("reg U=8", T * 8 * 16 / 1024);
      cudaEventRecord(e0); reg_loads<16><<<148, T>>>((const float4*)x, per / 16, sink); report("reg U=16", T * 16 * 16 / 1024);
    }
    cudaFuncSetAttribute(bulk_loads, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    for (int ch : {8192, 16384, 32768})
      for (int nst : {2, 4, 6}) {
        if ((size_t)ch * nst > 190 * 1024) continue;
        cudaEventRecord(e0); bulk_loads<<<148, 192, 256 + ch * nst>>>(x, per, nst, ch, sink);
        char nm[64]; snprintf(nm, 64, "bulk ch=%d nst=%d", ch, nst); report(nm, ch * nst / 1024);
      }
  }
  return 0;
}
