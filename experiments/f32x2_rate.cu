// Probe: issue rate of the packed fp32 instructions of sm_100 (FADD2 / FFMA2 = add.rn.f32x2 / fma.rn.f32x2) against the
// scalar FADD / FFMA: 16 independent dependency chains per thread, 8 / 16 / 32 warps per SM, 148 blocks.
// Prints lane-operations per clock per SM (128 = one scalar instruction per scheduler per clock).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O2 experiments/f32x2_rate.cu -o /tmp/f32x2_rate
#include <cstdio>
#include <cstdint>
typedef unsigned long long c2;
constexpr int ITERS = 2048, CH = 16;

template <int MODE>   // 0 FADD, 1 FFMA, 2 FADD2, 3 FFMA2
__global__ void rate(float* out, long long* clk, float seed) {
  float a[CH]; c2 p[CH];
  for (int i = 0; i < CH; ++i) { a[i] = seed * (i + threadIdx.x); asm("mov.b64 %0, {%1, %1};" : "=l"(p[i]) : "f"(a[i])); }
  c2 m; asm("mov.b64 %0, {%1, %1};" : "=l"(m) : "f"(seed));
  const long long t0 = clock64();
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CH; ++i) {
      if (MODE == 0) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(a[i]) : "f"(seed));
      if (MODE == 1) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(a[i]) : "f"(seed));
      if (MODE == 2) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(m));
      if (MODE == 3) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[i]) : "l"(m));
    }
  }
  const long long t1 = clock64();
  float s = 0.f;
  for (int i = 0; i < CH; ++i) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(p[i])); s += a[i] + lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  // slowest warp of the block: the scheduler is not fair, one warp's own loop time understates the block's
  __shared__ long long tmin, tmax;
  if (threadIdx.x == 0) { tmin = t0; tmax = t1; }
  __syncthreads();
  if ((threadIdx.x & 31) == 0) { atomicMin(&tmin, t0); atomicMax(&tmax, t1); }
  __syncthreads();
  if (threadIdx.x == 0) clk[blockIdx.x] = tmax - tmin;
}

template <int MODE>
static void run(const char* name, float* out, long long* clk) {
  for (int warps : {4, 8, 16, 32}) {
    rate<MODE><<<148, warps * 32>>>(out, clk, 1.0001f);
    rate<MODE><<<148, warps * 32>>>(out, clk, 1.0001f);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return; }
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const double lanes = (double)ITERS * CH * warps * 32 * (MODE >= 2 ? 2 : 1);
    printf("%-6s %2d warps/SM: %7.0f clk, %6.1f instr/clk/SM, %6.1f lane-ops/clk/SM\n", name, warps, avg,
           (double)ITERS * CH * warps / avg, lanes / avg);
  }
}
int main() {
  float* out; long long* clk;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&clk, 148 * 8);
  run<0>("FADD", out, clk); run<1>("FFMA", out, clk); run<2>("FADD2", out, clk); run<3>("FFMA2", out, clk);
  return 0;
}
