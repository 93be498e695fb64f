// Bidirectional LSTM layer - the recurrent half (mlx_audio/tts/models/kokoro/modules.py:93-286: `LSTM`, gate order
// i, f, g, o; the `shared` layer of ProsodyPredictor.F0Ntrain :355-358).  The input projection x Wx^T + b_ih + b_hh of
// all time steps is one GEMM (kk_conv1d_tf32_tc as a 1x1 convolution); this kernel runs the L dependent steps
//     z = xw[t] + h Wh^T,   c = sigmoid(z_f) c + sigmoid(z_i) tanh(z_g),   h = sigmoid(z_o) tanh(c).
// Layout of the work: a thread-block CLUSTER of 8 CTAs owns one direction of 8 sequences.  CTA j of the cluster holds
// the recurrent weights of hidden units [j H/8, (j+1) H/8) - all four gates, H x 4 (H/8) floats, 128 KB at H = 256 -
// in shared memory for the whole sequence, so a step moves nothing but h: every CTA writes its slice of h(t) to a
// double-buffered global scratch, the cluster meets at barrier.cluster (release / acquire), and every CTA reads the
// 8 x H block back (L2 hits, ld.global.cg).  One barrier per step; 16 clusters = 128 CTAs cover B = 64.
// Ragged batches: `lens` = valid steps per sequence; the backward direction starts at each sequence's own last step,
// finished sequences stop updating, outputs past a sequence's end are zeros.
#include "common.cuh"
#include <cuda_runtime.h>

namespace kk {

constexpr int LSTM_CS = 8;     // CTAs per cluster = slices of the hidden units
constexpr int LSTM_R = 8;      // sequences per cluster

__device__ __forceinline__ void cluster_sync_acqrel() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
// packed fp32 (sm_100 FFMA2): two rows of one column per instruction, each lane the plain IEEE fma
typedef unsigned long long f2;
__device__ __forceinline__ f2 pack2(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float2 unpack2(f2 v) { float2 r; asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
constexpr int LSTM_KQ = 4;     // the k range of the recurrent product is split over four threads per column pair

// grid (LSTM_CS, ceil(B / LSTM_R), 2), cluster (LSTM_CS, 1, 1), block LSTM_R * UC threads (UC = H / LSTM_CS).
__global__ void lstm_bidir_cluster_kernel(const float* __restrict__ xw, const float* __restrict__ whT, float* hbuf,
                                          float* __restrict__ out, int B, int L, int H, const int* __restrict__ lens) {
  extern __shared__ __align__(16) unsigned char smem_[];
  const int UC = H / LSTM_CS, NCOL = 4 * UC;
  float* ws = reinterpret_cast<float*>(smem_);          // [H][NCOL]   recurrent weights of this CTA's units
  float* hs = ws + (size_t)H * NCOL;                    // [H][LSTM_R] h(t-1) of the cluster's sequences, k-major
  float* pre = hs + (size_t)H * LSTM_R;                 // [LSTM_KQ][LSTM_R][NCOL] partial h Wh^T of this step
  const int j = (int)cluster_ctarank();                 // hidden-unit slice
  const int dir = blockIdx.z, b0 = blockIdx.y * LSTM_R;
  const int tid = threadIdx.x, nthr = blockDim.x;

  // resident weights: column (g, uu) of this CTA = row g H + j UC + uu of Wh = column of whT[dir]
  for (int i = tid; i < H * NCOL; i += nthr) {
    const int k = i / NCOL, c = i - k * NCOL, g = c / UC, uu = c - g * UC;
    ws[i] = whT[((size_t)dir * H + k) * (4 * H) + g * H + j * UC + uu];
  }
  for (int i = tid; i < H * LSTM_R; i += nthr) hs[i] = 0.f;

  // gate phase: thread -> (sequence r, unit uu); matvec phase: thread -> (column pair, quarter of the k range), all
  // eight sequences: per k one 64-bit load of two weights, two 128-bit broadcast loads of h and 8 FFMA2 (the first
  // version - one column x four sequences per thread, 2 loads per 4 FFMA - was bound by the shared-memory latency)
  const int r = tid / UC, uu = tid - r * UC;
  const int b = b0 + r;
  const bool row_ok = b < B;
  const int len = row_ok ? (lens ? min(max(lens[b], 0), L) : L) : 0;
  int lmax = 0;
  for (int i = 0; i < LSTM_R; ++i) {
    const int bi = b0 + i;
    if (bi < B) lmax = max(lmax, lens ? min(max(lens[bi], 0), L) : L);
  }
  const int cp = tid % (NCOL / 2), kq = tid / (NCOL / 2);
  const int kper = H / LSTM_KQ;
  float cstate = 0.f;
  const size_t xw_row = (size_t)2 * 4 * H;
  __syncthreads();

  for (int s = 0; s < lmax; ++s) {
    const bool active = s < len;
    const int t = dir == 0 ? s : len - 1 - s;
    // this step's input projections (four gates of this thread's cell): in flight during the matvec
    float zi = 0.f, zf = 0.f, zg = 0.f, zo = 0.f;
    if (active) {
      const float* p = xw + ((size_t)b * L + t) * xw_row + (size_t)dir * 4 * H + j * UC + uu;
      zi = __ldg(p); zf = __ldg(p + H); zg = __ldg(p + 2 * H); zo = __ldg(p + 3 * H);
    }
    // h(t-1) Wh^T for this CTA's columns: partial sums over this thread's quarter of k
    f2 acc[2][LSTM_R / 2];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
      for (int q = 0; q < LSTM_R / 2; ++q) acc[c][q] = 0ull;
    const float2* wp = reinterpret_cast<const float2*>(ws + (size_t)kq * kper * NCOL) + cp;
    const float4* hp = reinterpret_cast<const float4*>(hs + (size_t)kq * kper * LSTM_R);
#pragma unroll 4
    for (int k = 0; k < kper; ++k) {
      const float2 w = wp[(size_t)k * (NCOL / 2)];
      const float4 hA = hp[2 * k], hB = hp[2 * k + 1];
      const f2 w0 = pack2(w.x, w.x), w1 = pack2(w.y, w.y);
      const f2 h01 = pack2(hA.x, hA.y), h23 = pack2(hA.z, hA.w), h45 = pack2(hB.x, hB.y), h67 = pack2(hB.z, hB.w);
      acc[0][0] = fma2(h01, w0, acc[0][0]); acc[0][1] = fma2(h23, w0, acc[0][1]);
      acc[0][2] = fma2(h45, w0, acc[0][2]); acc[0][3] = fma2(h67, w0, acc[0][3]);
      acc[1][0] = fma2(h01, w1, acc[1][0]); acc[1][1] = fma2(h23, w1, acc[1][1]);
      acc[1][2] = fma2(h45, w1, acc[1][2]); acc[1][3] = fma2(h67, w1, acc[1][3]);
    }
    {
      float2* pp = reinterpret_cast<float2*>(pre + (size_t)kq * LSTM_R * NCOL) + cp;   // (row, column pair)
#pragma unroll
      for (int q = 0; q < LSTM_R / 2; ++q) {
        const float2 c0 = unpack2(acc[0][q]), c1 = unpack2(acc[1][q]);
        pp[(size_t)(2 * q) * (NCOL / 2)] = make_float2(c0.x, c1.x);
        pp[(size_t)(2 * q + 1) * (NCOL / 2)] = make_float2(c0.y, c1.y);
      }
    }
    __syncthreads();
    float* hw = hbuf + (((size_t)(s & 1) * 2 + dir) * B) * H;   // h(t) of every sequence, this parity
    if (active) {
      const float* pr = pre + r * NCOL + uu;
      float z[4] = {zi, zf, zg, zo};
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float hw_ = pr[g * UC];
#pragma unroll
        for (int q = 1; q < LSTM_KQ; ++q) hw_ += pr[(size_t)q * LSTM_R * NCOL + g * UC];   // fixed order
        z[g] += hw_;
      }
      const float ig = sigmoidf_(z[0]), fg = sigmoidf_(z[1]);
      const float gg = tanhf(z[2]), og = sigmoidf_(z[3]);
      cstate = fmaf(fg, cstate, ig * gg);
      const float h = og * tanhf(cstate);
      __stcg(hw + (size_t)b * H + j * UC + uu, h);
      out[((size_t)b * L + t) * (2 * H) + dir * H + j * UC + uu] = h;
    } else if (row_ok) {
      // past the sequence's end: its state is no longer used (rows are independent); the exchange slot gets a defined
      // value and the output row s is zero
      __stcg(hw + (size_t)b * H + j * UC + uu, 0.f);
      out[((size_t)b * L + s) * (2 * H) + dir * H + j * UC + uu] = 0.f;
    }
    cluster_sync_acqrel();   // every slice of h(t) is written and visible cluster-wide
    // reload the full h(t) of the cluster's sequences, k-major
    for (int i = tid; i < H * LSTM_R; i += nthr) {
      const int rr = i / H, k = i - rr * H;
      hs[k * LSTM_R + rr] = (b0 + rr < B) ? __ldcg(hw + (size_t)(b0 + rr) * H + k) : 0.f;
    }
    __syncthreads();
  }
  // sequences shorter than L inside a cluster whose longest is shorter than L too: zero the remaining rows
  if (row_ok)
    for (int s = max(lmax, len); s < L; ++s) out[((size_t)b * L + s) * (2 * H) + dir * H + j * UC + uu] = 0.f;
}

}  // namespace kk

extern "C" int kk_lstm_bidir_f32(const float* xw, const float* whT, float* h_ws, float* out, int B, int L, int H,
                                 const int* lens, void* stream) {
  KK_CHECK_ARG(xw && whT && h_ws && out && B > 0 && L > 0, "kk_lstm_bidir_f32: bad arguments");
  if (H < 64 || H > 256 || H % 64 != 0) {
    kk::set_error("kk_lstm_bidir_f32: hidden size %d (supported: 64, 128, 192, 256)", H);
    return KK_ERR_UNSUPPORTED;
  }
  const int UC = H / kk::LSTM_CS, NCOL = 4 * UC;
  const size_t smem = ((size_t)H * NCOL + (size_t)H * kk::LSTM_R + (size_t)kk::LSTM_KQ * kk::LSTM_R * NCOL) * sizeof(float);
  KK_CUDA(cudaFuncSetAttribute(kk::lstm_bidir_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kk::LSTM_CS, (unsigned)kk::cdiv(B, kk::LSTM_R), 2);
  cfg.blockDim = dim3(kk::LSTM_R * UC);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = (cudaStream_t)stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kk::LSTM_CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  KK_CUDA(cudaLaunchKernelEx(&cfg, kk::lstm_bidir_cluster_kernel, xw, whT, h_ws, out, B, L, H, lens));
  KK_LAUNCH_CHECK("lstm_bidir_cluster_kernel");
  return KK_OK;
}
