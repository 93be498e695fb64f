// Bidirectional LSTM layer - the recurrent half (mlx_audio/tts/models/kokoro/modules.py:93-286: `LSTM`, gate order
// i, f, g, o; the `shared` layer of ProsodyPredictor.F0Ntrain :355-358).  The input projection x Wx^T + b_ih + b_hh of
// all time steps is one GEMM (kk_conv1d_tf32_tc as a 1x1 convolution); this kernel runs the L dependent steps
//     z = xw[t] + h Wh^T,   c = sigmoid(z_f) c + sigmoid(z_i) tanh(z_g),   h = sigmoid(z_o) tanh(c).
// Layout of the work: a thread-block CLUSTER of 8 CTAs owns one direction of R = 8, 10 or 12 sequences (the host picks
// R so that every cluster of the grid is co-resident: a B200 holds 15 of them).  CTA j of the cluster holds
// the recurrent weights of hidden units [j H/8, (j+1) H/8) - all four gates, H x 4 (H/8) floats, 128 KB at H = 256 -
// in REGISTERS for the whole sequence (64 per thread), so a step moves nothing but h: every CTA stores its slice of h(t) straight
// into the (double-buffered) h block of all eight CTAs - distributed shared memory, st.async.shared::cluster with
// mbarrier complete_tx - and every CTA waits on its OWN mbarrier until the R x H floats of the step have arrived: no
// cluster-wide barrier and no fence in the loop (barrier.cluster's release / acquire compile to MEMBAR.ALL.GPU and an
// L1 invalidate, ~1.2 k clocks a step).  14 clusters of 10 sequences = 112 CTAs cover B = 64 in one wave.
// Ragged batches: `lens` = valid steps per sequence; the backward direction starts at each sequence's own last step,
// finished sequences stop updating, outputs past a sequence's end are zeros.
#include "common.cuh"
#include "ptx_sm100.cuh"
#include <cuda_runtime.h>

namespace kk {
using namespace ptx;

constexpr int LSTM_CS = 8;     // CTAs per cluster = slices of the hidden units

__device__ __forceinline__ void cluster_sync_acqrel() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
#ifdef KK_LSTM_DBG
__device__ unsigned long long g_lstm_clk[8];   // lab build: clocks of one thread per phase (load issue, matvec, sync, gate, barrier)
#define LSTM_MARK(i) do { if (dbg) { const long long c_ = clock64(); g_lstm_clk[i] += (unsigned long long)(c_ - tmark); tmark = c_; } } while (0)
#else
#define LSTM_MARK(i) do {} while (0)
#endif
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
// packed fp32 (sm_100 FFMA2): the two columns of a pair per instruction, each lane the plain IEEE fma
typedef unsigned long long f2;
__device__ __forceinline__ f2 pack2(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ float2 unpack2(f2 v) { float2 r; asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
#ifndef LSTM_KQ_
#define LSTM_KQ_ 8   // 4: 2.06 ms, 8: 1.88 ms at B = 64 x L = 320, H = 256 (twice the warps hide the FFMA2 / ld.shared latency)
#endif
constexpr int LSTM_KQ = LSTM_KQ_;   // the k range of the recurrent product is split over this many threads per column pair

// grid (LSTM_CS, ceil(B / R), 2), cluster (LSTM_CS, 1, 1), block 2 UC LSTM_KQ threads (UC = H / LSTM_CS), the first R UC of
// them own a cell.  R = sequences per cluster: the result of a sequence does not depend on it (its sums run over k in the
// same order whoever its neighbours are); the host picks it so that all clusters are co-resident.
template <int R, int KPER>
__global__ void __launch_bounds__(LSTM_KQ * 64, 1) lstm_bidir_cluster_kernel(const float* __restrict__ xw, const float* __restrict__ whT,
                                          float* __restrict__ out, int B, int L, int H, const int* __restrict__ lens) {
  extern __shared__ __align__(16) unsigned char smem_[];
  const int UC = H / LSTM_CS, NCOL = 4 * UC;
  float* hs = reinterpret_cast<float*>(smem_);          // [2][R][H]       h(t-1) / h(t) of the cluster's sequences
  float* pre = hs + (size_t)2 * R * H;             // [LSTM_KQ][R][NCOL] partial h Wh^T of this step
  const int j = (int)cluster_ctarank();                 // hidden-unit slice
  const int dir = blockIdx.z, b0 = blockIdx.y * R;
  const int tid = threadIdx.x, nthr = blockDim.x;

  for (int i = tid; i < 2 * R * H; i += nthr) hs[i] = 0.f;
  __shared__ __align__(8) uint64_t hbar[2];             // "h buffer [i] is complete" (same offset in every CTA of the cluster)
  if (tid == 0) {
    mbar_init(&hbar[0], 1);
    mbar_init(&hbar[1], 1);
    fence_barrier_init();
  }

  // gate phase: thread -> (sequence r, unit uu); matvec phase: thread -> (column pair, quarter of the k range), all
  // eight sequences: per four k one 128-bit broadcast load of h per sequence and four 64-bit loads of two weights, 32
  // FFMA2 (the two columns of a pair in one instruction, h as the scalar operand).
  const bool cell = tid < R * UC;                  // (half of the block at LSTM_KQ = 8)
  const int r = cell ? tid / UC : 0, uu = cell ? tid - r * UC : 0;
  const int b = b0 + r;
  const bool row_ok = cell && b < B;
  const int len = row_ok ? (lens ? min(max(lens[b], 0), L) : L) : 0;
  int lmax = 0;
  for (int i = 0; i < R; ++i) {
    const int bi = b0 + i;
    if (bi < B) lmax = max(lmax, lens ? min(max(lens[bi], 0), L) : L);
  }
  const int cp = tid % (NCOL / 2), kq = tid / (NCOL / 2);
  // resident weights, in REGISTERS: this thread's two columns (c = 2 cp, 2 cp + 1; column (g, uu) of this CTA = row
  // g H + j UC + uu of Wh = column of whT[dir]) over its eighth of k - 2 KPER = 64 registers at H = 256.  The 128 KB slice
  // of a CTA fills half of the SM's register file instead of its shared memory, and a step reads it at no cost: the
  // shared-memory data pipe (1024 of ~3300 wavefronts per step were the weights) no longer bounds the matvec.
  f2 wreg[KPER];
  {
    const int c0 = 2 * cp, g0 = c0 / UC, u0 = c0 - g0 * UC;        // UC is even: both columns in the same gate
    const float* wsrc = whT + ((size_t)dir * H + (size_t)kq * KPER) * (4 * H) + g0 * H + j * UC + u0;
#pragma unroll
    for (int k = 0; k < KPER; ++k) {
      const float2 w2 = *reinterpret_cast<const float2*>(wsrc + (size_t)k * (4 * H));
      wreg[k] = pack2(w2.x, w2.y);
    }
  }
  float cstate = 0.f;
  const size_t xw_row = (size_t)2 * 4 * H;
  // this thread's h goes to the same place of every CTA of the cluster (distributed shared memory)
  const uint32_t mine = (uint32_t)__cvta_generic_to_shared(hs + (size_t)r * H + j * UC + uu);
  const uint32_t bar_rel = (uint32_t)__cvta_generic_to_shared(&hbar[0]) - mine;   // a peer's window has the same layout
  cluster_sync_acqrel();   // every CTA has cleared its h buffers and set up its mbarriers before a peer writes into them
#ifdef KK_LSTM_DBG
  const bool dbg = tid == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
  long long tmark = clock64();
#endif

  for (int s = 0; s < lmax; ++s) {
    const bool active = s < len;
    const int t = dir == 0 ? s : len - 1 - s;
    const int cur = s & 1;
    const bool last = s == lmax - 1;                    // (lmax is the same in every CTA of the cluster)
    if (tid == 0 && !last) mbar_expect_tx(&hbar[cur ^ 1], (uint32_t)(R * H * sizeof(float)));   // h(t) of this step
    // this step's input projections (four gates of this thread's cell): in flight during the matvec
    float zi = 0.f, zf = 0.f, zg = 0.f, zo = 0.f;
    if (active) {
      const float* p = xw + ((size_t)b * L + t) * xw_row + (size_t)dir * 4 * H + j * UC + uu;
      zi = __ldg(p); zf = __ldg(p + H); zg = __ldg(p + 2 * H); zo = __ldg(p + 3 * H);
    }
    LSTM_MARK(0);
    if (s > 0) mbar_wait(&hbar[cur], (uint32_t)(((s - 1) >> 1) & 1));   // all of h(t-1) is here
    LSTM_MARK(4);
    // h(t-1) Wh^T for this CTA's columns: partial sums over this thread's quarter of k
    f2 acc[R];
#pragma unroll
    for (int q = 0; q < R; ++q) acc[q] = 0ull;
    const float* hp = hs + (size_t)cur * R * H + kq * KPER;
#pragma unroll
    for (int k = 0; k < KPER; k += 4) {
#pragma unroll
      for (int q = 0; q < R; ++q) {
        const float4 hv = *reinterpret_cast<const float4*>(hp + (size_t)q * H + k);
        acc[q] = fma2(pack2(hv.x, hv.x), wreg[k + 0], acc[q]);
        acc[q] = fma2(pack2(hv.y, hv.y), wreg[k + 1], acc[q]);
        acc[q] = fma2(pack2(hv.z, hv.z), wreg[k + 2], acc[q]);
        acc[q] = fma2(pack2(hv.w, hv.w), wreg[k + 3], acc[q]);
      }
    }
    {
      f2* pp = reinterpret_cast<f2*>(pre + (size_t)kq * R * NCOL) + cp;   // (row, column pair)
#pragma unroll
      for (int q = 0; q < R; ++q) pp[(size_t)q * (NCOL / 2)] = acc[q];
    }
    LSTM_MARK(1);
    __syncthreads();
    LSTM_MARK(2);
    float hout = 0.f;
    if (active) {
      const float* pr = pre + r * NCOL + uu;
      float z[4] = {zi, zf, zg, zo};
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float hw_ = pr[g * UC];
#pragma unroll
        for (int q = 1; q < LSTM_KQ; ++q) hw_ += pr[(size_t)q * R * NCOL + g * UC];   // fixed order
        z[g] += hw_;
      }
      const float ig = sigmoidf_(z[0]), fg = sigmoidf_(z[1]);
      const float gg = tanhf(z[2]), og = sigmoidf_(z[3]);
      cstate = fmaf(fg, cstate, ig * gg);
      const float h = og * tanhf(cstate);
      hout = h;
    }
    // Every cell thread sends (zeros for a finished or absent sequence: the byte count per step is fixed), except on
    // the last step, whose h nobody reads.  WAR on the buffers: a peer can only send h(t+1) into the buffer this CTA
    // reads now after it has received this CTA's h(t), which is sent below - after the matvec of this step.
    if (cell && !last) {
      const uint32_t off = (uint32_t)((cur ^ 1) * R * H) * 4u, boff = bar_rel + 8u * (cur ^ 1);
#pragma unroll
      for (int c = 0; c < LSTM_CS; ++c) {
        uint32_t pc;                                    // (mapa per send: eight registers are worth more than eight instructions)
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(pc) : "r"(mine), "r"(c));
        asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(pc + off),
                     "r"(__float_as_uint(hout)), "r"(pc + boff) : "memory");
      }
    }
    // (past the sequence's end the state is no longer used - rows are independent - and output row s is zero)
    if (row_ok) out[((size_t)b * L + (active ? t : s)) * (2 * H) + dir * H + j * UC + uu] = hout;
    LSTM_MARK(3);
  }
  cluster_sync_acqrel();   // nobody leaves while a peer could still be sending to it
  // sequences shorter than L inside a cluster whose longest is shorter than L too: zero the remaining rows
  if (row_ok)
    for (int s = max(lmax, len); s < L; ++s) out[((size_t)b * L + s) * (2 * H) + dir * H + j * UC + uu] = 0.f;
}

template <int R, int KPER>
static int lstm_launch_k(const float* xw, const float* whT, float* out, int B, int L, int H, const int* lens, cudaStream_t stream,
                         int* fit) {
  const int UC = H / LSTM_CS, NCOL = 4 * UC;
  const size_t smem = ((size_t)2 * H * R + (size_t)LSTM_KQ * R * NCOL) * sizeof(float);   // h (two buffers), partial sums
  KK_CUDA(cudaFuncSetAttribute(lstm_bidir_cluster_kernel<R, KPER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(LSTM_CS, (unsigned)cdiv(B, R), 2);
  cfg.blockDim = dim3(2 * UC * LSTM_KQ);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = LSTM_CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (fit) {
    KK_CUDA(cudaOccupancyMaxActiveClusters(fit, lstm_bidir_cluster_kernel<R, KPER>, &cfg));
    return KK_OK;
  }
  KK_CUDA(cudaLaunchKernelEx(&cfg, lstm_bidir_cluster_kernel<R, KPER>, xw, whT, out, B, L, H, lens));
  return KK_OK;
}

// Launch with R sequences per cluster; with `fit` != nullptr only report how many such clusters are co-resident.
template <int R>
static int lstm_launch(const float* xw, const float* whT, float* out, int B, int L, int H, const int* lens, cudaStream_t stream,
                       int* fit) {
  switch (H / LSTM_KQ) {
    case 8: return lstm_launch_k<R, 8>(xw, whT, out, B, L, H, lens, stream, fit);
    case 16: return lstm_launch_k<R, 16>(xw, whT, out, B, L, H, lens, stream, fit);
    case 24: return lstm_launch_k<R, 24>(xw, whT, out, B, L, H, lens, stream, fit);
    default: return lstm_launch_k<R, 32>(xw, whT, out, B, L, H, lens, stream, fit);
  }
}

// co-resident clusters for (R, H) on the current device, cached (a property of the device and the kernel image)
static int lstm_max_clusters(int R, int H) {
  static int cache[3][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}, {0, 0, 0, 0}};
  const int ri = R == 8 ? 0 : R == 10 ? 1 : 2, hi = H / 64 - 1;
  if (cache[ri][hi] == 0) {
    int fit = 0, rc;
    if (R == 8) rc = lstm_launch<8>(nullptr, nullptr, nullptr, 8, 1, H, nullptr, nullptr, &fit);
    else if (R == 10) rc = lstm_launch<10>(nullptr, nullptr, nullptr, 8, 1, H, nullptr, nullptr, &fit);
    else rc = lstm_launch<12>(nullptr, nullptr, nullptr, 8, 1, H, nullptr, nullptr, &fit);
    if (rc != KK_OK) return -1;
    cache[ri][hi] = fit > 0 ? fit : -2;     // -2: "none fit" is cached too
  }
  return cache[ri][hi] > 0 ? cache[ri][hi] : 0;
}

}  // namespace kk

#ifdef KK_LSTM_DBG
extern "C" int kk_lstm_dbg_read(unsigned long long* host8, int reset) {
  cudaMemcpyFromSymbol(host8, kk::g_lstm_clk, sizeof(unsigned long long) * 8);
  if (reset) { unsigned long long z[8] = {0}; cudaMemcpyToSymbol(kk::g_lstm_clk, z, sizeof(z)); }
  return 0;
}
#endif

extern "C" int kk_lstm_bidir_f32(const float* xw, const float* whT, float* out, int B, int L, int H, const int* lens,
                                 void* stream) {
  KK_CHECK_ARG(xw && whT && out && B > 0 && L > 0, "kk_lstm_bidir_f32: bad arguments");
  if (H < 64 || H > 256 || H % 64 != 0) {
    kk::set_error("kk_lstm_bidir_f32: hidden size %d (supported: 64, 128, 192, 256)", H);
    return KK_ERR_UNSUPPORTED;
  }
  // Sequences per cluster.  A cluster needs 8 SMs of one GPC (one CTA per SM: the weights fill shared memory), and a B200
  // holds 15 such clusters, not 148 / 8 = 18: with 8 sequences per cluster B = 64 is 16 clusters = TWO waves (1.88 ms at
  // L = 320); with 10 it is 14 clusters, one wave (1.11 ms).  Pick the R with the least waves x time per step, the
  // latter ~ (4 + R) by measurement (5.9 k clocks at R = 8, 6.8 k at R = 10).
  int rc = KK_ERR_UNSUPPORTED;
  const int cand[3] = {8, 10, 12};
  int pick = 0;
  long best = -1;
  for (int i = 0; i < 3; ++i) {
    const int fit = kk::lstm_max_clusters(cand[i], H);
    if (fit < 0) return KK_ERR_CUDA;
    if (fit == 0) continue;
    const long waves = kk::cdiv(2 * kk::cdiv(B, cand[i]), fit), cost = waves * (4 + cand[i]);
    if (best < 0 || cost < best) { best = cost; pick = i; }
  }
  if (best < 0) {
    kk::set_error("kk_lstm_bidir_f32: no cluster of %d CTAs fits this device", kk::LSTM_CS);
    return KK_ERR_UNSUPPORTED;
  }
  switch (cand[pick]) {
    case 8: rc = kk::lstm_launch<8>(xw, whT, out, B, L, H, lens, (cudaStream_t)stream, nullptr); break;
    case 10: rc = kk::lstm_launch<10>(xw, whT, out, B, L, H, lens, (cudaStream_t)stream, nullptr); break;
    default: rc = kk::lstm_launch<12>(xw, whT, out, B, L, H, lens, (cudaStream_t)stream, nullptr); break;
  }
  if (rc != KK_OK) return rc;
  KK_LAUNCH_CHECK("lstm_bidir_cluster_kernel");
  return KK_OK;
}
