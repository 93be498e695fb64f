// Load-time weight preparation: weight-norm folding and GEMM-operand packing.
#include "common.cuh"

namespace kk {

// One block per output channel: n = sqrt(sum v^2) (double accumulation), w = v / (n + 1e-7) * g.
// Reference: weight_norm istftnet.py:87-93 with dim=0 (:130).
__global__ void fold_weight_norm_kernel(const float* __restrict__ v, const float* __restrict__ g,
                                        float* __restrict__ w, int KI, float eps, int g_first) {
  int o = blockIdx.x;
  const float* vo = v + (size_t)o * KI;
  double acc = 0.0;
  for (int i = threadIdx.x; i < KI; i += blockDim.x) {
    float x = vo[i];
    acc += (double)x * (double)x;
  }
  __shared__ double sh[32];
  for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    double t = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.0;
    for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
    if (threadIdx.x == 0) sh[0] = t;
  }
  __syncthreads();
  float n = sqrtf((float)sh[0]);
  float denom = n + eps;
  float go = g[o];
  // g_first: (g * v) / n, the order of the DAC layers (descript/nn/layers.py:57); else (v / (n + eps)) * g (istftnet.py:92-93)
  if (g_first) for (int i = threadIdx.x; i < KI; i += blockDim.x) w[(size_t)o * KI + i] = __fdiv_rn(__fmul_rn(go, vo[i]), denom);
  else for (int i = threadIdx.x; i < KI; i += blockDim.x) w[(size_t)o * KI + i] = (vo[i] / denom) * go;
}

template <typename OutT>
__device__ __forceinline__ OutT cvt(float x);
template <>
__device__ __forceinline__ float cvt<float>(float x) { return x; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt<__nv_bfloat16>(float x) { return __float2bfloat16(x); }

// out (N, taps*Cp).  mode 0: N=O, taps=K.  mode 1: N=up*I(small), taps=2, source w (O=big C_in, K=2up, I=small C_out).
template <typename OutT>
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, OutT* __restrict__ out, int mode, int O, int K,
                                        int I, int up, int Cp, long long total) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int taps = mode == 0 ? K : 2;
  int row_len = taps * Cp;
  int n = (int)(idx / row_len);
  int r = (int)(idx - (long long)n * row_len);
  int t = r / Cp;
  int c = r - t * Cp;
  float val = 0.f;
  if (mode == 0) {
    if (c < I) val = w[((size_t)n * K + t) * I + c];
  } else {
    int cout = I, cin = O;
    int ph = n / cout, co = n - ph * cout;
    int k = ph + up * (1 - t);
    if (c < cin) val = w[((size_t)c * K + k) * I + co];
  }
  out[idx] = cvt<OutT>(val);
}

}  // namespace kk

extern "C" int kk_fold_weight_norm(const float* v, const float* g, float* w, int O, int KI, void* stream) {
  KK_CHECK_ARG(v && g && w && O > 0 && KI > 0, "kk_fold_weight_norm: bad arguments O=%d KI=%d", O, KI);
  kk::fold_weight_norm_kernel<<<O, 256, 0, (cudaStream_t)stream>>>(v, g, w, KI, 1e-7f, 0);
  KK_LAUNCH_CHECK("fold_weight_norm_kernel");
  return KK_OK;
}

extern "C" int kk_fold_weight_norm_ex(const float* v, const float* g, float* w, int O, int KI, float eps, int g_first,
                                      void* stream) {
  KK_CHECK_ARG(v && g && w && O > 0 && KI > 0 && eps >= 0.f, "kk_fold_weight_norm_ex: bad arguments O=%d KI=%d", O, KI);
  kk::fold_weight_norm_kernel<<<O, 256, 0, (cudaStream_t)stream>>>(v, g, w, KI, eps, g_first);
  KK_LAUNCH_CHECK("fold_weight_norm_kernel");
  return KK_OK;
}

extern "C" int kk_pack_conv_weight(const float* w, void* out, int out_is_bf16, int mode, int O, int K, int I, int up,
                                   int Cp, void* stream) {
  KK_CHECK_ARG(w && out && O > 0 && K > 0 && I > 0, "kk_pack_conv_weight: bad arguments");
  KK_CHECK_ARG(mode == 0 || mode == 1, "kk_pack_conv_weight: mode %d", mode);
  int N, taps, cin;
  if (mode == 0) { N = O; taps = K; cin = I; }
  else {
    KK_CHECK_ARG(up > 0 && K == 2 * up, "kk_pack_conv_weight: transposed conv needs K == 2*up (K=%d up=%d)", K, up);
    N = up * I; taps = 2; cin = O;
  }
  KK_CHECK_ARG(Cp >= cin, "kk_pack_conv_weight: Cp=%d < C_in=%d", Cp, cin);
  long long total = (long long)N * taps * Cp;
  int blocks = kk::cdiv(total, 256);
  if (out_is_bf16)
    kk::pack_conv_weight_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>(
        w, (__nv_bfloat16*)out, mode, O, K, I, up, Cp, total);
  else
    kk::pack_conv_weight_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(w, (float*)out, mode, O, K, I, up,
                                                                                   Cp, total);
  KK_LAUNCH_CHECK("pack_conv_weight_kernel");
  return KK_OK;
}
