// Elementwise / depthwise kernels of the codec generators that reuse the decoder's convolution and Snake kernels
// (SURVEY 8(f) row 4): SNAC's depthwise WNConv1d (groups = channels), NoiseBlock and the Tanh head
// (mlx_audio/codec/models/snac/layers.py:18-61, 217-228, 196-201).  Channels-last (B, L, C) fp32, HBM-bound:
// one thread = 4 consecutive channels (16-byte accesses), a warp = 128 consecutive channels of one row.
#include "common.cuh"

namespace kk {

// y[b][l][c] = bias[c] + sum_t w[c][t] * x[b][l + t*dil - pad][c] (+ res[b][l][c]); rows outside [0, L) are zero padding
// (mx.conv1d with groups = C, layers.py:57).  w is (C, K) (the folded (C, K, 1) weight).
template <int VEC>
__global__ void __launch_bounds__(256)
dwconv1d_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                const float* __restrict__ res, float* __restrict__ out, int L, int C, int K, int dil, int pad,
                long long total_vec) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_vec) return;
  const int cv = C / VEC;
  const int c = (int)(idx % cv) * VEC;
  const long long row = idx / cv;             // b * L + l
  const int l = (int)(row % L);
  const float* xr = x + (row - l) * C + c;    // utterance base + channel
  float acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = bias ? bias[c + i] : 0.f;
  for (int t = 0; t < K; ++t) {
    const int li = l + t * dil - pad;
    if (li < 0 || li >= L) continue;
    float xv[VEC];
    if (VEC == 4) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(xr + (long long)li * C));
      xv[0] = v.x; xv[1] = v.y; xv[2] = v.z; xv[3] = v.w;
    } else {
      xv[0] = __ldg(xr + (long long)li * C);
    }
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] = fmaf(__ldg(w + (size_t)(c + i) * K + t), xv[i], acc[i]);
  }
  const long long o = row * C + c;
  if (res) {
#pragma unroll
    for (int i = 0; i < VEC; ++i) acc[i] += res[o + i];
  }
  if (VEC == 4) *reinterpret_cast<float4*>(out + o) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  else out[o] = acc[0];
}

// NoiseBlock (layers.py:222-228): out = x + noise * h with noise of shape (B, 1, C) - the reference unpacks
// ``B, C, T = x.shape`` on a channels-last tensor, so its noise varies per CHANNEL and is constant over time.
__global__ void __launch_bounds__(256)
noise_mix_kernel(const float* __restrict__ x, const float* __restrict__ h, const float* __restrict__ noise,
                 float* __restrict__ out, long long LC, int C, long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const long long b = idx / LC;
  const int c = (int)(idx % C);
  out[idx] = fmaf(__ldg(noise + b * C + c), h[idx], x[idx]);
}

__global__ void __launch_bounds__(256) tanh_kernel(const float* __restrict__ x, float* __restrict__ out, long long n) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) out[idx] = tanhf(x[idx]);
}

}  // namespace kk

extern "C" int kk_dwconv1d_f32(const float* x, const float* w, const float* bias, const float* res, float* out, int B,
                               int L, int C, int K, int dil, int pad, void* stream) {
  KK_CHECK_ARG(x && w && out && B > 0 && L > 0 && C > 0 && K > 0 && dil > 0 && pad >= 0, "kk_dwconv1d_f32: bad arguments");
  KK_CHECK_ARG((K - 1) * dil == 2 * pad, "kk_dwconv1d_f32: 'same' padding only ((K-1)*dil = %d, 2*pad = %d)", (K - 1) * dil, 2 * pad);
  cudaStream_t st = (cudaStream_t)stream;
  const bool v4 = C % 4 == 0 && (((uintptr_t)x | (uintptr_t)out | (uintptr_t)res) & 15) == 0;
  const long long total = (long long)B * L * (v4 ? C / 4 : C);
  if (v4) kk::dwconv1d_kernel<4><<<kk::cdiv(total, 256), 256, 0, st>>>(x, w, bias, res, out, L, C, K, dil, pad, total);
  else kk::dwconv1d_kernel<1><<<kk::cdiv(total, 256), 256, 0, st>>>(x, w, bias, res, out, L, C, K, dil, pad, total);
  KK_LAUNCH_CHECK("dwconv1d_kernel");
  return KK_OK;
}

extern "C" int kk_noise_mix_f32(const float* x, const float* h, const float* noise, float* out, int B, long long L, int C,
                                void* stream) {
  KK_CHECK_ARG(x && h && noise && out && B > 0 && L > 0 && C > 0, "kk_noise_mix_f32: bad arguments");
  const long long total = (long long)B * L * C;
  kk::noise_mix_kernel<<<kk::cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(x, h, noise, out, L * C, C, total);
  KK_LAUNCH_CHECK("noise_mix_kernel");
  return KK_OK;
}

extern "C" int kk_tanh_f32(const float* x, float* out, long long n, void* stream) {
  KK_CHECK_ARG(x && out && n > 0, "kk_tanh_f32: bad arguments");
  kk::tanh_kernel<<<kk::cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, out, n);
  KK_LAUNCH_CHECK("tanh_kernel");
  return KK_OK;
}
