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

// out = a x + b y (BigVGAN's sum over the resblocks of a stage and its division by their number, bigvgan.py:113-117)
__global__ void __launch_bounds__(256)
axpby_kernel(const float* __restrict__ x, const float* __restrict__ y, float* __restrict__ out, float a, float b, long long n) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) out[idx] = fmaf(a, x[idx], b * (y ? y[idx] : 0.f));
}

__global__ void __launch_bounds__(256) tanh_kernel(const float* __restrict__ x, float* __restrict__ out, long long n) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) out[idx] = tanhf(x[idx]);
}

// mx.clip(x, lo, hi) (BigVGAN's head without tanh, bigvgan.py:125-126)
__global__ void __launch_bounds__(256) clip_kernel(const float* __restrict__ x, float* __restrict__ out, float lo, float hi, long long n) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx < n) out[idx] = fminf(fmaxf(x[idx], lo), hi);
}

// ---------------------------------------------------------------- BigVGAN anti-aliased activation
// Activation1d (mlx_audio/codec/models/bigvgan/resample.py:152-177) with its default ratios and kernel sizes, fused:
//   UpSample1d(2, 12)  :104-134  edge-pad 5, depthwise transposed conv stride 2 with the kaiser-sinc filter f, x 2, crop 15:
//                                u[2q]   = 2 sum_r f[11 - 2r] x[clamp(q - 3 + r)],  u[2q+1] = 2 sum_r f[10 - 2r] x[clamp(q - 2 + r)]
//   SnakeBeta / Snake  activation.py:27-51   v = u + sin^2(alpha u) / (beta + 1e-9)   (alpha, beta through exp() if logscale)
//   DownSample1d(2,12) :137-149  edge-pad (5, 6), depthwise conv stride 2 with g:   out[t] = sum_j g[j] v[clamp(2t + j - 5)]
// One thread = one channel x AA_TT consecutive outputs: the AA_TT + 10 inputs and the 2 AA_TT + 10 activated samples of
// the tile live in registers (every activated sample is computed once per tile; 10 of them are the next tile's too).
// HBM traffic is the input once (+ halo) and the output once; the 2x-rate signal never exists in memory.
constexpr int AA_TT = 8;
struct AaFilters { float f[12], g[12]; };

__device__ __forceinline__ float aa_snake(float u, float alpha, float inv_beta) {
  const float s = sinf(u * alpha);
  return fmaf(inv_beta, s * s, u);
}

__global__ void __launch_bounds__(256)
aa_act_kernel(const float* __restrict__ x, float* __restrict__ out, const float* __restrict__ alpha,
              const float* __restrict__ beta, const AaFilters k, int L, int C, int logscale, float eps) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const int t0 = (blockIdx.y * blockDim.y + threadIdx.y) * AA_TT;
  if (t0 >= L) return;
  const int b = blockIdx.z;
  const float* xb = x + (size_t)b * L * C + c;
  float* ob = out + (size_t)b * L * C + c;
  float al = alpha[c], be = beta ? beta[c] : al;
  if (logscale) { al = expf(al); be = expf(be); }
  const float inv_beta = 1.0f / (be + eps);
  // inputs x[t0 - 5 .. t0 + AA_TT + 4], replicated at the ends (mode="edge")
  float xv[AA_TT + 10];
#pragma unroll
  for (int i = 0; i < AA_TT + 10; ++i) xv[i] = __ldg(xb + (size_t)min(max(t0 - 5 + i, 0), L - 1) * C);
  // activated 2x-rate samples v[m], m = 2 t0 - 5 + i.  m = 2q + 1 with q = t0 - 3 + i/2 for even i, m = 2q with
  // q = t0 - 2 + (i-1)/2 for odd i; x[q - 2 + r] = xv[i/2 + r], x[q - 3 + r] = xv[(i-1)/2 + r].
  float v[2 * AA_TT + 10];
#pragma unroll
  for (int i = 0; i < 2 * AA_TT + 10; ++i) {
    float u = 0.f;
    if ((i & 1) == 0) {
#pragma unroll
      for (int r = 0; r < 6; ++r) u = fmaf(k.f[10 - 2 * r], xv[i / 2 + r], u);
    } else {
#pragma unroll
      for (int r = 0; r < 6; ++r) u = fmaf(k.f[11 - 2 * r], xv[(i - 1) / 2 + r], u);
    }
    v[i] = aa_snake(2.0f * u, al, inv_beta);
  }
  // the 2x-rate signal is edge-replicated too: samples with m < 0 are v[m = 0] (i = 5 when t0 = 0 .. i = 1 when t0 = 2),
  // samples with m > 2L - 1 are v[m = 2L - 1]
  const int m_lo = 2 * t0 - 5, m_hi = 2 * L - 1;
  if (m_lo < 0 || m_lo + 2 * AA_TT + 9 > m_hi) {
    float v0 = 0.f, vl = 0.f;
#pragma unroll
    for (int i = 0; i < 2 * AA_TT + 10; ++i) {
      if (m_lo + i == 0) v0 = v[i];
      if (m_lo + i == m_hi) vl = v[i];
    }
#pragma unroll
    for (int i = 0; i < 2 * AA_TT + 10; ++i) {
      if (m_lo + i < 0) v[i] = v0;
      if (m_lo + i > m_hi) v[i] = vl;
    }
  }
#pragma unroll
  for (int tt = 0; tt < AA_TT; ++tt) {
    if (t0 + tt >= L) break;
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < 12; ++j) acc = fmaf(k.g[j], v[2 * tt + j], acc);
    ob[(size_t)(t0 + tt) * C] = acc;
  }
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

extern "C" int kk_aa_act_f32(const float* x, float* out, const float* alpha, const float* beta, const float* up_filter,
                             const float* down_filter, int B, int L, int C, int logscale, float eps, void* stream) {
  KK_CHECK_ARG(x && out && alpha && up_filter && down_filter && B > 0 && B <= 65535 && L > 0 && C > 0,
               "kk_aa_act_f32: bad arguments");
  kk::AaFilters k;   // host pointers: 12 taps each (resample.py:18-50, kernel_size 12)
  for (int i = 0; i < 12; ++i) { k.f[i] = up_filter[i]; k.g[i] = down_filter[i]; }
  const int bx = C >= 128 ? 128 : (C >= 64 ? 64 : 32);
  dim3 block(bx, 256 / bx);
  dim3 grid(kk::cdiv(C, bx), kk::cdiv(kk::cdiv(L, kk::AA_TT), (int)block.y), B);
  KK_CHECK_ARG(grid.y <= 65535, "kk_aa_act_f32: sequence too long (%d)", L);
  kk::aa_act_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(x, out, alpha, beta, k, L, C, logscale, eps);
  KK_LAUNCH_CHECK("aa_act_kernel");
  return KK_OK;
}

extern "C" int kk_axpby_f32(const float* x, const float* y, float* out, float a, float b, long long n, void* stream) {
  KK_CHECK_ARG(x && out && n > 0, "kk_axpby_f32: bad arguments");
  kk::axpby_kernel<<<kk::cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, y, out, a, b, n);
  KK_LAUNCH_CHECK("axpby_kernel");
  return KK_OK;
}

extern "C" int kk_clip_f32(const float* x, float* out, float lo, float hi, long long n, void* stream) {
  KK_CHECK_ARG(x && out && n > 0 && lo <= hi, "kk_clip_f32: bad arguments");
  kk::clip_kernel<<<kk::cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, out, lo, hi, n);
  KK_LAUNCH_CHECK("clip_kernel");
  return KK_OK;
}
