// iSTFT head of the generator: exp / sin of conv_post's 22 channels, 11 -> 20 inverse real DFT, periodic
// Hann window, gather-form overlap-add (hop 5), division by the window sum, centre trim - one kernel.
// Reference: Generator.__call__ istftnet.py:804-806, MLXSTFT.inverse :497-523 (mlx_unwrap is the identity
// because |sin(.)| <= 1 < pi, :442), utils.istft utils.py:104-158 (normaliser is sum(w), not sum(w^2)).
#include "common.cuh"
#include <math.h>

namespace kk {

constexpr int IH_FT = 61;          // frames of output per block -> 305 samples
constexpr int IH_NF = IH_FT + 3;   // frames touched per block

struct Idft20Tables { float win[20]; float cs[20]; float sn[20]; };

__global__ void __launch_bounds__(256)
istft_head_n20_kernel(const float* __restrict__ y, int y_pitch, float* __restrict__ audio, int F, int S,
                      const Idft20Tables tb) {
  __shared__ float s_re[IH_NF][11];
  __shared__ float s_im[IH_NF][11];
  __shared__ float s_fr[IH_NF][21];
  const int b = blockIdx.y;
  const int g0 = blockIdx.x * IH_FT;       // first output sample = 5*g0 (trimmed index)
  const int fbase = g0 + 2 - 3;            // u0 = 5*g0 + 10 -> u0/5 - 3
  const float* yb = y + (size_t)b * F * y_pitch;

  // 1a. spectrum: Re = exp(a) cos(sin(p)), Im = exp(a) sin(sin(p))
  for (int it = threadIdx.x; it < IH_NF * 11; it += blockDim.x) {
    int fi = it / 11, k = it - fi * 11;
    int f = fbase + fi;
    float re = 0.f, im = 0.f;
    if (f >= 0 && f < F) {
      const float* yr = yb + (size_t)f * y_pitch;
      float m = expf(yr[k]);                 // :804
      float p = sinf(yr[11 + k]);            // :805
      float s, c;
      sincosf(p, &s, &c);
      re = m * c;                            // :505
      im = m * s;                            // :506
    }
    s_re[fi][k] = re;
    s_im[fi][k] = im;
  }
  __syncthreads();
  // 1b. windowed time frames: irfft (c2r, imag of DC / Nyquist ignored, 1/n) * periodic hann
  for (int it = threadIdx.x; it < IH_NF * 20; it += blockDim.x) {
    int fi = it / 20, n = it - fi * 20;
    float acc = 0.f;
#pragma unroll
    for (int k = 1; k <= 9; ++k) {
      int j = (k * n) % 20;
      acc = fmaf(s_re[fi][k], tb.cs[j], acc);
      acc = fmaf(-s_im[fi][k], tb.sn[j], acc);
    }
    float v = s_re[fi][0] + ((n & 1) ? -s_re[fi][10] : s_re[fi][10]) + 2.0f * acc;
    s_fr[fi][n] = (v * 0.05f) * tb.win[n];   // utils.py:135, :145
  }
  __syncthreads();
  // 2. gather overlap-add: untrimmed position u = i + 10 receives frames f with 5f <= u < 5f + 20
  for (int it = threadIdx.x; it < IH_FT * 5; it += blockDim.x) {
    int i = 5 * g0 + it;
    if (i >= S) break;
    int u = i + 10;
    int fhi = u / 5;
    float acc = 0.f, ws = 0.f;
#pragma unroll
    for (int d = 3; d >= 0; --d) {           // increasing frame index, like the reference's scatter order
      int f = fhi - d;
      if (f < 0 || f >= F) continue;
      int n = u - 5 * f;
      acc += s_fr[f - fbase][n];
      ws += tb.win[n];
    }
    audio[(size_t)b * S + i] = (ws != 0.f) ? acc / ws : acc;   // utils.py:150
  }
}

}  // namespace kk

extern "C" int kk_istft_head_n20(const float* y, int y_pitch, float* audio, int B, int F, void* stream) {
  KK_CHECK_ARG(y && audio && B > 0 && F > 1 && y_pitch >= 22, "kk_istft_head_n20: bad arguments");
  kk::Idft20Tables tb;
  for (int n = 0; n < 20; ++n) {
    tb.win[n] = (float)(0.5 * (1.0 - cos(2.0 * M_PI * n / 20.0)));  // periodic: hanning(21)[:-1] utils.py:121
    tb.cs[n] = (float)cos(2.0 * M_PI * n / 20.0);
    tb.sn[n] = (float)sin(2.0 * M_PI * n / 20.0);
  }
  int S = 5 * (F - 1);
  dim3 grid(kk::cdiv(F - 1, kk::IH_FT), B);
  KK_CHECK_ARG(grid.y <= 65535, "kk_istft_head_n20: batch too large");
  kk::istft_head_n20_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(y, y_pitch, audio, F, S, tb);
  KK_LAUNCH_CHECK("istft_head_n20_kernel");
  return KK_OK;
}
