// iSTFT head of the generator: exp / sin of conv_post's 22 channels, 11 -> 20 inverse real DFT, periodic
// Hann window, gather-form overlap-add (hop 5), division by the window sum, centre trim - one kernel.
// Reference: Generator.__call__ istftnet.py:804-806, MLXSTFT.inverse :497-523 (mlx_unwrap is the identity
// because |sin(.)| <= 1 < pi, :442), utils.istft utils.py:104-158 (normaliser is sum(w), not sum(w^2)).
#include "common.cuh"
#include <math.h>

namespace kk {

constexpr int IH_THREADS = 256;
constexpr int IH_FT = IH_THREADS - 3;   // output frames per block (253 -> 1265 samples); 3 halo frames

// cos / sin (2 pi j / 20) and the periodic Hann window hanning(21)[:-1] (utils.py:121), evaluated in double
__device__ constexpr float IH_CS[20] = {1.000000000e+00f, 9.510565163e-01f, 8.090169944e-01f, 5.877852523e-01f, 3.090169944e-01f, 6.123233996e-17f, -3.090169944e-01f, -5.877852523e-01f, -8.090169944e-01f, -9.510565163e-01f, -1.000000000e+00f, -9.510565163e-01f, -8.090169944e-01f, -5.877852523e-01f, -3.090169944e-01f, -1.836970199e-16f, 3.090169944e-01f, 5.877852523e-01f, 8.090169944e-01f, 9.510565163e-01f};
__device__ constexpr float IH_SN[20] = {0.000000000e+00f, 3.090169944e-01f, 5.877852523e-01f, 8.090169944e-01f, 9.510565163e-01f, 1.000000000e+00f, 9.510565163e-01f, 8.090169944e-01f, 5.877852523e-01f, 3.090169944e-01f, 1.224646799e-16f, -3.090169944e-01f, -5.877852523e-01f, -8.090169944e-01f, -9.510565163e-01f, -1.000000000e+00f, -9.510565163e-01f, -8.090169944e-01f, -5.877852523e-01f, -3.090169944e-01f};
__device__ constexpr float IH_WIN[20] = {0.000000000e+00f, 2.447174185e-02f, 9.549150281e-02f, 2.061073739e-01f, 3.454915028e-01f, 5.000000000e-01f, 6.545084972e-01f, 7.938926261e-01f, 9.045084972e-01f, 9.755282581e-01f, 1.000000000e+00f, 9.755282581e-01f, 9.045084972e-01f, 7.938926261e-01f, 6.545084972e-01f, 5.000000000e-01f, 3.454915028e-01f, 2.061073739e-01f, 9.549150281e-02f, 2.447174185e-02f};

// One thread = one STFT frame: 22 conv_post channels -> spectrum -> 20 windowed samples (fully unrolled real
// inverse DFT with compile-time twiddles, cos even / sin odd symmetry) -> shared memory; then a gather-form
// overlap-add of the 4 frames covering every output sample.
__global__ void __launch_bounds__(IH_THREADS)
istft_head_n20_kernel(const float* __restrict__ y, int y_pitch, float* __restrict__ audio, const int F_pad,
                      const int S_pad, const int* __restrict__ lens, const int len_mul, const int len_add) {
  __shared__ float s_fr[IH_THREADS][21];
  const int b = blockIdx.y;
  // ragged batch: this utterance has F spectrogram frames / S samples inside buffers pitched for F_pad / S_pad
  const int F = ragged_len(lens, len_mul, len_add, b, F_pad);
  const int S = 5 * (F - 1);
  const int g0 = blockIdx.x * IH_FT;       // first output frame; first output sample = 5*g0 (trimmed index)
  const int fbase = g0 - 1;                // u0 = 5*g0 + 10 -> u0/5 - 3
  const int f = fbase + (int)threadIdx.x;
  float fr[20];
#pragma unroll
  for (int n = 0; n < 20; ++n) fr[n] = 0.f;
  if (f >= 0 && f < F) {
    const float* yr = y + ((size_t)b * F_pad + f) * y_pitch;
    float v[22];
    if ((y_pitch & 3) == 0 && (reinterpret_cast<uintptr_t>(y) & 15) == 0) {
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(yr) + i);
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
      }
      const float2 t2 = __ldg(reinterpret_cast<const float2*>(yr + 20));
      v[20] = t2.x; v[21] = t2.y;
    } else {
#pragma unroll
      for (int i = 0; i < 22; ++i) v[i] = __ldg(yr + i);
    }
    float re[11], im[11];
#pragma unroll
    for (int k = 0; k < 11; ++k) {
      const float m = expf(v[k]);            // :804
      const float p = sinf(v[11 + k]);       // :805
      float sp, cp;
      sincosf(p, &sp, &cp);
      re[k] = m * cp;                        // :505
      im[k] = m * sp;                        // :506
    }
    // irfft (c2r: imag of DC / Nyquist ignored, 1/n), utils.py:135, then the window, :145
#pragma unroll
    for (int n = 0; n <= 10; ++n) {
      float c = 0.f, s_ = 0.f;
#pragma unroll
      for (int k = 1; k <= 9; ++k) {
        c = fmaf(re[k], IH_CS[(k * n) % 20], c);
        s_ = fmaf(im[k], IH_SN[(k * n) % 20], s_);
      }
      const float base = re[0] + ((n & 1) ? -re[10] : re[10]);
      fr[n] = ((base + 2.0f * (c - s_)) * 0.05f) * IH_WIN[n];
      if (n >= 1 && n <= 9) fr[20 - n] = ((base + 2.0f * (c + s_)) * 0.05f) * IH_WIN[20 - n];
    }
  }
#pragma unroll
  for (int n = 0; n < 20; ++n) s_fr[threadIdx.x][n] = fr[n];
  __syncthreads();
  // gather overlap-add: untrimmed position u = i + 10 receives frames f with 5f <= u < 5f + 20
  for (int it = threadIdx.x; it < IH_FT * 5; it += IH_THREADS) {
    const int i = 5 * g0 + it;
    if (i >= S_pad) break;
    if (i >= S) { audio[(size_t)b * S_pad + i] = 0.f; continue; }   // padding of a ragged batch
    const int u = i + 10;
    const int fhi = u / 5;
    float acc = 0.f, ws = 0.f;
#pragma unroll
    for (int dd = 3; dd >= 0; --dd) {         // increasing frame index, like the reference's scatter order
      const int ff = fhi - dd;
      if (ff < 0 || ff >= F) continue;
      const int n = u - 5 * ff;
      acc += s_fr[ff - fbase][n];
      ws += IH_WIN[n];
    }
    audio[(size_t)b * S_pad + i] = (ws != 0.f) ? acc / ws : acc;   // utils.py:150
  }
}

}  // namespace kk

extern "C" int kk_istft_head_n20(const float* y, int y_pitch, float* audio, int B, int F, const int* lens, int len_mul,
                                 int len_add, void* stream) {
  KK_CHECK_ARG(y && audio && B > 0 && F > 1 && y_pitch >= 22, "kk_istft_head_n20: bad arguments");
  int S = 5 * (F - 1);
  dim3 grid(kk::cdiv(F - 1, kk::IH_FT), B);
  KK_CHECK_ARG(grid.y <= 65535, "kk_istft_head_n20: batch too large");
  kk::istft_head_n20_kernel<<<grid, kk::IH_THREADS, 0, (cudaStream_t)stream>>>(y, y_pitch, audio, F, S, lens, len_mul, len_add);
  KK_LAUNCH_CHECK("istft_head_n20_kernel");
  return KK_OK;
}
