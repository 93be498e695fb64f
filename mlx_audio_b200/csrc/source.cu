// Harmonic-plus-noise source (SineGen / SourceModuleHnNSF) and the n_fft=20 / hop=5 STFT that feeds the
// generator's noise branch.  The phase chain is bit-faithful to the reference's float32 op order:
// every multiply / add is an explicit round-to-nearest intrinsic so nvcc cannot contract them into FMAs
// (the reference materialises each op; phases reach 1e5..1e6 rad where one ulp is ~0.03 rad).
#include "common.cuh"
#include <math.h>

namespace kk {

constexpr int NH = 9;  // harmonic_num + 1 (istftnet.py:715, :546)

// ---- per (b, harmonic): r_j = (f0_j*k/24000) mod 1; c_j = cumsum; p_j = ((c_j*2)*pi)*upsample
// SineGen._f02sine istftnet.py:561-575.  The /300 linear down-sample (:568-572) is the identity on the
// nearest-upsampled f0 (it reads samples 300j+149 and 300j+150 with weight 0.5 each), and the rand_ini
// perturbation of sample 0 (:563-565) is never read.
__global__ void hnsf_phase_kernel(const float* __restrict__ f0, float* __restrict__ phase, int B, int T2,
                                  float upsample_f) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * NH) return;
  int b = idx / NH, k = idx - b * NH;
  const float* f = f0 + (size_t)b * T2;
  float* p = phase + (size_t)idx * T2;
  const float harm = (float)(k + 1);
  const float pi_f = 3.14159265358979323846f;
  float c = 0.f;
  // the cumulative sum is sequential by definition; the per-step fractions are not: compute them eight at a time so
  // that the loads and the divisions overlap instead of sitting on the add chain
  int j = 0;
  for (; j + 8 <= T2; j += 8) {
    float r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      float fk = __fmul_rn(f[j + u], harm);              // :610
      float q = __fdiv_rn(fk, 24000.0f);                 // :561
      r[u] = __fsub_rn(q, floorf(q));                    // python-sign "% 1"
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      c = __fadd_rn(c, r[u]);                            // sequential float32 cumsum :573
      p[j + u] = __fmul_rn(__fmul_rn(__fmul_rn(c, 2.0f), pi_f), upsample_f);  // *2, *pi (:573), *300 (:575)
    }
  }
  for (; j < T2; ++j) {
    float fk = __fmul_rn(f[j], harm);
    float q = __fdiv_rn(fk, 24000.0f);
    float r = __fsub_rn(q, floorf(q));
    c = __fadd_rn(c, r);
    p[j] = __fmul_rn(__fmul_rn(__fmul_rn(c, 2.0f), pi_f), upsample_f);
  }
}

// ---- Philox4x32-10 counter RNG + Box-Muller (only used when the caller passes no noise tensor)
__device__ __forceinline__ uint4 philox4x32(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

__device__ __forceinline__ void box_muller(uint32_t a, uint32_t b, float& n0, float& n1) {
  float u1 = ((float)(a >> 8) + 0.5f) * (1.0f / 16777216.0f);
  float u2 = ((float)(b >> 8) + 0.5f) * (1.0f / 16777216.0f);
  float r = sqrtf(-2.0f * __logf(u1));      // internal generator only: hardware log / sin / cos are ample
  float s, c;
  __sincosf(6.28318530717958647692f * u2, &s, &c);
  n0 = r * c;
  n1 = r * s;
}

// sin of a float32 phase of up to ~1e6 rad, correctly rounded to within an ulp like the reference's libm: the
// argument is reduced modulo 2 pi in double precision (B200 runs fp64 at half the fp32 rate), which keeps sinf on its
// short polynomial path instead of the Payne-Hanek one it takes above 105615 rad.
__device__ __forceinline__ float sin_large(float ph) {
  const double x = (double)ph;
  const double n = rint(x * 0.15915494309189533577);
  const double r = fma(n, -6.283185307179586232, x);      // |r| <= pi; 2 pi split error ~2.4e-16 * n: negligible
  return sinf((float)fma(n, -2.449293598294706414e-16, r));
}

// ---- per sample: linear up-sample of the phase with the reference's float32 coordinates and its
// negative-index wrap (interpolate.py:88-106), sin, uv gate, noise, 9->1 linear, tanh.
__global__ void hnsf_sample_kernel(const float* __restrict__ f0, const float* __restrict__ phase,
                                   const float* __restrict__ noise, unsigned long long seed,
                                   const unsigned long long* __restrict__ seed_dev,
                                   const float* __restrict__ lin_w, const float* __restrict__ lin_b,
                                   float* __restrict__ out, int B, int T2_pad, int upsample, float inv_up, float half_inv_up,
                                   const int* __restrict__ lens) {
  const long long S = (long long)T2_pad * upsample;
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)B * S) return;
  int b = (int)(idx / S);
  int i = (int)(idx - (long long)b * S);
  // ragged batch: this utterance has T2 f0 samples (the interpolation clamps / wraps at ITS last frame); the rest of the
  // padded row is silence
  const int T2 = ragged_len(lens, 2, 0, b, T2_pad);
  if (i >= T2 * upsample) { out[idx] = 0.f; return; }
  // x = ((float(i) * fl32(1/300)) + fl32(1/600)) - 0.5
  float x = __fsub_rn(__fadd_rn(__fmul_rn((float)i, inv_up), half_inv_up), 0.5f);
  float lo_f = floorf(x);
  float fr = __fsub_rn(x, lo_f);
  int lo = (int)lo_f;
  int hi = min(lo + 1, T2 - 1);
  if (lo < 0) lo += T2;  // MLX gather wraps index -1 to the last frame
  float one_m = __fsub_rn(1.0f, fr);
  float f0v = f0[(size_t)b * T2_pad + i / upsample];
  bool voiced = f0v > 10.0f;                          // voiced_threshold 10 (:716, :553)
  float amp = voiced ? 0.003f : __fdiv_rn(0.1f, 3.0f);  // uv*noise_std + (1-uv)*sine_amp/3 (:619)

  float nz[12];
  if (noise) {
    const float* np_ = noise + ((size_t)b * S + i) * NH;
#pragma unroll
    for (int k = 0; k < NH; ++k) nz[k] = np_[k];
  } else {
    if (seed_dev) seed = __ldg(seed_dev);   // device-resident seed: a captured CUDA graph draws fresh noise per replay
    uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32));
#pragma unroll
    for (int g = 0; g < 3; ++g) {
      uint4 r = philox4x32(make_uint4((uint32_t)i, (uint32_t)b, (uint32_t)g, 0x4b4bu), key);
      box_muller(r.x, r.y, nz[4 * g], nz[4 * g + 1]);
      box_muller(r.z, r.w, nz[4 * g + 2], nz[4 * g + 3]);
    }
  }
  float acc = 0.f;
  const float* pb = phase + (size_t)b * NH * T2_pad;
#pragma unroll
  for (int k = 0; k < NH; ++k) {
    float hk = __fmul_rn(amp, nz[k]);
    if (voiced) {   // sine * uv: unvoiced samples never see the sine (uv = 0, :622) - and f0 is constant over 300
                    // consecutive samples, so the branch is warp-uniform but for the segment edges
      const float* p = pb + (size_t)k * T2_pad;
      float ph = __fadd_rn(__fmul_rn(p[lo], one_m), __fmul_rn(p[hi], fr));  // interpolate.py:106
      float sine = __fmul_rn(sin_large(ph), 0.1f);                           // :579, :613
      hk = __fadd_rn(sine, hk);
    }
    acc = fmaf(hk, lin_w[k], acc);
  }
  out[idx] = tanhf(acc + lin_b[0]);  // l_linear + tanh (:666, :677)
}

// ---- STFT n_fft=20, hop=5, symmetric Hann, reflect pad 10, direct 20 -> 11 real DFT per frame.
struct Dft20Tables { float win[20]; float cs[20]; float sn[20]; };

constexpr int ST_FR = 128;   // frames per block (= threads)

__global__ void __launch_bounds__(ST_FR)
stft_n20_kernel(const float* __restrict__ sig, float* __restrict__ har, int har_pitch,
                __nv_bfloat16* __restrict__ har16, int pitch16, int rows16, int B, const int S_pad, const int F_pad,
                const Dft20Tables tb, const int* __restrict__ lens, const int len_mul) {
  // block = 128 consecutive frames of one utterance: the 655 samples they cover are staged in shared memory
  // (coalesced reads) and, when the bf16 rows are 24 wide, so is the output (128 x 48 B = 6 KB contiguous)
  __shared__ float s_in[ST_FR * 5 + 16];
  __shared__ __align__(16) __nv_bfloat16 s_out[ST_FR * 24];
  const int blocks_per_b = (F_pad + ST_FR - 1) / ST_FR;
  const int b = blockIdx.x / blocks_per_b;
  const int f0 = (blockIdx.x - b * blocks_per_b) * ST_FR;
  const int f = f0 + threadIdx.x;
  // ragged batch: S samples / F frames of this utterance inside rows pitched for S_pad / F_pad; the reflect pad mirrors
  // at S, frames in [F, F_pad) are written as zeros (the convolutions that read them expect zero padding)
  const int S = ragged_len(lens, len_mul, 0, b, S_pad);   // len_mul = samples per frame
  const int F = S / 5 + 1;
  const float* x = sig + (size_t)b * S_pad;
  for (int i = threadIdx.x; i < ST_FR * 5 + 15; i += ST_FR) {
    int p = 5 * f0 + i - 10;
    if (p < 0) p = -p;                       // reflect without edge repeat (utils.py:81-84)
    if (p >= S) p = 2 * (S - 1) - p;
    s_in[i] = (p >= 0 && p < S) ? x[p] : 0.f;   // (frames past F are not stored)
  }
  __syncthreads();
  const bool live = f < F_pad;          // row exists in the output tensors
  const bool real = f < F;              // ... and is a frame of this utterance
  float v[20];
#pragma unroll
  for (int n = 0; n < 20; ++n) v[n] = real ? s_in[5 * threadIdx.x + n] * tb.win[n] : 0.f;
  const bool staged = har16 && pitch16 == 24;
  float* o = (har && live) ? har + ((size_t)b * F_pad + f) * har_pitch : nullptr;
  __nv_bfloat16* o16 = (har16 && live && !staged) ? har16 + ((size_t)b * rows16 + f) * pitch16 : nullptr;
  __nv_bfloat16* os = s_out + threadIdx.x * 24;
#pragma unroll
  for (int k = 0; k <= 10; ++k) {
    float re = 0.f, im = 0.f;
#pragma unroll
    for (int n = 0; n < 20; ++n) {
      int j = (k * n) % 20;
      re = fmaf(v[n], tb.cs[j], re);
      im = fmaf(-v[n], tb.sn[j], im);
    }
    if (k == 0 || k == 10) im = 0.0f;        // r2c: DC / Nyquist imaginary parts are +0
    const float mag = sqrtf(re * re + im * im);   // mx.abs      istftnet.py:484
    const float ang = atan2f(im, re);             // mlx_angle   istftnet.py:409
    if (o) { o[k] = mag; o[11 + k] = ang; }
    if (o16) { o16[k] = __float2bfloat16(mag); o16[11 + k] = __float2bfloat16(ang); }
    if (staged) { os[k] = __float2bfloat16(mag); os[11 + k] = __float2bfloat16(ang); }
  }
  if (o) for (int c = 22; c < har_pitch; ++c) o[c] = 0.f;
  if (o16) for (int c = 22; c < pitch16; ++c) o16[c] = __float2bfloat16(0.f);
  if (staged) {
    os[22] = os[23] = __float2bfloat16(0.f);
    __syncthreads();
    const int nf = min(ST_FR, F_pad - f0);           // live frames of this block
    const int n16 = nf * 3;                          // 16-byte chunks (48 B per frame)
    uint4* dst = reinterpret_cast<uint4*>(har16 + ((size_t)b * rows16 + f0) * 24);   // 16-byte aligned: f0 % 128 == 0
    const uint4* src = reinterpret_cast<const uint4*>(s_out);
    for (int i = threadIdx.x; i < n16; i += ST_FR) dst[i] = src[i];
  }
}

}  // namespace kk

extern "C" int kk_hnsf_source(const float* f0, const float* noise, unsigned long long seed,
                              const unsigned long long* seed_dev, const float* lin_w, const float* lin_b,
                              float* phase_ws, float* har_source, int B, int T2, int upsample, const int* lens,
                              void* stream) {
  KK_CHECK_ARG(f0 && lin_w && lin_b && phase_ws && har_source, "kk_hnsf_source: null pointer");
  KK_CHECK_ARG(B > 0 && T2 > 1 && upsample > 1, "kk_hnsf_source: bad shape B=%d T2=%d up=%d", B, T2, upsample);
  KK_CHECK_ARG((long long)T2 * upsample < (1LL << 24), "kk_hnsf_source: utterance too long for float32 sample index");
  cudaStream_t st = (cudaStream_t)stream;
  kk::hnsf_phase_kernel<<<kk::cdiv(B * kk::NH, 64), 64, 0, st>>>(f0, phase_ws, B, T2, (float)upsample);
  KK_LAUNCH_CHECK("hnsf_phase_kernel");
  // Coordinates as the reference computes them: arange * (in/size) in float32, + 0.5*(in/size) as float32
  float inv_up = (float)(1.0 / (double)upsample);
  float half_inv_up = (float)(0.5 * (1.0 / (double)upsample));
  long long total = (long long)B * T2 * upsample;
  kk::hnsf_sample_kernel<<<kk::cdiv(total, 256), 256, 0, st>>>(f0, phase_ws, noise, seed, seed_dev, lin_w, lin_b, har_source, B,
                                                              T2, upsample, inv_up, half_inv_up, lens);
  KK_LAUNCH_CHECK("hnsf_sample_kernel");
  return KK_OK;
}

static int stft_n20_launch(const float* har_source, float* har, int har_pitch, void* har16, int pitch16, int rows16,
                           int B, int S, const int* lens, int len_mul, void* stream) {
  kk::Dft20Tables tb;
  for (int n = 0; n < 20; ++n) {
    tb.win[n] = (float)(0.5 * (1.0 - cos(2.0 * M_PI * n / 19.0)));  // symmetric hanning(20) utils.py:11-14
    tb.cs[n] = (float)cos(2.0 * M_PI * n / 20.0);
    tb.sn[n] = (float)sin(2.0 * M_PI * n / 20.0);
  }
  int F = S / 5 + 1;
  const int blocks = B * ((F + kk::ST_FR - 1) / kk::ST_FR);
  kk::stft_n20_kernel<<<blocks, kk::ST_FR, 0, (cudaStream_t)stream>>>(
      har_source, har, har_pitch, (__nv_bfloat16*)har16, pitch16, rows16, B, S, F, tb, lens, len_mul);
  KK_LAUNCH_CHECK("stft_n20_kernel");
  return KK_OK;
}

extern "C" int kk_stft_n20(const float* har_source, float* har, int har_pitch, int B, int S, const int* lens, int len_mul,
                           void* stream) {
  KK_CHECK_ARG(har_source && har && B > 0 && S > 10 && har_pitch >= 22 && S % 5 == 0, "kk_stft_n20: bad arguments");
  return stft_n20_launch(har_source, har, har_pitch, nullptr, 0, 0, B, S, lens, len_mul, stream);
}

extern "C" int kk_stft_n20_bf16(const float* har_source, float* har, int har_pitch, void* har_bf16, int bf_pitch,
                                int bf_rows, int B, int S, const int* lens, int len_mul, void* stream) {
  KK_CHECK_ARG(har_source && har_bf16 && B > 0 && S > 10 && S % 5 == 0, "kk_stft_n20_bf16: bad arguments");
  KK_CHECK_ARG(bf_pitch != 24 || ((uintptr_t)har_bf16 & 15) == 0, "kk_stft_n20_bf16: har_bf16 must be 16-byte aligned");
  KK_CHECK_ARG((!har || har_pitch >= 22) && bf_pitch >= 22 && bf_rows >= S / 5 + 1,
               "kk_stft_n20_bf16: pitch / rows too small (bf_rows %d < frames %d)", bf_rows, S / 5 + 1);
  return stft_n20_launch(har_source, har, har_pitch, har_bf16, bf_pitch, bf_rows, B, S, lens, len_mul, stream);
}

// ---- kk_hnsf_source_noalloc: the same harmonic source with its phase scratch in a file-scope device array, for callers
// that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.  Every GPU that loads the library holds the
// array (4 MiB) until exit.  All calls share it and it keeps the last call's contents; nothing in it has to start at
// zero (hnsf_phase_kernel writes every element hnsf_sample_kernel reads), but calls that overlap on the GPU (different
// streams, or graphs replayed concurrently) race on it.
namespace kk {
static __device__ __align__(256) float g_hnsf_phase_scratch[KK_HNSF_NOALLOC_MAX_FLOATS];
}

extern "C" int kk_hnsf_source_noalloc(const float* f0, const float* noise, unsigned long long seed,
                                      const unsigned long long* seed_dev, const float* lin_w, const float* lin_b,
                                      float* har_source, int B, int T2, int upsample, const int* lens, void* stream) {
  KK_CHECK_ARG(B > 0 && T2 > 1, "kk_hnsf_source_noalloc: bad shape B=%d T2=%d", B, T2);
  const long long need = (long long)B * kk::NH * T2;
  KK_CHECK_ARG(need <= (long long)KK_HNSF_NOALLOC_MAX_FLOATS,
               "kk_hnsf_source_noalloc: B*9*T2 = %lld floats of phase scratch exceed the static maximum of %lld "
               "(KK_HNSF_NOALLOC_MAX_FLOATS); use kk_hnsf_source with caller-provided scratch",
               need, (long long)KK_HNSF_NOALLOC_MAX_FLOATS);
  // In host code the array's name is a host-side stand-in, not a device address: resolve it once per device, with that
  // device current (cudaGetSymbolAddress allocates nothing).
  constexpr int MAX_DEV = 64;
  static float* scratch_of[MAX_DEV] = {nullptr};
  int dev = 0;
  KK_CUDA(cudaGetDevice(&dev));
  KK_CHECK_ARG(dev >= 0 && dev < MAX_DEV, "kk_hnsf_source_noalloc: device ordinal %d out of range", dev);
  float* ws = scratch_of[dev];
  if (!ws) {
    KK_CUDA(cudaGetSymbolAddress((void**)&ws, kk::g_hnsf_phase_scratch));
    scratch_of[dev] = ws;   // (two threads racing here store the same value)
  }
  return kk_hnsf_source(f0, noise, seed, seed_dev, lin_w, lin_b, ws, har_source, B, T2, upsample, lens, stream);
}
