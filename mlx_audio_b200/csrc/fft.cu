// General STFT / log-mel / iSTFT (mlx_audio/utils.py:52-158) as shared-memory staged mixed-radix FFTs.
// A real frame of n_fft samples is packed into an n_fft/2-point complex sequence, transformed with an
// out-of-place Stockham autosort FFT (radices 4, 2, 3, 5; twiddles from a per-block shared-memory table) and
// unpacked to the n_fft/2+1 half spectrum.  Windowing, centre (reflect / zero) padding, |X| or |X|^2, the
// mel projection + log, and on the inverse side the periodic window, gather-form overlap-add, window-sum
// normalisation and trimming are fused into the same kernels: the framed signal and (for log-mel) the
// spectrogram never touch HBM.
#include "common.cuh"
#include "fft_fast.cuh"
#ifdef KK_FFT_R32   // lab build: the single-exchange forward kernel (measured slower, profiles/r02_fft_experiments.txt)
#include "../../experiments/fft_r32.cuh"
#endif
#include <math.h>

namespace kk {

struct FftPlan {
  int n_fft, nc;       // real size, complex size n_fft/2
  int nrad;
  int rad[16];
};

static int make_fft_plan(int n_fft, FftPlan* p) {
  if (n_fft < 4 || (n_fft & 1) || n_fft > 4096) return -1;
  p->n_fft = n_fft;
  p->nc = n_fft / 2;
  p->nrad = 0;
  int n = p->nc;
  while (n % 4 == 0) { p->rad[p->nrad++] = 4; n /= 4; }
  while (n % 2 == 0) { p->rad[p->nrad++] = 2; n /= 2; }
  while (n % 3 == 0) { p->rad[p->nrad++] = 3; n /= 3; }
  while (n % 5 == 0) { p->rad[p->nrad++] = 5; n /= 5; }
  return n == 1 ? 0 : -1;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }

// One Stockham stage of radix R over `batch` independent sequences of length nc stored back to back.
// tw[j] = exp(-2 pi i j / nc).
template <int R>
__device__ __forceinline__ void stockham_stage(const float2* __restrict__ in, float2* __restrict__ out,
                                               const float2* __restrict__ tw, int nc, int ns, int batch) {
  const int T = nc / R;
  const int tw_step = nc / (ns * R);
  for (int it = threadIdx.x; it < batch * T; it += blockDim.x) {
    const int g = it / T, j = it - g * T;
    const float2* a = in + g * nc;
    float2* o = out + g * nc;
    const int k = j % ns;
    float2 v[R];
#pragma unroll
    for (int t = 0; t < R; ++t) {
      float2 x = a[j + t * T];
      v[t] = (t == 0 || k == 0) ? x : cmul(x, tw[(t * k * tw_step) % nc]);
    }
    const int j0 = (j / ns) * ns * R + k;
    if (R == 2) {
      o[j0] = cadd(v[0], v[1]);
      o[j0 + ns] = csub(v[0], v[1]);
    } else if (R == 4) {
      float2 s0 = cadd(v[0], v[2]), d0 = csub(v[0], v[2]);
      float2 s1 = cadd(v[1], v[3]), d1 = csub(v[1], v[3]);
      float2 md1 = make_float2(d1.y, -d1.x);  // -i * d1
      o[j0] = cadd(s0, s1);
      o[j0 + ns] = cadd(d0, md1);
      o[j0 + 2 * ns] = csub(s0, s1);
      o[j0 + 3 * ns] = csub(d0, md1);
    } else {
      const int rs = nc / R;  // tw[q*t*rs] = exp(-2 pi i q t / R)
#pragma unroll
      for (int q = 0; q < R; ++q) {
        float2 acc = v[0];
#pragma unroll
        for (int t = 1; t < R; ++t) acc = cadd(acc, cmul(v[t], tw[((q * t) % R) * rs]));
        o[j0 + q * ns] = acc;
      }
    }
  }
}

// Runs every stage; returns the buffer holding the result.  All threads of the block must call it.
__device__ float2* block_fft(float2* a, float2* b, const float2* tw, const FftPlan& p, int batch) {
  int ns = 1;
  for (int s = 0; s < p.nrad; ++s) {
    const int R = p.rad[s];
    if (R == 4) stockham_stage<4>(a, b, tw, p.nc, ns, batch);
    else if (R == 2) stockham_stage<2>(a, b, tw, p.nc, ns, batch);
    else if (R == 3) stockham_stage<3>(a, b, tw, p.nc, ns, batch);
    else stockham_stage<5>(a, b, tw, p.nc, ns, batch);
    __syncthreads();
    ns *= R;
    float2* t = a; a = b; b = t;
  }
  return a;
}

__device__ __forceinline__ void fill_tables(float2* tw, float2* tw2, int nc, int n_fft) {
  for (int j = threadIdx.x; j < nc; j += blockDim.x) {
    float s, c;
    sincospif(-2.0f * (float)j / (float)nc, &s, &c);
    tw[j] = make_float2(c, s);
  }
  for (int j = threadIdx.x; j <= nc; j += blockDim.x) {
    float s, c;
    sincospif(-2.0f * (float)j / (float)n_fft, &s, &c);
    tw2[j] = make_float2(c, s);
  }
}

// ---------------------------------------------------------------- forward: STFT (MODE 0) / log-mel (MODE 1)
// Ragged batches: `lens` (device, B ints or null) holds the valid samples of every signal (forward) / the valid frames
// (inverse).  A signal is transformed as if it ended there (its own reflect padding, its own last frame); frames /
// samples past its end are written as zeros.  ragged_samples / ragged_frames clamp to the padded extent.
struct StftArgs {
  const float* x; const float* window; float* out; const int* lens;
  int n, hop, pad_mode, frames, frames_per_block, group;
  // log-mel only
  const float* fb; int n_mels, power, log_mode; float floor_;
};

// valid samples of signal b and the number of frames they give (0 if the signal is too short for a frame or, with
// reflect padding, for the reflection)
__device__ __forceinline__ int ragged_samples(const StftArgs& a, int b) { return a.lens ? min(max(a.lens[b], 0), a.n) : a.n; }
__device__ __forceinline__ int ragged_frames(const StftArgs& a, int n_b, int n_fft) {
  const int pad = a.pad_mode == 0 ? 0 : n_fft / 2;
  if (n_b <= 0 || n_b + 2 * pad < n_fft || (a.pad_mode == 1 && n_b <= pad)) return 0;
  return min(a.frames, 1 + (n_b + 2 * pad - n_fft) / a.hop);
}

template <int MODE>
__global__ void __launch_bounds__(256) stft_kernel(const StftArgs a, const FftPlan p) {
  extern __shared__ __align__(16) unsigned char smem_[];
  const int nc = p.nc, n_fft = p.n_fft, bins = nc + 1;
  float2* tw = reinterpret_cast<float2*>(smem_);
  float2* tw2 = tw + nc;
  float* win = reinterpret_cast<float*>(tw2 + bins);
  float2* buf0 = reinterpret_cast<float2*>(win + n_fft);
  float2* buf1 = buf0 + a.group * nc;
  int* mel_lo = reinterpret_cast<int*>(buf1 + a.group * nc);
  int* mel_hi = mel_lo + (MODE == 1 ? a.n_mels : 0);

  fill_tables(tw, tw2, nc, n_fft);
  for (int i = threadIdx.x; i < n_fft; i += blockDim.x) win[i] = a.window[i];
  if (MODE == 1) {
    for (int m = threadIdx.x; m < a.n_mels; m += blockDim.x) {
      const float* row = a.fb + (size_t)m * bins;
      int lo = bins, hi = 0;
      for (int k = 0; k < bins; ++k)
        if (row[k] != 0.f) { if (k < lo) lo = k; hi = k + 1; }
      mel_lo[m] = lo; mel_hi[m] = hi;
    }
  }
  __syncthreads();

  const int b = blockIdx.y;
  const float* xb = a.x + (size_t)b * a.n;
  const int pad = a.pad_mode == 0 ? 0 : n_fft / 2;
  const int n_b = ragged_samples(a, b), frames_b = ragged_frames(a, n_b, n_fft);
  const int f_begin = blockIdx.x * a.frames_per_block;
  const int f_end = min(a.frames, f_begin + a.frames_per_block);
  for (int f0 = f_begin; f0 < f_end; f0 += a.group) {
    const int ng = min(a.group, f_end - f0);
    // load + window + pack even/odd samples into a complex sequence
    for (int it = threadIdx.x; it < ng * nc; it += blockDim.x) {
      const int g = it / nc, i = it - g * nc;
      const int base = (f0 + g) * a.hop - pad;
      float v[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        int q = base + 2 * i + e;
        float s = 0.f;
        if (f0 + g >= frames_b) {        // past the signal's last frame: zeros
        } else if (a.pad_mode == 1) {    // reflect without edge repeat (utils.py:81-84)
          if (q < 0) q = -q;
          if (q >= n_b) q = 2 * (n_b - 1) - q;
          s = xb[q];
        } else if (q >= 0 && q < n_b) {
          s = xb[q];
        }
        v[e] = s * win[2 * i + e];
      }
      buf0[g * nc + i] = make_float2(v[0], v[1]);
    }
    __syncthreads();
    float2* Z = block_fft(buf0, buf1, tw, p, ng);
    float2* other = (Z == buf0) ? buf1 : buf0;
    // unpack: X[k] = (Z[k]+conj(Z[nc-k]))/2 + e^{-2 pi i k/n} (Z[k]-conj(Z[nc-k]))/(2i)
    if (MODE == 0) {
      for (int it = threadIdx.x; it < ng * bins; it += blockDim.x) {
        const int g = it / bins, k = it - g * bins;
        float2 zk = Z[g * nc + (k == nc ? 0 : k)];
        float2 zc = cconj(Z[g * nc + ((nc - k) % nc)]);
        float2 e = make_float2(0.5f * (zk.x + zc.x), 0.5f * (zk.y + zc.y));
        float2 d = csub(zk, zc);
        float2 o = make_float2(0.5f * d.y, -0.5f * d.x);  // d / (2i)
        float2 X = cadd(e, cmul(tw2[k], o));
        if (k == 0 || k == nc) X.y = 0.f;
        reinterpret_cast<float2*>(a.out)[((size_t)b * a.frames + f0 + g) * bins + k] = X;
      }
      __syncthreads();
    } else {
      float* P = reinterpret_cast<float*>(other);  // ng * bins floats <= ng * nc float2 (bins <= 2 nc)
      for (int it = threadIdx.x; it < ng * bins; it += blockDim.x) {
        const int g = it / bins, k = it - g * bins;
        float2 zk = Z[g * nc + (k == nc ? 0 : k)];
        float2 zc = cconj(Z[g * nc + ((nc - k) % nc)]);
        float2 e = make_float2(0.5f * (zk.x + zc.x), 0.5f * (zk.y + zc.y));
        float2 d = csub(zk, zc);
        float2 o = make_float2(0.5f * d.y, -0.5f * d.x);
        float2 X = cadd(e, cmul(tw2[k], o));
        if (k == 0 || k == nc) X.y = 0.f;
        float m2 = X.x * X.x + X.y * X.y;
        P[g * bins + k] = a.power == 2 ? m2 : sqrtf(m2);
      }
      __syncthreads();
      for (int it = threadIdx.x; it < ng * a.n_mels; it += blockDim.x) {
        const int g = it / a.n_mels, m = it - g * a.n_mels;
        const float* row = a.fb + (size_t)m * bins;
        float acc = 0.f;
        for (int k = mel_lo[m]; k < mel_hi[m]; ++k) acc = fmaf(P[g * bins + k], __ldg(row + k), acc);
        if (a.log_mode == 1) acc = logf(fmaxf(acc, a.floor_));
        else if (a.log_mode == 2) acc = log10f(fmaxf(acc, a.floor_));
        else if (a.log_mode == 3) acc = logf(acc + a.floor_);
        a.out[((size_t)b * a.frames + f0 + g) * a.n_mels + m] = f0 + g < frames_b ? acc : 0.f;
      }
      __syncthreads();
    }
  }
}


// ---------------------------------------------------------------- inverse: iSTFT with gather overlap-add
// lens: valid frames per signal (or null); `tail` = samples the caller cuts from the end of the full overlap-add
// (n_fft/2 with centre trimming, 0 with an explicit length), so that a shorter signal ends where it would alone.
struct IstftArgs {
  const float* spec; const float* window; float* out; const int* lens;
  int frames, hop, trim, tail, out_len, hops_per_block, group, max_frames;
};
template <class Args>
__device__ __forceinline__ void ragged_istft(const Args& a, int b, int n_fft, int hop, int& frames_b, long long& total_b,
                                             long long& out_len_b) {
  frames_b = a.lens ? min(max(a.lens[b], 0), a.frames) : a.frames;
  total_b = frames_b > 0 ? (long long)(frames_b - 1) * hop + n_fft : 0;
  out_len_b = a.lens ? min((long long)a.out_len, max(0LL, total_b - a.trim - a.tail)) : a.out_len;
}

__global__ void __launch_bounds__(256) istft_kernel(const IstftArgs a, const FftPlan p) {
  extern __shared__ __align__(16) unsigned char smem_[];
  const int nc = p.nc, n_fft = p.n_fft, bins = nc + 1;
  float2* tw = reinterpret_cast<float2*>(smem_);
  float2* tw2 = tw + nc;
  float* win = reinterpret_cast<float*>(tw2 + bins);
  float2* buf0 = reinterpret_cast<float2*>(win + n_fft);
  float2* buf1 = buf0 + a.group * nc;
  float* tf = reinterpret_cast<float*>(buf1 + a.group * nc);   // (max_frames, n_fft) windowed time frames

  fill_tables(tw, tw2, nc, n_fft);
  for (int i = threadIdx.x; i < n_fft; i += blockDim.x) win[i] = a.window[i];
  __syncthreads();

  const int b = blockIdx.y;
  const long long total = (long long)(a.frames - 1) * a.hop + n_fft;   // untrimmed length (utils.py:126)
  int frames_b; long long total_b, out_len_b;
  ragged_istft(a, b, n_fft, a.hop, frames_b, total_b, out_len_b);
  const long long u0 = (long long)blockIdx.x * a.hops_per_block * a.hop;
  const long long u1 = min(total, u0 + (long long)a.hops_per_block * a.hop);
  if (u0 >= total) return;
  const int f_lo = u0 >= n_fft ? (int)((u0 - n_fft) / a.hop) + 1 : 0;
  const int f_hi = min(frames_b - 1, (int)((u1 - 1) / a.hop));
  const int nf = max(f_hi - f_lo + 1, 0);
  const float2* sb = reinterpret_cast<const float2*>(a.spec) + (size_t)b * a.frames * bins;
  const float inv_nc = 1.0f / (float)nc;

  for (int g0 = 0; g0 < nf; g0 += a.group) {
    const int ng = min(a.group, nf - g0);
    // pack: Z[k] = E[k] + i O[k], E = (X[k]+conj(X[nc-k]))/2, O = (X[k]-conj(X[nc-k]))/2 * e^{+2 pi i k/n};
    // the inverse transform is conj(FFT(conj(Z)))/nc, so conj(Z) is stored.  irfft ignores Im X[0], Im X[nc].
    for (int it = threadIdx.x; it < ng * nc; it += blockDim.x) {
      const int g = it / nc, k = it - g * nc;
      const float2* X = sb + (size_t)(f_lo + g0 + g) * bins;
      float2 xk = X[k], xc = X[nc - k];
      if (k == 0) { xk.y = 0.f; xc.y = 0.f; }
      xc = cconj(xc);
      float2 e = make_float2(0.5f * (xk.x + xc.x), 0.5f * (xk.y + xc.y));
      float2 d = make_float2(0.5f * (xk.x - xc.x), 0.5f * (xk.y - xc.y));
      float2 o = cmul(d, cconj(tw2[k]));
      float2 z = make_float2(e.x - o.y, e.y + o.x);   // e + i*o
      buf0[g * nc + k] = cconj(z);
    }
    __syncthreads();
    float2* Z = block_fft(buf0, buf1, tw, p, ng);
    for (int it = threadIdx.x; it < ng * nc; it += blockDim.x) {
      const int g = it / nc, i = it - g * nc;
      float2 z = Z[g * nc + i];
      float* dst = tf + (size_t)(g0 + g) * n_fft + 2 * i;
      // conj and the 1/n normalisation of irfft (1/nc here, the factor 1/2 is already inside E and O)
      dst[0] = (z.x * inv_nc) * win[2 * i];
      dst[1] = (-z.y * inv_nc) * win[2 * i + 1];
    }
    __syncthreads();
  }
  // gather overlap-add in increasing frame order (utils.py:138-150)
  for (long long u = u0 + threadIdx.x; u < u1; u += blockDim.x) {
    const long long oi = u - a.trim;
    if (oi < 0 || oi >= a.out_len) continue;
    int fa = u >= n_fft ? (int)((u - n_fft) / a.hop) + 1 : 0;
    int fb_ = min(frames_b - 1, (int)(u / a.hop));
    float acc = 0.f, ws = 0.f;
    for (int f = fa; f <= fb_; ++f) {
      const int n = (int)(u - (long long)f * a.hop);
      acc += tf[(size_t)(f - f_lo) * n_fft + n];
      ws += win[n];
    }
    a.out[(size_t)b * a.out_len + oi] = oi >= out_len_b ? 0.f : (ws != 0.f) ? acc / ws : acc;
  }
}


// ================================================================== n_fft = 1024 fast path (fft_fast.cuh)
// Four frames per block iteration, 64 threads each.  Forward: pass 1 reads the windowed samples straight from
// global memory (float2, coalesced; reflect / zero padding only on the edge frames), the half spectrum is unpacked
// from the last exchange buffer.  MODE 1 (log-mel) keeps |X|^p in shared memory and projects it there.
constexpr int MEL_ITEMS = 384;   // <= 8-bin pieces of the filterbank's non-zero spans (80 htk mels at 1024: ~170)
constexpr int MEL_MAX = 128;     // sizes the flag array: GROUPS * MEL_MAX >= MEL_ITEMS words

template <int MODE, bool RT>
__global__ void __launch_bounds__(ff::THREADS, RT ? 2 : 4) stft_fast_kernel(const StftArgs a) {
  extern __shared__ __align__(16) unsigned char smem_[];
  constexpr int NC = ff::NC, NT = ff::NT, n_fft = 2 * NC, bins = NC + 1;
  float2* tw = reinterpret_cast<float2*>(smem_);
  float2* tw2 = tw + ff::TW;
  float2* win = tw2 + bins + 1;                       // n_fft floats as NC float2
  float2* bufs = win + NC;
  int* mel_lo = reinterpret_cast<int*>(bufs + 2 * ff::GROUPS * ff::BUF);
  const int nm2 = (a.n_mels + 3) & ~3;                      // keeps the float4 table below 16-byte aligned
  int* mel_hi = mel_lo + (MODE == 1 ? nm2 : 0);
  // balanced sparse mel projection: the filterbank's non-zero spans cut into items of <= 8 bins
  int* it_m = mel_hi + (MODE == 1 ? nm2 : 0);               // [MEL_ITEMS] mel row of the item
  int* it_k = it_m + (MODE == 1 ? MEL_ITEMS : 0);           // [MEL_ITEMS] first bin
  float* fbc = reinterpret_cast<float*>(it_k + (MODE == 1 ? MEL_ITEMS : 0));   // [MEL_ITEMS][8] weights, zero padded
  int* it_f = reinterpret_cast<int*>(fbc + (MODE == 1 ? MEL_ITEMS * 8 : 0));   // [MEL_ITEMS] reduction flags (<= GROUPS * MEL_MAX words)
  float* melacc = reinterpret_cast<float*>(it_f);
  int* n_items_s = reinterpret_cast<int*>(melacc + (MODE == 1 ? ff::GROUPS * MEL_MAX : 0));
  const int g = threadIdx.x / NT, j = threadIdx.x % NT;
  using ff::c2; using ff::pk; using ff::up;
  c2* A = reinterpret_cast<c2*>(bufs + (2 * g) * ff::BUF);
  c2* Bf = A + ff::BUF;
  const c2* twc = reinterpret_cast<const c2*>(tw);
  const ff::SwzIdx sz(j);

  // RT: window, twiddles and unpack factors of this thread live in registers (two blocks per SM instead of four)
  c2 rw[8], rt2[8], rt3[8], ru[4];
  if (RT) {
    ff::fill_reg_twiddles(j, rt2, rt3);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      rw[t] = pk(a.window[2 * (j + NT * t)], a.window[2 * (j + NT * t) + 1]);
      if (t < 4) {
        float sn, cs_;
        sincospif(-(float)(j + NT * t) / (float)NC, &sn, &cs_);
        ru[t] = pk(cs_, sn);
      }
    }
  } else {
    ff::fill_tables(tw, tw2);
    for (int i = threadIdx.x; i < NC; i += blockDim.x) win[i] = make_float2(a.window[2 * i], a.window[2 * i + 1]);
  }
  // log-mel: the unpack factor of point j + 64 q as u0 * W16^q (u0 = exp(-i pi j / 512) in two registers, second factor
  // a compile-time constant) instead of a table read
  c2 u0 = 0;
  if (MODE == 1) {
    float sn, cs_;
    sincospif(-(float)j / (float)NC, &sn, &cs_);
    u0 = pk(cs_, sn);
  }
  if (MODE == 1) {   // non-zero span of every mel filter: one warp per row
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int m = warp; m < a.n_mels; m += ff::THREADS / 32) {
      const float* row = a.fb + (size_t)m * bins;
      int lo = bins, hi = 0;
      for (int k = lane; k < bins; k += 32)
        if (row[k] != 0.f) { lo = min(lo, k); hi = max(hi, k + 1); }
      for (int off = 16; off > 0; off >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, off));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, off));
      }
      if (lane == 0) { mel_lo[m] = lo; mel_hi[m] = hi; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {   // item list (serial over the rows: a few hundred cheap steps, once per block)
      // A row's items are consecutive and never straddle a group of 32 (dummy items pad the group), so that a warp
      // reduces them with shuffles and the row's first item owns the result: no atomics, no accumulator array.
      // it_f: bit d (d < 4) = item + 2^d belongs to the same row; bit 4 = first item of its row.
      int n = 0;
      auto dummy = [&]() { it_m[n] = -1; it_k[n] = 0; it_f[n] = 0; ++n; };
      for (int m = 0; m < a.n_mels && n >= 0; ++m) {
        const int lo = mel_lo[m], hi = mel_hi[m];
        const int cnt = lo < hi ? (hi - lo + 7) / 8 : 1;     // an all-zero row still reports log(floor)
        if (cnt > 16) { n = -1; break; }                     // four shuffle steps cover 16 items
        if ((n & 31) + cnt > 32) while ((n & 31) && n < MEL_ITEMS) dummy();
        if (n + cnt > MEL_ITEMS) { n = -1; break; }          // too dense for the table: row-loop fallback
        for (int c = 0; c < cnt; ++c) {
          int fl = c == 0 ? 16 : 0;
          for (int d = 0; d < 4; ++d) if (c + (1 << d) < cnt) fl |= 1 << d;
          it_m[n] = m; it_k[n] = lo < hi ? lo + 8 * c : 0; it_f[n] = fl; ++n;
        }
      }
      if (n > 0) while ((n & 31) && n < MEL_ITEMS) dummy();
      if (n > 0 && (n & 31)) n = -1;
      *n_items_s = n;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < max(*n_items_s, 0) * 8; i += blockDim.x) {
      const int it = i >> 3, k = it_k[it] + (i & 7), m = it_m[it];
      fbc[i] = (m >= 0 && k >= mel_lo[m] && k < mel_hi[m]) ? a.fb[(size_t)m * bins + k] : 0.f;
    }
  }
  __syncthreads();
  const int n_items = MODE == 1 ? *n_items_s : 0;

  const int b = blockIdx.y;
  const float* xb = a.x + (size_t)b * a.n;
  const int pad = a.pad_mode == 0 ? 0 : NC;
  const bool vec_ok = ((a.hop | pad) & 1) == 0 && ((reinterpret_cast<uintptr_t>(xb) & 7) == 0);
  const int n_b = ragged_samples(a, b), frames_b = ragged_frames(a, n_b, n_fft);
  const int f_begin = blockIdx.x * a.frames_per_block;
  const int f_end = min(a.frames, f_begin + a.frames_per_block);
  const int f_live = min(f_end, frames_b);       // frames past the signal's last one transform zeros
  // samples 2(j + 64 t), +1 of frame f (zero for a frame past the block's range or the signal's end)
  auto load_frame = [&](int f, float2 (&v)[8]) {
    const int base = f * a.hop - pad;
    if (f < f_live && vec_ok && base >= 0 && base + n_fft <= n_b) {
      const float2* src = reinterpret_cast<const float2*>(xb + base);
#pragma unroll
      for (int t = 0; t < 8; ++t) v[t] = __ldg(src + j + NT * t);
    } else {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        float sv[2] = {0.f, 0.f};
        if (f < f_live) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            int q = base + 2 * (j + NT * t) + e;
            if (a.pad_mode == 1) {           // reflect without edge repeat (utils.py:81-84)
              if (q < 0) q = -q;
              if (q >= n_b) q = 2 * (n_b - 1) - q;
              sv[e] = xb[q];
            } else if (q >= 0 && q < n_b) {
              sv[e] = xb[q];
            }
          }
        }
        v[t] = make_float2(sv[0], sv[1]);
      }
    }
  };
  // the four transforms of a block are independent: each pair of warps synchronises on its own named barrier
  auto gsync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(NT) : "memory"); };
  float2 nx[8];
  load_frame(f_begin + g, nx);
  for (int f0 = f_begin; f0 < f_end; f0 += ff::GROUPS) {
    const int f = f0 + g;
    const bool live = f < f_end;
    // ---- pass 1: windowed samples packed as complex points; the next frame's samples are requested first
    {
      c2 v[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) v[t] = ff::mul2(pk(nx[t]), RT ? rw[t] : pk(win[j + NT * t]));   // rounded products: no contraction into the first butterfly
      load_frame(f + ff::GROUPS, nx);
      ff::pass1_store(v, A, sz);
    }
    gsync();
    constexpr bool POW = MODE == 1;   // log-mel: no bit-for-bit twin, twiddles as products (see ff::tw_powers)
    if (RT) ff::pass2_r(A, Bf, rt2, sz); else ff::pass2<POW>(A, Bf, twc, j, sz);
    gsync();
    c2 z[8];                                      // z[q] = Z[j + 64 q]: kept in registers for the unpack
    if (RT) ff::pass3_r(Bf, rt3, sz, z); else ff::pass3<POW>(Bf, twc, j, sz, z);
    // ---- unpack in conjugate pairs: with E = (Z[k] + conj Z[nc-k]) / 2, O = (Z[k] - conj Z[nc-k]) / (2i),
    // u = e^{-2 pi i k / n}:   X[k] = E + u O   and   X[nc-k] = conj(E - u O).
    // Thread j owns the pairs of its lower four points k = j + 64 q (q < 4); their partners are the upper four points
    // of thread 64 - j, the only values that cross shared memory (thread 0 pairs with itself, and owns k = nc / 2).
#pragma unroll
    for (int q = 4; q < 8; ++q) A[sz.cons(q)] = z[q];
    gsync();
    float* P = reinterpret_cast<float*>(Bf);   // MODE 1: |X|^p of this frame (513 floats <= BUF float2)
    auto emit = [&](int k, c2 Xp) {
      if (MODE == 0) {
        if (live) reinterpret_cast<c2*>(a.out)[((size_t)b * a.frames + f) * bins + k] = Xp;
      } else {
        const float2 X = up(Xp);
        const float m2 = X.x * X.x + X.y * X.y;
        float mg = m2;
        if (a.power != 2) asm("sqrt.approx.f32 %0, %1;" : "=f"(mg) : "f"(m2));   // 1 ulp, no slow-path branch
        P[k] = mg;
      }
    };
    auto unpack_pair = [&](int k, c2 zk, c2 zc, c2 u, bool self) {
      // with zc' = conj(zc):  e = (zk + zc') / 2,  o = (zk - zc') / (2i),  t = u o
      const c2 s = ff::fma2(zc, pk(1.f, -1.f), zk);
      const c2 d = ff::fma2(zc, pk(-1.f, 1.f), zk);
      const c2 e = ff::mul2(s, pk(0.5f, 0.5f)), ec = ff::mul2(s, pk(0.5f, -0.5f));
      const c2 o = ff::mul2(ff::sw(d), pk(0.5f, -0.5f));
      const c2 t = ff::cm(u, o);
      c2 X1 = ff::add2(e, t);
      c2 X2 = ff::fma2(t, pk(-1.f, 1.f), ec);                  // conj(e - t) = (e.x - t.x, t.y - e.y)
      if (k == 0) { X1 = pk(up(X1).x, 0.f); X2 = pk(up(X2).x, 0.f); }   // rfft: DC and Nyquist are real
      emit(k, X1);
      if (!self) emit(NC - k, X2);
    };
    {
      c2 zc[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) zc[q] = A[sz.mirror(q)];
      if (j == 0) zc[0] = z[0];                                // k = 0 pairs with nc = 0 (mod nc): its own point
#pragma unroll
      for (int q = 0; q < 4; ++q)
        unpack_pair(j + NT * q, z[q], zc[q],
                    RT ? ru[q] : MODE == 1 ? (q ? ff::cmc(u0, ff::w32c(2 * q), -ff::w32s(2 * q)) : u0) : pk(tw2[j + NT * q]), false);
      if (j == 0) unpack_pair(NC / 2, z[4], z[4], pk(0.f, -1.f), true);   // e^{-i pi / 2}
    }
    gsync();
    if (MODE == 1) {
      auto finish = [&](float acc, int m) {
        if (a.log_mode == 1) acc = logf(fmaxf(acc, a.floor_));
        else if (a.log_mode == 2) acc = log10f(fmaxf(acc, a.floor_));
        else if (a.log_mode == 3) acc = logf(acc + a.floor_);
        if (live) a.out[((size_t)b * a.frames + f) * a.n_mels + m] = f < frames_b ? acc : 0.f;
      };
      if (n_items >= 0) {
        // 32 items per warp step; lane l reads element (i + (l >> 2)) & 7 of its 8-bin piece at step i: consecutive
        // items of a row sit 8 floats apart, the skew spreads the 32 lanes over all banks (P and weights alike)
        const int lane = j & 31, skew = lane >> 2;
        for (int it = (j & 32) + lane; it < n_items; it += NT) {
          const float* pk = P + it_k[it];                 // (bins past 512 read exchange-buffer leftovers x 0)
          const float* wk = fbc + it * 8;
          const int fl = it_f[it];
          float acc = 0.f;
#pragma unroll
          for (int i = 0; i < 8; ++i) { const int e = (i + skew) & 7; acc = fmaf(pk[e], wk[e], acc); }
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const float o = __shfl_down_sync(0xffffffffu, acc, 1 << d);
            if (fl & (1 << d)) acc += o;
          }
          if (fl & 16) finish(acc, it_m[it]);
        }
      } else {
        for (int m = j; m < a.n_mels; m += NT) {
          const float* row = a.fb + (size_t)m * bins;
          float acc = 0.f;
          for (int k = mel_lo[m]; k < mel_hi[m]; ++k) acc = fmaf(P[k], __ldg(row + k), acc);
          finish(acc, m);
        }
      }
      gsync();
    }
  }
}

#ifdef KK_FFT_R32   // lab build: the single-exchange forward kernel (measured slower; experiments/)
#include "../../experiments/fft_r32_kernel.cuh"
#endif

// Inverse, streaming form (hop divides n_fft, R = n_fft / hop in {1, 2, 4}): a block walks its frames four at a time;
// every group inverse-transforms one frame straight from the global half spectrum into its (scaled, not yet windowed)
// time frame in shared memory.  Overlap-add is a gather with the running sums in REGISTERS: thread `tid` owns the
// output offsets tid + 256 o (o < hop / 256) of every hop, keeps the synthesis-window taps that can touch them
// (R x hop/256 = 4 values) and the R - 1 partial sums that later frames still add to; per iteration it reads the 16
// samples of the four new frames that land on its offsets, adds them in increasing frame order (the reference's
// scatter-add order, utils.py:138-147), emits the four hops that are now complete divided by the window sum
// (utils.py:150, same order) and carries the rest.  No ring buffer, no atomics, no second pass over the output.
// A block first replays the frames before its range (warm-up, nothing emitted); the last block drains the tail.
struct IstftStreamArgs {
  const float* spec; const float* window; float* out; const int* lens;
  int frames, hop, trim, tail, out_len, frames_per_block, R;
};

template <int R, bool RT>
__global__ void __launch_bounds__(ff::THREADS, RT ? 2 : 4) istft_stream_kernel(const IstftStreamArgs a) {
  extern __shared__ __align__(16) unsigned char smem_[];
  constexpr int NC = ff::NC, NT = ff::NT, n_fft = 2 * NC, bins = NC + 1, G = ff::GROUPS;
  constexpr int hop = n_fft / R, OFFS = hop / ff::THREADS;     // R * OFFS == 4
  static_assert(R * OFFS == 4 && R - 1 <= G, "streaming iSTFT: hop must be 256, 512 or 1024");
  float2* tw = reinterpret_cast<float2*>(smem_);
  float2* tw2 = tw + ff::TW;
  float2* bufs = tw2 + bins + 1;
  const int g = threadIdx.x / NT, j = threadIdx.x % NT;
  using ff::c2; using ff::pk; using ff::up;
  c2* A = reinterpret_cast<c2*>(bufs + (2 * g) * ff::BUF);
  c2* Bf = A + ff::BUF;
  const c2* twc = reinterpret_cast<const c2*>(tw);
  const ff::SwzIdx sz(j);

  c2 rt2[8], rt3[8], ru[8];         // RT: twiddles and pack factors in registers (see stft_fast_kernel)
  if (RT) {
    ff::fill_reg_twiddles(j, rt2, rt3);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      float sn, cs_;
      sincospif(-(float)(j + NT * t) / (float)NC, &sn, &cs_);
      ru[t] = pk(cs_, sn);
    }
  } else {
    ff::fill_tables(tw, tw2);
  }
  // table variant: the pack factor of point j + 64 t is u0 * W16^t, u0 = exp(-i pi j / 512) in two registers and the
  // second factor a compile-time constant (8 fewer 64-bit shared loads per thread and frame: the L1 / shared data pipe
  // is what bounds this kernel)
  c2 u0;
  {
    float sn, cs_;
    sincospif(-(float)j / (float)NC, &sn, &cs_);
    u0 = pk(cs_, sn);
  }
  // synthesis-window taps of this thread's offsets, and the partial sums carried between iterations
  float w[OFFS][R], carry[OFFS][R > 1 ? R - 1 : 1], ws_full[OFFS];
#pragma unroll
  for (int o = 0; o < OFFS; ++o) {
#pragma unroll
    for (int r = 0; r < R; ++r) w[o][r] = a.window[r * hop + threadIdx.x + ff::THREADS * o];
#pragma unroll
    for (int c = 0; c < (R > 1 ? R - 1 : 1); ++c) carry[o][c] = 0.f;
    ws_full[o] = 0.f;
#pragma unroll
    for (int r = R - 1; r >= 0; --r) ws_full[o] = __fadd_rn(ws_full[o], w[o][r]);   // all R frames present
  }
  __syncthreads();

  const int b = blockIdx.y;
  const long long total = (long long)(a.frames - 1) * hop + n_fft;   // untrimmed length (utils.py:126)
  int frames_b; long long total_b, out_len_b;                        // this signal's own extent (ragged batches)
  ragged_istft(a, b, n_fft, hop, frames_b, total_b, out_len_b);
  const int hops_total = (int)((total + hop - 1) / hop);
  const int H0 = blockIdx.x * a.frames_per_block;
  const int H1 = (blockIdx.x == gridDim.x - 1) ? hops_total : min(hops_total, H0 + a.frames_per_block);
  const float2* sb = reinterpret_cast<const float2*>(a.spec) + (size_t)b * a.frames * bins;
  const float inv_nc = 1.0f / (float)NC;
  auto gsync = [&]() { asm volatile("bar.sync %0, %1;" ::"r"(g + 1), "n"(NT) : "memory"); };

  for (int fi = H0 >= G ? H0 - G : 0; fi < H1; fi += G) {
    const int f = fi + g;
    const bool live = f < frames_b;
    {
      // L2 prefetch of the next iteration's four spectrum rows (contiguous: 4 * 4104 bytes = 129 lines)
      const long long nf = (long long)fi + G;
      if (nf < a.frames) {
        const char* nx = reinterpret_cast<const char*>(sb + (size_t)nf * bins);
        const long long nbytes = (long long)min((long long)G, (long long)a.frames - nf) * bins * 8;
        const long long o = (long long)threadIdx.x * 128;
        if (o < nbytes) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + o));
      }
      // pack: Z[k] = E[k] + i O[k], E = (X[k]+conj(X[nc-k]))/2, O = (X[k]-conj(X[nc-k]))/2 * e^{+2 pi i k/n};
      // the inverse transform is conj(FFT(conj(Z)))/nc, so conj(Z) goes in.  irfft ignores Im X[0], Im X[nc].
      c2 v[8];
      const float2* X = sb + (size_t)(live ? f : 0) * bins;
      float2 xk[8], xc[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) { xk[t] = __ldg(X + j + NT * t); xc[t] = __ldg(X + NC - (j + NT * t)); }
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int k = j + NT * t;
        float2 a_ = xk[t], c_ = xc[t];
        if (k == 0) { a_.y = 0.f; c_.y = 0.f; }
        // with c' = conj(c):  e = (a + c') / 2,  dd = (a - c') / 2
        const c2 e = ff::mul2(ff::fma2(pk(c_), pk(1.f, -1.f), pk(a_)), pk(0.5f, 0.5f));
        const c2 dd = ff::mul2(ff::fma2(pk(c_), pk(-1.f, 1.f), pk(a_)), pk(0.5f, 0.5f));
        const float2 w2 = up(RT ? ru[t] : (t ? ff::cmc(u0, ff::w32c(2 * t), -ff::w32s(2 * t)) : u0));
        const c2 o = ff::cm(dd, pk(w2.x, -w2.y));
        const float2 z = up(ff::fma2(ff::sw(o), pk(-1.f, 1.f), e));   // e + i*o = (e.x - o.y, e.y + o.x)
        v[t] = pk(z.x, -z.y);                                         // conj
      }
      ff::pass1_store(v, A, sz);
    }
    gsync();
    if (RT) ff::pass2_r(A, Bf, rt2, sz); else ff::pass2<true>(A, Bf, twc, j, sz);
    gsync();
    {
      c2 v[8];
      if (RT) ff::pass3_r(Bf, rt3, sz, v); else ff::pass3<true>(Bf, twc, j, sz, v);
      c2* tfg = A;       // the group's time frame (n_fft floats), reusing its first exchange buffer
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        // conj only (the factor 1/2 of 1/n is already inside E and O): the 1/nc of irfft is a power of two, so it commutes exactly with the window products and the
        // overlap-add sums and is applied once per output sample below instead of twice per transform point
        const float2 vq = up(v[q]);
        tfg[j + NT * q] = live ? pk(vq.x, -vq.y) : pk(0.f, 0.f);
      }
    }
    __syncthreads();
#pragma unroll
    for (int o = 0; o < OFFS; ++o) {
      const int off = threadIdx.x + ff::THREADS * o;
      // the 16 (frame, hop-of-frame) samples on this offset: frame gg's r-th hop lands on hop fi + gg + r
      float tv[G][R];
#pragma unroll
      for (int gg = 0; gg < G; ++gg)
#pragma unroll
        for (int r = 0; r < R; ++r)
          tv[gg][r] = __fmul_rn(reinterpret_cast<const float*>(bufs + (2 * gg) * ff::BUF)[r * hop + off], w[o][r]);
      // interior iterations (every hop inside this block's range, every covering frame present, every sample inside
      // the trimmed output): precomputed window sum, 32-bit addressing
      const bool interior = fi >= H0 && fi + G <= H1 && fi >= R - 1 && fi + G <= frames_b &&
                            (long long)fi * hop >= a.trim && (long long)(fi + G) * hop <= total_b &&
                            (long long)(fi + G) * hop - a.trim <= out_len_b;
      float* orow = a.out + (size_t)b * a.out_len + ((long long)fi * hop - a.trim) + off;
#pragma unroll
      for (int hh = 0; hh < G; ++hh) {
        float acc = (hh < R - 1) ? carry[o][hh < R - 1 ? hh : 0] : 0.f;
#pragma unroll
        for (int gg = 0; gg < G; ++gg) {
          const int r = hh - gg;
          if (r >= 0 && r < R) acc = __fadd_rn(acc, tv[gg][r]);
        }
        if (interior) {
          acc = __fmul_rn(acc, inv_nc);
          orow[hh * hop] = (ws_full[o] != 0.f) ? acc / ws_full[o] : acc;
          continue;
        }
        const int hcur = fi + hh;
        const long long u = (long long)hcur * hop + off;
        const long long oi = u - a.trim;
        if (hcur >= H0 && hcur < H1 && u < total && oi >= 0 && oi < a.out_len) {
          float ws = 0.f;
#pragma unroll
          for (int r = R - 1; r >= 0; --r) {             // increasing frame order, like the reference's loop
            const int fr = hcur - r;                     // frames covering this sample
            if (fr >= 0 && fr < frames_b) ws = __fadd_rn(ws, w[o][r]);
          }
          acc = __fmul_rn(acc, inv_nc);
          a.out[(size_t)b * a.out_len + oi] = oi >= out_len_b ? 0.f : (ws != 0.f) ? acc / ws : acc;   // past the signal's own end: zeros
        }
      }
      // partial sums of the next R - 1 hops (frames of this iteration that reach past it)
#pragma unroll
      for (int c = 0; c < R - 1; ++c) {
        float acc = 0.f;
#pragma unroll
        for (int gg = 0; gg < G; ++gg) {
          const int r = G + c - gg;
          if (r >= 0 && r < R) acc = __fadd_rn(acc, tv[gg][r]);
        }
        carry[o][c] = acc;
      }
    }
    __syncthreads();
  }
}

}  // namespace kk

// ================================================================== C ABI
namespace kk {
static int stft_group(int nc) { int g = 2048 / nc; return g < 1 ? 1 : (g > 8 ? 8 : g); }

template <int MODE>
static int launch_stft(const StftArgs& a0, const FftPlan& p, int B, cudaStream_t st) {
  StftArgs a = a0;
#ifdef KK_FFT_R32
  if (p.n_fft == 2 * fr::NC && MODE == 0) {   // single-exchange radix-32 x 16 path
    int fpb = 32;
    while (fpb < 256 && (long long)cdiv(a.frames, fpb * 2) * B >= 148 * 16) fpb *= 2;
    a.frames_per_block = fpb;
    const size_t smem = (size_t)(fr::NC + fr::TG + fr::SLOTS * fr::XBUF) * 8;
    KK_CUDA(cudaFuncSetAttribute(stft_r32_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    stft_r32_kernel<MODE><<<dim3(cdiv(a.frames, fpb), B), fr::THREADS, smem, st>>>(a);
    KK_LAUNCH_CHECK("stft_r32_kernel");
    return KK_OK;
  }
#endif
  if (p.n_fft == 2 * ff::NC) {   // register-resident radix-8 path
    int fpb = 32;
    while (fpb < 256 && (long long)cdiv(a.frames, fpb * 2) * B >= 148 * 16) fpb *= 2;
    a.frames_per_block = fpb;
    size_t smem = (size_t)ff::TW * 8 + (size_t)(ff::NC + 2) * 8 + (size_t)ff::NC * 8 +
                  (size_t)2 * ff::GROUPS * ff::BUF * 8 +
                  (MODE == 1 ? (size_t)2 * (a.n_mels + 4) * 4 + (size_t)MEL_ITEMS * (8 + 32) + (size_t)ff::GROUPS * MEL_MAX * 4 + 16 : 0) + 16;
    dim3 grid(cdiv(a.frames, fpb), B);
    // register-table variant once a thread runs enough frames to amortise its 24 sincospif (A/B knob KK_FFT_RT=0/1 in -DKK_TUNE builds)
#ifdef KK_TUNE
    static const int rt_env = [] { const char* e = getenv("KK_FFT_RT"); return e ? atoi(e) : -1; }();
#else
    constexpr int rt_env = -1;                               // the product library reads no environment
#endif
    const bool rt = MODE == 0 && (rt_env >= 0 ? rt_env != 0 : fpb >= 64);   // log-mel: measured slower (2.26 -> 2.66 ms)
    if (rt) {
      KK_CUDA(cudaFuncSetAttribute(stft_fast_kernel<MODE, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      stft_fast_kernel<MODE, true><<<grid, ff::THREADS, smem, st>>>(a);
    } else {
      KK_CUDA(cudaFuncSetAttribute(stft_fast_kernel<MODE, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      stft_fast_kernel<MODE, false><<<grid, ff::THREADS, smem, st>>>(a);
    }
    KK_LAUNCH_CHECK("stft_fast_kernel");
    return KK_OK;
  }
  a.group = stft_group(p.nc);
  // enough blocks for several waves, but amortise the per-block table setup over >= 4 groups
  int fpb = a.group * 4;
  while (fpb < 256 && (long long)cdiv(a.frames, fpb * 2) * B >= 148 * 8) fpb *= 2;
  a.frames_per_block = fpb;
  size_t smem = (size_t)p.nc * 8 + (size_t)(p.nc + 1) * 8 + (size_t)p.n_fft * 4 + (size_t)2 * a.group * p.nc * 8 +
                (MODE == 1 ? (size_t)2 * a.n_mels * 4 : 0) + 16;
  KK_CUDA(cudaFuncSetAttribute(stft_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  dim3 grid(cdiv(a.frames, fpb), B);
  stft_kernel<MODE><<<grid, 256, smem, st>>>(a, p);
  KK_LAUNCH_CHECK("stft_kernel");
  return KK_OK;
}
}  // namespace kk

static int check_stft_args(const char* who, const float* x, const float* window, const void* out, int B, int n,
                           int n_fft, int hop, int pad_mode, int frames, kk::FftPlan* p) {
  KK_CHECK_ARG(x && window && out, "%s: null pointer", who);
  KK_CHECK_ARG(B > 0 && B <= 65535 && n > 0 && hop > 0 && frames > 0, "%s: bad shape", who);
  KK_CHECK_ARG(pad_mode >= 0 && pad_mode <= 2, "%s: pad_mode %d", who, pad_mode);
  if (kk::make_fft_plan(n_fft, p) != 0) {
    kk::set_error("%s: n_fft=%d must be even, <= 4096 and factor into 2, 3, 5", who, n_fft);
    return KK_ERR_UNSUPPORTED;
  }
  int padded = n + (pad_mode ? 2 * (n_fft / 2) : 0);
  KK_CHECK_ARG(padded >= n_fft && frames <= 1 + (padded - n_fft) / hop, "%s: frames=%d exceeds the signal", who, frames);
  KK_CHECK_ARG(pad_mode != 1 || n > n_fft / 2, "%s: reflect padding needs n > n_fft/2", who);
  return KK_OK;
}

extern "C" int kk_stft_f32(const float* x, const float* window, float* spec, int B, int n, int n_fft, int hop,
                           int pad_mode, int frames, const int* lens, void* stream) {
  kk::FftPlan p;
  int rc = check_stft_args("kk_stft_f32", x, window, spec, B, n, n_fft, hop, pad_mode, frames, &p);
  if (rc) return rc;
  kk::StftArgs a = {};
  a.x = x; a.window = window; a.out = spec; a.n = n; a.hop = hop; a.pad_mode = pad_mode; a.frames = frames;
  a.lens = lens;
  return kk::launch_stft<0>(a, p, B, (cudaStream_t)stream);
}

extern "C" int kk_logmel_f32(const float* x, const float* window, const float* fb, float* out, int B, int n,
                             int n_fft, int hop, int pad_mode, int frames, int n_mels, int power, int log_mode,
                             float floor_, const int* lens, void* stream) {
  kk::FftPlan p;
  int rc = check_stft_args("kk_logmel_f32", x, window, out, B, n, n_fft, hop, pad_mode, frames, &p);
  if (rc) return rc;
  KK_CHECK_ARG(fb && n_mels > 0 && n_mels <= 1024 && (power == 1 || power == 2) && log_mode >= 0 && log_mode <= 3,
               "kk_logmel_f32: bad mel arguments");
  kk::StftArgs a = {};
  a.x = x; a.window = window; a.out = out; a.n = n; a.hop = hop; a.pad_mode = pad_mode; a.frames = frames;
  a.fb = fb; a.n_mels = n_mels; a.power = power; a.log_mode = log_mode; a.floor_ = floor_; a.lens = lens;
  return kk::launch_stft<1>(a, p, B, (cudaStream_t)stream);
}

extern "C" int kk_istft_f32(const float* spec, const float* window, float* out, int B, int frames, int n_fft, int hop,
                            int trim, int out_len, const int* lens, int tail, void* stream) {
  KK_CHECK_ARG(spec && window && out && B > 0 && B <= 65535 && frames > 0 && hop > 0 && trim >= 0 && out_len > 0 && tail >= 0,
               "kk_istft_f32: bad arguments");
  kk::FftPlan p;
  if (kk::make_fft_plan(n_fft, &p) != 0) {
    kk::set_error("kk_istft_f32: n_fft=%d must be even, <= 4096 and factor into 2, 3, 5", n_fft);
    return KK_ERR_UNSUPPORTED;
  }
  kk::IstftArgs a = {};
  a.spec = spec; a.window = window; a.out = out; a.frames = frames; a.hop = hop; a.trim = trim; a.out_len = out_len;
  a.lens = lens; a.tail = tail;
  a.group = kk::stft_group(p.nc) > 4 ? 4 : kk::stft_group(p.nc);
  const int overlap = kk::cdiv(n_fft, hop);
  if (n_fft == 2 * kk::ff::NC && (hop == 256 || hop == 512 || hop == 1024)) {   // register-resident radix-8, streaming OLA
    kk::IstftStreamArgs sa = {};
    sa.spec = spec; sa.window = window; sa.out = out; sa.frames = frames; sa.hop = hop; sa.trim = trim;
    sa.lens = lens; sa.tail = tail;
    sa.out_len = out_len; sa.R = n_fft / hop;
    long long total = (long long)(frames - 1) * hop + n_fft;
    const int hops_total = (int)((total + hop - 1) / hop);
    int fpb = 32;   // multiple of GROUPS; the R - 1 warm-up frames cost (GROUPS / fpb) extra
    while (fpb < 256 && (long long)kk::cdiv(hops_total, fpb * 2) * B >= 148 * 16) fpb *= 2;
    sa.frames_per_block = fpb;
    size_t smem = (size_t)kk::ff::TW * 8 + (size_t)(kk::ff::NC + 2) * 8 + (size_t)2 * kk::ff::GROUPS * kk::ff::BUF * 8 + 16;
    dim3 grid(kk::cdiv(hops_total, fpb), B);
    auto launch = [&](auto kern) -> int {
      KK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      kern<<<grid, kk::ff::THREADS, smem, (cudaStream_t)stream>>>(sa);
      return KK_OK;
    };
    // register-table variant: measured slower here (1.35 -> 1.50 ms: the inverse leans on occupancy to cover its
    // spectrum loads); kept behind KK_IFFT_RT=1 (-DKK_TUNE builds) for A/B
#ifdef KK_TUNE
    static const bool rt = [] { const char* e = getenv("KK_IFFT_RT"); return e && atoi(e) != 0; }();
#else
    constexpr bool rt = false;                               // the product library reads no environment
#endif
    int lrc = sa.R == 4 ? (rt ? launch(kk::istft_stream_kernel<4, true>) : launch(kk::istft_stream_kernel<4, false>))
            : sa.R == 2 ? (rt ? launch(kk::istft_stream_kernel<2, true>) : launch(kk::istft_stream_kernel<2, false>))
                        : (rt ? launch(kk::istft_stream_kernel<1, true>) : launch(kk::istft_stream_kernel<1, false>));
    if (lrc) return lrc;
    KK_LAUNCH_CHECK("istft_stream_kernel");
    return KK_OK;
  }
  size_t fixed = (size_t)p.nc * 8 + (size_t)(p.nc + 1) * 8 + (size_t)n_fft * 4 + (size_t)2 * a.group * p.nc * 8 + 16;
  int hpb = 16;
  while (hpb > 1 && fixed + (size_t)(hpb + overlap) * n_fft * 4 > 160 * 1024) hpb >>= 1;
  a.hops_per_block = hpb;
  a.max_frames = hpb + overlap;
  size_t smem = fixed + (size_t)a.max_frames * n_fft * 4;
  if (smem > 200 * 1024) {
    kk::set_error("kk_istft_f32: hop=%d too small for n_fft=%d (overlap %d frames does not fit shared memory)", hop,
                  n_fft, overlap);
    return KK_ERR_UNSUPPORTED;
  }
  KK_CUDA(cudaFuncSetAttribute(kk::istft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  long long total = (long long)(frames - 1) * hop + n_fft;
  dim3 grid(kk::cdiv(total, (long long)hpb * hop), B);
  kk::istft_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(a, p);
  KK_LAUNCH_CHECK("istft_kernel");
  return KK_OK;
}
