// LAB CODE (not part of the product build; csrc/fft.cu includes it only under -DKK_FFT_R32): measured slower than the
// radix-8 kernels, kept for the record (profiles/r02_fft_experiments.txt 2).
// Single-exchange 512-point complex FFT (= the 1024-point real transform of the general STFT / iSTFT / log-mel kernels,
// mlx_audio/utils.py:52-158): one transform = 16 threads x 32 points, 512 = 32 x 16.
//   forward:  pass 1  thread j holds x[j + 16 t] (t < 32): radix-32 in registers, twiddle W512^(j k1), ONE exchange
//             pass 2  thread m takes the rows k1 = m and 32 - m (thread 0: 0 and 16): two radix-16, giving
//                     Z[k1 + 32 k2] (k2 < 16) - the conjugate partner of every point (512 - k) is in the same thread,
//                     so the real-transform unpack needs no second exchange (thread 0's two rows pair up inside
//                     themselves and are unpacked cooperatively from a 32-entry scratch)
//   inverse:  the mirror image (pack in the thread that holds both partners, two radix-16, exchange, radix-32).
// A warp runs two transforms side by side and synchronises with __syncwarp only.  Against the former 64 threads x 8
// points / three radix-8 passes (two exchanges + a third for the unpack partners) this moves ~100 instead of ~186
// shared-memory wavefronts per frame: ncu had the L1 / shared data pipe as the limiter of all three kernels (80-87 %).
// Exchange layout: row k1 at pitch 17 (Y[k1][j] at k1 * 17 + j): stores (fixed k1, lanes j) are consecutive, loads
// (fixed j, lane m at row m or 32 - m) hit bank (j +- m) mod 16: both conflict free, every address is base + immediate.
#pragma once
#include "../mlx_audio_b200/csrc/fft_fast.cuh"

namespace kk {
namespace fr {
using ff::c2; using ff::pk; using ff::up; using ff::sw; using ff::add2; using ff::sub2; using ff::mul2; using ff::fma2;
using ff::add_mi; using ff::sub_mi; using ff::cm; using ff::fft8; using ff::w32c; using ff::w32s; using ff::cmc;

constexpr int NC = 512;             // complex points
constexpr int LANES = 16;           // threads per transform
constexpr int PITCH = 17;           // exchange row pitch (c2 elements)
constexpr int XBUF = 32 * PITCH;    // exchange buffer of one transform (c2 elements)
#ifndef R32_THREADS
#define R32_THREADS 256
#endif
constexpr int THREADS = R32_THREADS;   // 8 warps = 16 transforms per block iteration
constexpr int SLOTS = THREADS / LANES;

__device__ __forceinline__ c2 mul_mi(c2 a) { const float2 t = up(a); return pk(t.y, -t.x); }   // -i * a
__device__ __forceinline__ c2 conj2(c2 a) { const float2 t = up(a); return pk(t.x, -t.y); }

// forward 4-point DFT, natural order in and out
__device__ __forceinline__ void fft4(c2& a0, c2& a1, c2& a2, c2& a3) {
  const c2 s02 = add2(a0, a2), d02 = sub2(a0, a2), s13 = add2(a1, a3), d13 = sub2(a1, a3);
  a0 = add2(s02, s13); a2 = sub2(s02, s13);
  a1 = add_mi(d02, d13); a3 = sub_mi(d02, d13);
}

// forward 32-point DFT of v[t] (natural order in); out: v[8 ka + kb] = V[ka + 4 kb]   (o32() maps k -> register)
__device__ __forceinline__ void dft32(c2 (&v)[32]) {
#pragma unroll
  for (int tb = 0; tb < 8; ++tb) fft4(v[tb], v[8 + tb], v[16 + tb], v[24 + tb]);
#pragma unroll
  for (int ka = 1; ka < 4; ++ka)
#pragma unroll
    for (int tb = 1; tb < 8; ++tb) {
      const int e = ka * tb;
      v[8 * ka + tb] = (e == 8) ? mul_mi(v[8 * ka + tb]) : cmc(v[8 * ka + tb], w32c(e), -w32s(e));
    }
#pragma unroll
  for (int ka = 0; ka < 4; ++ka) {
    c2 w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = v[8 * ka + i];
    fft8(w);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[8 * ka + i] = w[i];
  }
}
__host__ __device__ constexpr int o32(int k) { return 8 * (k & 3) + (k >> 2); }

// forward 16-point DFT of v[t]; out: v[4 ka + kb] = V[ka + 4 kb]   (o16() maps k -> register)
__device__ __forceinline__ void dft16(c2 (&v)[16]) {
#pragma unroll
  for (int tb = 0; tb < 4; ++tb) fft4(v[tb], v[4 + tb], v[8 + tb], v[12 + tb]);
#pragma unroll
  for (int ka = 1; ka < 4; ++ka)
#pragma unroll
    for (int tb = 1; tb < 4; ++tb) {
      const int e = 2 * ka * tb;       // W16^(ka tb) = W32^(2 ka tb)
      v[4 * ka + tb] = (e == 8) ? mul_mi(v[4 * ka + tb]) : cmc(v[4 * ka + tb], w32c(e), -w32s(e));
    }
#pragma unroll
  for (int ka = 0; ka < 4; ++ka) fft4(v[4 * ka], v[4 * ka + 1], v[4 * ka + 2], v[4 * ka + 3]);
}
__host__ __device__ constexpr int o16(int k) { return 4 * (k & 3) + (k >> 2); }

// Twiddles.  The inter-pass factor W512^(m k1), k1 = ka + 4 kb, is applied as the product of ga[ka] = W512^(m ka) and
// gb[kb] = W512^(4 m kb): 10 values per thread instead of a 31-entry row (a 64-bit shared load costs one wavefront per
// half-warp even when both halves read the same words, and the L1 / shared data pipe is what bounds these kernels).
// They live in a shared table tg[i * 16 + m] (i < 3: ga[i + 1], i >= 3: gb[i - 2]) and are read right where they are
// used: held in registers they cost 20 of a budget of 128 that the 32 points + the next frame's loads need.
// u0 = exp(-2 pi i m / 1024) (registers) is the base of the real-transform factors: u(m + 32 k2) = u0 * W32^k2, the
// second factor a compile-time constant.
constexpr int TG = 10 * LANES;
__device__ __forceinline__ void fill_tables(c2* tg) {
  for (int i = threadIdx.x; i < TG; i += blockDim.x) {
    const int r = i >> 4, m = i & 15;
    float s, c;
    sincospif(-2.0f * (float)(r < 3 ? m * (r + 1) : 4 * m * (r - 2)) / (float)NC, &s, &c);
    tg[i] = pk(c, s);
  }
}
struct Twid {
  c2 u0;
  const c2* tg;      // + m
  __device__ __forceinline__ Twid(const c2* table, int m) : tg(table + m) {
    float s, c;
    sincospif(-(float)m / (float)NC, &s, &c);
    u0 = pk(c, s);
  }
  __device__ __forceinline__ c2 ga(int ka) const { return tg[(ka - 1) * LANES]; }
  __device__ __forceinline__ c2 gb(int kb) const { return tg[(kb + 2) * LANES]; }
  // exp(-2 pi i (m + 32 k2) / 1024) scaled by `scale`
  __device__ __forceinline__ c2 u(int k2, float scale) const {
    return k2 == 0 && scale == 1.f ? u0 : cmc(u0, scale * w32c(k2), -scale * w32s(k2));
  }
};

// forward transform of one frame held as v[t] = z[m + 16 t] by its 16 threads (m = lane & 15): exchange through X,
// result rows A = Z[ka_row + 32 k2] at A[o16(k2)], B likewise for the second row (see top).  Ends with every thread's
// loads from X complete but NOT synchronised: the caller issues __syncwarp() before X is written again.
__device__ __forceinline__ void forward512(c2 (&v)[32], c2* X, const Twid& tw, int m, c2 (&A)[16], c2 (&B)[16]) {
  dft32(v);
#pragma unroll
  for (int k1 = 0; k1 < 32; ++k1) {
    c2 y = v[o32(k1)];
    if (k1 & 3) y = cm(y, tw.ga(k1 & 3));
    if (k1 >> 2) y = cm(y, tw.gb(k1 >> 2));
    X[k1 * PITCH + m] = y;
  }
  __syncwarp();
  const c2* ra = X + m * PITCH;
  const c2* rb = X + (m ? 32 - m : 16) * PITCH;
#pragma unroll
  for (int j = 0; j < 16; ++j) { A[j] = ra[j]; B[j] = rb[j]; }
  dft16(A);
  dft16(B);
}

}  // namespace fr
}  // namespace kk
