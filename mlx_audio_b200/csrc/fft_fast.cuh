// Register-resident 512-point complex FFT (= 1024-point real transform, the n_fft of the reference's 24 kHz mel / Vocos
// front ends: indextts/mel.py:20-37, codec/models/vocos/mel.py:22-33) for the general STFT / log-mel / iSTFT kernels.
// One transform = 64 threads x 8 points: three radix-8 Stockham passes, butterflies in registers, two exchanges
// through shared memory (XOR-swizzled, see swz(): every exchange access pattern is bank-conflict free); a block runs
// four transforms side by side.  The first pass reads global memory directly (windowed samples / packed half spectrum).
#pragma once
#include "common.cuh"

namespace kk {
namespace ff {

constexpr int NC = 512;            // complex points
constexpr int NT = NC / 8;         // threads per transform
constexpr int GROUPS = 4;          // transforms per block
constexpr int THREADS = NT * GROUPS;
constexpr int BUF = NC;            // exchange buffer (float2 elements), swizzled, no padding

// Exchange-buffer swizzle.  A 64-bit shared access is served per half-warp: 16 float2 = the 32 banks once.  Patterns:
//   consecutive runs  i = j + 64 t          (pass-2 / pass-3 loads, pass-3 stores, unpack)  - any XOR of the low 4
//                                             bits by a function of i >> 4 keeps an aligned run a permutation;
//   stride 8          i = 8 j + q           (pass-1 stores): low4 = 8 (j & 1) + q, i >> 4 = j >> 1  -> XOR (i >> 4) & 7
//                                             spreads the eight j >> 1 over the low three bits;
//   two octets        i = 64 h + k + 8 q    (pass-2 stores, h = 0 / 1, k < 8): XOR bit 3 with bit 6 (= h) separates them.
// (The former i + (i >> 3) padding left a 2-way conflict on every consecutive access: 16 elements span 17 slots - ncu
// counted 3.0 wavefronts per 64-bit load request instead of 2.)
__device__ __forceinline__ int pad(int i) { return i ^ (((i >> 4) & 7) | ((i >> 3) & 8)); }

// The swizzled addresses of one thread (j < 64) in closed form, so that an access costs an immediate offset or one
// logic op instead of the five of pad() (address arithmetic was ~30 % of the issued instructions):
//   cons(t)   = pad(j + 64 t)                        consecutive runs
//   st1(q)    = pad(8 j + q)                          pass-1 stores
//   st2(q)    = pad(64 (j >> 3) + (j & 7) + 8 q)      pass-2 stores
//   mirror(q) = pad((512 - (j + 64 q)) & 511)         the conjugate-symmetric partner of point j + 64 q
struct SwzIdx {
  int c0, c1, p1, m7, e, o, kk, d0, d1, zoff;
  __device__ __forceinline__ explicit SwzIdx(int j) {
    c0 = j ^ (j >> 4); c1 = c0 ^ 12;
    p1 = (8 * j) ^ (j & 8); m7 = (j >> 1) & 7;
    const int h = j >> 3, hb = h & 1;
    kk = (j & 7) ^ (4 * hb); e = 64 * h + 8 * hb; o = 64 * h + 8 - 8 * hb;
    const int jm = (64 - j) & 63, m0 = jm ^ (jm >> 4), m1 = m0 ^ 12;
    zoff = j == 0 ? 64 : 0;
    d0 = j == 0 ? m1 : m0; d1 = j == 0 ? m0 : m1;
  }
  __device__ __forceinline__ int cons(int t) const { return 64 * t + ((t & 1) ? c1 : c0); }
  __device__ __forceinline__ int st1(int q) const { return p1 + (q ^ m7); }
  __device__ __forceinline__ int st2(int q) const { return ((q & 1) ? o : e) + 16 * (q >> 1) + (kk ^ (q >> 1)); }
  __device__ __forceinline__ int mirror(int q) const { return ((64 * (7 - q) + zoff) & 511) + (((7 - q) & 1) ? d1 : d0); }
};
// Packed arithmetic: a complex value is one 64-bit register pair and every complex add / subtract / scale is ONE
// FADD2 / FFMA2 (sm_100 f32x2: both lanes IEEE round-to-nearest, so the bits equal the scalar forms').  ptxas folds
// the lane swap, a per-lane negation and a scalar broadcast into operand modifiers (R.F32x2.LO_HI, .NP, R.F32), so a
// multiply by -i costs nothing and a complex product is FMUL2 + FFMA2.  These kernels are issue bound (ncu: 70 % of
// the issue slots, 25 % occupancy): the scalar forms spent 2-4 slots where these spend 1-2.
typedef unsigned long long c2;
__device__ __forceinline__ c2 pk(float lo, float hi) { c2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ c2 pk(float2 v) { return pk(v.x, v.y); }
__device__ __forceinline__ float2 up(c2 v) { float2 r; asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v)); return r; }
__device__ __forceinline__ c2 sw(c2 v) { const float2 t = up(v); return pk(t.y, t.x); }
__device__ __forceinline__ c2 add2(c2 a, c2 b) { c2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c2 sub2(c2 a, c2 b) { c2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c2 mul2(c2 a, c2 b) { c2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ c2 fma2(c2 a, c2 b, c2 c) { c2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
// a + (-i) d  and  a - (-i) d  (d's lanes swapped, signs from the constant: the products by +-1 are exact)
__device__ __forceinline__ c2 add_mi(c2 a, c2 d) { return fma2(sw(d), pk(1.f, -1.f), a); }
__device__ __forceinline__ c2 sub_mi(c2 a, c2 d) { return fma2(sw(d), pk(-1.f, 1.f), a); }
// complex product with the contraction spelled out, so that every kernel variant (shared-memory or register tables,
// any batch size) produces the same bits for the same frame:
//   (fma(a.x, b.x, -(a.y * b.y)), fma(a.x, b.y, a.y * b.x))
__device__ __forceinline__ c2 cm(c2 a, c2 b) {
  const float2 af = up(a);
  float2 t = up(mul2(pk(af.y, af.y), sw(b)));
  t.x = -t.x;
  return fma2(pk(af.x, af.x), b, pk(t));
}
__device__ __forceinline__ float2 cm(float2 a, float2 b) { return up(cm(pk(a), pk(b))); }
__device__ __forceinline__ float2 ca(float2 a, float2 b) { return up(add2(pk(a), pk(b))); }
__device__ __forceinline__ float2 cs(float2 a, float2 b) { return up(sub2(pk(a), pk(b))); }

__host__ __device__ constexpr float w32c(int e) {   // cos(2 pi e / 32), correctly rounded
  switch (e & 31) {
    case 0: return 0x1.0000000000000p+0f;   // 1
    case 1: return 0x1.f6297c0000000p-1f;   // 0.980785251
    case 2: return 0x1.d906bc0000000p-1f;   // 0.923879504
    case 3: return 0x1.a9b6620000000p-1f;   // 0.831469595
    case 4: return 0x1.6a09e60000000p-1f;   // 0.707106769
    case 5: return 0x1.1c73b40000000p-1f;   // 0.555570245
    case 6: return 0x1.87de2a0000000p-2f;   // 0.382683426
    case 7: return 0x1.8f8b840000000p-3f;   // 0.195090324
    case 8: return 0x0.0p+0f;   // 0
    case 9: return -0x1.8f8b840000000p-3f;   // -0.195090324
    case 10: return -0x1.87de2a0000000p-2f;   // -0.382683426
    case 11: return -0x1.1c73b40000000p-1f;   // -0.555570245
    case 12: return -0x1.6a09e60000000p-1f;   // -0.707106769
    case 13: return -0x1.a9b6620000000p-1f;   // -0.831469595
    case 14: return -0x1.d906bc0000000p-1f;   // -0.923879504
    case 15: return -0x1.f6297c0000000p-1f;   // -0.980785251
    case 16: return -0x1.0000000000000p+0f;   // -1
    case 17: return -0x1.f6297c0000000p-1f;   // -0.980785251
    case 18: return -0x1.d906bc0000000p-1f;   // -0.923879504
    case 19: return -0x1.a9b6620000000p-1f;   // -0.831469595
    case 20: return -0x1.6a09e60000000p-1f;   // -0.707106769
    case 21: return -0x1.1c73b40000000p-1f;   // -0.555570245
    case 22: return -0x1.87de2a0000000p-2f;   // -0.382683426
    case 23: return -0x1.8f8b840000000p-3f;   // -0.195090324
    case 24: return 0x0.0p+0f;   // 0
    case 25: return 0x1.8f8b840000000p-3f;   // 0.195090324
    case 26: return 0x1.87de2a0000000p-2f;   // 0.382683426
    case 27: return 0x1.1c73b40000000p-1f;   // 0.555570245
    case 28: return 0x1.6a09e60000000p-1f;   // 0.707106769
    case 29: return 0x1.a9b6620000000p-1f;   // 0.831469595
    case 30: return 0x1.d906bc0000000p-1f;   // 0.923879504
    case 31: return 0x1.f6297c0000000p-1f;   // 0.980785251
  }
  return 0.f;
}
__host__ __device__ constexpr float w32s(int e) {   // sin(2 pi e / 32), correctly rounded
  switch (e & 31) {
    case 0: return 0x0.0p+0f;   // 0
    case 1: return 0x1.8f8b840000000p-3f;   // 0.195090324
    case 2: return 0x1.87de2a0000000p-2f;   // 0.382683426
    case 3: return 0x1.1c73b40000000p-1f;   // 0.555570245
    case 4: return 0x1.6a09e60000000p-1f;   // 0.707106769
    case 5: return 0x1.a9b6620000000p-1f;   // 0.831469595
    case 6: return 0x1.d906bc0000000p-1f;   // 0.923879504
    case 7: return 0x1.f6297c0000000p-1f;   // 0.980785251
    case 8: return 0x1.0000000000000p+0f;   // 1
    case 9: return 0x1.f6297c0000000p-1f;   // 0.980785251
    case 10: return 0x1.d906bc0000000p-1f;   // 0.923879504
    case 11: return 0x1.a9b6620000000p-1f;   // 0.831469595
    case 12: return 0x1.6a09e60000000p-1f;   // 0.707106769
    case 13: return 0x1.1c73b40000000p-1f;   // 0.555570245
    case 14: return 0x1.87de2a0000000p-2f;   // 0.382683426
    case 15: return 0x1.8f8b840000000p-3f;   // 0.195090324
    case 16: return 0x0.0p+0f;   // 0
    case 17: return -0x1.8f8b840000000p-3f;   // -0.195090324
    case 18: return -0x1.87de2a0000000p-2f;   // -0.382683426
    case 19: return -0x1.1c73b40000000p-1f;   // -0.555570245
    case 20: return -0x1.6a09e60000000p-1f;   // -0.707106769
    case 21: return -0x1.a9b6620000000p-1f;   // -0.831469595
    case 22: return -0x1.d906bc0000000p-1f;   // -0.923879504
    case 23: return -0x1.f6297c0000000p-1f;   // -0.980785251
    case 24: return -0x1.0000000000000p+0f;   // -1
    case 25: return -0x1.f6297c0000000p-1f;   // -0.980785251
    case 26: return -0x1.d906bc0000000p-1f;   // -0.923879504
    case 27: return -0x1.a9b6620000000p-1f;   // -0.831469595
    case 28: return -0x1.6a09e60000000p-1f;   // -0.707106769
    case 29: return -0x1.1c73b40000000p-1f;   // -0.555570245
    case 30: return -0x1.87de2a0000000p-2f;   // -0.382683426
    case 31: return -0x1.8f8b840000000p-3f;   // -0.195090324
  }
  return 0.f;
}

// a * (c + i s), contraction as ff::cm
__device__ __forceinline__ c2 cmc(c2 a, float c, float s) {
  const float2 af = up(a);
  return fma2(pk(af.x, af.x), pk(c, s), mul2(pk(af.y, af.y), pk(-s, c)));
}
// forward 8-point DFT (e^{-2 pi i t q / 8}), natural order in and out: 26 packed instructions
__device__ __forceinline__ void fft8(c2 (&v)[8]) {
  const float h = 0.70710678118654752440f;
  const c2 a0 = add2(v[0], v[4]), a4 = sub2(v[0], v[4]);
  const c2 a2 = add2(v[2], v[6]), d26 = sub2(v[2], v[6]);
  const c2 a1 = add2(v[1], v[5]), a5 = sub2(v[1], v[5]);
  const c2 a3 = add2(v[3], v[7]), d37 = sub2(v[3], v[7]);
  const c2 b0 = add2(a0, a2), b2 = sub2(a0, a2), b4 = add_mi(a4, d26), b6 = sub_mi(a4, d26);
  const c2 b1 = add2(a1, a3), d13 = sub2(a1, a3);
  const c2 s5 = add_mi(a5, d37), d7 = sub_mi(a5, d37);
  // b5 = s5 (1 - i)/sqrt 2 and b7 = d7 (-1 - i)/sqrt 2 folded into the last butterflies as explicit FMAs:
  // pq = (s5.x + s5.y, s5.y - s5.x), ru = (d7.y - d7.x, d7.x + d7.y)
  const c2 pq = fma2(sw(s5), pk(1.f, -1.f), s5);
  const c2 ru = fma2(d7, pk(-1.f, 1.f), sw(d7));
  v[0] = add2(b0, b1); v[4] = sub2(b0, b1);
  v[2] = add_mi(b2, d13); v[6] = sub_mi(b2, d13);
  v[1] = fma2(pk(h, h), pq, b4); v[5] = fma2(pk(-h, -h), pq, b4);
  v[3] = fma2(pk(h, -h), ru, b6); v[7] = fma2(pk(-h, h), ru, b6);
}

// Twiddle tables laid out per pass so that a warp reads consecutive addresses (a single exp(-2 pi i j/512) table
// indexed t*k*8 / t*j puts 8 distinct addresses on one bank):
//   tw[t*8 + k]        = exp(-2 pi i t k / 64)    pass 2 (t, k < 8)           - 64 entries
//   tw[64 + t*64 + j]  = exp(-2 pi i t j / 512)   pass 3 (t < 8, j < 64)      - 512 entries
//   tw2[k]             = exp(-2 pi i k / 1024)    real-transform unpack, k <= 512
constexpr int TW = 64 + NC;
__device__ __forceinline__ void fill_tables(float2* tw, float2* tw2) {
  for (int i = threadIdx.x; i < TW; i += blockDim.x) {
    float s, c;
    if (i < 64) sincospif(-2.0f * (float)((i >> 3) * (i & 7)) / 64.0f, &s, &c);
    else sincospif(-2.0f * (float)(((i - 64) >> 6) * ((i - 64) & 63)) / (float)NC, &s, &c);
    tw[i] = make_float2(c, s);
  }
  for (int j = threadIdx.x; j <= NC; j += blockDim.x) {
    float s, c;
    sincospif(-(float)j / (float)NC, &s, &c);
    tw2[j] = make_float2(c, s);
  }
}

// The seven twiddles w^t (t = 1..7) of a radix-8 pass from the three table values w, w^2, w^4: four complex products
// instead of four more 64-bit shared loads per pass (the L1 / shared data pipe bounds these kernels).  Used (POW) by the
// kernels whose output is not compared bit for bit with a register-table variant: the inverse and the log-mel.
__device__ __forceinline__ void tw_powers(c2 w1, c2 w2, c2 w4, c2 (&w)[8]) {
  w[1] = w1; w[2] = w2; w[4] = w4;
  w[3] = cm(w1, w2); w[5] = cm(w1, w4); w[6] = cm(w2, w4); w[7] = cm(w[3], w4);
}

// pass 1 (ns = 1): v holds points j + 64 t; result to A
__device__ __forceinline__ void pass1_store(c2 (&v)[8], c2* A, const SwzIdx& z) {
  fft8(v);
#pragma unroll
  for (int q = 0; q < 8; ++q) A[z.st1(q)] = v[q];
}
// pass 2 (ns = 8): A -> B
template <bool POW = false>
__device__ __forceinline__ void pass2(const c2* A, c2* B, const c2* tw, int j, const SwzIdx& z) {
  c2 v[8], w[8];
  const int k = j & 7;
  if (POW) tw_powers(tw[8 + k], tw[16 + k], tw[32 + k], w);
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    v[t] = A[z.cons(t)];
    if (t) v[t] = cm(v[t], POW ? w[t] : tw[t * 8 + k]);
  }
  fft8(v);
#pragma unroll
  for (int q = 0; q < 8; ++q) B[z.st2(q)] = v[q];
}
// pass 3 (ns = 64): B -> registers; v[q] = point j + 64 q of the transform
template <bool POW = false>
__device__ __forceinline__ void pass3(const c2* B, const c2* tw, int j, const SwzIdx& z, c2 (&v)[8]) {
  c2 w[8];
  if (POW) tw_powers(tw[64 + 64 + j], tw[64 + 128 + j], tw[64 + 256 + j], w);
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    v[t] = B[z.cons(t)];
    if (t) v[t] = cm(v[t], POW ? w[t] : tw[64 + t * 64 + j]);
  }
  fft8(v);
}


// Register-table variants: the twiddles a thread needs are the same for every frame (they depend on j only), so a
// kernel that can afford ~60 more registers keeps them out of shared memory altogether (30 of the 78 64-bit shared
// requests per thread and frame are table reads).  Same sincospif arguments as fill_tables: bit-identical values.
__device__ __forceinline__ void fill_reg_twiddles(int j, c2 (&t2)[8], c2 (&t3)[8]) {
  const int k = j & 7;
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    float s, c;
    sincospif(-2.0f * (float)(t * k) / 64.0f, &s, &c);
    t2[t] = pk(c, s);
    sincospif(-2.0f * (float)(t * j) / (float)NC, &s, &c);
    t3[t] = pk(c, s);
  }
}
__device__ __forceinline__ void pass2_r(const c2* A, c2* B, const c2 (&t2)[8], const SwzIdx& z) {
  c2 v[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    v[t] = A[z.cons(t)];
    if (t) v[t] = cm(v[t], t2[t]);
  }
  fft8(v);
#pragma unroll
  for (int q = 0; q < 8; ++q) B[z.st2(q)] = v[q];
}
__device__ __forceinline__ void pass3_r(const c2* B, const c2 (&t3)[8], const SwzIdx& z, c2 (&v)[8]) {
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    v[t] = B[z.cons(t)];
    if (t) v[t] = cm(v[t], t3[t]);
  }
  fft8(v);
}

}  // namespace ff
}  // namespace kk
