// Coalesced elementwise / reduction kernels: AdaIN style projection, instance-norm statistics,
// AdaIN-apply + Snake/LeakyReLU (+ depthwise transposed-conv pool), decoder front glue.
#include "common.cuh"

namespace kk {

// ---------------------------------------------------------------- style projection (AdaIN1d.fc)
// h (B, R) = s (B, S) W^T (R, S) + bias: block = 32 output rows x 32 style vectors, both staged in shared memory;
// warp = 4 rows, lane = style vector (no cross-lane reduction; W is read as broadcast float4).
constexpr int SA_ROWS = 32;
__global__ void __launch_bounds__(256)
style_affine_kernel(const float* __restrict__ s, const float* __restrict__ W, const float* __restrict__ bias,
                    float* __restrict__ h, int B, int R, int S) {
  extern __shared__ __align__(16) float sh[];
  float* sh_w = sh;                       // [SA_ROWS][S]
  float* sh_s = sh + SA_ROWS * S;         // [S][33] (transposed, padded)
  const int b0 = blockIdx.y * 32, r0 = blockIdx.x * SA_ROWS;
  const int nb = min(32, B - b0), nr = min(SA_ROWS, R - r0);
  for (int i = threadIdx.x; i < nr * S; i += blockDim.x) sh_w[i] = W[(size_t)r0 * S + i];
  for (int i = threadIdx.x; i < 32 * S; i += blockDim.x) {
    const int b = i / S, k = i - b * S;
    sh_s[k * 33 + b] = b < nb ? s[(size_t)(b0 + b) * S + k] : 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rw = warp * 4;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k = 0; k + 4 <= S; k += 4) {
    const float s0 = sh_s[(k + 0) * 33 + lane], s1 = sh_s[(k + 1) * 33 + lane];
    const float s2 = sh_s[(k + 2) * 33 + lane], s3 = sh_s[(k + 3) * 33 + lane];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 w = *reinterpret_cast<const float4*>(sh_w + (rw + i) * S + k);
      acc[i] = fmaf(s0, w.x, fmaf(s1, w.y, fmaf(s2, w.z, fmaf(s3, w.w, acc[i]))));
    }
  }
  for (int k = S & ~3; k < S; ++k) {
    const float sv = sh_s[k * 33 + lane];
#pragma unroll
    for (int i = 0; i < 4; ++i) acc[i] = fmaf(sv, sh_w[(rw + i) * S + k], acc[i]);
  }
  if (lane < nb) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = r0 + rw + i;
      if (r < R) h[(size_t)(b0 + lane) * R + r] = acc[i] + (bias ? bias[r] : 0.f);
    }
  }
}

// ---------------------------------------------------------------- instance-norm statistics
// block (32 channels, 8 row lanes); rows [l0, l0+rows_per_block).  fp32 partials -> double atomics.
__global__ void instnorm_stats_kernel(const float* __restrict__ x, double* __restrict__ stats, int L, int C,
                                      int pitch, int rows_per_block, const int* __restrict__ lens, int len_mul,
                                      int len_add) {
  int b = blockIdx.z;
  int c = blockIdx.x * 32 + threadIdx.x;
  int l0 = blockIdx.y * rows_per_block;
  int l1 = min(ragged_len(lens, len_mul, len_add, b, L), l0 + rows_per_block);   // padding rows are not statistics
  // fp64 partial sums: the row split (rows_per_block) depends on the batch size, and fp32 partials would make an
  // utterance's statistics - hence, after bf16 rounding downstream, its audio - depend on what it is batched with
  double s = 0.0, ss = 0.0;
  if (c < C) {
    const float* xb = x + (size_t)b * L * pitch + c;
    for (int l = l0 + threadIdx.y; l < l1; l += 8) {
      const double v = xb[(size_t)l * pitch];
      s += v;
      ss = fma(v, v, ss);
    }
  }
  __shared__ double sh[2][8][33];
  sh[0][threadIdx.y][threadIdx.x] = s;
  sh[1][threadIdx.y][threadIdx.x] = ss;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    double ds = 0.0, dss = 0.0;
    for (int i = 0; i < 8; ++i) { ds += sh[0][i][threadIdx.x]; dss += sh[1][i][threadIdx.x]; }
    atomicAdd(stats + ((size_t)b * C + c) * 2, ds);
    atomicAdd(stats + ((size_t)b * C + c) * 2 + 1, dss);
  }
}

// ---------------------------------------------------------------- AdaIN apply + activation
template <typename OutT>
__device__ __forceinline__ void store_out(OutT* p, float v);
template <>
__device__ __forceinline__ void store_out<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void store_out<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

struct AdaCoef { float mean, rstd, g1, beta, alpha, inv_alpha; };

__device__ __forceinline__ float ada_act(float x, const AdaCoef& k, bool norm, int act, float slope) {
  float y = x;
  if (norm) y = k.g1 * ((x - k.mean) * k.rstd) + k.beta;      // (1+gamma)*xn + beta   istftnet.py:337
  if (act == KK_ACT_SNAKE) {                                    // x + (1/a) sin(a x)^2  istftnet.py:382
    float t = sinf(k.alpha * y);
    y = y + k.inv_alpha * (t * t);
  } else if (act == KK_ACT_SNAKE_FAST) {                        // same, MUFU sine (bf16 pipelines)
    float t = __sinf(k.alpha * y);
    y = y + k.inv_alpha * (t * t);
  } else if (act == KK_ACT_LEAKY) {
    y = y > 0.f ? y : y * slope;
  }
  return y;
}

template <typename OutT, bool POOL>
__global__ void adain_act_kernel(const float* __restrict__ x, int x_pitch, const double* __restrict__ stats,
                                 const float* __restrict__ h, int h_stride, const float* __restrict__ alpha, int act,
                                 float slope, OutT* __restrict__ out, int out_pitch, int L, int C,
                                 const float* __restrict__ pool_w, const float* __restrict__ pool_b,
                                 int rows_per_block, const int* __restrict__ lens, int len_mul, int len_add) {
  int b = blockIdx.z;
  int c = blockIdx.x * 32 + threadIdx.x;
  if (c >= C) return;
  const int Lb = ragged_len(lens, len_mul, len_add, b, L);   // rows of this utterance; [Lb, L) is padding -> zeros
  AdaCoef k = {0.f, 1.f, 1.f, 0.f, 1.f, 1.f};
  bool norm = stats != nullptr;
  if (norm) {
    double s = stats[((size_t)b * C + c) * 2], ss = stats[((size_t)b * C + c) * 2 + 1];
    double mean = s / Lb;
    double var = ss / Lb - mean * mean;
    if (var < 0.0) var = 0.0;
    k.mean = (float)mean;
    k.rstd = 1.0f / sqrtf((float)var + 1e-5f);
    k.g1 = 1.0f + h[(size_t)b * h_stride + c];
    k.beta = h[(size_t)b * h_stride + C + c];
  }
  if (act == KK_ACT_SNAKE || act == KK_ACT_SNAKE_FAST) { k.alpha = alpha[c]; k.inv_alpha = 1.0f / k.alpha; }
  int l0 = blockIdx.y * rows_per_block;
  int l1 = min(L, l0 + rows_per_block);
  const float* xb = x + (size_t)b * L * x_pitch + c;
  if (!POOL) {
    OutT* ob = out + (size_t)b * L * out_pitch + c;
    for (int l = l0 + threadIdx.y; l < l1; l += 8)
      store_out<OutT>(ob + (size_t)l * out_pitch, l < Lb ? ada_act(xb[(size_t)l * x_pitch], k, norm, act, slope) : 0.f);
  } else {
    // depthwise ConvTranspose1d k3 s2 p1 + bias, then left zero pad (istftnet.py:880-881):
    // P[0]=0, P[2i+1] = a[i] w1 + b, P[2i+2] = a[i] w2 + a[i+1] w0 + b   (i+1 < L)
    float w0 = pool_w[c * 3 + 0], w1 = pool_w[c * 3 + 1], w2 = pool_w[c * 3 + 2], pb = pool_b[c];
    OutT* ob = out + (size_t)b * (2 * L) * out_pitch + c;
    for (int l = l0 + threadIdx.y; l < l1; l += 8) {
      if (l == 0) store_out<OutT>(ob, 0.f);
      if (l >= Lb) {   // padding rows of a ragged batch: the convolution that reads this tensor expects zeros
        store_out<OutT>(ob + (size_t)(2 * l + 1) * out_pitch, 0.f);
        if (l + 1 < L) store_out<OutT>(ob + (size_t)(2 * l + 2) * out_pitch, 0.f);
        continue;
      }
      float a0 = ada_act(xb[(size_t)l * x_pitch], k, norm, act, slope);
      store_out<OutT>(ob + (size_t)(2 * l + 1) * out_pitch, a0 * w1 + pb);
      if (l + 1 < Lb) {
        float a1 = ada_act(xb[(size_t)(l + 1) * x_pitch], k, norm, act, slope);
        store_out<OutT>(ob + (size_t)(2 * l + 2) * out_pitch, a0 * w2 + a1 * w0 + pb);
      } else if (l + 1 < L) {
        store_out<OutT>(ob + (size_t)(2 * l + 2) * out_pitch, 0.f);   // row 2 Lb: first padding row
      }
    }
  }
}


// Vectorised variant: one thread owns 4 consecutive channels (float4 loads, 8/16-byte stores); a warp covers
// 128 channels = one or more whole 512-byte row segments.  Needs C, pitches % 4 == 0 and aligned bases.
template <typename OutT>
__global__ void __launch_bounds__(256)
adain_act_vec4_kernel(const float* __restrict__ x, int x_pitch, const double* __restrict__ stats,
                      const float* __restrict__ h, int h_stride, const float* __restrict__ alpha, int act, float slope,
                      OutT* __restrict__ out, int out_pitch, int L, int C, int rows_per_block,
                      const int* __restrict__ lens, int len_mul, int len_add) {
  const int b = blockIdx.z;
  const int c = (blockIdx.x * 32 + threadIdx.x) * 4;
  if (c >= C) return;
  const int Lb = ragged_len(lens, len_mul, len_add, b, L);
  AdaCoef k[4];
  const bool norm = stats != nullptr;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    k[i] = AdaCoef{0.f, 1.f, 1.f, 0.f, 1.f, 1.f};
    if (norm) {
      double s = stats[((size_t)b * C + c + i) * 2], ss = stats[((size_t)b * C + c + i) * 2 + 1];
      double mean = s / Lb;
      double var = ss / Lb - mean * mean;
      if (var < 0.0) var = 0.0;
      k[i].mean = (float)mean;
      k[i].rstd = 1.0f / sqrtf((float)var + 1e-5f);
      k[i].g1 = 1.0f + h[(size_t)b * h_stride + c + i];
      k[i].beta = h[(size_t)b * h_stride + C + c + i];
    }
    if (act == KK_ACT_SNAKE || act == KK_ACT_SNAKE_FAST) { k[i].alpha = alpha[c + i]; k[i].inv_alpha = 1.0f / k[i].alpha; }
  }
  const int l0 = blockIdx.y * rows_per_block;
  const int l1 = min(L, l0 + rows_per_block);
  const float* xb = x + (size_t)b * L * x_pitch + c;
  OutT* ob = out + (size_t)b * L * out_pitch + c;
#pragma unroll 4
  for (int l = l0 + threadIdx.y; l < l1; l += 8) {
    float4 v = __ldg(reinterpret_cast<const float4*>(xb + (size_t)l * x_pitch));
    float y0 = ada_act(v.x, k[0], norm, act, slope), y1 = ada_act(v.y, k[1], norm, act, slope);
    float y2 = ada_act(v.z, k[2], norm, act, slope), y3 = ada_act(v.w, k[3], norm, act, slope);
    if (l >= Lb) y0 = y1 = y2 = y3 = 0.f;   // padding rows of a ragged batch read as the next conv's zero padding
    if (sizeof(OutT) == 4) {
      *reinterpret_cast<float4*>(ob + (size_t)l * out_pitch) = make_float4(y0, y1, y2, y3);
    } else {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(y0, y1), p1 = __floats2bfloat162_rn(y2, y3);
      uint2 q;
      q.x = *reinterpret_cast<uint32_t*>(&p0);
      q.y = *reinterpret_cast<uint32_t*>(&p1);
      *reinterpret_cast<uint2*>(ob + (size_t)l * out_pitch) = q;
    }
  }
}

// ---------------------------------------------------------------- decoder front glue
// asr (B, C, T) channels-first -> out[b][t][off + c]  (32x32 shared-memory transpose)
__global__ void transpose_cf_to_cl_kernel(const float* __restrict__ src, float* __restrict__ dst, int C, int T,
                                          int dst_pitch, int dst_off) {
  __shared__ float tile[32][33];
  int b = blockIdx.z;
  int t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    int c = c0 + i, t = t0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && t < T) ? src[((size_t)b * C + c) * T + t] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    int t = t0 + i, c = c0 + threadIdx.x;
    if (t < T && c < C) dst[((size_t)b * T + t) * dst_pitch + dst_off + c] = tile[threadIdx.x][i];
  }
}

// F0_conv / N_conv: ConvWeighted(1,1,k3,stride 2,padding 1) on a (B, 2T) curve (istftnet.py:933-934, :949-950)
__global__ void f0n_conv_kernel(const float* __restrict__ f0, const float* __restrict__ nn,
                                const float* __restrict__ wf, const float* __restrict__ wn, float* __restrict__ out,
                                int out_pitch, int f0_off, int n_off, int B, int T, const int* __restrict__ lens) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * T) return;
  int b = idx / T, t = idx - b * T;
  const float* f = f0 + (size_t)b * 2 * T;
  const float* g = nn + (size_t)b * 2 * T;
  const int n2 = 2 * ragged_len(lens, 1, 0, b, T);          // curve samples of this utterance; beyond: zero padding
  auto at = [&](const float* p, int j) { return (j >= 0 && j < n2) ? p[j] : 0.f; };
  float vf = at(f, 2 * t - 1) * wf[0] + at(f, 2 * t) * wf[1] + at(f, 2 * t + 1) * wf[2] + wf[3];
  float vn = at(g, 2 * t - 1) * wn[0] + at(g, 2 * t) * wn[1] + at(g, 2 * t + 1) * wn[2] + wn[3];
  out[(size_t)idx * out_pitch + f0_off] = vf;
  out[(size_t)idx * out_pitch + n_off] = vn;
}

__global__ void copy_channels_kernel(const float* __restrict__ src, int src_pitch, int src_off,
                                     float* __restrict__ dst, int dst_pitch, int dst_off, long long rows, int C) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * C) return;
  long long r = idx / C;
  int c = (int)(idx - r * C);
  dst[r * dst_pitch + dst_off + c] = src[r * src_pitch + src_off + c];
}

template <typename OutT>
__global__ void convert_act_kernel(const float* __restrict__ x, int x_pitch, OutT* __restrict__ out, int out_pitch,
                                   long long rows, int C, float slope) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= rows * C) return;
  long long r = idx / C;
  int c = (int)(idx - r * C);
  float v = x[r * x_pitch + c];
  store_out<OutT>(out + r * out_pitch + c, v > 0.f ? v : v * slope);
}

// Same, four channels per thread (16-byte loads, 8- or 16-byte stores): block = 64 channel quads x 4 rows, no division.
template <typename OutT>
__global__ void __launch_bounds__(256) convert_act_vec4_kernel(const float* __restrict__ x, int x_pitch,
                                                               OutT* __restrict__ out, int out_pitch, long long rows,
                                                               int C, float slope) {
  const int c = 4 * (blockIdx.y * 64 + threadIdx.x);
  const long long r = (long long)blockIdx.x * 4 + threadIdx.y;
  if (c >= C || r >= rows) return;
  const float* xr = x + r * x_pitch + c;
  OutT* orow = out + r * out_pitch + c;
  if (c + 4 <= C) {
    const float4 v = *reinterpret_cast<const float4*>(xr);
    const float a0 = v.x > 0.f ? v.x : v.x * slope, a1 = v.y > 0.f ? v.y : v.y * slope;
    const float a2 = v.z > 0.f ? v.z : v.z * slope, a3 = v.w > 0.f ? v.w : v.w * slope;
    if constexpr (sizeof(OutT) == 2) {
      const __nv_bfloat162 lo = __floats2bfloat162_rn(a0, a1), hi = __floats2bfloat162_rn(a2, a3);
      uint2 pk;
      pk.x = *reinterpret_cast<const uint32_t*>(&lo);
      pk.y = *reinterpret_cast<const uint32_t*>(&hi);
      *reinterpret_cast<uint2*>(orow) = pk;
    } else {
      *reinterpret_cast<float4*>(orow) = make_float4(a0, a1, a2, a3);
    }
  } else {
    for (int i = 0; c + i < C; ++i) { const float v = xr[i]; store_out<OutT>(orow + i, v > 0.f ? v : v * slope); }
  }
}

}  // namespace kk

// ================================================================== C ABI
extern "C" int kk_style_affine(const float* s, const float* W, const float* bias, float* h, int B, int R, int S,
                               void* stream) {
  KK_CHECK_ARG(s && W && h && B > 0 && R > 0 && S > 0 && S <= 384, "kk_style_affine: bad arguments B=%d R=%d S=%d", B, R, S);
  KK_CHECK_ARG(S % 4 == 0 && ((uintptr_t)W & 15) == 0, "kk_style_affine: S must be a multiple of 4 and W 16-byte aligned");
  dim3 grid(kk::cdiv(R, kk::SA_ROWS), kk::cdiv(B, 32));
  size_t smem = ((size_t)kk::SA_ROWS * S + (size_t)S * 33) * sizeof(float);
  kk::style_affine_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(s, W, bias, h, B, R, S);
  KK_LAUNCH_CHECK("style_affine_kernel");
  return KK_OK;
}

static int rows_per_block_for(int L, int B, int cblocks) {
  // aim for >= ~4 waves of 148 SMs without making blocks tiny
  int rpb = 512;
  while (rpb > 64 && (long long)kk::cdiv(L, rpb) * B * cblocks < 148 * 8) rpb >>= 1;
  return rpb;
}

extern "C" int kk_instnorm_stats(const float* x, double* stats, int B, int L, int C, int pitch, const int* lens,
                                 int len_mul, int len_add, void* stream) {
  KK_CHECK_ARG(x && stats && B > 0 && L > 0 && C > 0 && pitch >= C, "kk_instnorm_stats: bad arguments");
  int cb = kk::cdiv(C, 32);
  int rpb = rows_per_block_for(L, B, cb);
  dim3 grid(cb, kk::cdiv(L, rpb), B), block(32, 8);
  kk::instnorm_stats_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(x, stats, L, C, pitch, rpb, lens, len_mul, len_add);
  KK_LAUNCH_CHECK("instnorm_stats_kernel");
  return KK_OK;
}

extern "C" int kk_adain_act(const float* x, int x_pitch, const double* stats, const float* h, int h_stride,
                            const float* alpha, int act, float slope, void* out, int out_is_bf16, int out_pitch,
                            int B, int L, int C, int pool, const float* pool_w, const float* pool_b, const int* lens,
                            int len_mul, int len_add, void* stream) {
  KK_CHECK_ARG(x && out && B > 0 && L > 0 && C > 0 && x_pitch >= C && out_pitch >= C, "kk_adain_act: bad shape");
  KK_CHECK_ARG(act >= KK_ACT_NONE && act <= KK_ACT_SNAKE_FAST, "kk_adain_act: act=%d", act);
  KK_CHECK_ARG((act != KK_ACT_SNAKE && act != KK_ACT_SNAKE_FAST) || alpha, "kk_adain_act: snake needs alpha");
  KK_CHECK_ARG(!stats || h, "kk_adain_act: normalisation needs the style projection h");
  KK_CHECK_ARG(!pool || (pool_w && pool_b), "kk_adain_act: pool needs weights");
  int cb = kk::cdiv(C, 32);
  int rpb = rows_per_block_for(L, B, cb);
  dim3 grid(cb, kk::cdiv(L, rpb), B), block(32, 8);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t osz = out_is_bf16 ? 2 : 4;
  if (!pool && C % 4 == 0 && x_pitch % 4 == 0 && out_pitch % 4 == 0 && ((uintptr_t)x & 15) == 0 &&
      ((uintptr_t)out & (4 * osz - 1)) == 0) {
    int cb4 = kk::cdiv(C, 128);
    int rpb4 = rows_per_block_for(L, B, cb4);
    dim3 grid4(cb4, kk::cdiv(L, rpb4), B);
    if (out_is_bf16)
      kk::adain_act_vec4_kernel<__nv_bfloat16><<<grid4, block, 0, st>>>(x, x_pitch, stats, h, h_stride, alpha, act, slope,
                                                                        (__nv_bfloat16*)out, out_pitch, L, C, rpb4, lens,
                                                                        len_mul, len_add);
    else
      kk::adain_act_vec4_kernel<float><<<grid4, block, 0, st>>>(x, x_pitch, stats, h, h_stride, alpha, act, slope,
                                                                (float*)out, out_pitch, L, C, rpb4, lens, len_mul, len_add);
    KK_LAUNCH_CHECK("adain_act_vec4_kernel");
    return KK_OK;
  }
#define KK_ADA(T, P) kk::adain_act_kernel<T, P><<<grid, block, 0, st>>>(x, x_pitch, stats, h, h_stride, alpha, act, \
    slope, (T*)out, out_pitch, L, C, pool_w, pool_b, rpb, lens, len_mul, len_add)
  if (out_is_bf16) { if (pool) KK_ADA(__nv_bfloat16, true); else KK_ADA(__nv_bfloat16, false); }
  else { if (pool) KK_ADA(float, true); else KK_ADA(float, false); }
#undef KK_ADA
  KK_LAUNCH_CHECK("adain_act_kernel");
  return KK_OK;
}

extern "C" int kk_front_assemble(const float* asr, const float* F0_curve, const float* N, const float* f0_w3b,
                                 const float* n_w3b, float* out, int out_pitch, int asr_off, int f0_off, int n_off,
                                 int B, int C_asr, int T, const int* lens, void* stream) {
  KK_CHECK_ARG(F0_curve && N && f0_w3b && n_w3b && out && B > 0 && T > 0, "kk_front_assemble: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (asr) {
    dim3 grid(kk::cdiv(T, 32), kk::cdiv(C_asr, 32), B), block(32, 8);
    kk::transpose_cf_to_cl_kernel<<<grid, block, 0, st>>>(asr, out, C_asr, T, out_pitch, asr_off);
    KK_LAUNCH_CHECK("transpose_cf_to_cl_kernel");
  }
  kk::f0n_conv_kernel<<<kk::cdiv((long long)B * T, 256), 256, 0, st>>>(F0_curve, N, f0_w3b, n_w3b, out, out_pitch,
                                                                      f0_off, n_off, B, T, lens);
  KK_LAUNCH_CHECK("f0n_conv_kernel");
  return KK_OK;
}

extern "C" int kk_copy_channels(const float* src, int src_pitch, int src_off, float* dst, int dst_pitch, int dst_off,
                                int rows, int C, void* stream) {
  KK_CHECK_ARG(src && dst && rows > 0 && C > 0, "kk_copy_channels: bad arguments");
  long long total = (long long)rows * C;
  kk::copy_channels_kernel<<<kk::cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(src, src_pitch, src_off, dst,
                                                                                   dst_pitch, dst_off, rows, C);
  KK_LAUNCH_CHECK("copy_channels_kernel");
  return KK_OK;
}

namespace kk {
// Vocos ISTFTHead (codec/models/vocos/vocos.py:126-133): rows of [log-magnitude (bins) | phase (bins)] ->
// S = min(exp(m), clip) * (cos p + i sin p), written frame-major (rows, bins, 2) - the layout kk_istft_f32 reads.
__global__ void __launch_bounds__(256) polar_to_spec_kernel(const float* __restrict__ x, int x_pitch,
                                                            float2* __restrict__ spec, long long rows, int bins,
                                                            float clip) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * bins) return;
  const long long r = i / bins;
  const int k = (int)(i - r * bins);
  const float* xr = x + r * x_pitch;
  const float mag = fminf(expf(xr[k]), clip);
  float sn, cs;
  sincosf(xr[bins + k], &sn, &cs);
  spec[i] = make_float2(mag * cs, mag * sn);
}
}  // namespace kk

extern "C" int kk_polar_to_spec(const float* x, int x_pitch, float* spec, long long rows, int bins, float clip,
                                void* stream) {
  KK_CHECK_ARG(x && spec && rows > 0 && bins > 0 && x_pitch >= 2 * bins, "kk_polar_to_spec: bad arguments");
  const long long total = rows * bins;
  KK_CHECK_ARG(kk::cdiv(total, 256) <= 0x7fffffffLL, "kk_polar_to_spec: too many elements");
  kk::polar_to_spec_kernel<<<(unsigned)kk::cdiv(total, 256), 256, 0, (cudaStream_t)stream>>>(
      x, x_pitch, reinterpret_cast<float2*>(spec), rows, bins, clip);
  KK_LAUNCH_CHECK("polar_to_spec_kernel");
  return KK_OK;
}

namespace kk {
// Whisper log-mel dynamic-range step (stt/models/whisper/audio.py:80-81): per signal m = max(x),
// x <- (max(x, m - range) + add) * scale.  One block per signal (<= 3000 x 128 values): two passes over L2-resident data.
__global__ void __launch_bounds__(1024) clamp_below_max_affine_kernel(float* __restrict__ x, long long n_per,
                                                                      float range, float add, float scale) {
  __shared__ float red[32];
  float* xb = x + (size_t)blockIdx.x * n_per;
  float m = -INFINITY;
  for (long long i = threadIdx.x; i < n_per; i += blockDim.x) m = fmaxf(m, xb[i]);
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  m = red[threadIdx.x & 31];
  for (int off = 16; off > 0; off >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, off));
  const float lo = __fadd_rn(m, -range);
  for (long long i = threadIdx.x; i < n_per; i += blockDim.x)
    xb[i] = __fmul_rn(__fadd_rn(fmaxf(xb[i], lo), add), scale);
}
}  // namespace kk

extern "C" int kk_clamp_below_max_affine(float* x, int B, long long n_per, float range, float add, float scale,
                                         void* stream) {
  KK_CHECK_ARG(x && B > 0 && n_per > 0, "kk_clamp_below_max_affine: bad arguments");
  kk::clamp_below_max_affine_kernel<<<B, 1024, 0, (cudaStream_t)stream>>>(x, n_per, range, add, scale);
  KK_LAUNCH_CHECK("clamp_below_max_affine_kernel");
  return KK_OK;
}

namespace kk {
// Parakeet normalisation: one block per signal, thread (ty, f) walks frames ty, ty + 8, ... of feature f (coalesced
// across features), fp64 partial sums combined through shared memory; then the same walk applies the normalisation.
__global__ void __launch_bounds__(1024) normalize_features_kernel(float* __restrict__ x, int frames, int n_feat,
                                                                  int per_feature, float eps) {
  extern __shared__ double sm[];           // [2][rows][n_feat]
  const int rows = blockDim.y, ty = threadIdx.y;
  float* xb = x + (size_t)blockIdx.x * frames * n_feat;
  double* ssum = sm;
  double* ssq = sm + (size_t)rows * n_feat + 2;
  const int gslot = rows * n_feat;          // per-signal totals (per_feature == 0)
  for (int f = threadIdx.x; f < n_feat; f += blockDim.x) {
    double s = 0.0, q = 0.0;
    for (int t = ty; t < frames; t += rows) { const double v = xb[(size_t)t * n_feat + f]; s += v; q += v * v; }
    ssum[ty * n_feat + f] = s; ssq[ty * n_feat + f] = q;
  }
  __syncthreads();
  if (ty == 0) {
    for (int f = threadIdx.x; f < n_feat; f += blockDim.x) {
      double s = 0.0, q = 0.0;
      for (int r = 0; r < rows; ++r) { s += ssum[r * n_feat + f]; q += ssq[r * n_feat + f]; }
      ssum[f] = s; ssq[f] = q;
    }
  }
  __syncthreads();
  if (!per_feature) {
    if (ty == 0 && threadIdx.x == 0) {
      double s = 0.0, q = 0.0;
      for (int f = 0; f < n_feat; ++f) { s += ssum[f]; q += ssq[f]; }
      ssum[gslot] = s; ssq[gslot] = q;
    }
    __syncthreads();
  }
  const double cnt = per_feature ? (double)frames : (double)frames * n_feat;
  for (int f = threadIdx.x; f < n_feat; f += blockDim.x) {
    const double s = per_feature ? ssum[f] : ssum[gslot], q = per_feature ? ssq[f] : ssq[gslot];
    const double mean = s / cnt;
    const double var = fmax(q / cnt - mean * mean, 0.0);
    const float m = (float)mean, inv = 1.0f / ((float)sqrt(var) + eps);
    for (int t = ty; t < frames; t += rows) {
      const size_t i = (size_t)t * n_feat + f;
      xb[i] = (xb[i] - m) * inv;
    }
  }
}
}  // namespace kk

extern "C" int kk_normalize_features(float* x, int B, int frames, int n_feat, int per_feature, float eps, void* stream) {
  KK_CHECK_ARG(x && B > 0 && frames > 0 && n_feat > 0 && n_feat <= 1024, "kk_normalize_features: bad arguments");
  int tx = n_feat >= 128 ? 128 : (n_feat >= 64 ? 64 : 32);
  int rows = 1024 / tx;
  while (rows > 1 && (size_t)2 * (rows * n_feat + 2) * 8 > 96 * 1024) rows >>= 1;
  const size_t smem = (size_t)2 * ((size_t)rows * n_feat + 2) * 8;
  KK_CUDA(cudaFuncSetAttribute(kk::normalize_features_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  kk::normalize_features_kernel<<<B, dim3(tx, rows), smem, (cudaStream_t)stream>>>(x, frames, n_feat, per_feature, eps);
  KK_LAUNCH_CHECK("normalize_features_kernel");
  return KK_OK;
}

namespace kk {
// Snake activations of the codec generators (BigVGAN Snake / SnakeBeta, codec/models/bigvgan/activation.py:5-51; DAC
// Snake1d, codec/models/descript/nn/layers.py:116-129):  y = x + 1 / (b_c + eps) * sin(a_c x)^2,  b = a unless a
// separate beta is given, a / b = exp(.) of the parameters in log-scale mode.  Element (o, c, i) of an (outer, C, inner)
// tensor: inner = 1 is channels-last, inner = L channels-first.  Reference op order, no contraction.
__global__ void __launch_bounds__(256) snake_kernel(const float* __restrict__ x, float* __restrict__ out,
                                                    const float* __restrict__ alpha, const float* __restrict__ beta,
                                                    long long total, int C, long long inner, int logscale, float eps) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)((i / inner) % C);
  float a = alpha[c], b = beta ? beta[c] : a;
  if (logscale) { a = expf(a); b = beta ? expf(b) : a; }
  const float v = x[i];
  const float s = sinf(__fmul_rn(v, a));
  out[i] = __fadd_rn(v, __fmul_rn(__fdiv_rn(1.0f, __fadd_rn(b, eps)), __fmul_rn(s, s)));
}
}  // namespace kk

extern "C" int kk_snake_f32(const float* x, float* out, const float* alpha, const float* beta, long long outer, int C,
                            long long inner, int logscale, float eps, void* stream) {
  KK_CHECK_ARG(x && out && alpha && outer > 0 && C > 0 && inner > 0, "kk_snake_f32: bad arguments");
  const long long total = outer * C * inner;
  KK_CHECK_ARG((total + 255) / 256 <= 0x7fffffffLL, "kk_snake_f32: too many elements");
  kk::snake_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(x, out, alpha, beta, total, C, inner,
                                                                                     logscale, eps);
  KK_LAUNCH_CHECK("snake_kernel");
  return KK_OK;
}

extern "C" int kk_convert_act(const float* x, int x_pitch, void* out, int out_is_bf16, int out_pitch, long long rows,
                              int C, float slope, void* stream) {
  KK_CHECK_ARG(x && out && rows > 0 && C > 0, "kk_convert_act: bad arguments");
  long long total = rows * C;
  int blocks = kk::cdiv(total, 256);
  const bool vec = x_pitch % 4 == 0 && out_pitch % 4 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)out & 15) == 0 &&
                   (rows + 3) / 4 <= 0x7fffffffLL;
  if (vec) {
    dim3 grid((unsigned)((rows + 3) / 4), kk::cdiv(kk::cdiv(C, 4), 64)), block(64, 4);
    if (out_is_bf16)
      kk::convert_act_vec4_kernel<__nv_bfloat16><<<grid, block, 0, (cudaStream_t)stream>>>(
          x, x_pitch, (__nv_bfloat16*)out, out_pitch, rows, C, slope);
    else
      kk::convert_act_vec4_kernel<float><<<grid, block, 0, (cudaStream_t)stream>>>(x, x_pitch, (float*)out, out_pitch,
                                                                                  rows, C, slope);
    KK_LAUNCH_CHECK("convert_act_vec4_kernel");
    return KK_OK;
  }
  if (out_is_bf16)
    kk::convert_act_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>(x, x_pitch, (__nv_bfloat16*)out,
                                                                                     out_pitch, rows, C, slope);
  else
    kk::convert_act_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(x, x_pitch, (float*)out, out_pitch, rows,
                                                                             C, slope);
  KK_LAUNCH_CHECK("convert_act_kernel");
  return KK_OK;
}
