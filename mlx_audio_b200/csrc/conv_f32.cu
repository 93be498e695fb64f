// fp32 CUDA-core implicit-GEMM convolution (strict-fp32 parity path; also the tiny / strided layers).
// Tile 128(M) x 64(N) x 16(K), 256 threads, 8x4 outputs per thread, shared-memory staged.
#include "conv_epilogue.cuh"

namespace kk {

constexpr int F_BM = 128, F_BN = 64, F_BK = 16;

__global__ void __launch_bounds__(256)
conv1d_f32_kernel(const float* __restrict__ x, const float* __restrict__ w, const kk_conv_desc d) {
  __shared__ __align__(16) float As[F_BK][F_BM + 4];
  __shared__ __align__(16) float Bs[F_BK][F_BN + 4];
  __shared__ float s_sum[16][F_BN], s_sq[16][F_BN];  // one slot per row group: deterministic combine

  const int b = blockIdx.z;
  const int m0 = blockIdx.x * F_BM, n0 = blockIdx.y * F_BN;
  const int tid = threadIdx.x;
  const int ty = tid >> 4, tx = tid & 15;
  const float* xb = x + (size_t)b * d.L_in * d.x_pitch;
  const int wrow = d.taps * d.Cw;

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // loader mapping: A: 128 rows x 16 k = 2048 elements, 8 per thread; B: 64 x 16 = 1024, 4 per thread
  const int a_row = tid >> 1;          // 0..127
  const int a_k0 = (tid & 1) * 8;      // 0 or 8
  const int b_row = tid >> 2;          // 0..63
  const int b_k0 = (tid & 3) * 4;      // 0,4,8,12

  for (int t = 0; t < d.taps; ++t) {
    const int in_row = (m0 + a_row) * d.stride + t * d.dil - d.pad_left;
    const bool row_ok = (m0 + a_row) < d.M && in_row >= 0 && in_row < d.L_in;
    const float* xr = xb + (size_t)(row_ok ? in_row : 0) * d.x_pitch;
    const bool n_ok = (n0 + b_row) < d.N;
    const float* wr = w + (size_t)(n_ok ? n0 + b_row : 0) * wrow + (size_t)t * d.Cw;
    for (int c0 = 0; c0 < d.C_in; c0 += F_BK) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int c = c0 + a_k0 + i;
        As[a_k0 + i][a_row] = (row_ok && c < d.C_in) ? xr[c] : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int c = c0 + b_k0 + i;
        Bs[b_k0 + i][b_row] = (n_ok && c < d.C_in) ? wr[c] : 0.f;
      }
      __syncthreads();
#pragma unroll
      for (int kk_ = 0; kk_ < F_BK; ++kk_) {
        float4 a0 = *reinterpret_cast<const float4*>(&As[kk_][ty * 8]);
        float4 a1 = *reinterpret_cast<const float4*>(&As[kk_][ty * 8 + 4]);
        float4 bv = *reinterpret_cast<const float4*>(&Bs[kk_][tx * 4]);
        float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
      }
      __syncthreads();
    }
  }

  // ---- epilogue
  const bool do_stats = d.stats != nullptr;
  const EpiLim lim = epi_lim(d, b);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    int n = n0 + tx * 4 + j;
    float cs = 0.f, cq = 0.f;
    if (n < d.N) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        int m = m0 + ty * 8 + i;
        if (m >= d.M) continue;
        EpiCoord e = epi_coord(d, lim, m, n);
        if (e.out_row < 0) continue;
        float v = epi_value(d, b, e, acc[i][j]);   // 0 on the padding rows of a ragged batch (stored, not counted)
        epi_store(d, b, e, v);
        cs += v;
        cq = fmaf(v, v, cq);
      }
    }
    if (do_stats) {
      s_sum[ty][tx * 4 + j] = cs;
      s_sq[ty][tx * 4 + j] = cq;
    }
  }
  if (do_stats) {
    __syncthreads();
    if (tid < F_BN && n0 + tid < d.N) {
      int n = n0 + tid;
      int c = d.up > 1 ? n % d.C_out : n;
      float ts = 0.f, tq = 0.f;
#pragma unroll
      for (int i = 0; i < 16; ++i) { ts += s_sum[i][tid]; tq += s_sq[i][tid]; }
      atomicAdd(d.stats + ((size_t)b * d.C_out + c) * 2, (double)ts);
      atomicAdd(d.stats + ((size_t)b * d.C_out + c) * 2 + 1, (double)tq);
    }
  }
}

int validate_conv_desc(const kk_conv_desc* d, const char* who) {
  KK_CHECK_ARG(d, "%s: null descriptor", who);
  KK_CHECK_ARG(d->B > 0 && d->L_in > 0 && d->C_in > 0 && d->x_pitch >= d->C_in, "%s: bad input shape", who);
  KK_CHECK_ARG(d->taps > 0 && d->dil > 0 && d->stride > 0 && d->Cw >= d->C_in, "%s: bad taps/dil/stride/Cw", who);
  KK_CHECK_ARG(d->M > 0 && d->N > 0 && d->C_out > 0 && d->up > 0 && d->N == d->C_out * d->up,
               "%s: N (%d) must equal C_out*up (%d*%d)", who, d->N, d->C_out, d->up);
  KK_CHECK_ARG(d->L_out > 0 && d->L_true > 0, "%s: bad output length", who);
  KK_CHECK_ARG(d->out || d->out_act || d->out_add, "%s: no output buffer", who);
  KK_CHECK_ARG(!d->out_add || d->add_pitch >= d->C_out, "%s: add_pitch < C_out", who);
  KK_CHECK_ARG(!d->out || d->out_pitch >= d->C_out, "%s: out_pitch < C_out", who);
  KK_CHECK_ARG(!d->out_act || d->act_pitch >= d->C_out, "%s: act_pitch < C_out", who);
  KK_CHECK_ARG(!d->res || (d->res_pitch >= d->C_out && d->L_res > 0 && d->res_shift >= 0), "%s: bad residual", who);
  KK_CHECK_ARG(!d->res2 || d->res2_pitch >= d->C_out, "%s: bad residual 2", who);
  return KK_OK;
}

}  // namespace kk

extern "C" int kk_conv1d_f32(const float* x, const float* w, const kk_conv_desc* d, void* stream) {
  KK_CHECK_ARG(x && w, "kk_conv1d_f32: null tensor");
  int rc = kk::validate_conv_desc(d, "kk_conv1d_f32");
  if (rc) return rc;
  if (d->res && d->res_is_bf16) {
    kk::set_error("kk_conv1d_f32: a bf16 residual is implemented by kk_conv1d_adain_bf16_tc only");
    return KK_ERR_UNSUPPORTED;
  }
  if (d->out_add) {
    kk::set_error("kk_conv1d_f32: out_add is implemented by kk_conv1d_adain_bf16_tc only");
    return KK_ERR_UNSUPPORTED;
  }
  dim3 grid(kk::cdiv(d->M, kk::F_BM), kk::cdiv(d->N, kk::F_BN), d->B);
  KK_CHECK_ARG(grid.y <= 65535 && grid.z <= 65535, "kk_conv1d_f32: grid too large");
  kk::conv1d_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x, w, *d);
  KK_LAUNCH_CHECK("conv1d_f32_kernel");
  return KK_OK;
}
