// Epilogue of the tcgen05 convolution kernels, shared by conv_tc.cu (TMA-fed A operand) and conv_fused.cu
// (AdaIN + activation applied while staging the A operand).
#pragma once
#include "conv_epilogue.cuh"
#include "ptx_sm100.cuh"

namespace kk {

// One 32-column chunk of the epilogue for the warp that owns TMEM lane group `lg`: tensor memory -> registers
// -> (bias, residuals, scale) -> coalesced global stores through the warp-private staging tile; returns the
// per-column (sum, sum of squares) partials of the warp's 32 rows (lane i = column i of the chunk).
template <int BN>
__device__ __forceinline__ void epilogue_chunk(const kk_conv_desc& d, const int b, const int m0, const int n0,
                                               const int ch, const int lg, const int lane, const uint32_t tmem_base,
                                               float* stage_w, const bool do_stats, float& col_sum, float& col_sq) {
  const int m = m0 + lg * 32 + lane;  // GEMM row of this thread
  {
      float v[32];
      ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * 32), v);
      ptx::tmem_ld_wait();
      const int nb = n0 + ch * 32;           // first GEMM column of this chunk
      // phase is constant inside a 32-column chunk (host guarantees C_out % 32 == 0 when up > 1)
      int ph = 0, cbase = nb;
      if (d.up > 1) { ph = nb / d.C_out; cbase = nb - ph * d.C_out; }
      const int j = m * d.up + ph - d.up_off;
      const bool acc_valid = (j >= 0) && (j < d.L_true);
      const int out_row = j + d.left_pad;
      const bool row_ok = (m < d.M) && (out_row >= 0) && (out_row < d.L_out);
      const size_t orow = (size_t)b * d.L_out + (row_ok ? out_row : 0);
      const float* resp = d.res ? d.res + ((size_t)b * d.L_res + ((row_ok ? out_row : 0) >> d.res_shift)) * d.res_pitch + cbase : nullptr;
      const float* res2p = d.res2 ? d.res2 + orow * d.res2_pitch + cbase : nullptr;
      // Fast path: full 32-column chunk, 16-byte aligned rows.  Global traffic goes through a warp-private
      // 32x32 fp32 staging tile (XOR-swizzled 16-byte chunks) so that every LDG/STG covers whole 128-byte
      // row segments (lane = column group) while TMEM hands out lane = row.
      const bool fast = (nb + 32 <= d.N) &&
                        (!d.out || (((reinterpret_cast<uintptr_t>(d.out) | (size_t)d.out_pitch * 4 | (size_t)cbase * 4) & 15) == 0)) &&
                        (!d.res || (((reinterpret_cast<uintptr_t>(d.res) | (size_t)d.res_pitch * 4 | (size_t)cbase * 4) & 15) == 0)) &&
                        (!d.res2 || (((reinterpret_cast<uintptr_t>(d.res2) | (size_t)d.res2_pitch * 4 | (size_t)cbase * 4) & 15) == 0)) &&
                        (!d.out_act || !d.act_is_bf16 ||
                         (((reinterpret_cast<uintptr_t>(d.out_act) | (size_t)d.act_pitch * 2 | (size_t)cbase * 2) & 15) == 0)) &&
                        (!d.out_act || d.act_is_bf16);
      if (fast) {
        float* st = stage_w;                       // [32 rows][8 chunks of 4 floats], chunk ^= row & 7
        const int mw = m0 + lg * 32;               // first GEMM row of this warp
        if (d.res || d.res2) {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), cc = lane & 7;
            const int jr = (mw + r) * d.up + ph - d.up_off;
            const int orr = jr + d.left_pad;
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if ((mw + r) < d.M && orr >= 0 && orr < d.L_out) {
              if (d.res) t = __ldg(reinterpret_cast<const float4*>(
                             d.res + ((size_t)b * d.L_res + (orr >> d.res_shift)) * d.res_pitch + cbase + cc * 4));
              if (d.res2) {
                float4 u = __ldg(reinterpret_cast<const float4*>(
                    d.res2 + ((size_t)b * d.L_out + orr) * d.res2_pitch + cbase + cc * 4));
                t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
              }
            }
            *reinterpret_cast<float4*>(st + r * 32 + ((cc ^ (r & 7)) << 2)) = t;
          }
          __syncwarp();
        }
        {
          float bsv[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) bsv[i] = d.bias ? __ldg(d.bias + cbase + i) : 0.f;
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
            if (d.res || d.res2) t = *reinterpret_cast<const float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2));
            float x0 = acc_valid ? v[c4 * 4 + 0] + bsv[c4 * 4 + 0] : 0.f;
            float x1 = acc_valid ? v[c4 * 4 + 1] + bsv[c4 * 4 + 1] : 0.f;
            float x2 = acc_valid ? v[c4 * 4 + 2] + bsv[c4 * 4 + 2] : 0.f;
            float x3 = acc_valid ? v[c4 * 4 + 3] + bsv[c4 * 4 + 3] : 0.f;
            x0 = row_ok ? (x0 + t.x) * d.scale : 0.f;
            x1 = row_ok ? (x1 + t.y) * d.scale : 0.f;
            x2 = row_ok ? (x2 + t.z) * d.scale : 0.f;
            x3 = row_ok ? (x3 + t.w) * d.scale : 0.f;
            v[c4 * 4 + 0] = x0; v[c4 * 4 + 1] = x1; v[c4 * 4 + 2] = x2; v[c4 * 4 + 3] = x3;
          }
        }
        __syncwarp();
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4)
          *reinterpret_cast<float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2)) =
              make_float4(v[c4 * 4], v[c4 * 4 + 1], v[c4 * 4 + 2], v[c4 * 4 + 3]);
        __syncwarp();
        if (d.out) {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + (lane >> 3), cc = lane & 7;
            const int orr = (mw + r) * d.up + ph - d.up_off + d.left_pad;
            if ((mw + r) < d.M && orr >= 0 && orr < d.L_out)
              *reinterpret_cast<float4*>(d.out + ((size_t)b * d.L_out + orr) * d.out_pitch + cbase + cc * 4) =
                  *reinterpret_cast<const float4*>(st + r * 32 + ((cc ^ (r & 7)) << 2));
          }
        }
        if (d.out_act) {
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int r = it * 8 + (lane >> 2), c8 = lane & 3;
            const int orr = (mw + r) * d.up + ph - d.up_off + d.left_pad;
            if ((mw + r) < d.M && orr >= 0 && orr < d.L_out) {
              float4 t0 = *reinterpret_cast<const float4*>(st + r * 32 + (((2 * c8) ^ (r & 7)) << 2));
              float4 t1 = *reinterpret_cast<const float4*>(st + r * 32 + (((2 * c8 + 1) ^ (r & 7)) << 2));
              __nv_bfloat162 p0 = __floats2bfloat162_rn(leaky(t0.x, d.act_slope), leaky(t0.y, d.act_slope));
              __nv_bfloat162 p1 = __floats2bfloat162_rn(leaky(t0.z, d.act_slope), leaky(t0.w, d.act_slope));
              __nv_bfloat162 p2 = __floats2bfloat162_rn(leaky(t1.x, d.act_slope), leaky(t1.y, d.act_slope));
              __nv_bfloat162 p3 = __floats2bfloat162_rn(leaky(t1.z, d.act_slope), leaky(t1.w, d.act_slope));
              uint4 q4;
              q4.x = *reinterpret_cast<uint32_t*>(&p0); q4.y = *reinterpret_cast<uint32_t*>(&p1);
              q4.z = *reinterpret_cast<uint32_t*>(&p2); q4.w = *reinterpret_cast<uint32_t*>(&p3);
              *reinterpret_cast<uint4*>((__nv_bfloat16*)d.out_act + ((size_t)b * d.L_out + orr) * d.act_pitch + cbase + c8 * 8) = q4;
            }
          }
        }
        __syncwarp();
      } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const bool col_ok = (nb + i) < d.N;
        float x = 0.f;
        if (row_ok && col_ok) {
          if (acc_valid) x = v[i] + (d.bias ? __ldg(d.bias + cbase + i) : 0.f);
          if (resp) x += __ldg(resp + i);
          if (res2p) x += __ldg(res2p + i);
          x *= d.scale;
        }
        v[i] = x;
      }
      if (row_ok) {
        if (d.out) {
          float* op = d.out + orow * d.out_pitch + cbase;
          if ((reinterpret_cast<uintptr_t>(op) & 15) == 0 && nb + 32 <= d.N) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) *reinterpret_cast<float4*>(op + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) if (nb + i < d.N) op[i] = v[i];
          }
        }
        if (d.out_act) {
          if (d.act_is_bf16) {
            __nv_bfloat16* ap = (__nv_bfloat16*)d.out_act + orow * d.act_pitch + cbase;
            if ((reinterpret_cast<uintptr_t>(ap) & 15) == 0 && nb + 32 <= d.N) {
#pragma unroll
              for (int i = 0; i < 32; i += 8) {
                __nv_bfloat162 p0 = __floats2bfloat162_rn(leaky(v[i], d.act_slope), leaky(v[i + 1], d.act_slope));
                __nv_bfloat162 p1 = __floats2bfloat162_rn(leaky(v[i + 2], d.act_slope), leaky(v[i + 3], d.act_slope));
                __nv_bfloat162 p2 = __floats2bfloat162_rn(leaky(v[i + 4], d.act_slope), leaky(v[i + 5], d.act_slope));
                __nv_bfloat162 p3 = __floats2bfloat162_rn(leaky(v[i + 6], d.act_slope), leaky(v[i + 7], d.act_slope));
                uint4 q;
                q.x = *reinterpret_cast<uint32_t*>(&p0); q.y = *reinterpret_cast<uint32_t*>(&p1);
                q.z = *reinterpret_cast<uint32_t*>(&p2); q.w = *reinterpret_cast<uint32_t*>(&p3);
                *reinterpret_cast<uint4*>(ap + i) = q;
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) if (nb + i < d.N) ap[i] = __float2bfloat16(leaky(v[i], d.act_slope));
            }
          } else {
            float* ap = (float*)d.out_act + orow * d.act_pitch + cbase;
#pragma unroll
            for (int i = 0; i < 32; ++i) if (nb + i < d.N) ap[i] = leaky(v[i], d.act_slope);
          }
        }
      }
      }  // fast / generic
      if (do_stats) {
        // column sums over the warp's 32 rows, read back from the staging tile (lane i = column i; the XOR
        // swizzle keeps every row read conflict-free).  Rows / columns that were not stored hold 0.
        float* st = stage_w;
        if (!fast) {
          __syncwarp();
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4)
            *reinterpret_cast<float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2)) =
                make_float4(v[c4 * 4], v[c4 * 4 + 1], v[c4 * 4 + 2], v[c4 * 4 + 3]);
          __syncwarp();
        }
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
        const int cc = lane >> 2, ci = lane & 3;
#pragma unroll
        for (int r = 0; r < 32; r += 2) {
          const float a0 = st[r * 32 + ((cc ^ (r & 7)) << 2) + ci];
          const float a1 = st[(r + 1) * 32 + ((cc ^ ((r + 1) & 7)) << 2) + ci];
          s0 += a0; q0 = fmaf(a0, a0, q0);
          s1 += a1; q1 = fmaf(a1, a1, q1);
        }
        col_sum = s0 + s1;
        col_sq = q0 + q1;
        __syncwarp();
      }
  }
}

}  // namespace kk
