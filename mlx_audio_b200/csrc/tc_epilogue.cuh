// Epilogue of the tcgen05 convolution kernels, shared by conv_tc.cu (TMA-fed A operand) and conv_fused.cu
// (AdaIN + activation applied while staging the A operand).
#pragma once
#include "conv_epilogue.cuh"
#include "ptx_sm100.cuh"

namespace kk {

// Geometry of one 32-column chunk: first GEMM column, polyphase index, first output channel, and whether the
// vectorised staging path applies (full chunk, 16-byte aligned rows everywhere).
struct EpiGeom { int nb, ph, cbase; bool fast; };

__device__ __forceinline__ EpiGeom epi_geom(const kk_conv_desc& d, const int n0, const int ch) {
  EpiGeom g;
  g.nb = n0 + ch * 32;
  g.ph = 0; g.cbase = g.nb;
  // phase is constant inside a 32-column chunk (host guarantees C_out % 32 == 0 when up > 1)
  if (d.up > 1) { g.ph = g.nb / d.C_out; g.cbase = g.nb - g.ph * d.C_out; }
  const size_t cb4 = (size_t)g.cbase * 4;
  g.fast = (g.nb + 32 <= d.N) &&
           (!d.out || (((reinterpret_cast<uintptr_t>(d.out) | (size_t)d.out_pitch * 4 | cb4) & 15) == 0)) &&
           (!d.res || (((reinterpret_cast<uintptr_t>(d.res) | (size_t)d.res_pitch * 4 | cb4) & 15) == 0)) &&
           (!d.res2 || (((reinterpret_cast<uintptr_t>(d.res2) | (size_t)d.res2_pitch * 4 | cb4) & 15) == 0)) &&
           (!d.bias || (((reinterpret_cast<uintptr_t>(d.bias) | cb4) & 15) == 0)) &&
           (!d.out_act || !d.act_is_bf16 ||
            (((reinterpret_cast<uintptr_t>(d.out_act) | (size_t)d.act_pitch * 2 | (size_t)g.cbase * 2) & 15) == 0)) &&
           (!d.out_act || d.act_is_bf16);
  return g;
}

// Residual prefetch for one chunk of the warp whose first GEMM row is mw: lane (r = it*4 + lane/8, cc = lane%8)
// loads 16 bytes of row r; t[] stays in flight while the previous chunk is being processed.
__device__ __forceinline__ void epi_prefetch_res(const kk_conv_desc& d, const EpiGeom& g, const int b, const int mw,
                                                 const int lane, float4 (&t)[8]) {
#pragma unroll
  for (int it = 0; it < 8; ++it) {
    const int r = it * 4 + (lane >> 3), cc = lane & 7;
    const int orr = (mw + r) * d.up + g.ph - d.up_off + d.left_pad;
    t[it] = make_float4(0.f, 0.f, 0.f, 0.f);
    if ((mw + r) < d.M && orr >= 0 && orr < d.L_out) {
      if (d.res) t[it] = __ldg(reinterpret_cast<const float4*>(
                     d.res + ((size_t)b * d.L_res + (orr >> d.res_shift)) * d.res_pitch + g.cbase + cc * 4));
      if (d.res2) {
        const float4 u = __ldg(reinterpret_cast<const float4*>(
            d.res2 + ((size_t)b * d.L_out + orr) * d.res2_pitch + g.cbase + cc * 4));
        t[it].x += u.x; t[it].y += u.y; t[it].z += u.z; t[it].w += u.w;
      }
    }
  }
}

// Issue the residual prefetch of chunk `ch` (first GEMM row of the tile m0, column origin n0) when that chunk will
// take the fast path and has a residual.  Call it for the first chunk of a tile BEFORE waiting on the accumulator.
__device__ __forceinline__ void epi_prefetch_chunk(const kk_conv_desc& d, const int b, const int m0, const int n0,
                                                   const int ch, const int lg, const int lane, float4 (&t)[8]) {
  if (!d.res && !d.res2) return;
  const EpiGeom g = epi_geom(d, n0, ch);
  if (g.fast) epi_prefetch_res(d, g, b, m0 + lg * 32, lane, t);
}

// One 32-column chunk of the epilogue for the warp that owns TMEM lane group `lg`: tensor memory -> registers
// -> (bias, residuals, scale) -> coalesced global stores through the warp-private staging tile; returns the
// per-column (sum, sum of squares) partials of the warp's 32 rows (lane i = column i of the chunk).
// t[] holds this chunk's prefetched residual (epi_prefetch_chunk); once it has been parked in the staging tile
// the prefetch of the NEXT chunk (has_next: tile origin (nx_m0, nx_n0) of utterance nx_b, chunk nx_ch) is issued so
// that its latency hides behind this chunk's TMEM read, arithmetic and stores.
template <int BN>
__device__ __forceinline__ void epilogue_chunk(const kk_conv_desc& d, const EpiLim lim, const int b, const int m0, const int n0,
                                               const int ch, const int lg, const int lane, const uint32_t tmem_base,
                                               float* stage_w, const bool do_stats, float& col_sum, float& col_sq,
                                               float4 (&t)[8], const bool has_next, const int nx_b, const int nx_m0,
                                               const int nx_n0, const int nx_ch) {
  const int m = m0 + lg * 32 + lane;  // GEMM row of this thread
  {
      const EpiGeom g = epi_geom(d, n0, ch);
      const int nb = g.nb, ph = g.ph, cbase = g.cbase;
      const bool fast = g.fast;
      float* st = stage_w;                       // [32 rows][8 chunks of 4 floats], chunk ^= row & 7
      if (fast && (d.res || d.res2)) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3), cc = lane & 7;
          *reinterpret_cast<float4*>(st + r * 32 + ((cc ^ (r & 7)) << 2)) = t[it];
        }
        __syncwarp();
      }
      if (has_next) epi_prefetch_chunk(d, nx_b, nx_m0, nx_n0, nx_ch, lg, lane, t);
      float v[32];
      ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * 32), v);
      ptx::tmem_ld_wait();
      const int j = m * d.up + ph - d.up_off;
      const bool acc_valid = (j >= 0) && (j < lim.lt);
      const int out_row = j + d.left_pad;
      const bool row_ok = (m < d.M) && (out_row >= 0) && (out_row < d.L_out);   // row exists in the tensors
      const bool row_live = row_ok && out_row < lim.lo;                         // ... and belongs to the utterance
      const size_t orow = (size_t)b * d.L_out + (row_ok ? out_row : 0);
      const float* resp = d.res ? d.res + ((size_t)b * d.L_res + ((row_ok ? out_row : 0) >> d.res_shift)) * d.res_pitch + cbase : nullptr;
      const float* res2p = d.res2 ? d.res2 + orow * d.res2_pitch + cbase : nullptr;
      // Fast path: full 32-column chunk, 16-byte aligned rows.  Global traffic goes through a warp-private
      // 32x32 fp32 staging tile (XOR-swizzled 16-byte chunks) so that every LDG/STG covers whole 128-byte
      // row segments (lane = column group) while TMEM hands out lane = row.
      if (fast) {
        const int mw = m0 + lg * 32;               // first GEMM row of this warp
        {
#pragma unroll
          for (int c4 = 0; c4 < 8; ++c4) {
            float4 rr = make_float4(0.f, 0.f, 0.f, 0.f);
            if (d.res || d.res2) rr = *reinterpret_cast<const float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2));
            float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);
            if (d.bias) bs = __ldg(reinterpret_cast<const float4*>(d.bias + cbase + c4 * 4));
            float x0 = acc_valid ? v[c4 * 4 + 0] + bs.x : 0.f;
            float x1 = acc_valid ? v[c4 * 4 + 1] + bs.y : 0.f;
            float x2 = acc_valid ? v[c4 * 4 + 2] + bs.z : 0.f;
            float x3 = acc_valid ? v[c4 * 4 + 3] + bs.w : 0.f;
            x0 = row_live ? (x0 + rr.x) * d.scale : 0.f;
            x1 = row_live ? (x1 + rr.y) * d.scale : 0.f;
            x2 = row_live ? (x2 + rr.z) * d.scale : 0.f;
            x3 = row_live ? (x3 + rr.w) * d.scale : 0.f;
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
        if (row_live && col_ok) {
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

// ---------------------------------------------------------------------------------------------------------------
// TMA epilogue (every chunk full and 16-byte aligned, no second residual - decided on the host).  Register
// loads / stores cannot keep enough bytes in flight per warp (profiles/r01_inflight_bw_probe.txt), so both the
// residual and the result travel through a warp-private 32x32 fp32 staging tile by TMA:
//   * interior chunks (all 32 rows of the warp map to rows inside the utterance): ONE 4 KB tensor-map box per
//     transfer.  The tensor map is the rank-4 view  V[b][m][phase][c] = T[b][m*up + phase - up_off + left_pad][c]
//     of the (B, L_out, pitch) tensor, so that the polyphase (transposed-conv) row interleave is just a
//     coordinate; the box is written with the 128-byte swizzle (16-byte chunk ^= row & 7), which makes the
//     lane = row float4 accesses and the lane = column scalar reads (instance-norm sums) bank-conflict free;
//   * edge chunks (first / last rows of an utterance, or a row-shifted shortcut): one 128-byte linear copy per
//     valid row, unswizzled (rare, so the bank conflicts do not matter).  A 128-byte request costs the TMA unit
//     about as much as a 4 KB one (~20 clk), which is why this is the fallback and not the rule.
// The bf16 activation copy (64-byte row segments) goes straight from registers.
#ifdef KK_TP_DBG
// lab build: clocks one thread per role of CTA 0 spends waiting (persistent conv; read by kk_tp_dbg_read)
static __device__ unsigned long long g_tp_clk[16];
#define TP_T0 const long long t0_ = clock64()
#define TP_ACC(i, who) do { if (blockIdx.x == 0 && threadIdx.x == (who)) g_tp_clk[i] += (unsigned long long)(clock64() - t0_); } while (0)
#define TP_MARK0 long long tm_ = clock64()
#define TP_MARK(i) do { const long long c_ = clock64(); if (blockIdx.x == 0 && threadIdx.x == 64) g_tp_clk[i] += (unsigned long long)(c_ - tm_); tm_ = c_; } while (0)
#else
#define TP_MARK0 do {} while (0)
#define TP_MARK(i) do {} while (0)
#define TP_T0 do {} while (0)
#define TP_ACC(i, who) do {} while (0)
#endif
constexpr int EPI_TILE_BYTES = 4096;
constexpr int EPI_NBUF = 2;        // staging tiles per epilogue warp = residual prefetch distance

struct EpiRow { int ph, cbase, out_row; bool acc_valid, row_ok, live; };

// row_ok: the row exists in the tensors; live: ... and belongs to the utterance (ragged batches: rows in
// [lim.lo, L_out) are padding - stored as zero, never counted)
__device__ __forceinline__ EpiRow epi_row(const kk_conv_desc& d, const EpiLim& lim, const int m, const int nb) {
  EpiRow e;
  e.ph = 0; e.cbase = nb;
  if (d.up > 1) { e.ph = nb / d.C_out; e.cbase = nb - e.ph * d.C_out; }
  const int j = m * d.up + e.ph - d.up_off;
  e.acc_valid = (j >= 0) && (j < lim.lt);
  e.out_row = j + d.left_pad;
  // nb >= N: a chunk of the last column tile past the GEMM's N (N a multiple of 32 but not of the tile width).  Its
  // "phase" / channel offset lie outside the tensors: the box path is clipped by the tensor map, the per-row path must
  // not store at all (phase = up lands on the NEXT row block's phase 0: it raced with the owner of those rows)
  e.row_ok = (nb < d.N) && (m < d.M) && (e.out_row >= 0) && (e.out_row < d.L_out);
  e.live = e.row_ok && e.out_row < lim.lo;
  return e;
}

// warp-uniform: do rows m0w .. m0w+31 of the chunk starting at column nb all land inside the utterance (and inside
// the true convolution output, so that no element needs masking)?
__device__ __forceinline__ bool epi_interior(const kk_conv_desc& d, const EpiLim& lim, const int m0w, const int nb) {
  const int ph = d.up > 1 ? nb / d.C_out : 0;
  const int j0 = m0w * d.up + ph - d.up_off, r0 = j0 + d.left_pad;
  return nb < d.N && m0w + 31 < d.M && r0 >= 0 && r0 + 31 * d.up < lim.lo && j0 >= 0 && j0 + 31 * d.up < lim.lt;
}

// all lanes: start the residual transfer of the chunk whose first column is nb, rows m0w..m0w+31 (lane = row)
__device__ __forceinline__ void epi_issue_res(const kk_conv_desc& d, const CUtensorMap* tmRes, const int b, const int m0w,
                                              const int nb, const int lane, uint8_t* st, uint64_t* bar) {
  const EpiLim lim = epi_lim(d, b);
  const EpiRow e = epi_row(d, lim, m0w + lane, nb);
  // bf16 residual (res_is_bf16): 64-byte rows - a 2 KB box with the 64-byte swizzle, or 64-byte row copies
  const uint32_t rowb = d.res_is_bf16 ? 64u : 128u;
  if (d.res_shift == 0 && epi_interior(d, lim, m0w, nb)) {
    if (lane == 0) {
      ptx::mbar_expect_tx(bar, 32u * rowb);
      ptx::tma_load_4d(st, tmRes, bar, e.cbase, e.ph, m0w, b);
    }
    return;
  }
  const unsigned mask = __ballot_sync(0xffffffffu, e.row_ok);
  if (lane == 0) ptx::mbar_expect_tx(bar, (uint32_t)__popc(mask) * rowb);
  __syncwarp();
  if (e.row_ok) {
    const size_t off = ((size_t)b * d.L_res + (e.out_row >> d.res_shift)) * d.res_pitch + e.cbase;
    const void* src = d.res_is_bf16 ? (const void*)(reinterpret_cast<const __nv_bfloat16*>(d.res) + off) : (const void*)(d.res + off);
    ptx::bulk_load_g2s(st + lane * rowb, src, rowb, bar);
  }
}

// INTERIOR: no masking anywhere, swizzled tile, box transfers.  Otherwise: per-element masks, plain tile, row copies.
template <bool INTERIOR>
__device__ __forceinline__ void epilogue_chunk_direct_t(const kk_conv_desc& d, const EpiLim lim, const CUtensorMap* tmOut,
                                                        const CUtensorMap* tmAct, uint8_t* act_tile, const int b,
                                                        const int m0w, const int nb, const int lane,
                                                        const uint32_t tmem_addr, uint8_t* st, uint64_t* rbar,
                                                        const uint32_t rparity, const bool do_stats, float& col_sum,
                                                        float& col_sq) {
  TP_MARK0;
  const EpiRow e = epi_row(d, lim, m0w + lane, nb);
  constexpr int WSW = INTERIOR ? 7 : 0;                                  // swizzle mask of the result tile
  const int rsw = (INTERIOR && d.res_shift == 0) ? 7 : 0;               // ... and of the tile as the residual arrived
  const bool acc_valid = INTERIOR || e.acc_valid, row_ok = INTERIOR || e.row_ok;
  float v[32];
  ptx::tmem_ld_32x32(tmem_addr, v);
  { TP_T0; if (d.res) ptx::mbar_wait(rbar, rparity); TP_ACC(4, 64); }
  ptx::tmem_ld_wait();
  TP_MARK(10);
  float* row = reinterpret_cast<float*>(st) + lane * 32;
  float rres[32];                                     // this lane's residual row (kept: the non-interior masks re-use it)
#pragma unroll
  for (int i = 0; i < 32; ++i) rres[i] = 0.f;
  if (d.res && row_ok) {
    if (d.res_is_bf16) {                              // 64-byte rows, 16-byte chunk ^= (row >> 1) & 3 when it came as a box
      const uint8_t* r16 = st + lane * 64;
      const int sw = rsw ? ((lane >> 1) & 3) : 0;
#pragma unroll
      for (int c8 = 0; c8 < 4; ++c8) {
        const uint4 q = *reinterpret_cast<const uint4*>(r16 + ((c8 ^ sw) << 4));
        const uint32_t w4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          rres[c8 * 8 + 2 * i] = __uint_as_float(w4[i] << 16);
          rres[c8 * 8 + 2 * i + 1] = __uint_as_float(w4[i] & 0xffff0000u);
        }
      }
    } else {
#pragma unroll
      for (int c4 = 0; c4 < 8; ++c4) {
        const float4 rr = *reinterpret_cast<const float4*>(row + ((c4 ^ (lane & rsw)) << 2));
        rres[c4 * 4 + 0] = rr.x; rres[c4 * 4 + 1] = rr.y; rres[c4 * 4 + 2] = rr.z; rres[c4 * 4 + 3] = rr.w;
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] += rres[i];
  }
  if (d.bias && nb < d.N) {
#pragma unroll
    for (int c4 = 0; c4 < 8; ++c4) {
      const float4 bs = __ldg(reinterpret_cast<const float4*>(d.bias + e.cbase + c4 * 4));
      v[c4 * 4 + 0] += bs.x; v[c4 * 4 + 1] += bs.y; v[c4 * 4 + 2] += bs.z; v[c4 * 4 + 3] += bs.w;
    }
  }
  if (!INTERIOR) {
    // accumulator rows outside the true output contribute nothing (the residual still does); rows outside the
    // utterance are zero
#pragma unroll
    for (int c4 = 0; c4 < 8; ++c4) {
      if (!acc_valid) {
        v[c4 * 4 + 0] = rres[c4 * 4 + 0]; v[c4 * 4 + 1] = rres[c4 * 4 + 1];
        v[c4 * 4 + 2] = rres[c4 * 4 + 2]; v[c4 * 4 + 3] = rres[c4 * 4 + 3];
      }
      if (!e.live) v[c4 * 4 + 0] = v[c4 * 4 + 1] = v[c4 * 4 + 2] = v[c4 * 4 + 3] = 0.f;
    }
  }
  if (d.scale != 1.0f) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] *= d.scale;
  }
  TP_MARK(11);
  const size_t orow = (size_t)b * d.L_out + (row_ok ? e.out_row : 0);
  // bf16 activation copy.  Interior chunks with a staging tile: 32 rows x 64 bytes written with the 64-byte swizzle
  // (16-byte chunk ^= (row >> 1) & 3: conflict-free for lane = row) and stored as ONE tensor-map box together with the
  // fp32 tile below (row-strided 16-byte register stores reached 1.9 TB/s, the box store 4.4).  Otherwise from registers.
  const bool act_box = INTERIOR && act_tile != nullptr && d.out_act != nullptr;
  if (row_ok && d.out_act) {
    uint4 q[4];
#pragma unroll
    for (int c8 = 0; c8 < 4; ++c8) {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(leaky(v[c8 * 8 + 0], d.act_slope), leaky(v[c8 * 8 + 1], d.act_slope));
      __nv_bfloat162 p1 = __floats2bfloat162_rn(leaky(v[c8 * 8 + 2], d.act_slope), leaky(v[c8 * 8 + 3], d.act_slope));
      __nv_bfloat162 p2 = __floats2bfloat162_rn(leaky(v[c8 * 8 + 4], d.act_slope), leaky(v[c8 * 8 + 5], d.act_slope));
      __nv_bfloat162 p3 = __floats2bfloat162_rn(leaky(v[c8 * 8 + 6], d.act_slope), leaky(v[c8 * 8 + 7], d.act_slope));
      q[c8].x = *reinterpret_cast<uint32_t*>(&p0); q[c8].y = *reinterpret_cast<uint32_t*>(&p1);
      q[c8].z = *reinterpret_cast<uint32_t*>(&p2); q[c8].w = *reinterpret_cast<uint32_t*>(&p3);
    }
    if (act_box) {
      TP_T0;
      ptx::bulk_wait_read<0>();                        // the previous box store has read the (single) bf16 tile
      TP_ACC(5, 64);
      __syncwarp();
#pragma unroll
      for (int c8 = 0; c8 < 4; ++c8)
        *reinterpret_cast<uint4*>(act_tile + lane * 64 + ((c8 ^ ((lane >> 1) & 3)) << 4)) = q[c8];
    } else {
      uint4* ap = reinterpret_cast<uint4*>((__nv_bfloat16*)d.out_act + orow * d.act_pitch + e.cbase);
#pragma unroll
      for (int c8 = 0; c8 < 4; ++c8) ap[c8] = q[c8];
    }
  }
  TP_MARK(12);
  if (d.out || do_stats) {
    if (!d.res) ptx::bulk_wait_read<EPI_NBUF - 1>();   // the store that left this tile EPI_NBUF items ago has read it
    __syncwarp();                                      // ... for every lane; and every lane holds its residual row
#pragma unroll
    for (int c4 = 0; c4 < 8; ++c4)
      *reinterpret_cast<float4*>(row + ((c4 ^ (lane & WSW)) << 2)) = make_float4(v[c4 * 4], v[c4 * 4 + 1], v[c4 * 4 + 2], v[c4 * 4 + 3]);
    if (d.out) {
      ptx::fence_proxy_async();                        // generic writes -> visible to the TMA engine
      if (INTERIOR) {
        __syncwarp();
        if (lane == 0) ptx::tma_store_4d(tmOut, st, e.cbase, e.ph, m0w, b);
      } else if (row_ok) {
        ptx::bulk_store_s2g(d.out + orow * d.out_pitch + e.cbase, row, 128u);
      }
      if (!act_box) ptx::bulk_commit();
    }
  }
  TP_MARK(13);
  if (act_box) {                                       // one bulk group per chunk: fp32 box (if any) + bf16 box
    ptx::fence_proxy_async();
    __syncwarp();
    if (lane == 0) ptx::tma_store_4d(tmAct, act_tile, e.cbase, e.ph, m0w, b);
    ptx::bulk_commit();
  }
  TP_MARK(14);
  if (do_stats) {
    __syncwarp();
    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;   // lane = column: sums over the warp's 32 rows
    const int cc = lane >> 2, ci = lane & 3;
#pragma unroll
    for (int r = 0; r < 32; r += 2) {
      const float a0 = reinterpret_cast<const float*>(st)[r * 32 + ((cc ^ (r & WSW)) << 2) + ci];
      const float a1 = reinterpret_cast<const float*>(st)[(r + 1) * 32 + ((cc ^ ((r + 1) & WSW)) << 2) + ci];
      s0 += a0; q0 = fmaf(a0, a0, q0);
      s1 += a1; q1 = fmaf(a1, a1, q1);
    }
    col_sum = s0 + s1;
    col_sq = q0 + q1;
  }
  TP_MARK(15);
  __syncwarp();                                    // the tile may be overwritten by the next residual copy
}

__device__ __forceinline__ void epilogue_chunk_direct(const kk_conv_desc& d, const CUtensorMap* tmOut,
                                                      const CUtensorMap* tmAct, uint8_t* act_tile, const int b,
                                                      const int m0w, const int nb, const int lane,
                                                      const uint32_t tmem_addr, uint8_t* st, uint64_t* rbar,
                                                      const uint32_t rparity, const bool do_stats, float& col_sum,
                                                      float& col_sq) {
  const EpiLim lim = epi_lim(d, b);
  if (epi_interior(d, lim, m0w, nb))
    epilogue_chunk_direct_t<true>(d, lim, tmOut, tmAct, act_tile, b, m0w, nb, lane, tmem_addr, st, rbar, rparity, do_stats, col_sum, col_sq);
  else
    epilogue_chunk_direct_t<false>(d, lim, tmOut, tmAct, act_tile, b, m0w, nb, lane, tmem_addr, st, rbar, rparity, do_stats, col_sum, col_sq);
}

}  // namespace kk
