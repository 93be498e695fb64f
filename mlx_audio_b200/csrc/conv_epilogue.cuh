// Epilogue element mapping shared by the fp32 (FFMA) and the tcgen05 convolution kernels.
// GEMM element (b, m, n) with accumulator `acc` -> output tensors, see kk_conv_desc in kokoro_b200.h.
#pragma once
#include "common.cuh"

namespace kk {

struct EpiCoord {
  int out_row;   // row in the output tensor, <0 when nothing is to be written
  int c;         // output channel
  bool valid;    // accumulator contributes (0 <= j < L_true of this utterance)
  bool pad;      // row exists in the tensors but lies beyond this utterance (ragged batch): stored as zero, no stats
};

// Per-utterance row limits of a ragged batch (kokoro_b200.h): lt = true conv output rows, lo = output rows.
struct EpiLim { int lt, lo; };
__device__ __forceinline__ EpiLim epi_lim(const kk_conv_desc& d, int b) {
  EpiLim l;
  l.lt = ragged_len(d.lens, d.len_mul, d.len_add, b, d.L_true);
  l.lo = l.lt + (d.L_out - d.L_true);
  return l;
}

__device__ __forceinline__ EpiCoord epi_coord(const kk_conv_desc& d, const EpiLim& lim, int m, int n) {
  EpiCoord e;
  int ph = 0;
  e.c = n;
  if (d.up > 1) { ph = n / d.C_out; e.c = n - ph * d.C_out; }
  int j = m * d.up + ph - d.up_off;
  e.valid = (j >= 0) && (j < lim.lt);
  int r = j + d.left_pad;
  e.out_row = (r >= 0 && r < d.L_out) ? r : -1;
  e.pad = r >= lim.lo;
  return e;
}

// Value to store for one element (everything except the stats reduction).  Returns the stored value.
__device__ __forceinline__ float epi_value(const kk_conv_desc& d, int b, const EpiCoord& e, float acc) {
  float v = 0.f;
  if (e.pad) return 0.f;
  if (e.valid) v = acc + (d.bias ? __ldg(d.bias + e.c) : 0.f);
  if (d.res) v += __ldg(d.res + ((size_t)b * d.L_res + (e.out_row >> d.res_shift)) * d.res_pitch + e.c);
  if (d.res2) v += __ldg(d.res2 + ((size_t)b * d.L_out + e.out_row) * d.res2_pitch + e.c);
  return v * d.scale;
}

__device__ __forceinline__ void epi_store(const kk_conv_desc& d, int b, const EpiCoord& e, float v) {
  size_t row = (size_t)b * d.L_out + e.out_row;
  if (d.out) d.out[row * d.out_pitch + e.c] = v;
  if (d.out_act) {
    float a = leaky(v, d.act_slope);
    if (d.act_is_bf16) ((__nv_bfloat16*)d.out_act)[row * d.act_pitch + e.c] = __float2bfloat16(a);
    else ((float*)d.out_act)[row * d.act_pitch + e.c] = a;
  }
}

}  // namespace kk
