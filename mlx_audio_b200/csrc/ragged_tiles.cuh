// Ragged batches in the persistent tcgen05 convolution kernels: the (utterance, row-tile) list is built from the
// per-utterance lengths, so that padding rows cost nothing.  Every CTA derives the same table in shared memory at kernel
// start (one warp, B loads + a shuffle scan): ts[b] = index of utterance b's first tile, ts[B] = number of tiles.
#pragma once
#include "conv_epilogue.cuh"

namespace kk {

constexpr int RG_MAX_B = 256;   // utterances per ragged launch (table size in shared memory)
// Rows past an utterance's end that are still processed - and therefore WRITTEN AS ZERO in the activation outputs: the
// next convolution reads them through TMA, unmasked, as its zero padding.  Generator levels (20 T, 120 T rows): up to
// (taps - 1) * dil / 2 <= 28 rows (k = 11, dilation 5) -> 64.  Front levels (T and 2 T rows, len_mul <= 2): the
// consumers are k = 3 convolutions, the depthwise pool and the first polyphase upsampler (one or two rows) -> 8; with
// 128-row tiles over a few hundred rows the larger margin often costs a whole tile.
constexpr int RG_ZPAD = 64;
constexpr int RG_ZPAD_FRONT = 8;

// GEMM rows of utterance b that need processing (rows with an output row inside the utterance, plus the margin)
__device__ __forceinline__ int rg_rows(const kk_conv_desc& d, int b) {
  const int lt = ragged_len(d.lens, d.len_mul, d.len_add, b, d.L_true);
  const int m = (lt + d.up_off + d.up - 1) / d.up + (d.len_mul <= 2 ? RG_ZPAD_FRONT : RG_ZPAD);
  return m < d.M ? m : d.M;
}

// one full warp
__device__ __forceinline__ void rg_build(int* ts, const kk_conv_desc& d, int tile_rows, int lane) {
  int carry = 0;
  for (int b0 = 0; b0 < d.B; b0 += 32) {
    const int b = b0 + lane;
    const int n = b < d.B ? (rg_rows(d, b) + tile_rows - 1) / tile_rows : 0;
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (b < d.B) ts[b] = carry + incl - n;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) ts[d.B] = carry;
}

// tile index (over utterances x row tiles) -> (utterance, row tile).  Uniform batches: a division; ragged: a monotone
// walk of the table (every role visits its tiles in increasing order; look-ahead users keep their own map).
struct RgMap {
  const int* ts;
  int m_tiles;
  int b;
  __device__ __forceinline__ RgMap(const int* ts_, int m_tiles_) : ts(ts_), m_tiles(m_tiles_), b(0) {}
  __device__ __forceinline__ void locate(int tile, int& bb, int& mt) {
#ifdef KK_NO_RAGGED
    bb = tile / m_tiles; mt = tile - bb * m_tiles; return;
#endif
    if (!ts) { bb = tile / m_tiles; mt = tile - bb * m_tiles; return; }
    while (tile >= ts[b + 1]) ++b;
    bb = b;
    mt = tile - ts[b];
  }
};

}  // namespace kk
