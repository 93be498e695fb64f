// Implicit-GEMM Conv1d / polyphase ConvTranspose1d on the 5th-gen tensor cores (sm_100a):
//   TMA (cp.async.bulk.tensor, 128B swizzle) -> shared-memory ring -> tcgen05.mma (bf16 x bf16 -> fp32 in
//   tensor memory) -> tcgen05.ld epilogue (bias, residuals, scale, instance-norm partial sums, optional
//   activated bf16 copy).
// A tile = 128 (positions) x BN (output channels) of one utterance.  The K loop walks (tap, 64-channel block): the A
// tile of a tap is the SAME activation tensor shifted by tap*dil - pad rows; TMA's out-of-bounds zero fill provides
// the per-utterance conv zero padding for free (tensor map dims are (C, L, B), so a box never bleeds into the
// neighbouring utterance).
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected lane), warps 2-9 = epilogue
// (warp w owns TMEM lanes 32*(w%4)..+31); warp 2 also allocates / frees tensor memory.
#include "tc_epilogue.cuh"
#include "ragged_tiles.cuh"
#include <stdlib.h>

namespace kk {

constexpr int TC_BM = 128;       // UMMA M
constexpr int TC_BK = 64;        // bf16 elements per K block = one 128-byte swizzle row
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;  // 16 KB

// Persistent kernel: one CTA per SM loops over output tiles.  Tensor memory holds TWO accumulators so the
// epilogue of tile i (8 warps) overlaps the TMA/MMA main loop of tile i+1; barrier set-up, TMEM allocation and
// tensor-map prefetch are paid once per SM instead of once per tile.

// fp32 -> nearest value with 10 explicit mantissa bits (exactly representable in TF32): round half away from zero
__device__ __forceinline__ float tf32_rn(float v) { return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u); }

constexpr int TP_EPI_WARPS = 8;
constexpr int TP_THREADS = 64 + TP_EPI_WARPS * 32;  // 320
constexpr int TP_SPLIT_WARPS = 4;                   // TF32x3 only: hi / lo operand split of the A tile
constexpr int TP_THREADS_TF32 = TP_THREADS + TP_SPLIT_WARPS * 32;

// TF32 = false: bf16 operands (K block = 64 channels).  TF32 = true: fp32 operands on the kind::tf32 path with the
// error-compensated 3xTF32 split (the strict-fp32 mode of the reference arithmetic, istftnet.py:128-166): the K block
// is 32 channels (the same 128-byte swizzle rows), every operand exists as hi = x with the 13 low mantissa bits
// rounded away (exactly representable in TF32, so the tensor core's own conversion cannot change it) and lo = x - hi
// (exact in fp32, then rounded to TF32 as well); three MMAs per K step accumulate lo*hi + hi*lo + hi*hi into the same fp32 accumulator.  The
// weights are split at pack time (kk_split_tf32), the activation tile in shared memory by four extra warps, so the
// fp32 activation tensor crosses HBM once.
template <int BN, bool TF32 = false>
struct TpCfg {
  static constexpr int BK = TF32 ? 32 : 64;                  // channels per K block (128 bytes per row)
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = TF32 ? 2 * TC_A_BYTES + 2 * B_BYTES : TC_A_BYTES + B_BYTES;   // [A | A_lo | B | B_lo]
  static constexpr int TX_BYTES = TF32 ? TC_A_BYTES + 2 * B_BYTES : STAGE_BYTES;                  // bytes TMA delivers
  static constexpr int STAGES = TF32 ? 2 : (BN >= 256 ? 3 : (BN >= 128 ? 4 : (BN >= 64 ? 5 : 6)));
  static constexpr int NBUF = EPI_NBUF;                      // residual staging tiles per epilogue warp
  static constexpr int STAGING_BYTES = TP_EPI_WARPS * NBUF * EPI_TILE_BYTES;
  static constexpr int STATS_BYTES = 2 * 8 * BN * 4;  // double-buffered [4 lane groups][sum, sumsq][BN]
  // one 32 x 32 bf16 tile per epilogue warp for the box store of the activation copy (no room next to BN = 256 stages)
  static constexpr int ACT_BYTES = (BN <= 128 && !TF32) ? TP_EPI_WARPS * 2048 : 0;
  static constexpr int OFF_ACT = STAGES * STAGE_BYTES + STAGING_BYTES + STATS_BYTES;
  static constexpr int OFF_BAR = OFF_ACT + ACT_BYTES;
  static constexpr int OFF_RG = OFF_BAR + 448;               // ragged tile table (RG_MAX_B + 1 ints)
  static constexpr int SMEM_BYTES = OFF_RG + (RG_MAX_B + 1) * 4 + 12 + 1024;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
  static constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;
};

template <int BN, bool TF32>
__global__ void __launch_bounds__(TF32 ? TP_THREADS_TF32 : TP_THREADS, 1)
conv1d_tc_persist_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmB2, const __grid_constant__ CUtensorMap tmRes,
                         const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmAct,
                         const kk_conv_desc d, const int m_tiles, const int n_tiles, const int total_tiles_uniform,
                         const int direct_epi, const int terms) {
  using Cfg = TpCfg<BN, TF32>;
  constexpr int TC_BK = Cfg::BK;
#ifdef KK_TP_DBG
  const long long t_start_ = clock64();
#endif
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float* staging = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES);
  float* s_stats_all = reinterpret_cast<float*>(smem + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::STAGING_BYTES);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* empty_bar = full_bar + Cfg::STAGES;
  uint64_t* tmem_full_bar = empty_bar + Cfg::STAGES;   // [2]
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;        // [2]
  uint64_t* res_bar = tmem_empty_bar + 2;              // [TP_EPI_WARPS * NBUF]
  uint64_t* split_bar = res_bar + TP_EPI_WARPS * Cfg::NBUF;   // [STAGES] TF32x3: the stage's A tile has been split
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(split_bar + 6);
  int* rg_ts = reinterpret_cast<int*>(smem + Cfg::OFF_RG);   // ragged batches: first row tile of each utterance
  const int* ts = d.lens ? rg_ts : nullptr;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cblocks = (d.C_in + TC_BK - 1) / TC_BK;
  const int num_kb = d.taps * cblocks;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < Cfg::STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(&tmem_full_bar[a], 1);
      ptx::mbar_init(&tmem_empty_bar[a], TP_EPI_WARPS * 32);
    }
    for (int i = 0; i < TP_EPI_WARPS * Cfg::NBUF; ++i) ptx::mbar_init(&res_bar[i], 1);
    for (int s = 0; s < Cfg::STAGES; ++s) ptx::mbar_init(&split_bar[s], TP_SPLIT_WARPS);
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  if (warp == 3 && d.lens) rg_build(rg_ts, d, TC_BM, lane);   // ragged batch: the tile list follows the lengths
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // tile = (utterance x row tile) * n_tiles + column tile
  const int total_tiles = ts ? ts[d.B] * n_tiles : total_tiles_uniform;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      uint32_t kbg = 0;
      RgMap map(ts, m_tiles);
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int nt = tile % n_tiles;
        const int rest = tile / n_tiles;
        int mt, b;
        map.locate(rest, b, mt);
        const int m0 = mt * TC_BM, n0 = nt * BN;
        for (int kb = 0; kb < num_kb; ++kb, ++kbg) {
          const int s = kbg % Cfg::STAGES;
          const uint32_t ph = (kbg / Cfg::STAGES) & 1;
          { TP_T0; ptx::mbar_wait(&empty_bar[s], ph ^ 1); TP_ACC(0, 0); }
          const int t = kb / cblocks, cb = kb - t * cblocks;
          uint8_t* sa = smem + s * Cfg::STAGE_BYTES;
          ptx::mbar_expect_tx(&full_bar[s], Cfg::TX_BYTES);
          ptx::tma_load_3d(sa, &tmA, &full_bar[s], cb * TC_BK, m0 + t * d.dil - d.pad_left, b);
          if constexpr (TF32) {
            ptx::tma_load_2d(sa + 2 * TC_A_BYTES, &tmB, &full_bar[s], t * d.Cw + cb * TC_BK, n0);
            ptx::tma_load_2d(sa + 2 * TC_A_BYTES + Cfg::B_BYTES, &tmB2, &full_bar[s], t * d.Cw + cb * TC_BK, n0);
          } else {
            ptx::tma_load_2d(sa + TC_A_BYTES, &tmB, &full_bar[s], t * d.Cw + cb * TC_BK, n0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = TF32 ? ptx::umma_idesc_tf32(TC_BM, BN) : ptx::umma_idesc_bf16(TC_BM, BN);
      uint32_t kbg = 0, it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const uint32_t acc = it & 1, aph = (it >> 1) & 1;
        { TP_T0; ptx::mbar_wait(&tmem_empty_bar[acc], aph ^ 1); TP_ACC(1, 32); }  // epilogue has drained this accumulator
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb, ++kbg) {
          const int s = kbg % Cfg::STAGES;
          const uint32_t ph = (kbg / Cfg::STAGES) & 1;
          const uint32_t sa = ptx::smem_u32(smem + s * Cfg::STAGE_BYTES);
          if constexpr (TF32) {
            ptx::mbar_wait(&split_bar[s], ph);           // TMA landed AND the A tile is split into hi / lo
            ptx::tc_fence_after();
            const uint32_t sal = sa + TC_A_BYTES, sb = sa + 2 * TC_A_BYTES, sbl = sb + Cfg::B_BYTES;
            // descriptors once per stage; +32 bytes per K step = +2 in the (address >> 4) field
            const uint64_t ah0 = ptx::umma_desc(sa, 16, 1024, ptx::UMMA_LAYOUT_SW128), al0 = ptx::umma_desc(sal, 16, 1024, ptx::UMMA_LAYOUT_SW128);
            const uint64_t bh0 = ptx::umma_desc(sb, 16, 1024, ptx::UMMA_LAYOUT_SW128), bl0 = ptx::umma_desc(sbl, 16, 1024, ptx::UMMA_LAYOUT_SW128);
#pragma unroll
            for (int k = 0; k < 4; ++k) {                // K = 8 fp32 = 32 bytes per MMA
              const uint64_t ah = ah0 + 2 * k, al = al0 + 2 * k, bh = bh0 + 2 * k, bl = bl0 + 2 * k;
              if (terms == 3) {                          // small terms first, the dominant product last
                ptx::umma_tf32(tacc, al, bh, idesc, (kb | k) != 0 ? 1u : 0u);
                ptx::umma_tf32(tacc, ah, bl, idesc, 1u);
                ptx::umma_tf32(tacc, ah, bh, idesc, 1u);
              } else {
                ptx::umma_tf32(tacc, ah, bh, idesc, (kb | k) != 0 ? 1u : 0u);
              }
            }
          } else {
            { TP_T0; ptx::mbar_wait(&full_bar[s], ph); TP_ACC(2, 32); }
            ptx::tc_fence_after();
            const uint32_t sb = sa + TC_A_BYTES;
            const uint64_t ad0 = ptx::umma_desc(sa, 16, 1024, ptx::UMMA_LAYOUT_SW128), bd0 = ptx::umma_desc(sb, 16, 1024, ptx::UMMA_LAYOUT_SW128);
#pragma unroll
            for (int k = 0; k < 4; ++k)                  // K = 16 bf16 = 32 bytes per MMA: +2 in the (address >> 4) field
              ptx::umma_bf16(tacc, ad0 + 2 * k, bd0 + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(&empty_bar[s]);
        }
        ptx::umma_commit(&tmem_full_bar[acc]);
      }
    }
  } else if (TF32 && warp >= 2 + TP_EPI_WARPS) {
    // ===================== TF32x3: split the landed A tile into hi / lo (warps 10..13) =====================
    // Elementwise and position-preserving, so the 128B swizzle TMA applied is irrelevant: A[i] -> (hi, lo) at the same
    // byte offset of the hi and the lo tile.  hi = x rounded to TF32, lo = x - hi rounded to TF32.
    const int st = threadIdx.x - (2 + TP_EPI_WARPS) * 32;          // 0..127
    uint32_t kbg = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < num_kb; ++kb, ++kbg) {
        const int s = kbg % Cfg::STAGES;
        const uint32_t ph = (kbg / Cfg::STAGES) & 1;
        ptx::mbar_wait(&full_bar[s], ph);
        float4* hi = reinterpret_cast<float4*>(smem + s * Cfg::STAGE_BYTES);
        float4* lo = reinterpret_cast<float4*>(smem + s * Cfg::STAGE_BYTES + TC_A_BYTES);
        if (terms == 3) {
#pragma unroll
          for (int i = 0; i < TC_A_BYTES / 16 / (TP_SPLIT_WARPS * 32); ++i) {
            const int j = i * (TP_SPLIT_WARPS * 32) + st;
            const float4 v = hi[j];
            float4 h, l;
            h.x = tf32_rn(v.x); l.x = tf32_rn(v.x - h.x);      // hi + lo = x to 2^-24 |x|; both exact TF32 values
            h.y = tf32_rn(v.y); l.y = tf32_rn(v.y - h.y);
            h.z = tf32_rn(v.z); l.z = tf32_rn(v.z - h.z);
            h.w = tf32_rn(v.w); l.w = tf32_rn(v.w - h.w);
            hi[j] = h;
            lo[j] = l;
          }
          ptx::fence_proxy_async();                      // generic-proxy stores -> visible to the tensor core
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&split_bar[s]);
      }
    }
  } else {
    // ===================== epilogue (warps 2..9) =====================
    const int ew = warp - 2;
    const int lg = warp & 3;            // TMEM lane group of this warp
    const int half = ew >> 2;           // which half of the tile's column chunks
    constexpr int CHUNKS = BN / 32;
    constexpr int CH_PER = (CHUNKS + 1) / 2;
    float* stage_w = staging + ew * 1024;
    const bool do_stats = d.stats != nullptr;
    uint32_t it = 0;
    const int ch_first = half * CH_PER;
    const int n_mine = (CHUNKS - ch_first) < CH_PER ? (CHUNKS - ch_first) : CH_PER;   // chunks of this warp (may be <= 0)
    float4 tres[8];
    // direct epilogue: item n = (n / n_mine)-th tile of this CTA, chunk ch_first + n % n_mine
    constexpr int NBUF = Cfg::NBUF;
    uint8_t* stg = reinterpret_cast<uint8_t*>(staging) + ew * (NBUF * EPI_TILE_BYTES);
    uint64_t* rbar = res_bar + ew * NBUF;
    uint32_t item = 0;
    RgMap map(ts, m_tiles), map_res(ts, m_tiles), map_next(ts, m_tiles);
    auto issue_res = [&](uint32_t n) {
      const int tl = blockIdx.x + (int)(n / (uint32_t)n_mine) * gridDim.x;
      if (tl >= total_tiles) return;
      int rb, rmt;
      map_res.locate(tl / n_tiles, rb, rmt);
      epi_issue_res(d, &tmRes, rb, rmt * TC_BM + lg * 32,
                    (tl % n_tiles) * BN + (ch_first + (int)(n % (uint32_t)n_mine)) * 32, lane,
                    stg + (n % NBUF) * EPI_TILE_BYTES, &rbar[n % NBUF]);
    };
    if (direct_epi) {
      if (d.res && n_mine > 0)
        for (int i = 0; i < NBUF; ++i) issue_res(i);
    } else if (n_mine > 0 && (int)blockIdx.x < total_tiles) {
      int fb, fmt;
      map_next.locate(blockIdx.x / n_tiles, fb, fmt);
      epi_prefetch_chunk(d, fb, fmt * TC_BM, (blockIdx.x % n_tiles) * BN, ch_first, lg, lane, tres);
    }
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int nt = tile % n_tiles;
      const int rest = tile / n_tiles;
      int mt, b;
      map.locate(rest, b, mt);
      const int m0 = mt * TC_BM, n0 = nt * BN;
      const int ntile = tile + gridDim.x;
      const bool more = ntile < total_tiles;
      int nx_b = 0, nx_mt = 0;
      if (more) map_next.locate(ntile / n_tiles, nx_b, nx_mt);
      const int nx_m0 = nx_mt * TC_BM, nx_n0 = (ntile % n_tiles) * BN;
      const uint32_t acc = it & 1, aph = (it >> 1) & 1;
      float* s_stats = s_stats_all + acc * (8 * BN);
      { TP_T0; ptx::mbar_wait(&tmem_full_bar[acc], aph); TP_ACC(3, 64); }
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < n_mine; ++c) {
        TP_T0;
        const int ch = ch_first + c;
        const bool last = c + 1 == n_mine;
        float cs = 0.f, cq = 0.f;
        if (direct_epi) {
          const uint32_t bufi = item % NBUF;
          epilogue_chunk_direct(d, &tmOut, &tmAct, Cfg::ACT_BYTES ? smem + Cfg::OFF_ACT + ew * 2048 : nullptr, b,
                                m0 + lg * 32, n0 + ch * 32, lane,
                                tmem_base + acc * BN + ((uint32_t)(lg * 32) << 16) + (uint32_t)(ch * 32),
                                stg + bufi * EPI_TILE_BYTES, &rbar[bufi], (item / NBUF) & 1, do_stats, cs, cq);
          if (d.res) {
            { TP_T0; if (d.out) ptx::bulk_wait_read<0>(); TP_ACC(6, 64); }      // the lane's row store has read the tile
            ptx::fence_proxy_async();                 // generic accesses to the tile precede the next TMA write
            __syncwarp();
            issue_res(item + NBUF);
          }
          ++item;
        } else {
          epilogue_chunk<BN>(d, epi_lim(d, b), b, m0, n0, ch, lg, lane, tmem_base + acc * BN, stage_w, do_stats, cs, cq, tres,
                             last ? more : true, last ? nx_b : b, last ? nx_m0 : m0, last ? nx_n0 : n0,
                             last ? ch_first : ch + 1);
        }
        if (do_stats) { s_stats[(lg * 2 + 0) * BN + ch * 32 + lane] = cs; s_stats[(lg * 2 + 1) * BN + ch * 32 + lane] = cq; }
        TP_ACC(8, 64);
      }
      // all tcgen05.ld of this thread have completed (wait::ld inside the chunk): hand the accumulator back
      ptx::tc_fence_before();
      ptx::mbar_arrive(&tmem_empty_bar[acc]);
      if (do_stats) {
        asm volatile("bar.sync 1, %0;" ::"n"(TP_EPI_WARPS * 32) : "memory");
        for (int i = threadIdx.x - 64; i < BN; i += TP_EPI_WARPS * 32) {
          const int n = n0 + i;
          if (n < d.N) {
            const int cc = d.up > 1 ? n % d.C_out : n;
            float ts = (s_stats[0 * BN + i] + s_stats[2 * BN + i]) + (s_stats[4 * BN + i] + s_stats[6 * BN + i]);
            float tq = (s_stats[1 * BN + i] + s_stats[3 * BN + i]) + (s_stats[5 * BN + i] + s_stats[7 * BN + i]);
            atomicAdd(d.stats + ((size_t)b * d.C_out + cc) * 2, (double)ts);
            atomicAdd(d.stats + ((size_t)b * d.C_out + cc) * 2 + 1, (double)tq);
          }
        }
      }
    }
    if (direct_epi) ptx::bulk_wait<0>();   // every TMA store of this lane has completed before the CTA retires
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
#ifdef KK_TP_DBG
  if (blockIdx.x == 0 && threadIdx.x == 0) { g_tp_clk[7] += (unsigned long long)(clock64() - t_start_); g_tp_clk[9] += 1; }
#endif
}

}  // namespace kk

// ================================================================== host side
namespace kk {

int validate_conv_desc(const kk_conv_desc* d, const char* who);  // conv_f32.cu

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  }
  return fn;
}

// Tensor map over a (rank-D) global tensor; dims/strides innermost first, strides in bytes for dims 1..rank-1.
int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
              const cuuint64_t* strides, const cuuint32_t* box, CUtensorMapSwizzle swz) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return KK_ERR_CUDA; }
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(m, dt, (cuuint32_t)rank, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return KK_ERR_CUDA; }
  return KK_OK;
}

// bf16 tensor map with 128B swizzle
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides,
                   const cuuint32_t* box) {
  return make_tmap(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, base, rank, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
}

int num_sms() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int BN, bool TF32>
static int launch_tc_persist(const void* x, const void* w, const void* w_lo, const kk_conv_desc* d, int terms,
                             cudaStream_t st) {
  using Cfg = TpCfg<BN, TF32>;
  constexpr int ESZ = TF32 ? 4 : 2;
  const CUtensorMapDataType dt = TF32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  const char* who = TF32 ? "kk_conv1d_tf32_tc" : "kk_conv1d_bf16_tc";
  CUtensorMap tmA, tmB, tmB2;
  {
    cuuint64_t dims[3] = {(cuuint64_t)d->C_in, (cuuint64_t)d->L_in, (cuuint64_t)d->B};
    cuuint64_t strides[2] = {(cuuint64_t)d->x_pitch * ESZ, (cuuint64_t)d->L_in * d->x_pitch * ESZ};
    cuuint32_t box[3] = {Cfg::BK, TC_BM, 1};
    int rc = make_tmap(&tmA, dt, x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)d->taps * d->Cw, (cuuint64_t)d->N};
    cuuint64_t strides[1] = {(cuuint64_t)d->taps * d->Cw * ESZ};
    cuuint32_t box[2] = {Cfg::BK, BN};
    int rc = make_tmap(&tmB, dt, w, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc) return rc;
    tmB2 = tmB;
    if (TF32 && (rc = make_tmap(&tmB2, dt, w_lo, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    KK_CUDA(cudaFuncSetAttribute(conv1d_tc_persist_kernel<BN, TF32>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Cfg::SMEM_BYTES));
    attr_set = true;
  }
  const int m_tiles = cdiv(d->M, TC_BM), n_tiles = cdiv(d->N, BN);
  const long long total = (long long)m_tiles * n_tiles * d->B;
  KK_CHECK_ARG(total < (1LL << 31), "%s: too many tiles", who);
  KK_CHECK_ARG(!d->lens || d->B <= RG_MAX_B, "%s: a ragged batch holds at most %d utterances", who, RG_MAX_B);
  const int grid = (int)(total < num_sms() ? total : num_sms());
  // direct epilogue (tc_epilogue.cuh): every 32-column chunk full, 16-byte aligned rows, one residual at most
  auto al16 = [](const void* p, long long pitch_bytes) { return (((uintptr_t)p | (uintptr_t)pitch_bytes) & 15) == 0; };
  const int direct = d->N % 32 == 0 && (d->up == 1 || d->C_out % 32 == 0) && !d->res2 &&
                     (!d->out || al16(d->out, (long long)d->out_pitch * 4)) &&
                     (!d->res || al16(d->res, (long long)d->res_pitch * (d->res_is_bf16 ? 2 : 4))) && (!d->bias || al16(d->bias, 0)) &&
                     (!d->out_act || (d->act_is_bf16 && al16(d->out_act, (long long)d->act_pitch * 2)));
  // rank-4 polyphase views (tc_epilogue.cuh) of the residual and output tensors; only coordinates whose rows lie
  // inside the utterance are ever issued, so the (possibly negative) row offset folded into the base is harmless
  CUtensorMap tmRes = tmB, tmOut = tmB, tmAct = tmB;
  if (direct) {
    const long long off = (long long)d->left_pad - d->up_off;
    cuuint32_t box[4] = {32, 1, 32, 1};
    if (d->res && d->res_shift == 0) {
      const cuuint64_t rsz = d->res_is_bf16 ? 2 : 4;     // bf16 residual: 32 x 32 box of 64-byte rows, 64-byte swizzle
      cuuint64_t dims[4] = {(cuuint64_t)d->C_out, (cuuint64_t)d->up, (cuuint64_t)d->M, (cuuint64_t)d->B};
      cuuint64_t strides[3] = {(cuuint64_t)d->res_pitch * rsz, (cuuint64_t)d->up * d->res_pitch * rsz,
                               (cuuint64_t)d->L_res * d->res_pitch * rsz};
      const char* base = reinterpret_cast<const char*>(d->res) + off * d->res_pitch * (long long)rsz;
      int rc = make_tmap(&tmRes, d->res_is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, 4,
                         dims, strides, box, d->res_is_bf16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
    if (d->out) {
      cuuint64_t dims[4] = {(cuuint64_t)d->C_out, (cuuint64_t)d->up, (cuuint64_t)d->M, (cuuint64_t)d->B};
      cuuint64_t strides[3] = {(cuuint64_t)d->out_pitch * 4, (cuuint64_t)d->up * d->out_pitch * 4,
                               (cuuint64_t)d->L_out * d->out_pitch * 4};
      int rc = make_tmap(&tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d->out + off * d->out_pitch, 4, dims, strides, box,
                         CU_TENSOR_MAP_SWIZZLE_128B);
      if (rc) return rc;
    }
  }
  if (direct && d->out_act && Cfg::ACT_BYTES) {
    const long long off = (long long)d->left_pad - d->up_off;
    cuuint32_t box[4] = {32, 1, 32, 1};
    cuuint64_t dims[4] = {(cuuint64_t)d->C_out, (cuuint64_t)d->up, (cuuint64_t)d->M, (cuuint64_t)d->B};
    cuuint64_t strides[3] = {(cuuint64_t)d->act_pitch * 2, (cuuint64_t)d->up * d->act_pitch * 2,
                             (cuuint64_t)d->L_out * d->act_pitch * 2};
    int rc = make_tmap(&tmAct, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (const __nv_bfloat16*)d->out_act + off * d->act_pitch, 4,
                       dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B);
    if (rc) return rc;
  }
  if (d->res && d->res_is_bf16 && !(direct && d->res_shift == 0)) {
    set_error("%s: a bf16 residual needs the TMA epilogue (N %% 32 == 0, 16-byte aligned rows, no second residual, "
              "res_shift 0)", who);
    return KK_ERR_UNSUPPORTED;
  }
  conv1d_tc_persist_kernel<BN, TF32><<<grid, TF32 ? TP_THREADS_TF32 : TP_THREADS, Cfg::SMEM_BYTES, st>>>(
      tmA, tmB, tmB2, tmRes, tmOut, tmAct, *d, m_tiles, n_tiles, (int)total, direct, terms);
  KK_LAUNCH_CHECK(TF32 ? "conv1d_tc_persist_kernel<tf32>" : "conv1d_tc_persist_kernel<bf16>");
  return KK_OK;
}

#ifdef KK_TP_DBG
}  // namespace kk
extern "C" int kk_tp_dbg_read(unsigned long long* host16, int reset) {
  cudaMemcpyFromSymbol(host16, kk::g_tp_clk, sizeof(unsigned long long) * 16);
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(kk::g_tp_clk, z, sizeof(z)); }
  return 0;
}
namespace kk {
#endif

// fp32 -> (hi, lo) with hi = x rounded to TF32's 10 mantissa bits, lo = x - hi (exact) rounded the same way, so that
// the tensor core's own conversion cannot change either operand
__global__ void split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float v = x[i];
  const float h = tf32_rn(v);
  hi[i] = h;
  lo[i] = tf32_rn(v - h);
}

}  // namespace kk

extern "C" int kk_conv1d_bf16_tc(const void* x, const void* w, const kk_conv_desc* d, int block_n, void* stream) {
  KK_CHECK_ARG(x && w, "kk_conv1d_bf16_tc: null tensor");
  int rc = kk::validate_conv_desc(d, "kk_conv1d_bf16_tc");
  if (rc) return rc;
  if (!kk_device_is_sm100()) {
    kk::set_error("kk_conv1d_bf16_tc: needs an sm_100 (Blackwell) device");
    return KK_ERR_UNSUPPORTED;
  }
  if (d->out_add) {
    kk::set_error("kk_conv1d_bf16_tc: out_add is implemented by kk_conv1d_adain_bf16_tc only");
    return KK_ERR_UNSUPPORTED;
  }
  KK_CHECK_ARG(d->stride == 1, "kk_conv1d_bf16_tc: stride must be 1 (got %d)", d->stride);
  KK_CHECK_ARG(d->Cw % 64 == 0, "kk_conv1d_bf16_tc: Cw (%d) must be a multiple of 64", d->Cw);
  KK_CHECK_ARG(d->x_pitch % 8 == 0, "kk_conv1d_bf16_tc: x_pitch (%d) must be a multiple of 8", d->x_pitch);
  KK_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0, "kk_conv1d_bf16_tc: operands must be 16-byte aligned");
  KK_CHECK_ARG(d->up == 1 || d->C_out % 32 == 0, "kk_conv1d_bf16_tc: polyphase needs C_out %% 32 == 0");
  KK_CHECK_ARG(d->B <= 65535, "kk_conv1d_bf16_tc: batch too large");
  if (block_n == 0) block_n = d->N > 64 ? 128 : (d->N > 32 ? 64 : 32);
  cudaStream_t st = (cudaStream_t)stream;
  switch (block_n) {
    case 32: return kk::launch_tc_persist<32, false>(x, w, nullptr, d, 1, st);
    case 64: return kk::launch_tc_persist<64, false>(x, w, nullptr, d, 1, st);
    case 128: return kk::launch_tc_persist<128, false>(x, w, nullptr, d, 1, st);
    case 256: return kk::launch_tc_persist<256, false>(x, w, nullptr, d, 1, st);
    default: kk::set_error("kk_conv1d_bf16_tc: block_n %d not in {32,64,128,256}", block_n); return KK_ERR_INVALID_ARG;
  }
}

extern "C" int kk_split_tf32(const float* x, float* hi, float* lo, long long n, void* stream) {
  KK_CHECK_ARG(x && hi && lo && n > 0, "kk_split_tf32: bad arguments");
  kk::split_tf32_kernel<<<kk::cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(x, hi, lo, n);
  KK_LAUNCH_CHECK("split_tf32_kernel");
  return KK_OK;
}

extern "C" int kk_conv1d_tf32_tc(const float* x, const float* w_hi, const float* w_lo, const kk_conv_desc* d, int terms,
                                 void* stream) {
  KK_CHECK_ARG(x && w_hi && w_lo, "kk_conv1d_tf32_tc: null tensor");
  int rc = kk::validate_conv_desc(d, "kk_conv1d_tf32_tc");
  if (rc) return rc;
  if (!kk_device_is_sm100()) {
    kk::set_error("kk_conv1d_tf32_tc: needs an sm_100 (Blackwell) device");
    return KK_ERR_UNSUPPORTED;
  }
  if (d->out_add) {
    kk::set_error("kk_conv1d_tf32_tc: out_add is implemented by kk_conv1d_adain_bf16_tc only");
    return KK_ERR_UNSUPPORTED;
  }
  KK_CHECK_ARG(terms == 1 || terms == 3, "kk_conv1d_tf32_tc: terms must be 1 (plain TF32) or 3 (3xTF32), got %d", terms);
  KK_CHECK_ARG(d->stride == 1, "kk_conv1d_tf32_tc: stride must be 1 (got %d)", d->stride);
  KK_CHECK_ARG(d->Cw % 32 == 0, "kk_conv1d_tf32_tc: Cw (%d) must be a multiple of 32", d->Cw);
  KK_CHECK_ARG(d->x_pitch % 4 == 0, "kk_conv1d_tf32_tc: x_pitch (%d) must be a multiple of 4", d->x_pitch);
  KK_CHECK_ARG((((uintptr_t)x | (uintptr_t)w_hi | (uintptr_t)w_lo) & 15) == 0, "kk_conv1d_tf32_tc: operands must be 16-byte aligned");
  KK_CHECK_ARG(d->up == 1 || d->C_out % 32 == 0, "kk_conv1d_tf32_tc: polyphase needs C_out %% 32 == 0");
  KK_CHECK_ARG(d->B <= 65535, "kk_conv1d_tf32_tc: batch too large");
  cudaStream_t st = (cudaStream_t)stream;
  if (d->N > 64) return kk::launch_tc_persist<128, true>(x, w_hi, w_lo, d, terms, st);
  if (d->N > 32) return kk::launch_tc_persist<64, true>(x, w_hi, w_lo, d, terms, st);
  return kk::launch_tc_persist<32, true>(x, w_hi, w_lo, d, terms, st);
}
