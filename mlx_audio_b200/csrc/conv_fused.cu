// Fused AdaIN + Snake/LeakyReLU + Conv1d on the 5th-gen tensor cores (sm_100a), TMA-fed end to end.
//
// Reference ops replaced (mlx_audio/tts/models/kokoro/istftnet.py): AdaINResBlock1's
//   xt = adain(x, s); xt = x + (1/a) sin(a xt)^2; xt = conv(xt) [; x = xt + x]      (:379-395, AdaIN1d :327-338)
// i.e. per reference call: instance-norm apply, style affine, Snake, swapaxes, conv, residual add - six
// materialised tensors.  Here the raw tensor x (B, L, C) crosses HBM once per output tile:
//   * an "X producer" thread streams tile + dilation-halo rows of one 64-channel block with TMA:
//       - bf16 input: straight into the A-operand slot (128B swizzle), transformed IN PLACE;
//       - fp32 input: through a ring of 32-row raw stages, transformed into the slot;
//     (register loads cannot keep enough bytes in flight - profiles/r01_inflight_bw_probe.txt);
//   * "transform" warps apply a_c x + b_c and the activation and leave bf16 in the canonical K-major
//     128B-swizzle layout (row = 128 B, 16-byte chunk index XOR (row & 7));
//   * one thread issues tcgen05.mma: for tap t the A operand is the SAME slot read through a descriptor whose
//     start address is shifted by t*dil rows (the swizzle is a function of the absolute smem address, so a row
//     shift that is not a multiple of the 8-row atom is legal - profiles/r01_umma_rowshift_probe.txt);
//   * the weights (B operand) stream through a second TMA ring, one (channel-block, tap) slab per stage;
//   * MT (1 or 2) 128-row sub-tiles share every B stage; accumulators are double-buffered in tensor memory
//     (2 x MT x BN = 512 columns) so the epilogue of tile i overlaps the main loop of tile i+1;
//   * epilogue warps own 32x32 fp32 staging tiles: the residual arrives in them by TMA (two tiles ahead), the
//     result is written back in place and leaves by TMA store; instance-norm sums are read off the same tile
//     and accumulated in fp64 registers across the CTA's contiguous run of tiles (one atomic flush per
//     utterance change instead of one per tile).
// Warp roles: 0 = TMA weights, 1 = MMA issuer, 2 = TMA input (+ TMEM allocation), 3 = spare, 4..11 = transform,
// 12..19 = epilogue; registers are re-balanced per warpgroup with setmaxnreg (control 40, transform 72, epilogue 144)
// as the first statement of every role branch (ptxas then allocates each branch against its own budget).  Persistent: CTA i walks a contiguous range of the (utterance, row-tile) list.
#include "conv_epilogue.cuh"
#include "ptx_sm100.cuh"
#include "ragged_tiles.cuh"
#include <stdlib.h>

namespace kk {

constexpr int FV_XF_WARPS = 8;
constexpr int FV_EPI_WARPS = 8;
constexpr int FV_XF_THREADS = FV_XF_WARPS * 32;
constexpr int FV_XF_WARP0 = 4;
constexpr int FV_EPI_WARP0 = 12;
constexpr int FV_THREADS = (FV_EPI_WARP0 + FV_EPI_WARPS) * 32;  // 640 = 5 warpgroups, launched with 96 registers
// per-warpgroup register reallocation (setmaxnreg): freed by dec 128*(56 + 2*24) = 13312 >= needed by inc 256*48
constexpr int FV_REGS_CTRL = 40;
// transform / epilogue budgets: <SWAP> 96 / 120 (lean register-store epilogue, the transform is the critical path);
// otherwise 72 / 144.  Freed by dec: 128*56 (+ 256*24) >= needed by inc: 256*24 (256*48).
template <bool SWAP> struct FvRegs { static constexpr int XF = SWAP ? 96 : 72, EPI = SWAP ? 120 : 144; };
constexpr int FV_HALO_MAX = 56;
constexpr int FV_BOX_ROWS = 32;            // rows per input TMA box
constexpr int FV_RAW_BYTES = 32 * 256;     // fp32 raw stage: 32 rows x 64 channels
constexpr int FV_XBOX_BYTES = 32 * 128;    // bf16 in-place box: 32 rows x 64 channels
constexpr int FV_MAX_RAW = 8;

template <int MT, int BN, int CB>
struct FvCfg {
  static constexpr int SLOT_ROWS = MT == 2 ? 320 : 192;          // >= 128*MT + halo, multiple of 32
  static constexpr int SLOT_BYTES = SLOT_ROWS * 128;
  static constexpr int NSLOT = 2;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGES = BN >= 256 ? 3 : 4;
  static constexpr int STG_HALF = FV_EPI_WARPS * 4096;           // one 32x32 fp32 tile per epilogue warp
  static constexpr int NRAW = STG_HALF / FV_RAW_BYTES;           // fp32 raw ring = staging half 1 (no residual then)
  static constexpr int COEF_BYTES = CB * 64 * 16;
  static constexpr int OFF_B = NSLOT * SLOT_BYTES;
  static constexpr int OFF_STG = OFF_B + STAGES * B_BYTES;       // [half 0][half 1]
  static constexpr int OFF_RAW = OFF_STG + STG_HALF;             // aliases staging half 1
  static constexpr int OFF_COEF = OFF_STG + 2 * STG_HALF;
  static constexpr int OFF_BAR = OFF_COEF + COEF_BYTES;
  static constexpr int OFF_RG = OFF_BAR + 640;                   // ragged tile table (RG_MAX_B + 1 ints)
  static constexpr int SMEM_BYTES = OFF_RG + (RG_MAX_B + 1) * 4 + 12 + 1024;
  static constexpr uint32_t TMEM_COLS = 512;
  static_assert(SLOT_ROWS >= 128 * MT + FV_HALO_MAX && SLOT_ROWS % 32 == 0, "slot too small");
  static_assert(2 * MT * BN <= 512, "accumulators exceed tensor memory");
  static_assert(NRAW <= FV_MAX_RAW && STAGES + (BN == 128 ? STG_HALF / B_BYTES : 0) <= 8, "barrier arrays");
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};

__device__ __forceinline__ uint32_t fv_pack(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

// coef = (a, b, -, -); Snake: (a, b + 1/(2 alpha), 2 alpha, -1/(2 alpha)) for the half-angle form
//   y + (1/alpha) sin^2(alpha y) = (y + 1/(2 alpha)) - cos(2 alpha y) / (2 alpha)
// (FFMA, FFMA, range-reduce, MUFU.COS, FFMA: one instruction fewer than the sine-squared form)
template <int ACT>
__device__ __forceinline__ float fv_apply(float x, const float4& k, float slope) {
  float y = fmaf(k.x, x, k.y);
  if (ACT == KK_ACT_SNAKE) {
    const float c = __cosf(fmaf(k.z, y, -1.0f));   // 2 alpha (y' - 1/(2 alpha)) = 2 alpha y
    y = fmaf(k.w, c, y);
  } else if (ACT == KK_ACT_LEAKY) {
    y = y > 0.f ? y : y * slope;
  }
  return y;
}

// bf16 input, in place: the slot holds rows [0, nrows) as TMA wrote them (128B swizzle).  Lane = (row%4, logical
// 16-byte chunk q): 8 channels cb*64 + 8q.., coefficients in registers; physical chunk = q ^ (row & 7).
// Batches of XB steps: all loads, then all arithmetic, then all stores (the shared-memory accesses are volatile
// asm, which the compiler will not reorder - the batching is what exposes the instruction-level parallelism).
__device__ __forceinline__ float4 fv_lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

// NB steps (4 rows each, FV_XF_WARPS steps apart) of the in-place transform: all loads, then all arithmetic, then all
// stores (the shared-memory accesses are volatile asm, which the compiler will not reorder - the batching is what
// exposes the instruction-level parallelism).
#ifdef KK_TUNE
__device__ int g_fv_xf_knock = 0;   // lab: 1 = no activation arithmetic (affine only), 2 = no arithmetic at all (load + store), 3 = nothing
#define FV_XF_KNOCK g_fv_xf_knock
#else
#define FV_XF_KNOCK 0
#endif
template <int ACT, int NB>
__device__ __forceinline__ void fv_xf_steps(uint32_t slot, const float4 (&k)[8], float slope, int s0, int nrows,
                                            int l_base, int L_in, int q, int sub) {
  if (FV_XF_KNOCK == 3) return;
  uint4 t[NB];
  uint32_t p[NB];
  bool ok[NB];
#pragma unroll
  for (int u = 0; u < NB; ++u) {
    const int r = 4 * (s0 + u * FV_XF_WARPS) + sub;
    const int l = l_base + r;
    p[u] = slot + (uint32_t)r * 128u + (uint32_t)((q ^ (r & 7)) << 4);
    ok[u] = r < nrows && l >= 0 && l < L_in;
    t[u] = make_uint4(0u, 0u, 0u, 0u);
    if (ok[u]) asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(t[u].x), "=r"(t[u].y), "=r"(t[u].z), "=r"(t[u].w) : "r"(p[u]));
  }
#pragma unroll
  for (int u = 0; u < NB; ++u) {   // branch-free so that the NB cosine chains interleave
    if (FV_XF_KNOCK == 2) continue;
    if (FV_XF_KNOCK == 1) {
      const uint4 v = t[u];
      t[u].x = fv_pack(fv_apply<KK_ACT_NONE>(__uint_as_float(v.x << 16), k[0], slope), fv_apply<KK_ACT_NONE>(__uint_as_float(v.x & 0xffff0000u), k[1], slope));
      t[u].y = fv_pack(fv_apply<KK_ACT_NONE>(__uint_as_float(v.y << 16), k[2], slope), fv_apply<KK_ACT_NONE>(__uint_as_float(v.y & 0xffff0000u), k[3], slope));
      t[u].z = fv_pack(fv_apply<KK_ACT_NONE>(__uint_as_float(v.z << 16), k[4], slope), fv_apply<KK_ACT_NONE>(__uint_as_float(v.z & 0xffff0000u), k[5], slope));
      t[u].w = fv_pack(fv_apply<KK_ACT_NONE>(__uint_as_float(v.w << 16), k[6], slope), fv_apply<KK_ACT_NONE>(__uint_as_float(v.w & 0xffff0000u), k[7], slope));
      continue;
    }
    const uint4 v = t[u];
    const uint32_t x0 = fv_pack(fv_apply<ACT>(__uint_as_float(v.x << 16), k[0], slope), fv_apply<ACT>(__uint_as_float(v.x & 0xffff0000u), k[1], slope));
    const uint32_t x1 = fv_pack(fv_apply<ACT>(__uint_as_float(v.y << 16), k[2], slope), fv_apply<ACT>(__uint_as_float(v.y & 0xffff0000u), k[3], slope));
    const uint32_t x2 = fv_pack(fv_apply<ACT>(__uint_as_float(v.z << 16), k[4], slope), fv_apply<ACT>(__uint_as_float(v.z & 0xffff0000u), k[5], slope));
    const uint32_t x3 = fv_pack(fv_apply<ACT>(__uint_as_float(v.w << 16), k[6], slope), fv_apply<ACT>(__uint_as_float(v.w & 0xffff0000u), k[7], slope));
    // rows outside the utterance are zero: the conv's zero padding (post-activation)
    t[u].x = ok[u] ? x0 : 0u; t[u].y = ok[u] ? x1 : 0u; t[u].z = ok[u] ? x2 : 0u; t[u].w = ok[u] ? x3 : 0u;
  }
#pragma unroll
  for (int u = 0; u < NB; ++u)
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(p[u]), "r"(t[u].x), "r"(t[u].y), "r"(t[u].z), "r"(t[u].w) : "memory");
}

// bf16 input, in place: the slot holds rows [0, nrows) as TMA wrote them (128B swizzle).  Lane = (row%4, logical
// 16-byte chunk q): 8 channels cb*64 + 8q.., coefficients in registers; physical chunk = q ^ (row & 7).
// Full batches of XB steps per warp, then single steps for the ragged tail (tile + halo is not a multiple of the
// batch).  Slot rows beyond nrows (up to the next multiple of 4) are written too: they are inside the slot and no
// MMA reads them.
template <int ACT>
__device__ __forceinline__ void fv_xf_inplace(uint32_t slot, const float4* __restrict__ coef, float slope, int cb,
                                              int nrows, int l_base, int L_in, int xw, int lane) {
  constexpr int XB = 4;
  const int q = lane & 7, sub = lane >> 3;
  float4 k[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) k[i] = fv_lds128(ptx::smem_u32(coef + cb * 64 + q * 8 + i));
  const int nsteps = (nrows + 3) >> 2;
  int s0 = xw;
  for (; s0 + (XB - 1) * FV_XF_WARPS < nsteps; s0 += FV_XF_WARPS * XB)
    fv_xf_steps<ACT, XB>(slot, k, slope, s0, nrows, l_base, L_in, q, sub);
  for (; s0 < nsteps; s0 += FV_XF_WARPS) fv_xf_steps<ACT, 1>(slot, k, slope, s0, nrows, l_base, L_in, q, sub);
}

// fp32 input: one raw stage (32 rows x 64 channels, row-major 256 B) -> slot rows [row0, row0 + 32).  A stage is
// owned by a PAIR of warps (wsub = 0/1): warp rows 4i + 2 wsub + {0, 1}, i = 0..7; lane = (row parity, 4-channel
// group q).  Two batches of four steps: loads, arithmetic, stores (see above).
template <int ACT>
__device__ __forceinline__ void fv_xf_raw(uint32_t raw, uint32_t slot, const float4 (&k)[4], float slope, int row0,
                                          int nrows, int l_base, int L_in, int wsub, int lane) {
  constexpr int NS = 4;
  const int q = lane & 15, half = lane >> 4;
  const uint32_t chunk = (uint32_t)(q >> 1), sub8 = (uint32_t)(q & 1) * 8u;
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    float4 v[NS];
    bool ok[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const int rr = 4 * (g * NS + i) + 2 * wsub + half;
      const int r = row0 + rr;
      const int l = l_base + r;
      ok[i] = r < nrows && l >= 0 && l < L_in;
      v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ok[i])
        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[i].x), "=f"(v[i].y), "=f"(v[i].z), "=f"(v[i].w)
                     : "r"(raw + (uint32_t)(rr * 256 + q * 16)));
    }
    uint2 o[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const uint32_t x0 = fv_pack(fv_apply<ACT>(v[i].x, k[0], slope), fv_apply<ACT>(v[i].y, k[1], slope));
      const uint32_t x1 = fv_pack(fv_apply<ACT>(v[i].z, k[2], slope), fv_apply<ACT>(v[i].w, k[3], slope));
      o[i].x = ok[i] ? x0 : 0u;   // branch-free so that the cosine chains of the NS steps interleave
      o[i].y = ok[i] ? x1 : 0u;
    }
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const uint32_t r = (uint32_t)(row0 + 4 * (g * NS + i) + 2 * wsub + half);
      asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(slot + r * 128u + ((chunk ^ (r & 7u)) << 4) + sub8), "r"(o[i].x),
                   "r"(o[i].y) : "memory");
    }
  }
}

// Debug instrumentation (KK_FUSED_DBG builds only): cycles spent waiting on each barrier class, per CTA.
#ifdef KK_FUSED_DBG
__device__ unsigned long long* g_fv_dbg = nullptr;
#define FV_WAIT(bar, par, acc) do { long long _t = clock64(); ptx::mbar_wait(bar, par); (acc) += clock64() - _t; } while (0)
#define FV_DBG_DECL(name) long long name = 0
#define FV_DBG_OUT(idx, v) do { if (g_fv_dbg) g_fv_dbg[(size_t)blockIdx.x * 16 + (idx)] = (unsigned long long)(v); } while (0)
#else
#define FV_WAIT(bar, par, acc) ptx::mbar_wait(bar, par)
#define FV_DBG_DECL(name)
#define FV_DBG_OUT(idx, v)
#endif

// Residual tile (32 rows x 32 channels at utterance b, row row0, channel ch0) -> staging tile, by 16-byte cp.async from
// all 32 lanes; completion = the caller's cp.async group (commit after this call, wait_group + __syncwarp before the
// tile is read; round 1's per-lane deferred mbarrier arrives cost 10 % of an epilogue-bound launch).  fp32: 128-byte rows in
// the 128B-swizzle layout (chunk ^= row & 7); bf16 (res_is_bf16): 64-byte rows, chunk ^= (row >> 1) & 3.
// Deliberately NOT a TMA load: residual boxes in the TMA queue delayed the weight and input loads behind them
// (C=256 k=11: weight-ring stalls 18 % -> 47 % of the issuer's time).
__device__ __forceinline__ void fv_issue_res(const kk_conv_desc& d, int b, int row0, int ch0, uint8_t* tile, int lane) {
  const uint32_t t = ptx::smem_u32(tile);
  if (d.res_is_bf16) {
    const int c = lane & 3;
    const __nv_bfloat16* base = reinterpret_cast<const __nv_bfloat16*>(d.res);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = i * 8 + (lane >> 2);
      const int row = row0 + r;
      const bool ok = row < d.L_res;
      const __nv_bfloat16* src = base + ((size_t)b * d.L_res + (ok ? row : 0)) * d.res_pitch + ch0 + c * 8;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(t + (uint32_t)(r * 64 + ((c ^ ((r >> 1) & 3)) << 4))),
                   "l"(src), "r"(ok ? 16u : 0u)
                   : "memory");
    }
  } else {
    const int c = lane & 7;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = i * 4 + (lane >> 3);
      const int row = row0 + r;
      const bool ok = row < d.L_res;
      const float* src = d.res + ((size_t)b * d.L_res + (ok ? row : 0)) * d.res_pitch + ch0 + c * 4;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(t + (uint32_t)(r * 128 + ((c ^ (r & 7)) << 4))),
                   "l"(src), "r"(ok ? 16u : 0u)
                   : "memory");
    }
  }
}

// contiguous share of the tile list for CTA `cta` of `ncta`
__device__ __forceinline__ void fv_tile_range(int total, int cta, int ncta, int& start, int& count) {
  const int base = total / ncta, rem = total - base * ncta;
  count = base + (cta < rem ? 1 : 0);
  start = cta * base + (cta < rem ? cta : rem);
}

// SWAP: the MMA runs "transposed" - weights as the M = 128 operand, the tile rows as N.  <MT = 2, BN = 128>: the 256
// tile rows are N = 256; <MT = 1, BN = 256> (round 2): 128-row tiles, two 128-channel M blocks per K step (same 256
// accumulator columns).  The accumulator is [channel lanes][row columns], so the epilogue is lane = channel:
// per-thread statistics, no staging round trip for the bf16 copy.  At 128 channels
// one 128x256x16 MMA replaces two 128x128x16 ones (12 KB instead of 16 KB of shared-memory
// operand reads per 128 tensor cycles).  The accumulator is then [channel lanes][row columns].
// CL = 2: CTAs run as cluster pairs that share the weight stream - each loads half of every weight slab and
// TMA-multicasts it into both shared memories (the slab sequence is the same for every tile), halving the L2 -> SM
// weight traffic, which is what bounds the 256-channel configuration (32 KB per 128 x 256 x 64 MMA block).
template <int MT, int BN, int CB, bool SWAP, int CL>
__global__ void __launch_bounds__(FV_THREADS, 1)
conv1d_adain_tc_kernel(const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmX,
                       const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmOut,
                       const __grid_constant__ CUtensorMap tmAdd, const kk_conv_desc d, const kk_adain_in a, const int m_tiles, const int total_tiles_uniform,
                       const int flags) {
  using Cfg = FvCfg<MT, BN, CB>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem + Cfg::OFF_B;
  float4* coef = reinterpret_cast<float4*>(smem + Cfg::OFF_COEF);
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);   // [3] slot transformed, MMA may read
  uint64_t* a_empty = a_full + 3;                                        // [3] MMAs on the slot completed
  uint64_t* x_full = a_empty + 3;                                        // [3] bf16 input landed in the slot
  uint64_t* raw_full = x_full + 3;                                       // [FV_MAX_RAW]
  uint64_t* raw_empty = raw_full + FV_MAX_RAW;                           // [FV_MAX_RAW]
  uint64_t* b_full = raw_empty + FV_MAX_RAW;                             // [8]
  uint64_t* b_empty = b_full + 8;                                        // [8]
  uint64_t* tmem_full = b_empty + 8;                                     // [2]
  uint64_t* tmem_empty = tmem_full + 2;                                  // [2]
  uint64_t* res_full = tmem_empty + 2;                                   // [FV_EPI_WARPS * 2] (unused since the residual copies complete by cp.async groups; kept for the layout)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_full + FV_EPI_WARPS * 2);
  int* rg_ts = reinterpret_cast<int*>(smem + Cfg::OFF_RG);              // ragged batches: first tile of each utterance
  const int* ts = d.lens ? rg_ts : nullptr;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int halo = (d.taps - 1) * d.dil;
  const int R = 128 * MT + halo;                         // slot rows a tile needs
  const int nbox = (R + FV_BOX_ROWS - 1) / FV_BOX_ROWS;  // 32-row TMA boxes per block
  const bool has_res = d.res != nullptr;
  const size_t res_esz = d.res_is_bf16 ? 2 : 4;
  const bool act_id = d.act_slope == 1.0f;   // out_act is a plain bf16 copy
  const bool has_st = d.out != nullptr || d.out_add != nullptr;   // the staging tile leaves by TMA (store and/or reduce-add)
  // fp32 input comes without a residual (host-checked): the raw ring lives in the second staging half
  // Staging half 0 is free when nothing leaves through it (no residual, no fp32 output: the "conv1" launches).  Short
  // kernels (taps <= 7) are bound by the input stream: the raw ring takes it (8 stages, two per warp pair, so that
  // one loads while the other is transformed); long kernels are bound by the MMA: the weight ring takes it.
  const bool half0_free = SWAP && d.res == nullptr && d.out == nullptr && d.out_add == nullptr;   // (the non-SWAP epilogue always stages)
  const bool deep_raw = half0_free && !a.x_is_bf16 && d.taps <= 7 && !(flags & 2);   // flags bit 1: A/B knob (off)
  const uint32_t nraw = deep_raw ? 2 * Cfg::NRAW : Cfg::NRAW;
  uint8_t* sRaw = smem + (deep_raw ? Cfg::OFF_STG : Cfg::OFF_RAW);
  const uint32_t nst = (half0_free && !deep_raw) ? Cfg::STAGES + Cfg::STG_HALF / Cfg::B_BYTES : Cfg::STAGES;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmB);
    ptx::prefetch_tmap(&tmX);
    if (has_res) ptx::prefetch_tmap(&tmRes);
    if (d.out) ptx::prefetch_tmap(&tmOut);
    if (d.out_add) ptx::prefetch_tmap(&tmAdd);
    for (int s = 0; s < 3; ++s) {
      ptx::mbar_init(&a_full[s], FV_XF_WARPS);
      ptx::mbar_init(&a_empty[s], 1);
      ptx::mbar_init(&x_full[s], 1);
    }
    for (int s = 0; s < 8; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], CL);   // every CTA of the pair has finished reading the stage
    }
    for (int s = 0; s < FV_MAX_RAW; ++s) {
      ptx::mbar_init(&raw_full[s], 1);
      ptx::mbar_init(&raw_empty[s], 2);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full[i], 1);
      ptx::mbar_init(&tmem_empty[i], FV_EPI_WARPS);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  if (warp == 3 && d.lens) rg_build(rg_ts, d, 128 * MT, lane);   // ragged batch: the tile list follows the lengths
  ptx::tc_fence_before();
  if constexpr (CL > 1) ptx::cluster_sync();   // the peer's barriers exist before anything is multicast at them
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int total_tiles = ts ? ts[d.B] : total_tiles_uniform;
  int t_start, t_count;
  fv_tile_range(total_tiles, blockIdx.x, gridDim.x, t_start, t_count);
  // the weight pipeline of a pair runs for the longer of the two tile lists (the other CTA passes phantom slabs)
  int t_pair = t_count;
  uint32_t crank = 0;
  if constexpr (CL > 1) {
    int s2, c2;
    fv_tile_range(total_tiles, blockIdx.x ^ 1, gridDim.x, s2, c2);
    t_pair = c2 > t_count ? c2 : t_count;
    crank = ptx::cluster_ctarank();
  }

  if (warp == 0) {
    // ===================== TMA producer: weight slabs, order (tile, channel block, tap) =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FV_REGS_CTRL));
    if (lane == 0) {
      uint32_t g = 0;
      FV_DBG_DECL(w_bempty);
      for (int j = 0; j < t_pair; ++j) {
        for (int cb = 0; cb < CB; ++cb) {
          for (int t = 0; t < d.taps; ++t, ++g) {
            const uint32_t s = g % nst, ph = (g / nst) & 1;
            FV_WAIT(&b_empty[s], ph ^ 1, w_bempty);
            ptx::mbar_expect_tx(&b_full[s], Cfg::B_BYTES);
            if constexpr (CL > 1)   // my half of the slab (output channels crank*BN/2 ..), delivered to both CTAs
              ptx::tma_load_2d_mc(sB + s * Cfg::B_BYTES + crank * (Cfg::B_BYTES / CL), &tmB, &b_full[s],
                                  t * d.Cw + cb * 64, (int)crank * (BN / CL), (uint16_t)((1u << CL) - 1));
            else
              ptx::tma_load_2d(sB + s * Cfg::B_BYTES, &tmB, &b_full[s], t * d.Cw + cb * 64, 0);
          }
        }
      }
      FV_DBG_OUT(0, w_bempty);
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FV_REGS_CTRL));
    if (lane == 0) {
      // SWAP: M = 128 output channels (BN / 128 such blocks), N = the tile's 128 * MT rows
      constexpr uint32_t idesc = SWAP ? ptx::umma_idesc_bf16(128, 128 * MT) : ptx::umma_idesc_bf16(128, BN);
      static_assert(!SWAP || (MT == 2 && BN == 128) || (MT == 1 && BN == 256),
                    "SWAP: 256 rows x 128 channels, or 128 rows x 2 blocks of 128 channels (256 TMEM columns per tile)");
      uint32_t gb = 0, ga = 0;
      FV_DBG_DECL(w_tempty); FV_DBG_DECL(w_afull); FV_DBG_DECL(w_bfull);
#ifdef KK_FUSED_DBG
      const long long t_begin = clock64();
#endif
      for (int j = 0; j < t_pair; ++j) {
        if (CL > 1 && j >= t_count) {   // phantom tile: keep the pair's weight ring turning, no MMAs
          for (int i = 0; i < CB * d.taps; ++i, ++gb) {
            const uint32_t s = gb % nst, ph = (gb / nst) & 1;
            FV_WAIT(&b_full[s], ph, w_bfull);
            ptx::umma_commit_mc(&b_empty[s], (uint16_t)((1u << CL) - 1));
          }
          continue;
        }
        const uint32_t buf = j & 1, aph = (j >> 1) & 1;
        FV_WAIT(&tmem_empty[buf], aph ^ 1, w_tempty);   // epilogue has drained this accumulator set
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * (MT * BN);
        for (int cb = 0; cb < CB; ++cb, ++ga) {
          const uint32_t slot = ga % Cfg::NSLOT, sph = (ga / Cfg::NSLOT) & 1;
          FV_WAIT(&a_full[slot], sph, w_afull);
          ptx::tc_fence_after();
          const uint32_t sa0 = ptx::smem_u32(smem + slot * Cfg::SLOT_BYTES);
          for (int t = 0; t < d.taps; ++t, ++gb) {
            const uint32_t s = gb % nst, ph = (gb / nst) & 1;
            FV_WAIT(&b_full[s], ph, w_bfull);
            ptx::tc_fence_after();
            const uint32_t sb = ptx::smem_u32(sB + s * Cfg::B_BYTES);
            if (SWAP) {
              const uint32_t sx = sa0 + (uint32_t)(t * d.dil) * 128u;                // rows t*dil .. of the slot
              // descriptors once per slab; a K step of 16 elements = 32 bytes = +2 in the (address >> 4) field (one
              // integer add instead of rebuilding the descriptor; measured neutral - profiles/r02_fused_experiments.txt)
              const uint64_t xd0 = ptx::umma_desc(sx, 16, 1024, ptx::UMMA_LAYOUT_SW128);
              const uint64_t wd0 = ptx::umma_desc(sb, 16, 1024, ptx::UMMA_LAYOUT_SW128);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
#pragma unroll
                for (int mb = 0; mb < BN / 128; ++mb)     // 128-channel blocks of the slab: accumulator columns mb * rows
                  ptx::umma_bf16(tacc + mb * (128 * MT), wd0 + (uint64_t)(mb * 1024 + k * 2), xd0 + (uint64_t)(k * 2), idesc,
                                 (cb | t | k) != 0 ? 1u : 0u);
              }
            } else {
              const uint64_t bd0 = ptx::umma_desc(sb, 16, 1024, ptx::UMMA_LAYOUT_SW128);
              const uint64_t ad0 = ptx::umma_desc(sa0 + (uint32_t)(t * d.dil) * 128u, 16, 1024, ptx::UMMA_LAYOUT_SW128);
#pragma unroll
              for (int mt = 0; mt < MT; ++mt) {           // row-shifted views of the slot: +128 rows = +1024 in the field
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  ptx::umma_bf16(tacc + mt * BN, ad0 + (uint64_t)(mt * 1024 + k * 2), bd0 + (uint64_t)(k * 2), idesc,
                                 (cb | t | k) != 0 ? 1u : 0u);
              }
            }
            if constexpr (CL > 1) ptx::umma_commit_mc(&b_empty[s], (uint16_t)((1u << CL) - 1));
            else ptx::umma_commit(&b_empty[s]);
          }
          ptx::umma_commit(&a_empty[slot]);
        }
        ptx::umma_commit(&tmem_full[buf]);
      }
      FV_DBG_OUT(1, w_tempty); FV_DBG_OUT(2, w_afull); FV_DBG_OUT(3, w_bfull);
#ifdef KK_FUSED_DBG
      FV_DBG_OUT(10, clock64() - t_begin);
#endif
    }
  } else if (warp == 2) {
    // ===================== TMA producer: input rows (tile + halo), one 64-channel block at a time ============
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FV_REGS_CTRL));
    if (lane == 0) {
      uint32_t ga = 0, gr = 0;
      FV_DBG_DECL(w_xprod);
      RgMap map(ts, m_tiles), map_next(ts, m_tiles);
      for (int j = 0; j < t_count; ++j) {
        const int tile = t_start + j;
        int b, mtile;
        map.locate(tile, b, mtile);
        const int l_base = mtile * (128 * MT) - d.pad_left;
        if ((flags & 1) && j + 1 < t_count) {
          // pull the NEXT tile's rows (input and residual) into L2 now: the TMA loads one tile later then see
          // L2-hit latency instead of HBM latency
          int nb_, nmt;
          map_next.locate(tile + 1, nb_, nmt);
          const int nm0 = nmt * (128 * MT);
          const size_t xesz = a.x_is_bf16 ? 2 : 4;
          int l0 = nm0 - d.pad_left + (nb_ == b ? R - halo : 0), l1 = nm0 - d.pad_left + R;   // rows not already fetched
          l0 = l0 < 0 ? 0 : l0;
          l1 = l1 > d.L_in ? d.L_in : l1;
          if (l1 > l0)
            ptx::prefetch_l2_bulk(reinterpret_cast<const char*>(a.x) + ((size_t)nb_ * d.L_in + l0) * a.x_pitch * xesz,
                                  (uint32_t)((size_t)(l1 - l0) * a.x_pitch * xesz));
          int r1 = nm0 + 128 * MT;
          r1 = r1 > d.M ? d.M : r1;
          if (has_res && !(flags & 4))
            ptx::prefetch_l2_bulk(reinterpret_cast<const uint8_t*>(d.res) + ((size_t)nb_ * d.L_res + nm0) * d.res_pitch * res_esz,
                                  (uint32_t)((size_t)(r1 - nm0) * d.res_pitch * res_esz));
        }
        for (int cb = 0; cb < CB; ++cb, ++ga) {
          const uint32_t slot = ga % Cfg::NSLOT, sph = (ga / Cfg::NSLOT) & 1;
          if (a.x_is_bf16) {
            FV_WAIT(&a_empty[slot], sph ^ 1, w_xprod);     // the MMAs that read this slot have completed
            ptx::mbar_expect_tx(&x_full[slot], nbox * FV_XBOX_BYTES);
            for (int i = 0; i < nbox; ++i)
              ptx::tma_load_3d(smem + slot * Cfg::SLOT_BYTES + i * FV_XBOX_BYTES, &tmX, &x_full[slot], cb * 64,
                               l_base + i * FV_BOX_ROWS, b);
          } else {
            for (int i = 0; i < nbox; ++i, ++gr) {
              const uint32_t s = gr % nraw, ph = (gr / nraw) & 1;
              FV_WAIT(&raw_empty[s], ph ^ 1, w_xprod);
              ptx::mbar_expect_tx(&raw_full[s], FV_RAW_BYTES);
              ptx::tma_load_3d(sRaw + s * FV_RAW_BYTES, &tmX, &raw_full[s], cb * 64, l_base + i * FV_BOX_ROWS, b);
            }
          }
        }
      }
      FV_DBG_OUT(4, w_xprod);
    }
  } else if (warp >= FV_XF_WARP0 && warp < FV_XF_WARP0 + FV_XF_WARPS) {
    // ===================== transform warps: x -> act(a x + b) -> bf16 swizzled slot =====================
    if constexpr (FvRegs<SWAP>::XF < 96) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FvRegs<SWAP>::XF));
    const int xw = warp - FV_XF_WARP0;
    const int xt = threadIdx.x - FV_XF_WARP0 * 32;
    const bool snake = a.act == KK_ACT_SNAKE || a.act == KK_ACT_SNAKE_FAST;
    int cur_b = -1;
    uint32_t ga = 0, gr = 0;
    FV_DBG_DECL(w_xdata); FV_DBG_DECL(w_xslot);
    RgMap map(ts, m_tiles);
    for (int j = 0; j < t_count; ++j) {
      const int tile = t_start + j;
      int b, mtile;
      map.locate(tile, b, mtile);
      const int l_base = mtile * (128 * MT) - d.pad_left;
      // ragged batch: rows of this utterance (rows beyond are padding: zero after the activation, not in the divisor)
      const int Lb = ragged_len(d.lens, d.len_mul, d.len_add, b, d.L_in);
      if (b != cur_b) {
        asm volatile("bar.sync 2, %0;" ::"n"(FV_XF_THREADS) : "memory");   // everyone is done with the old table
        for (int c = xt; c < CB * 64; c += FV_XF_THREADS) {
          float4 k = make_float4(1.f, 0.f, 1.f, 1.f);
          if (a.stats) {
            const double s = a.stats[((size_t)b * d.C_in + c) * 2], ss = a.stats[((size_t)b * d.C_in + c) * 2 + 1];
            const double mean = s / Lb;
            double var = ss / Lb - mean * mean;
            if (var < 0.0) var = 0.0;
            const float rstd = 1.0f / sqrtf((float)var + 1e-5f);
            const float g1 = 1.0f + a.h[(size_t)b * a.h_stride + c];
            k.x = g1 * rstd;
            k.y = a.h[(size_t)b * a.h_stride + d.C_in + c] - k.x * (float)mean;
          }
          if (snake) { const float al = a.alpha[c]; k.y += 0.5f / al; k.z = 2.0f * al; k.w = -0.5f / al; }
          coef[c] = k;
        }
        asm volatile("bar.sync 2, %0;" ::"n"(FV_XF_THREADS) : "memory");
        cur_b = b;
      }
      for (int cb = 0; cb < CB; ++cb, ++ga) {
        const uint32_t slot = ga % Cfg::NSLOT, sph = (ga / Cfg::NSLOT) & 1;
        const uint32_t sp = ptx::smem_u32(smem + slot * Cfg::SLOT_BYTES);
        if (a.x_is_bf16) {
          FV_WAIT(&x_full[slot], sph, w_xdata);
          if (snake) fv_xf_inplace<KK_ACT_SNAKE>(sp, coef, a.slope, cb, R, l_base, Lb, xw, lane);
          else if (a.act == KK_ACT_LEAKY) fv_xf_inplace<KK_ACT_LEAKY>(sp, coef, a.slope, cb, R, l_base, Lb, xw, lane);
          else if (a.stats) fv_xf_inplace<KK_ACT_NONE>(sp, coef, a.slope, cb, R, l_base, Lb, xw, lane);
          // else: no normalisation and no activation - the operand is exactly what TMA wrote (rows outside the
          // utterance were zero-filled by the tensor map), so the slot is handed to the MMA as is
        } else {
          float4 k[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) k[i] = fv_lds128(ptx::smem_u32(coef + cb * 64 + (lane & 15) * 4 + i));
          FV_WAIT(&a_empty[slot], sph ^ 1, w_xslot);   // the MMAs that read this slot have completed
          for (int i = 0; i < nbox; ++i, ++gr) {
            if ((int)(gr & 3u) != (xw >> 1)) continue;   // raw stage gr belongs to warp pair gr % 4
            const uint32_t s = gr % nraw, ph = (gr / nraw) & 1;
            FV_WAIT(&raw_full[s], ph, w_xdata);
            const uint32_t rp = ptx::smem_u32(sRaw + s * FV_RAW_BYTES);
            if (snake) fv_xf_raw<KK_ACT_SNAKE>(rp, sp, k, a.slope, i * FV_BOX_ROWS, R, l_base, Lb, xw & 1, lane);
            else if (a.act == KK_ACT_LEAKY) fv_xf_raw<KK_ACT_LEAKY>(rp, sp, k, a.slope, i * FV_BOX_ROWS, R, l_base, Lb, xw & 1, lane);
            else fv_xf_raw<KK_ACT_NONE>(rp, sp, k, a.slope, i * FV_BOX_ROWS, R, l_base, Lb, xw & 1, lane);
            __syncwarp();                              // the warp's reads of the stage are complete
            if (lane == 0) ptx::mbar_arrive(&raw_empty[s]);
          }
        }
        ptx::fence_proxy_async();                      // generic-proxy stores -> visible to the tensor core
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&a_full[slot]);
      }
    }
    if (warp == FV_XF_WARP0 && lane == 0) { FV_DBG_OUT(5, w_xdata); FV_DBG_OUT(6, w_xslot); }
  } else if (warp >= FV_EPI_WARP0) {
    // ===================== epilogue warps =====================
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(FvRegs<SWAP>::EPI));
    if constexpr (SWAP) {
    // Transposed accumulator: TMEM lane = output channel, column = tile row.  Work item n = (tile j, 32-row chunk
    // rc) of this warp: channels 32*lg..+31 (lane = channel) x rows half*128 + rc*32..+31.  fp32 results are
    // written into the staging tile (in place over the residual) and leave by TMA (store / reduce-add; register
    // stores measured slower); the bf16 copy goes straight from registers, one 64-byte row segment per store.
    // Two shapes share this code: 256 rows x 128 channels (`half` = which 128 rows) and 128 rows x 256 channels (`half`
    // = which 128-channel block); either way the warp's accumulator columns are half * 128 + rc * 32.
    const int ew = warp - FV_EPI_WARP0;
    const int lg = warp & 3;
    const int half = ew >> 2;
    constexpr int TR = 128 * MT;                       // rows per tile
    constexpr bool CHB = BN == 256;                    // `half` selects a channel block (else a row half)
    constexpr int HR = CHB ? TR : TR / 2;              // rows a warp covers per tile, in NRC chunks of 32
    constexpr int NRC = HR / 32;
    const int row_half = CHB ? 0 : half * HR;
    uint8_t* stg0 = smem + Cfg::OFF_STG + ew * 4096;   // [32 rows][32 channels] fp32, 16-byte chunk ^= row & 7
    const bool do_stats = d.stats != nullptr;
    const bool via_smem = has_st;   // fp32 results leave by TMA (store and/or reduce-add) from the staging tile
    const bool res16 = d.res_is_bf16 != 0;
    const int ch0 = lg * 32 + (CHB ? half * 128 : 0), chan = ch0 + lane;
    const float bias_s = (d.bias ? __ldg(d.bias + chan) : 0.f) * d.scale;   // (v + bias) * scale = fma(v, scale, bias * scale)
    const int n_items = t_count * NRC;
#ifdef KK_TUNE   // lab knobs (KK_L2PF bits 4 / 5): residual fetched but not added / added (from stale staging) but not fetched
    const bool dbg_noadd = (flags & 16) != 0, dbg_nofetch = (flags & 32) != 0;
    const bool dbg_notmem = (flags & 256) != 0, dbg_nostore = (flags & 512) != 0, dbg_nostats = (flags & 1024) != 0;   // knock-outs
#else
    constexpr bool dbg_noadd = false, dbg_nofetch = false, dbg_notmem = false, dbg_nostore = false, dbg_nostats = false;
#endif
    RgMap map(ts, m_tiles), map_res(ts, m_tiles);
    // bf16 residual: a running per-lane cursor.  The NRC chunks of a tile are consecutive rows, and a lane's four
    // 16-byte pieces of a chunk are 8 rows apart, so within a tile the source pointer just advances by 8 rows per
    // copy; the tile lookup and the 64-bit address arithmetic happen once per tile.  (Measured: with the address of
    // every copy rebuilt from (b, row, pitch) the fetch cost an epilogue-bound launch 0.15 of 0.72 ms - instructions
    // on the critical warps, not bytes: prefetching the rows into L2 changed nothing.)
    const uint8_t* rcur = nullptr;
    int rleft = 0;
    const size_t rstep8 = (size_t)8 * d.res_pitch * 2;
    const uint32_t rdst = (uint32_t)((lane >> 2) * 64 + (((lane & 3) ^ ((lane >> 3) & 3)) << 4));   // + i * 512: (r >> 1) & 3 = (lane >> 3) & 3 for r = 8 i + (lane >> 2)
    // Completion by cp.async groups (one per item, committed even when there is nothing left to fetch, so that
    // "all but the newest group" always means "item n has landed"): the mbarrier form - 32 per-lane arrives and a
    // try_wait per chunk on the warps the launch waits for - cost 0.07 of 0.71 ms on the k = 3 conv2 of the 128-channel
    // stage with the copies themselves switched off.
    auto issue_res = [&](int n) {   // all lanes: fetch the residual tile of item n into staging buffer n & 1
      if (dbg_nofetch) return;
      if (n >= n_items) { asm volatile("cp.async.commit_group;" ::: "memory"); return; }
      if (res16) {
        if (n % NRC == 0) {
          int b, mtile;
          map_res.locate(t_start + n / NRC, b, mtile);
          const int row = mtile * TR + row_half + (lane >> 2);
          rcur = reinterpret_cast<const uint8_t*>(d.res) + (((size_t)b * d.L_res + row) * d.res_pitch + ch0 + (lane & 3) * 8) * 2;
          rleft = d.L_res - row;
        }
        const uint32_t t = ptx::smem_u32(stg0 + (n & 1) * Cfg::STG_HALF) + rdst;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const bool ok = rleft > 8 * i;
#ifdef KK_TUNE
          if ((flags & 64) && i) continue;          // lab: one copy of four
          if ((flags & 128)) continue;              // lab: no copies, barrier protocol only
#endif
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(t + (uint32_t)(i * 512)),
                       "l"(ok ? rcur + i * rstep8 : reinterpret_cast<const uint8_t*>(d.res)), "r"(ok ? 16u : 0u)
                       : "memory");
        }
        rcur += 4 * rstep8;
        rleft -= 32;
        asm volatile("cp.async.commit_group;" ::: "memory");
        return;
      }
      const int tile = t_start + n / NRC, rc = n % NRC;
      int b, mtile;
      map_res.locate(tile, b, mtile);
      const int row0 = mtile * TR + row_half + rc * 32;
      fv_issue_res(d, b, row0, ch0, stg0 + (n & 1) * Cfg::STG_HALF, lane);
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (has_res) { issue_res(0); issue_res(1); }
    double acc_s = 0.0, acc_q = 0.0;   // fp64 running instance-norm sums of this thread's channel
    int stat_b = -1;
    FV_DBG_DECL(w_tfull);
    auto flush_stats = [&](int b) {
      if (b < 0) return;
      atomicAdd(d.stats + ((size_t)b * d.C_out + chan) * 2, acc_s);
      atomicAdd(d.stats + ((size_t)b * d.C_out + chan) * 2 + 1, acc_q);
      acc_s = acc_q = 0.0;
    };
    // row-strided bf16 stores of the thread's channel; `pitch` is a literal on the hot path (immediate offsets)
    auto store_bf16 = [&](__nv_bfloat16* ap, const int pitch, const float (&o)[32], int nvalid) {
      if (nvalid >= 32) {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) ap[jj * pitch] = __float2bfloat16_rn(act_id ? o[jj] : leaky(o[jj], d.act_slope));
      } else {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj)
          if (jj < nvalid) ap[jj * pitch] = __float2bfloat16_rn(act_id ? o[jj] : leaky(o[jj], d.act_slope));
      }
    };
    for (int j = 0; j < t_count; ++j) {
      const int tile = t_start + j;
      int b, mtile;
      map.locate(tile, b, mtile);
      const int tm0 = mtile * TR + row_half;
      const int Mb = ragged_len(d.lens, d.len_mul, d.len_add, b, d.M);   // rows of this utterance (ragged batches)
      if (do_stats && b != stat_b) { flush_stats(stat_b); stat_b = b; }
      const uint32_t buf = j & 1, aph = (j >> 1) & 1;
      FV_WAIT(&tmem_full[buf], aph, w_tfull);
      ptx::tc_fence_after();
      const uint32_t tacc = tmem_base + buf * (MT * BN) + (uint32_t)(half * HR) + ((uint32_t)(lg * 32) << 16);
#pragma unroll 1
      for (int rc = 0; rc < NRC; ++rc) {
        const int n = j * NRC + rc;
        const int p = has_res ? (n & 1) : 0;
        float* st = reinterpret_cast<float*>(stg0 + p * Cfg::STG_HALF);
        const int row0 = tm0 + rc * 32;
        const int nvalid = Mb - row0;            // rows of this chunk inside the utterance (may be <= 0 or >= 32)
        const int nstore = d.M - row0;           // rows of this chunk inside the tensors: padding rows are stored as 0
        float v[32];
        if (!dbg_notmem) ptx::tmem_ld_32x32(tacc + (uint32_t)(rc * 32), v);
        else {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) v[jj] = 0.f;
        }
        if (has_res && !dbg_nofetch) {
          asm volatile("cp.async.wait_group 1;" ::: "memory");   // this lane's copies of item n (item n + 1 may be in flight)
          __syncwarp();                                           // ... and every other lane's
        } else if (via_smem) {
          if (lane == 0) ptx::bulk_wait_read<0>();       // the previous TMA operation has finished reading this tile
          __syncwarp();
        }
        ptx::tmem_ld_wait();
        if (rc == NRC - 1) {                             // last TMEM read of this tile: hand the accumulator back
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tmem_empty[buf]);
        }
        // lane = channel: element (row jj, channel) of the swizzled tile; conflict-free for every jj.
        const uint32_t sbase = ptx::smem_u32(st) + (uint32_t)(lane & 3) * 4u;
        const uint32_t lchunk = (uint32_t)(lane >> 2);
        if (has_res && res16 && !dbg_noadd) {
          // bf16 residual tile: 64-byte rows, lane = channel reads 2 bytes of row jj (one contiguous row per instruction)
          const uint32_t hbase = ptx::smem_u32(st) + (uint32_t)(lane & 7) * 2u;
          const uint32_t hchunk = (uint32_t)(lane >> 3);
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) {
            uint16_t r;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(hbase + (uint32_t)jj * 64u + ((hchunk ^ (uint32_t)((jj >> 1) & 3)) << 4)));
            v[jj] += __uint_as_float((uint32_t)r << 16);
          }
          if (via_smem) __syncwarp();   // the fp32 result tile overlays the (smaller) bf16 residual tile
        } else if (has_res) {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) {
            float r;
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(r) : "r"(sbase + (uint32_t)jj * 128u + ((lchunk ^ (uint32_t)(jj & 7)) << 4)));
            v[jj] += r;
          }
        }
        float s_ = 0.f, q_ = 0.f;
        if (nvalid >= 32) {   // whole chunk inside the utterance (all but the last tile of an utterance): no masks
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) v[jj] = fmaf(v[jj], d.scale, bias_s);
        } else {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) v[jj] = jj < nvalid ? fmaf(v[jj], d.scale, bias_s) : 0.f;
        }
        if (do_stats && !dbg_nostats) {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj) {
            s_ += v[jj];
            q_ = fmaf(v[jj], v[jj], q_);
          }
          acc_s += (double)s_;
          acc_q += (double)q_;
        }
        if (via_smem) {
#pragma unroll
          for (int jj = 0; jj < 32; ++jj)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(sbase + (uint32_t)jj * 128u + ((lchunk ^ (uint32_t)(jj & 7)) << 4)), "f"(v[jj]) : "memory");
          ptx::fence_proxy_async();                      // staging writes -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            if (d.out) ptx::tma_store_3d(&tmOut, st, ch0, row0, b);   // rows >= L_out are clipped by the tensor map
            if (d.out_add) ptx::tma_reduce_add_3d(&tmAdd, st, ch0, row0, b);
            ptx::bulk_commit();
          }
        }
        if (d.out_act && nstore > 0 && !dbg_nostore) {
          __nv_bfloat16* ap = (__nv_bfloat16*)d.out_act + ((size_t)b * d.L_out + row0) * d.act_pitch + chan;
          if (d.act_pitch == 128) store_bf16(ap, 128, v, nstore);
          else if (d.act_pitch == 256) store_bf16(ap, 256, v, nstore);
          else store_bf16(ap, d.act_pitch, v, nstore);
        }
        if (has_res) {
          if (via_smem && lane == 0) ptx::bulk_wait_read<0>();   // the TMA engine is done reading this staging tile
          __syncwarp();                                          // ... and so is every lane
          issue_res(n + 2);
        }
      }
    }
    if (do_stats) flush_stats(stat_b);
    if (ew == 0 && lane == 0) { FV_DBG_OUT(7, w_tfull); FV_DBG_OUT(8, 0); }
    if (via_smem && lane == 0) ptx::bulk_wait<0>();      // all TMA stores have completed before the CTA retires
    } else {
    // Work item n = (tile j, chunk c) of this warp: 32 rows (TMEM lane group lg of sub-tile mt) x 32 columns.
    const int ew = warp - FV_EPI_WARP0;
    const int lg = warp & 3;                 // TMEM lane group this warp may access
    const int half = ew >> 2;
    const int mt = MT == 2 ? half : 0;
    constexpr int NCH = MT == 2 ? BN / 32 : BN / 64;   // chunks per warp and tile
    const int chunk0 = MT == 2 ? 0 : half * NCH;
    uint8_t* stg0 = smem + Cfg::OFF_STG + ew * 4096;   // staging tile p at stg0 + p * STG_HALF (1024-aligned)
    const bool do_stats = d.stats != nullptr;
    const bool out_aligned_act = d.out_act != nullptr;
    const bool res16 = d.res_is_bf16 != 0;
    const int n_items = t_count * NCH;
    RgMap map(ts, m_tiles), map_res(ts, m_tiles);

    auto issue_res = [&](int n) {   // all lanes: fetch the residual tile of item n into staging buffer n & 1
      if (n < n_items) {            // completion by cp.async groups, one per item (see the transposed epilogue above)
        const int tile = t_start + n / NCH, c = n % NCH;
        int b, mtile;
        map_res.locate(tile, b, mtile);
        const int m0 = mtile * (128 * MT) + mt * 128 + lg * 32;
        fv_issue_res(d, b, m0, (chunk0 + c) * 32, stg0 + (n & 1) * Cfg::STG_HALF, lane);
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    if (has_res) { issue_res(0); issue_res(1); }

    double acc_s[NCH], acc_q[NCH];   // fp64 running instance-norm sums of column `lane` of each chunk
#pragma unroll
    for (int c = 0; c < NCH; ++c) acc_s[c] = acc_q[c] = 0.0;
    int stat_b = -1;
    FV_DBG_DECL(w_tfull);
    auto flush_stats = [&](int b) {
      if (b < 0) return;
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int col = (chunk0 + c) * 32 + lane;
        atomicAdd(d.stats + ((size_t)b * d.C_out + col) * 2, acc_s[c]);
        atomicAdd(d.stats + ((size_t)b * d.C_out + col) * 2 + 1, acc_q[c]);
        acc_s[c] = acc_q[c] = 0.0;
      }
    };

    for (int j = 0; j < t_count; ++j) {
      const int tile = t_start + j;
      int b, mtile;
      map.locate(tile, b, mtile);
      const int m0 = mtile * (128 * MT) + mt * 128 + lg * 32;   // first row of the warp
      const int Mb = ragged_len(d.lens, d.len_mul, d.len_add, b, d.M);   // rows of this utterance (ragged batches)
      if (do_stats && b != stat_b) { flush_stats(stat_b); stat_b = b; }
      const uint32_t buf = j & 1, aph = (j >> 1) & 1;
      FV_WAIT(&tmem_full[buf], aph, w_tfull);
      ptx::tc_fence_after();
      const uint32_t tacc = tmem_base + buf * (MT * BN) + mt * BN + ((uint32_t)(lg * 32) << 16);
      const int m = m0 + lane;
      const bool row_ok = m < Mb;              // rows beyond the utterance are stored as zeros
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int n = j * NCH + c;
        const int p = has_res ? (n & 1) : 0;
        float* st = reinterpret_cast<float*>(stg0 + p * Cfg::STG_HALF);   // [32 rows][8 chunks of 4 floats], chunk ^= row & 7
        const int cbase = (chunk0 + c) * 32;
        float v[32];
        ptx::tmem_ld_32x32(tacc + (uint32_t)cbase, v);
        if (has_res) {
          asm volatile("cp.async.wait_group 1;" ::: "memory");   // this lane's copies of item n (item n + 1 may be in flight)
          __syncwarp();                                           // ... and every other lane's
        } else if (has_st) {
          if (lane == 0) ptx::bulk_wait_read<0>();       // the previous TMA store has finished reading this tile
          __syncwarp();
        }
        ptx::tmem_ld_wait();
        if (c == NCH - 1) {                              // last TMEM read of this tile: hand the accumulators back
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&tmem_empty[buf]);
        }
        const bool rd = has_res && !res16;
        uint4 rb[4];
        if (has_res && res16) {   // bf16 residual: this lane's 64-byte row, read whole before the fp32 tile overlays it
#pragma unroll
          for (int i = 0; i < 4; ++i)
            rb[i] = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(st) + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4));
          __syncwarp();
        } else {
#pragma unroll
          for (int i = 0; i < 4; ++i) rb[i] = make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2) {   // loads batched ahead of the stores, four 16-byte chunks at a time
          float4 rr[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int c4 = h2 * 4 + i;
            const uint32_t lo = (c4 & 1) ? rb[c4 >> 1].z : rb[c4 >> 1].x, hi = (c4 & 1) ? rb[c4 >> 1].w : rb[c4 >> 1].y;
            rr[i] = make_float4(__uint_as_float(lo << 16), __uint_as_float(lo & 0xffff0000u), __uint_as_float(hi << 16),
                                __uint_as_float(hi & 0xffff0000u));
            if (rd) rr[i] = *reinterpret_cast<const float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2));
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int c4 = h2 * 4 + i;
            float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);
            if (d.bias) {
              if (cbase + c4 * 4 + 4 <= d.C_out) bs = __ldg(reinterpret_cast<const float4*>(d.bias + cbase + c4 * 4));
              else {   // ragged last channels (C_out not a multiple of 4 / 32)
                if (cbase + c4 * 4 + 0 < d.C_out) bs.x = __ldg(d.bias + cbase + c4 * 4 + 0);
                if (cbase + c4 * 4 + 1 < d.C_out) bs.y = __ldg(d.bias + cbase + c4 * 4 + 1);
                if (cbase + c4 * 4 + 2 < d.C_out) bs.z = __ldg(d.bias + cbase + c4 * 4 + 2);
              }
            }
            rr[i].x = row_ok ? (v[c4 * 4 + 0] + bs.x + rr[i].x) * d.scale : 0.f;
            rr[i].y = row_ok ? (v[c4 * 4 + 1] + bs.y + rr[i].y) * d.scale : 0.f;
            rr[i].z = row_ok ? (v[c4 * 4 + 2] + bs.z + rr[i].z) * d.scale : 0.f;
            rr[i].w = row_ok ? (v[c4 * 4 + 3] + bs.w + rr[i].w) * d.scale : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int c4 = h2 * 4 + i;
            *reinterpret_cast<float4*>(st + lane * 32 + ((c4 ^ (lane & 7)) << 2)) = rr[i];
          }
        }
        if (has_st) ptx::fence_proxy_async();            // staging writes -> visible to the TMA store
        __syncwarp();
        if (has_st && lane == 0) {
          if (d.out) ptx::tma_store_3d(&tmOut, st, cbase, m0, b);   // rows >= L_out are clipped by the tensor map
          if (d.out_add) ptx::tma_reduce_add_3d(&tmAdd, st, cbase, m0, b);
          ptx::bulk_commit();
        }
        if (out_aligned_act) {
#pragma unroll
          for (int it = 0; it < 4; ++it) {
            const int r = it * 8 + (lane >> 2), c8 = lane & 3;
            if (m0 + r < d.M) {
              const float4 t0 = *reinterpret_cast<const float4*>(st + r * 32 + (((2 * c8) ^ (r & 7)) << 2));
              const float4 t1 = *reinterpret_cast<const float4*>(st + r * 32 + (((2 * c8 + 1) ^ (r & 7)) << 2));
              uint4 q4;
              if (act_id) {
                q4.x = fv_pack(t0.x, t0.y); q4.y = fv_pack(t0.z, t0.w);
                q4.z = fv_pack(t1.x, t1.y); q4.w = fv_pack(t1.z, t1.w);
              } else {
                q4.x = fv_pack(leaky(t0.x, d.act_slope), leaky(t0.y, d.act_slope));
                q4.y = fv_pack(leaky(t0.z, d.act_slope), leaky(t0.w, d.act_slope));
                q4.z = fv_pack(leaky(t1.x, d.act_slope), leaky(t1.y, d.act_slope));
                q4.w = fv_pack(leaky(t1.z, d.act_slope), leaky(t1.w, d.act_slope));
              }
              *reinterpret_cast<uint4*>((__nv_bfloat16*)d.out_act + ((size_t)b * d.L_out + m0 + r) * d.act_pitch + cbase + c8 * 8) = q4;
            }
          }
        }
        if (do_stats) {
          // column sums over the warp's 32 rows (lane i = column i; the swizzle keeps every read conflict-free)
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
          const int cc = lane >> 2, ci = lane & 3;
#pragma unroll
          for (int r = 0; r < 32; r += 2) {
            const float a0 = st[r * 32 + ((cc ^ (r & 7)) << 2) + ci];
            const float a1 = st[(r + 1) * 32 + ((cc ^ ((r + 1) & 7)) << 2) + ci];
            s0 += a0; q0 = fmaf(a0, a0, q0);
            s1 += a1; q1 = fmaf(a1, a1, q1);
          }
          acc_s[c] += (double)(s0 + s1);
          acc_q[c] += (double)(q0 + q1);
        }
        if (has_res) {
          if (has_st && lane == 0) ptx::bulk_wait_read<0>();   // the TMA store is done reading this staging tile
          __syncwarp();                                        // ... and so is every lane
          issue_res(n + 2);
        } else {
          __syncwarp();
        }
      }
    }
    if (do_stats) flush_stats(stat_b);
    if (ew == 0 && lane == 0) { FV_DBG_OUT(7, w_tfull); FV_DBG_OUT(8, 0); }
    if (has_st && lane == 0) ptx::bulk_wait<0>();        // all TMA stores have completed before the CTA retires
    }  // !SWAP
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(FV_REGS_CTRL));   // spare warp of warpgroup 0
  }
  ptx::tc_fence_before();
  if constexpr (CL > 1) ptx::cluster_sync();   // nothing of the peer's is still in flight towards this CTA
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

}  // namespace kk

// ================================================================== host side
namespace kk {

int validate_conv_desc(const kk_conv_desc* d, const char* who);                            // conv_f32.cu
int make_tmap(CUtensorMap* m, CUtensorMapDataType dt, const void* base, int rank, const cuuint64_t* dims,
              const cuuint64_t* strides, const cuuint32_t* box, CUtensorMapSwizzle swz);   // conv_tc.cu
int num_sms();                                                                             // conv_tc.cu

// Experiment knob: the environment variable `name` in -DKK_TUNE builds (KK_TUNE=1 python -m mlx_audio_b200._build), the
// default otherwise - the product library reads no environment.
static inline int fv_tune_int(const char* name, int dflt) {
#ifdef KK_TUNE
  const char* v = getenv(name);
  return v ? atoi(v) : dflt;
#else
  (void)name;
  return dflt;
#endif
}

template <int MT, int BN, int CB, bool SWAP, int CL = 1>
static int launch_fused(const kk_adain_in* a, const void* w, const kk_conv_desc* d, cudaStream_t st) {
  using Cfg = FvCfg<MT, BN, CB>;
  CUtensorMap tmB, tmX, tmRes, tmOut, tmAdd;
  int rc;
  {
    cuuint64_t dims[2] = {(cuuint64_t)d->taps * d->Cw, (cuuint64_t)d->N};
    cuuint64_t strides[1] = {(cuuint64_t)d->taps * d->Cw * 2};
    cuuint32_t box[2] = {64, BN / CL};   // CL > 1: every CTA of the pair loads (and multicasts) its share of a slab
    if ((rc = make_tmap(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, w, 2, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  {
    const cuuint64_t esz = a->x_is_bf16 ? 2 : 4;
    cuuint64_t dims[3] = {(cuuint64_t)d->C_in, (cuuint64_t)d->L_in, (cuuint64_t)d->B};
    cuuint64_t strides[2] = {(cuuint64_t)a->x_pitch * esz, (cuuint64_t)d->L_in * a->x_pitch * esz};
    cuuint32_t box[3] = {64, FV_BOX_ROWS, 1};
    rc = a->x_is_bf16
             ? make_tmap(&tmX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, a->x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B)
             : make_tmap(&tmX, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, a->x, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc) return rc;
  }
  tmRes = tmB;
  tmOut = tmB;
  if (d->res && !d->res_is_bf16) {
    cuuint64_t dims[3] = {(cuuint64_t)d->C_out, (cuuint64_t)d->L_res, (cuuint64_t)d->B};
    cuuint64_t strides[2] = {(cuuint64_t)d->res_pitch * 4, (cuuint64_t)d->L_res * d->res_pitch * 4};
    cuuint32_t box[3] = {32, 32, 1};
    if ((rc = make_tmap(&tmRes, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d->res, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  if (d->out) {
    cuuint64_t dims[3] = {(cuuint64_t)d->C_out, (cuuint64_t)d->L_out, (cuuint64_t)d->B};
    cuuint64_t strides[2] = {(cuuint64_t)d->out_pitch * 4, (cuuint64_t)d->L_out * d->out_pitch * 4};
    cuuint32_t box[3] = {32, 32, 1};
    if ((rc = make_tmap(&tmOut, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d->out, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  tmAdd = tmB;
  if (d->out_add) {
    cuuint64_t dims[3] = {(cuuint64_t)d->C_out, (cuuint64_t)d->L_out, (cuuint64_t)d->B};
    cuuint64_t strides[2] = {(cuuint64_t)d->add_pitch * 4, (cuuint64_t)d->L_out * d->add_pitch * 4};
    cuuint32_t box[3] = {32, 32, 1};
    if ((rc = make_tmap(&tmAdd, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, d->out_add, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    KK_CUDA(cudaFuncSetAttribute(conv1d_adain_tc_kernel<MT, BN, CB, SWAP, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Cfg::SMEM_BYTES));
    attr_set = true;
  }
  const int m_tiles = cdiv(d->M, 128 * MT);
  const long long total = (long long)m_tiles * d->B;
  KK_CHECK_ARG(total < (1LL << 31), "kk_conv1d_adain_bf16_tc: too many tiles");
  KK_CHECK_ARG(!d->lens || d->B <= RG_MAX_B, "kk_conv1d_adain_bf16_tc: a ragged batch holds at most %d utterances", RG_MAX_B);
  int grid = (int)(total < num_sms() ? total : num_sms());
  // bit 0 = next-tile L2 prefetch (on), bit 1 = disable the deep raw ring, bit 2 = no residual prefetch (on).  The
  // environment is read in -DKK_TUNE builds only (experiments); the product build has no knobs.
  static const int flags = fv_tune_int("KK_L2PF", 5);
#ifdef KK_TUNE
  {
    static const int knock = fv_tune_int("KK_XF_KNOCK", 0);
    static bool set = false;
    if (!set) { cudaMemcpyToSymbol(g_fv_xf_knock, &knock, sizeof(int)); set = true; }
  }
#endif
  if constexpr (CL > 1) {
    grid -= grid % CL;                        // whole clusters (callers fall back to CL = 1 for tiny problems)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(FV_THREADS);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CL; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    KK_CUDA(cudaLaunchKernelEx(&cfg, conv1d_adain_tc_kernel<MT, BN, CB, SWAP, CL>, tmB, tmX, tmRes, tmOut, tmAdd, *d, *a,
                               m_tiles, (int)total, flags));
  } else {
    conv1d_adain_tc_kernel<MT, BN, CB, SWAP, CL><<<grid, FV_THREADS, Cfg::SMEM_BYTES, st>>>(tmB, tmX, tmRes, tmOut, tmAdd, *d, *a,
                                                                                    m_tiles, (int)total, flags);
  }
  KK_LAUNCH_CHECK("conv1d_adain_tc_kernel");
  return KK_OK;
}

}  // namespace kk

#ifdef KK_FUSED_DBG
extern "C" int kk_debug_fused_buffer(unsigned long long* buf) {
  KK_CUDA(cudaMemcpyToSymbol(kk::g_fv_dbg, &buf, sizeof(buf)));
  return KK_OK;
}
#endif

extern "C" int kk_conv1d_adain_bf16_tc(const kk_adain_in* a, const void* w, const kk_conv_desc* d, void* stream) {
  KK_CHECK_ARG(a && a->x && w, "kk_conv1d_adain_bf16_tc: null tensor");
  int rc = kk::validate_conv_desc(d, "kk_conv1d_adain_bf16_tc");
  if (rc) return rc;
  if (!kk_device_is_sm100()) {
    kk::set_error("kk_conv1d_adain_bf16_tc: needs an sm_100 (Blackwell) device");
    return KK_ERR_UNSUPPORTED;
  }
  KK_CHECK_ARG(d->stride == 1 && d->up == 1 && d->up_off == 0 && d->left_pad == 0,
               "kk_conv1d_adain_bf16_tc: plain stride-1 convolutions only");
  KK_CHECK_ARG(d->M == d->L_true && d->M == d->L_out, "kk_conv1d_adain_bf16_tc: M must equal L_true and L_out");
  KK_CHECK_ARG(d->x_pitch == a->x_pitch, "kk_conv1d_adain_bf16_tc: x_pitch mismatch between the descriptors");
  KK_CHECK_ARG((d->taps - 1) * d->dil <= kk::FV_HALO_MAX, "kk_conv1d_adain_bf16_tc: (taps-1)*dil = %d exceeds %d",
               (d->taps - 1) * d->dil, kk::FV_HALO_MAX);
  KK_CHECK_ARG(d->Cw == d->C_in, "kk_conv1d_adain_bf16_tc: Cw (%d) must equal C_in (%d)", d->Cw, d->C_in);
  const int xesz = a->x_is_bf16 ? 2 : 4;
  KK_CHECK_ARG((a->x_pitch * xesz) % 16 == 0 && ((uintptr_t)a->x & 15) == 0 && ((uintptr_t)w & 15) == 0,
               "kk_conv1d_adain_bf16_tc: x and w must be 16-byte aligned with 16-byte row pitch");
  KK_CHECK_ARG(!d->out || (((uintptr_t)d->out & 15) == 0 && d->out_pitch % 4 == 0),
               "kk_conv1d_adain_bf16_tc: out must be 16-byte aligned with a pitch multiple of 4");
  KK_CHECK_ARG(!d->out_add || (((uintptr_t)d->out_add & 15) == 0 && d->add_pitch % 4 == 0),
               "kk_conv1d_adain_bf16_tc: out_add must be 16-byte aligned with a pitch multiple of 4");
  KK_CHECK_ARG(!d->res || (((uintptr_t)d->res & 15) == 0 && d->res_pitch % (d->res_is_bf16 ? 8 : 4) == 0 &&
                           d->res_shift == 0 && d->L_res == d->L_out),
               "kk_conv1d_adain_bf16_tc: residual must be 16-byte aligned (pointer and row pitch), same length as the "
               "output, no shift");
  if (d->res2) {
    kk::set_error("kk_conv1d_adain_bf16_tc: a second residual is not supported here; accumulate with out_add instead");
    return KK_ERR_UNSUPPORTED;
  }
  KK_CHECK_ARG(!d->bias || ((uintptr_t)d->bias & 15) == 0, "kk_conv1d_adain_bf16_tc: bias must be 16-byte aligned");
  KK_CHECK_ARG(!d->out_act || (d->act_is_bf16 && ((uintptr_t)d->out_act & 15) == 0 && d->act_pitch % 8 == 0),
               "kk_conv1d_adain_bf16_tc: out_act must be bf16, 16-byte aligned, pitch multiple of 8");
  KK_CHECK_ARG(a->act >= KK_ACT_NONE && a->act <= KK_ACT_SNAKE_FAST, "kk_conv1d_adain_bf16_tc: act=%d", a->act);
  KK_CHECK_ARG((a->act != KK_ACT_SNAKE && a->act != KK_ACT_SNAKE_FAST) || a->alpha,
               "kk_conv1d_adain_bf16_tc: snake needs alpha");
  KK_CHECK_ARG(!a->stats || a->h, "kk_conv1d_adain_bf16_tc: normalisation needs the style projection h");
  if (!a->x_is_bf16 && d->res) {
    kk::set_error("kk_conv1d_adain_bf16_tc: fp32 input with a residual is not supported (the raw-input ring and the "
                  "residual staging share shared memory)");
    return KK_ERR_UNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  static const bool swap128 = kk::fv_tune_int("KK_FUSED_SWAP", 1) != 0;
  // cluster pairs (shared weight stream): KK_FUSED_CL bit 0 = 256-channel configuration, bit 1 = 128.  Default
  // off: measured neutral to slightly negative (C=256 k=7: 5.05 -> 5.20 ms per 12 launches), i.e. the weight stalls
  // are a latency / ring-depth effect, not L2 bandwidth.  Kept as an A/B knob.
  static const int cl_mask = kk::fv_tune_int("KK_FUSED_CL", 0);
  const long long tiles128 = (long long)kk::cdiv(d->M, 256) * d->B, tiles256 = (long long)kk::cdiv(d->M, 128) * d->B;
  if (d->C_in == 128 && d->N == 128) {
    if (!swap128) return kk::launch_fused<2, 128, 2, false>(a, w, d, st);
    return ((cl_mask & 2) && tiles128 >= 4) ? kk::launch_fused<2, 128, 2, true, 2>(a, w, d, st)
                                            : kk::launch_fused<2, 128, 2, true>(a, w, d, st);
  }
  if (d->C_in == 256 && d->N == 256) {
    // Transposed form (128-row tiles, two 128-channel M blocks, lane = channel epilogue) vs the row-major one, chosen per
    // launch from what was measured at B=64 x T=320 (profiles/r02_fused_experiments.txt, item 4): the transposed
    // epilogue is ~1.4x cheaper, its MMAs read 128 instead of 96 B/clk of operands - it wins where the CUDA-core roles
    // bound the launch (k = 3: 0.67-0.80x; k = 7 with a residual: 0.87x) and loses where the MMA does (k = 7 conv1
    // 1.08x, k = 11 1.03-1.27x).  Mask bits: 1 taps <= 3, 2 taps <= 7 with residual, 4 taps <= 7 without, 8 longer.
    static const int swap256 = kk::fv_tune_int("KK_FUSED_SWAP256", 3);
    const int bit = d->taps <= 3 ? 1 : (d->taps <= 7 ? (d->res ? 2 : 4) : 8);
    if (swap256 & bit) return kk::launch_fused<1, 256, 4, true>(a, w, d, st);
    return ((cl_mask & 1) && tiles256 >= 4) ? kk::launch_fused<1, 256, 4, false, 2>(a, w, d, st)
                                            : kk::launch_fused<1, 256, 4, false>(a, w, d, st);
  }
  if (d->C_in == 128 && d->N <= 32) {   // narrow head (conv_post, istftnet.py:798-803): no stats / bf16 copy / residual
    KK_CHECK_ARG(!d->stats && !d->out_act && !d->res && !d->out_add && d->out,
                 "kk_conv1d_adain_bf16_tc: the N <= 32 configuration writes `out` only");
    return kk::launch_fused<2, 32, 2, false>(a, w, d, st);
  }
  kk::set_error("kk_conv1d_adain_bf16_tc: unsupported shape C_in=%d N=%d (128->128, 256->256 and 128-><=32 only)", d->C_in, d->N);
  return KK_ERR_UNSUPPORTED;
}
