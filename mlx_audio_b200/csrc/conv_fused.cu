// Fused AdaIN + Snake/LeakyReLU + Conv1d on the 5th-gen tensor cores (sm_100a).
//
// Reference ops replaced (mlx_audio/tts/models/kokoro/istftnet.py): AdaINResBlock1's
//   xt = adain(x, s); xt = x + (1/a) sin(a xt)^2; xt = conv(xt)          (:379-393, AdaIN1d :327-338)
// i.e. per reference call: instance-norm mean/var apply, style affine, Snake, swapaxes, conv - five materialised
// tensors.  Here the raw tensor x (B, L, C) is read once per output tile:
//   * "transform" warps load tile + dilation halo rows of one 64-channel block, apply a_c x + b_c and the
//     activation in registers, and store bf16 into a shared-memory slot in the canonical K-major 128B-swizzle
//     layout (row = 128 B, 16-byte chunk index XOR (row & 7));
//   * one thread issues tcgen05.mma: for tap t the A operand is the SAME slot read through a descriptor whose
//     start address is shifted by t*dil rows (the swizzle is a function of the absolute smem address, so a row
//     shift that is not a multiple of the 8-row atom is legal - profiles/r01_umma_rowshift_probe.txt);
//   * the weights (B operand) stream through a TMA ring, one (tap, channel-block) slab of BN x 64 per stage;
//   * MT (1 or 2) 128-row sub-tiles share every B stage; accumulators are double-buffered in tensor memory
//     (2 x MT x BN = 512 columns) so the epilogue of tile i overlaps the main loop of tile i+1.
// Warp roles: 0 = TMA (weights), 1 = MMA issuer, 2..7 = transform (warp 2 owns the TMEM allocation),
// 8..15 = epilogue.  Persistent: one CTA per SM walks the (utterance, row-tile) list.
#include "tc_epilogue.cuh"
#include <stdlib.h>

namespace kk {

constexpr int FU_XF_WARPS = 6;
constexpr int FU_EPI_WARPS = 8;
constexpr int FU_XF_THREADS = FU_XF_WARPS * 32;
constexpr int FU_EPI_THREADS = FU_EPI_WARPS * 32;
constexpr int FU_THREADS = 64 + FU_XF_THREADS + FU_EPI_THREADS;  // 512
constexpr int FU_HALO_MAX = 56;
constexpr int FU_UNROLL = 8;

template <int MT, int BN, int CB>
struct FuCfg {
  static constexpr int SLOT_ROWS = 128 * MT + FU_HALO_MAX;
  static constexpr int SLOT_BYTES = SLOT_ROWS * 128;            // multiple of 1024 for MT in {1,2}
  static constexpr int NSLOT = MT == 2 ? 3 : 4;
  static constexpr int B_BYTES = BN * 128;
  static constexpr int STAGES = BN >= 256 ? 3 : 4;
  static constexpr int STAGING_BYTES = FU_EPI_WARPS * 4096;
  static constexpr int COEF_BYTES = CB * 64 * 16;
  static constexpr int OFF_B = NSLOT * SLOT_BYTES;
  static constexpr int OFF_STAGING = OFF_B + STAGES * B_BYTES;
  static constexpr int OFF_COEF = OFF_STAGING + STAGING_BYTES;
  static constexpr int OFF_BAR = OFF_COEF + COEF_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 256 + 1024;
  static constexpr uint32_t TMEM_COLS = 512;
  static_assert(SLOT_BYTES % 1024 == 0, "slots must keep the 1024-byte swizzle alignment");
  static_assert(2 * MT * BN <= 512, "accumulators exceed tensor memory");
  static_assert(SMEM_BYTES <= 232448, "shared memory budget");
};

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}

// coef = (a, b, alpha, 1/alpha)
template <int ACT>
__device__ __forceinline__ float xf_apply(float x, const float4& k, float slope) {
  float y = fmaf(k.x, x, k.y);
  if (ACT == KK_ACT_SNAKE) {
    float t = __sinf(k.z * y);
    y = fmaf(k.w, t * t, y);
  } else if (ACT == KK_ACT_LEAKY) {
    y = y > 0.f ? y : y * slope;
  }
  return y;
}

// Transform one 64-channel block of the tile (rows [0, R) of the slot <-> input rows l_base + r).
// A warp covers two rows per step: lanes 0-15 row 2p, lanes 16-31 row 2p+1, 4 channels per lane.
template <bool XBF16, int ACT>
__device__ __forceinline__ void xf_block(const kk_adain_in& a, const float4* __restrict__ coef, uint8_t* slot, int b,
                                         int L_in, int l_base, int R, int cb, int xw, int lane) {
  const int half = lane >> 4, q = lane & 15;
  const int c0 = cb * 64 + q * 4;
  const float4 k0 = coef[c0], k1 = coef[c0 + 1], k2 = coef[c0 + 2], k3 = coef[c0 + 3];
  const uint32_t chunk = (uint32_t)(q >> 1), sub = (uint32_t)(q & 1) * 8u;
  constexpr size_t ESZ = XBF16 ? 2 : 4;
  const char* xb = reinterpret_cast<const char*>(a.x) + ((size_t)b * L_in * a.x_pitch + c0) * ESZ;
  const size_t row_bytes = (size_t)a.x_pitch * ESZ;
  const int npairs = (R + 1) >> 1;
  for (int p0 = xw; p0 < npairs; p0 += FU_XF_WARPS * FU_UNROLL) {
    float4 v[FU_UNROLL];
#pragma unroll
    for (int u = 0; u < FU_UNROLL; ++u) {
      const int r = 2 * (p0 + u * FU_XF_WARPS) + half;
      const int l = l_base + r;
      v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (r < R && l >= 0 && l < L_in) {
        if (XBF16) {
          const uint2 t = __ldg(reinterpret_cast<const uint2*>(xb + (size_t)l * row_bytes));
          v[u].x = __uint_as_float(t.x << 16); v[u].y = __uint_as_float(t.x & 0xffff0000u);
          v[u].z = __uint_as_float(t.y << 16); v[u].w = __uint_as_float(t.y & 0xffff0000u);
        } else {
          v[u] = __ldg(reinterpret_cast<const float4*>(xb + (size_t)l * row_bytes));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < FU_UNROLL; ++u) {
      const int p = p0 + u * FU_XF_WARPS;
      if (p < npairs) {
        const int r = 2 * p + half;
        const int l = l_base + r;
        uint2 o = make_uint2(0u, 0u);   // rows outside the utterance: the conv's zero padding (post-activation)
        if (r < R && l >= 0 && l < L_in) {
          o.x = pack_bf16x2(xf_apply<ACT>(v[u].x, k0, a.slope), xf_apply<ACT>(v[u].y, k1, a.slope));
          o.y = pack_bf16x2(xf_apply<ACT>(v[u].z, k2, a.slope), xf_apply<ACT>(v[u].w, k3, a.slope));
        }
        *reinterpret_cast<uint2*>(slot + (uint32_t)r * 128u + ((chunk ^ ((uint32_t)r & 7u)) << 4) + sub) = o;
      }
    }
  }
}

template <int MT, int BN, int CB>
__global__ void __launch_bounds__(FU_THREADS, 1)
conv1d_adain_tc_kernel(const __grid_constant__ CUtensorMap tmB, const kk_conv_desc d, const kk_adain_in a,
                       const int m_tiles, const int total_tiles, const int flags) {
  using Cfg = FuCfg<MT, BN, CB>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sB = smem + Cfg::OFF_B;
  float* staging = reinterpret_cast<float*>(smem + Cfg::OFF_STAGING);
  float4* coef = reinterpret_cast<float4*>(smem + Cfg::OFF_COEF);
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + Cfg::OFF_BAR);
  uint64_t* a_empty = a_full + Cfg::NSLOT;
  uint64_t* b_full = a_empty + Cfg::NSLOT;
  uint64_t* b_empty = b_full + Cfg::STAGES;
  uint64_t* tmem_full = b_empty + Cfg::STAGES;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;          // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int halo = (d.taps - 1) * d.dil;
  const int R = 128 * MT + halo;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < Cfg::NSLOT; ++s) {
      ptx::mbar_init(&a_full[s], FU_XF_THREADS);
      ptx::mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < Cfg::STAGES; ++s) {
      ptx::mbar_init(&b_full[s], 1);
      ptx::mbar_init(&b_empty[s], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tmem_full[i], 1);
      ptx::mbar_init(&tmem_empty[i], FU_EPI_THREADS);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<Cfg::TMEM_COLS>(tmem_slot);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer: weight slabs, order (tile, channel block, tap) =====================
    if (lane == 0) {
      uint32_t g = 0;
      const size_t xesz = a.x_is_bf16 ? 2 : 4;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        // pull the NEXT tile's input rows and residual rows into L2 while this tile is being computed: the
        // transform / epilogue warps' loads then see L2-hit latency instead of HBM latency
        const int ntile = tile + gridDim.x;
        if ((flags & 1) && ntile < total_tiles) {
          const int nb_ = ntile / m_tiles, nm0 = (ntile - nb_ * m_tiles) * (128 * MT);
          int l0 = nm0 - d.pad_left, l1 = l0 + R;
          l0 = l0 < 0 ? 0 : l0;
          l1 = l1 > d.L_in ? d.L_in : l1;
          if (l1 > l0)
            ptx::prefetch_l2_bulk(reinterpret_cast<const char*>(a.x) + ((size_t)nb_ * d.L_in + l0) * a.x_pitch * xesz,
                                  (uint32_t)((size_t)(l1 - l0) * a.x_pitch * xesz));
          int r1 = nm0 + 128 * MT;
          r1 = r1 > d.M ? d.M : r1;
          if (d.res && d.res_shift == 0)
            ptx::prefetch_l2_bulk(d.res + ((size_t)nb_ * d.L_res + nm0) * d.res_pitch,
                                  (uint32_t)((size_t)(r1 - nm0) * d.res_pitch * 4));
          if (d.res2)
            ptx::prefetch_l2_bulk(d.res2 + ((size_t)nb_ * d.L_out + nm0) * d.res2_pitch,
                                  (uint32_t)((size_t)(r1 - nm0) * d.res2_pitch * 4));
        }
        for (int cb = 0; cb < CB; ++cb) {
          for (int t = 0; t < d.taps; ++t, ++g) {
            const uint32_t s = g % Cfg::STAGES, ph = (g / Cfg::STAGES) & 1;
            ptx::mbar_wait(&b_empty[s], ph ^ 1);
            ptx::mbar_expect_tx(&b_full[s], Cfg::B_BYTES);
            ptx::tma_load_2d(sB + s * Cfg::B_BYTES, &tmB, &b_full[s], t * d.Cw + cb * 64, 0);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_bf16(128, BN);
      uint32_t gb = 0, ga = 0, it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const uint32_t buf = it & 1, aph = (it >> 1) & 1;
        ptx::mbar_wait(&tmem_empty[buf], aph ^ 1);   // epilogue has drained this accumulator set
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + buf * (MT * BN);
        for (int cb = 0; cb < CB; ++cb, ++ga) {
          const uint32_t slot = ga % Cfg::NSLOT, sph = (ga / Cfg::NSLOT) & 1;
          ptx::mbar_wait(&a_full[slot], sph);
          ptx::tc_fence_after();
          const uint32_t sa0 = ptx::smem_u32(smem + slot * Cfg::SLOT_BYTES);
          for (int t = 0; t < d.taps; ++t, ++gb) {
            const uint32_t s = gb % Cfg::STAGES, ph = (gb / Cfg::STAGES) & 1;
            ptx::mbar_wait(&b_full[s], ph);
            ptx::tc_fence_after();
            const uint32_t sb = ptx::smem_u32(sB + s * Cfg::B_BYTES);
#pragma unroll
            for (int mt = 0; mt < MT; ++mt) {
              const uint32_t sa = sa0 + (uint32_t)(mt * 128 + t * d.dil) * 128u;   // row-shifted view of the slot
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint64_t ad = ptx::umma_desc(sa + k * 32, 16, 1024, ptx::UMMA_LAYOUT_SW128);
                const uint64_t bd = ptx::umma_desc(sb + k * 32, 16, 1024, ptx::UMMA_LAYOUT_SW128);
                ptx::umma_bf16(tacc + mt * BN, ad, bd, idesc, (cb | t | k) != 0 ? 1u : 0u);
              }
            }
            ptx::umma_commit(&b_empty[s]);
          }
          ptx::umma_commit(&a_empty[slot]);
        }
        ptx::umma_commit(&tmem_full[buf]);
      }
    }
  } else if (warp < 2 + FU_XF_WARPS) {
    // ===================== transform warps: x -> act(a x + b) -> bf16 swizzled slot =====================
    const int xw = warp - 2;
    const int xt = threadIdx.x - 64;
    int cur_b = -1;
    uint32_t ga = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
      const int b = tile / m_tiles, m0 = (tile - b * m_tiles) * (128 * MT);
      if (b != cur_b) {
        asm volatile("bar.sync 2, %0;" ::"n"(FU_XF_THREADS) : "memory");   // everyone is done with the old table
        for (int c = xt; c < CB * 64; c += FU_XF_THREADS) {
          float4 k = make_float4(1.f, 0.f, 1.f, 1.f);
          if (a.stats) {
            const double s = a.stats[((size_t)b * d.C_in + c) * 2], ss = a.stats[((size_t)b * d.C_in + c) * 2 + 1];
            const double mean = s / d.L_in;
            double var = ss / d.L_in - mean * mean;
            if (var < 0.0) var = 0.0;
            const float rstd = 1.0f / sqrtf((float)var + 1e-5f);
            const float g1 = 1.0f + a.h[(size_t)b * a.h_stride + c];
            k.x = g1 * rstd;
            k.y = a.h[(size_t)b * a.h_stride + d.C_in + c] - k.x * (float)mean;
          }
          if (a.act == KK_ACT_SNAKE || a.act == KK_ACT_SNAKE_FAST) { k.z = a.alpha[c]; k.w = 1.0f / k.z; }
          coef[c] = k;
        }
        asm volatile("bar.sync 2, %0;" ::"n"(FU_XF_THREADS) : "memory");
        cur_b = b;
      }
      const int l_base = m0 - d.pad_left;
      for (int cb = 0; cb < CB; ++cb, ++ga) {
        const uint32_t slot = ga % Cfg::NSLOT, sph = (ga / Cfg::NSLOT) & 1;
        ptx::mbar_wait(&a_empty[slot], sph ^ 1);     // the MMAs that read this slot have completed
        uint8_t* sp = smem + slot * Cfg::SLOT_BYTES;
        const bool snake = a.act == KK_ACT_SNAKE || a.act == KK_ACT_SNAKE_FAST;
        if (a.x_is_bf16) {
          if (snake) xf_block<true, KK_ACT_SNAKE>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
          else if (a.act == KK_ACT_LEAKY) xf_block<true, KK_ACT_LEAKY>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
          else xf_block<true, KK_ACT_NONE>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
        } else {
          if (snake) xf_block<false, KK_ACT_SNAKE>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
          else if (a.act == KK_ACT_LEAKY) xf_block<false, KK_ACT_LEAKY>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
          else xf_block<false, KK_ACT_NONE>(a, coef, sp, b, d.L_in, l_base, R, cb, xw, lane);
        }
        ptx::fence_proxy_async();                    // generic-proxy stores -> visible to the tensor core
        ptx::mbar_arrive(&a_full[slot]);
      }
    }
  } else {
    // ===================== epilogue warps =====================
    const int ew = warp - (2 + FU_XF_WARPS);
    const int et = threadIdx.x - (64 + FU_XF_THREADS);
    const int lg = warp & 3;                 // TMEM lane group this warp may access
    const int half = ew >> 2;
    const int mt = MT == 2 ? half : 0;
    const int chunk0 = MT == 2 ? 0 : half * (BN / 64);
    constexpr int NCH = MT == 2 ? BN / 32 : BN / 64;   // chunks per warp
    float* stage_w = staging + ew * 1024;
    const bool do_stats = d.stats != nullptr;
    uint32_t it = 0;
    float4 tres[8];
    if ((int)blockIdx.x < total_tiles) {
      const int b0 = blockIdx.x / m_tiles;
      epi_prefetch_chunk(d, b0, (blockIdx.x - b0 * m_tiles) * (128 * MT) + mt * 128, 0, chunk0, lg, lane, tres);
    }
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int b = tile / m_tiles, m0 = (tile - b * m_tiles) * (128 * MT) + mt * 128;
      const int ntile = tile + gridDim.x;
      const bool more = ntile < total_tiles;
      const int nx_b = ntile / m_tiles, nx_m0 = (ntile - nx_b * m_tiles) * (128 * MT) + mt * 128;
      const uint32_t buf = it & 1, aph = (it >> 1) & 1;
      ptx::mbar_wait(&tmem_full[buf], aph);
      ptx::tc_fence_after();
      const uint32_t tacc = tmem_base + buf * (MT * BN) + mt * BN;
      float ps[NCH], pq[NCH];
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        ps[c] = pq[c] = 0.f;
        const bool last = c + 1 == NCH;
        epilogue_chunk<BN>(d, b, m0, 0, chunk0 + c, lg, lane, tacc, stage_w, do_stats, ps[c], pq[c], tres,
                           last ? more : true, last ? nx_b : b, last ? nx_m0 : m0, 0, last ? chunk0 : chunk0 + c + 1);
      }
      ptx::tc_fence_before();
      ptx::mbar_arrive(&tmem_empty[buf]);     // all tcgen05.ld of this thread have completed
      if (do_stats) {
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          stage_w[(c * 2 + 0) * 32 + lane] = ps[c];
          stage_w[(c * 2 + 1) * 32 + lane] = pq[c];
        }
        asm volatile("bar.sync 1, %0;" ::"n"(FU_EPI_THREADS) : "memory");
        for (int i = et; i < 2 * BN; i += FU_EPI_THREADS) {
          const int q = i / BN, col = i - q * BN;
          int w0, w1, c;
          if (MT == 2) { w0 = 0; w1 = FU_EPI_WARPS; c = col >> 5; }
          else { const int hs = col / (BN / 2); w0 = hs * 4; w1 = w0 + 4; c = (col - hs * (BN / 2)) >> 5; }
          float acc = 0.f;
          for (int w = w0; w < w1; ++w) acc += staging[w * 1024 + (c * 2 + q) * 32 + (col & 31)];   // fixed order
          atomicAdd(d.stats + ((size_t)b * d.C_out + col) * 2 + q, (double)acc);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(FU_EPI_THREADS) : "memory");
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<Cfg::TMEM_COLS>(tmem_base);
}

}  // namespace kk

// ================================================================== host side
namespace kk {

int validate_conv_desc(const kk_conv_desc* d, const char* who);                       // conv_f32.cu
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const cuuint64_t* dims,  // conv_tc.cu
                   const cuuint64_t* strides, const cuuint32_t* box);
int num_sms();                                                                        // conv_tc.cu

template <int MT, int BN, int CB>
static int launch_fused(const kk_adain_in* a, const void* w, const kk_conv_desc* d, cudaStream_t st) {
  using Cfg = FuCfg<MT, BN, CB>;
  CUtensorMap tmB;
  {
    cuuint64_t dims[2] = {(cuuint64_t)d->taps * d->Cw, (cuuint64_t)d->N};
    cuuint64_t strides[1] = {(cuuint64_t)d->taps * d->Cw * 2};
    cuuint32_t box[2] = {64, BN};
    int rc = make_tmap_bf16(&tmB, w, 2, dims, strides, box);
    if (rc) return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    KK_CUDA(cudaFuncSetAttribute(conv1d_adain_tc_kernel<MT, BN, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 Cfg::SMEM_BYTES));
    attr_set = true;
  }
  const int m_tiles = cdiv(d->M, 128 * MT);
  const long long total = (long long)m_tiles * d->B;
  KK_CHECK_ARG(total < (1LL << 31), "kk_conv1d_adain_bf16_tc: too many tiles");
  const int grid = (int)(total < num_sms() ? total : num_sms());
  static const int flags = getenv("KK_NO_L2PF") ? 0 : 1;   // experiment knob: disable the next-tile L2 prefetch
  conv1d_adain_tc_kernel<MT, BN, CB><<<grid, FU_THREADS, Cfg::SMEM_BYTES, st>>>(tmB, *d, *a, m_tiles, (int)total, flags);
  KK_LAUNCH_CHECK("conv1d_adain_tc_kernel");
  return KK_OK;
}

}  // namespace kk

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
  KK_CHECK_ARG((d->taps - 1) * d->dil <= kk::FU_HALO_MAX, "kk_conv1d_adain_bf16_tc: (taps-1)*dil = %d exceeds %d",
               (d->taps - 1) * d->dil, kk::FU_HALO_MAX);
  KK_CHECK_ARG(d->Cw == d->C_in, "kk_conv1d_adain_bf16_tc: Cw (%d) must equal C_in (%d)", d->Cw, d->C_in);
  KK_CHECK_ARG(a->x_pitch % 4 == 0 && ((uintptr_t)a->x & 15) == 0 && ((uintptr_t)w & 15) == 0,
               "kk_conv1d_adain_bf16_tc: x must be 16-byte aligned with a pitch multiple of 4");
  KK_CHECK_ARG(a->act >= KK_ACT_NONE && a->act <= KK_ACT_SNAKE_FAST, "kk_conv1d_adain_bf16_tc: act=%d", a->act);
  KK_CHECK_ARG((a->act != KK_ACT_SNAKE && a->act != KK_ACT_SNAKE_FAST) || a->alpha,
               "kk_conv1d_adain_bf16_tc: snake needs alpha");
  KK_CHECK_ARG(!a->stats || a->h, "kk_conv1d_adain_bf16_tc: normalisation needs the style projection h");
  cudaStream_t st = (cudaStream_t)stream;
  if (d->C_in == 128 && d->N == 128) return kk::launch_fused<2, 128, 2>(a, w, d, st);
  if (d->C_in == 256 && d->N == 256) return kk::launch_fused<1, 256, 4>(a, w, d, st);
  kk::set_error("kk_conv1d_adain_bf16_tc: unsupported shape C_in=%d N=%d (128->128 and 256->256 only)", d->C_in, d->N);
  return KK_ERR_UNSUPPORTED;
}
