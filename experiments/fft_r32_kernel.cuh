// LAB CODE: the single-exchange forward STFT kernel (included by csrc/fft.cu inside namespace kk, under -DKK_FFT_R32 only;
// needs StftArgs / ragged helpers of fft.cu).  Measured slower than the radix-8 kernel: profiles/r02_fft_experiments.txt 2.
// ---------------------------------------------------------------- single-exchange forward kernel (fft_r32.cuh)
// 16 threads per frame, two frames per warp, 16 per block iteration; no block-level synchronisation after the tables.
// Pass 1 reads the samples straight from global memory (thread m: the float2 at 2 (m + 16 t), 128 contiguous bytes per
// frame and request; reflect / zero padding only on the edge frames) and multiplies by the window (shared table, the
// two frames of a warp read the same words: one wavefront).  The half spectrum leaves from the registers of pass 2.
#ifndef R32_BLOCKS
#define R32_BLOCKS 1
#endif
template <int MODE>
__global__ void __launch_bounds__(fr::THREADS, R32_BLOCKS) stft_r32_kernel(const StftArgs a) {
  extern __shared__ __align__(16) unsigned char smem_[];
  using fr::c2; using fr::pk; using fr::up;
  constexpr int NC = fr::NC, n_fft = 2 * NC, bins = NC + 1;
  c2* win = reinterpret_cast<c2*>(smem_);           // (w[2 i], w[2 i + 1]) at i
  c2* tg = win + NC;
  c2* xbuf = tg + fr::TG;
  const int lane = threadIdx.x & 31, m = lane & 15, slot = threadIdx.x >> 4;
  c2* X = xbuf + slot * fr::XBUF;
  const fr::Twid tw(tg, m);
  fr::fill_tables(tg);
  // real-transform factor of the pair this lane unpacks for lane 0 (rows 16 and 0, see below): k = 16 + 32 q or 32 q
  c2 u_coop;
  {
    float sn, cs_;
    sincospif(-(float)(m < 8 ? 16 + 32 * m : 32 * (m - 8)) / (float)NC, &sn, &cs_);
    u_coop = pk(cs_, sn);
  }
  for (int i = threadIdx.x; i < NC; i += blockDim.x) win[i] = pk(a.window[2 * i], a.window[2 * i + 1]);
  __syncthreads();

  const int b = blockIdx.y;
  const float* xb = a.x + (size_t)b * a.n;
  const int pad = a.pad_mode == 0 ? 0 : NC;
  const bool vec_ok = ((a.hop | pad) & 1) == 0 && ((reinterpret_cast<uintptr_t>(xb) & 7) == 0);
  const int f_begin = blockIdx.x * a.frames_per_block;
  const int f_end = min(a.frames, f_begin + a.frames_per_block);
  // Software pipeline: there is no register room for a second frame, and at 16 warps per SM the wait for a frame's 32
  // loads was 36 % of a warp's time (ncu stall_long_sb).  The loads of the NEXT frame are issued inside the unpack
  // loop of this one, two per pair, into the registers the consumed spectrum points free.
  struct Src { const float2* vec; int base; bool live, fast; };
  auto source = [&](int f) {
    Src s;
    s.base = f * a.hop - pad;
    s.live = f < f_end;
    s.fast = s.live && vec_ok && s.base >= 0 && s.base + n_fft <= a.n;
    s.vec = reinterpret_cast<const float2*>(xb + s.base) + m;                  // dereferenced only if fast
    return s;
  };
  // an ordered (not .nc) load behind a __syncwarp: stays behind the pairs whose registers it reuses (left to the
  // compiler, all 32 loads are hoisted above the unpack and 64 registers spill)
  auto ldg_ordered = [](const float2* p) -> c2 {
    c2 r;
    asm volatile("ld.global.b64 %0, [%1];" : "=l"(r) : "l"(p) : "memory");
    return r;
  };
  auto load_point = [&](const Src& s, int t) -> c2 {   // edge frames: padding resolved per sample
    float sv[2] = {0.f, 0.f};
    if (s.live) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        int q = s.base + 2 * (m + 16 * t) + e;
        if (a.pad_mode == 1) {           // reflect without edge repeat (utils.py:81-84)
          if (q < 0) q = -q;
          if (q >= a.n) q = 2 * (a.n - 1) - q;
          sv[e] = xb[q];
        } else if (q >= 0 && q < a.n) {
          sv[e] = xb[q];
        }
      }
    }
    return pk(sv[0], sv[1]);
  };
  // the loop bound is warp uniform (a warp's second frame may be past the end: it computes on zeros, stores nothing)
  c2 v[32];
  bool have = false;             // v already holds this iteration's samples (warp uniform)
  for (int f0 = f_begin + (slot & ~1); f0 < f_end; f0 += fr::SLOTS) {
    const int f = f0 + (slot & 1);
    const bool live = f < f_end;
    if (!have) {                 // first iteration, and frames that need padding: loaded here, not pipelined
      const Src cur = source(f);
      if (__all_sync(0xffffffffu, cur.fast || !cur.live)) {
        if (cur.fast) {
#pragma unroll
          for (int t = 0; t < 32; ++t) v[t] = pk(__ldg(cur.vec + 16 * t));
        }
      } else {
#pragma unroll
        for (int t = 0; t < 32; ++t) v[t] = load_point(cur, t);
      }
    }
#pragma unroll
    for (int t = 0; t < 32; ++t) v[t] = ff::mul2(v[t], win[m + 16 * t]);
    c2 A[16], B[16];
    fr::forward512(v, X, tw, m, A, B);
    // ---- unpack: with E = (Z[k] + conj Z[nc-k]) / 2, O = (Z[k] - conj Z[nc-k]) / (2i), u = e^{-2 pi i k / n}:
    //      X[k] = E + u O   and   X[nc-k] = conj(E - u O)
    c2* orow = reinterpret_cast<c2*>(a.out) + ((size_t)b * a.frames + (live ? f : 0)) * bins;
    const Src nxt = source(f + fr::SLOTS);
    // warp-uniform: the next frames of both halves load unpadded (a frame past the end loads nothing, computes on
    // whatever its registers hold and stores nothing)
    const bool nfast = __all_sync(0xffffffffu, nxt.fast || !nxt.live);
    // uh = u / 2
    auto unpack_pair = [&](c2 zk, c2 zc, c2 uh, c2& X1, c2& X2) {
      const c2 s = ff::fma2(zc, pk(1.f, -1.f), zk);            // zk + conj(zc) = 2 E
      const float2 d = up(ff::fma2(zc, pk(-1.f, 1.f), zk));    // zk - conj(zc) = 2i O
      const float2 t = up(ff::cm(uh, pk(d.y, -d.x)));          // u O
      X1 = ff::fma2(s, pk(0.5f, 0.5f), pk(t));
      X2 = ff::fma2(s, pk(0.5f, -0.5f), pk(-t.x, t.y));
    };
    if (MODE == 0) {
      if (nfast && nxt.fast) {
#pragma unroll
        for (int t = 0; t < 32; ++t) v[t] = pk(__ldg(nxt.vec + 16 * t));
      }
      // lane 0 holds rows 0 and 16, which pair up inside themselves: handed to the 16 lanes through a scratch
      __syncwarp();
      if (m == 0) {
#pragma unroll
        for (int k2 = 0; k2 < 16; ++k2) {   // volatile: merged into 128-bit stores, the rows get permuted through local memory
          *reinterpret_cast<volatile c2*>(X + k2) = A[fr::o16(k2)];
          *reinterpret_cast<volatile c2*>(X + 16 + k2) = B[fr::o16(k2)];
        }
        if (live) orow[NC / 2] = fr::conj2(A[fr::o16(8)]);     // k = 256 pairs with itself: X = conj(Z)
      }
      // rows m and 32 - m of lanes 1..15: point k = m + 32 k2 pairs with 512 - k = (32 - m) + 32 (15 - k2)
#pragma unroll
      for (int k2 = 0; k2 < 16; ++k2) {
        c2 X1, X2;
        unpack_pair(A[fr::o16(k2)], B[fr::o16(15 - k2)], tw.u(k2, 0.5f), X1, X2);
        if (live && m) { orow[m + 32 * k2] = X1; orow[NC - m - 32 * k2] = X2; }
      }
      __syncwarp();
      {
        // lanes 0..7: row 16, Z[16 + 32 q] with Z[16 + 32 (15 - q)]; lanes 8..15: row 0, Z[32 q] with Z[32 (16 - q)]
        // (q = 0: DC and Nyquist, both from Z[0])
        const int q = m & 7, rowb = m < 8;
        const c2 zk = X[rowb ? 16 + q : q], zc = X[rowb ? 31 - q : ((16 - q) & 15)];
        const int k = rowb ? 16 + 32 * q : 32 * q;
        c2 X1, X2;
        unpack_pair(zk, zc, ff::mul2(u_coop, pk(0.5f, 0.5f)), X1, X2);
        if (k == 0) { X1 = pk(up(X1).x, 0.f); X2 = pk(up(X2).x, 0.f); }   // rfft: DC and Nyquist are real
        if (live) { orow[k] = X1; orow[NC - k] = X2; }
      }
      __syncwarp();
    }
    have = nfast;
  }
}
