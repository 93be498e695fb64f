/* libkokoro_b200 - C ABI of the B200-native Kokoro ISTFTNet decoder + STFT/iSTFT/mel path.
 *
 * Every entry point: extern "C", POD arguments only (device pointers, sizes, a cudaStream_t passed as
 * void*), returns 0 on success or a negative KK_ERR_* code; kk_last_error() gives the thread-local message.
 * Functions enqueue work on the caller's stream and never synchronise or allocate (CUDA-graph capturable).
 * The caller owns every buffer.  Activations are channels-last (B, L, pitch) with `pitch` >= channels.
 *
 * Each group cites the reference interface it replaces (paths relative to the mlx-audio tree).
 */
#ifndef KOKORO_B200_H
#define KOKORO_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KK_OK 0
#define KK_ERR_INVALID_ARG (-1)
#define KK_ERR_UNSUPPORTED (-2)
#define KK_ERR_CUDA (-3)

#define KK_ACT_NONE 0
#define KK_ACT_LEAKY 1 /* LeakyReLU(slope)                      istftnet.py:692-693, :831 */
#define KK_ACT_SNAKE 2 /* x + (1/alpha) sin(alpha x)^2          istftnet.py:382, :389     */
#define KK_ACT_SNAKE_FAST 3 /* Snake with the hardware (MUFU) sine; for bf16-output pipelines */

/* ---- ragged batches (per-utterance lengths inside ONE launch) -------------------------------------------------
 * The reference decodes one utterance per call (kokoro.py:135-165), so every utterance has its own length; a batch here
 * is a PADDED tensor (B, L, pitch) plus `lens`, a device array of B ints giving each utterance's length in asr FRAMES
 * (T_b <= T).  A kernel working at a level with `mul*T + add` rows per utterance receives (lens, len_mul, len_add) and
 * treats rows >= len_mul*lens[b] + len_add as padding: they are excluded from instance-norm statistics (and from their
 * divisor), read as the convolution's zero padding, and activation tensors that a later convolution reads through TMA
 * are written as ZERO there.  lens == NULL: every utterance has the full length (the arguments are ignored).
 * Results of utterance b are then identical to decoding it alone (no op mixes utterances, istftnet.py:221-230). */

int kk_version(void);
const char* kk_last_error(void);
/* 1 when the current device is compute capability 10.x (tcgen05/TMEM/TMA paths usable). */
int kk_device_is_sm100(void);

/* ---- weights (load time) ------------------------------------------------------------------
 * weight_norm (istftnet.py:53-93, called from ConvWeighted.__call__ :130 on EVERY forward in the
 * reference; folded once here): w[o,:] = v[o,:] / (||v[o,:]||_2 + 1e-7) * g[o].  v is (O, KI). */
int kk_fold_weight_norm(const float* v, const float* g, float* w, int O, int KI, void* stream);
/* Re-lay a folded conv weight (O,K,I) into the GEMM "B operand" (N, taps*Cp) [K-major], fp32 or bf16.
 * mode 0: plain conv              out[n][t*Cp+c] = w[n][t][c]                      (mx.conv1d)
 * mode 1: transposed conv, polyphase (istftnet.py:161-166,:783; mx.conv_transpose1d with w.T):
 *         w is the folded (C_in_big, K, C_out_small) ConvWeighted tensor, K = 2*up, N = up*C_out,
 *         out[ph*C_out+co][t*Cp+ci] = w[ci][ph + up*(1-t)][co],  t in {0,1}. */
int kk_pack_conv_weight(const float* w, void* out, int out_is_bf16, int mode, int O, int K, int I, int up,
                        int Cp, void* stream);

/* ---- AdaIN style projection (AdaIN1d.fc, istftnet.py:331-336), all layers in one launch:
 * h[b][r] = sum_k s[b][k] W[r][k] + bias[r];  W is the row-concatenation of every fc.weight (R, S). */
int kk_style_affine(const float* s, const float* W, const float* bias, float* h, int B, int R, int S, void* stream);

/* ---- instance-norm statistics (InstanceNorm1d, istftnet.py:229-230): accumulates per (b,c) sum and
 * sum of squares over time into stats (B, C, 2) doubles (caller zeroes them). */
int kk_instnorm_stats(const float* x, double* stats, int B, int L, int C, int pitch, const int* lens, int len_mul,
                      int len_add, void* stream);

/* ---- AdaIN apply + activation (istftnet.py:337 then :382/:389 Snake or :876/:889 LeakyReLU):
 * y = act((1+gamma)*(x-mean)/sqrt(var+1e-5)+beta); gamma = h[b][c], beta = h[b][C+c] (h row stride h_stride).
 * stats==NULL skips the normalisation (plain activation).  Output fp32 or bf16, pitch out_pitch.
 * pool != 0: additionally applies AdainResBlk1d's depthwise ConvTranspose1d k3 s2 p1 (+bias) and the
 * left zero pad (istftnet.py:880-881): output has 2L rows; pool_w is the folded (C,3) weight. */
int kk_adain_act(const float* x, int x_pitch, const double* stats, const float* h, int h_stride,
                 const float* alpha, int act, float slope, void* out, int out_is_bf16, int out_pitch,
                 int B, int L, int C, int pool, const float* pool_w, const float* pool_b, const int* lens, int len_mul,
                 int len_add, void* stream);

/* ---- Decoder front glue (istftnet.py:949-953,:957): F0_conv / N_conv (k3 s2 p1 on one channel) and the
 * channels-first -> channels-last transpose of asr, written into out (B, T, out_pitch) at channel offsets
 * asr_off (512 channels; skipped when asr == NULL), f0_off, n_off.  w3b = {w0,w1,w2,bias} folded.
 * lens (frames per utterance, or NULL): curve samples at or beyond 2*lens[b] read as the conv's zero padding. */
int kk_front_assemble(const float* asr, const float* F0_curve, const float* N, const float* f0_w3b,
                      const float* n_w3b, float* out, int out_pitch, int asr_off, int f0_off, int n_off,
                      int B, int C_asr, int T, const int* lens, void* stream);
/* Strided channel-block copy between channels-last tensors (the concatenations of :957). */
int kk_copy_channels(const float* src, int src_pitch, int src_off, float* dst, int dst_pitch, int dst_off,
                     int rows, int C, void* stream);
/* fp32 -> bf16 (or fp32) channels-last conversion with optional LeakyReLU (shortcut / leaky_relu :777,:798). */
int kk_convert_act(const float* x, int x_pitch, void* out, int out_is_bf16, int out_pitch, long long rows, int C,
                   float slope, void* stream);

/* ---- convolution as implicit GEMM (mx.conv1d / mx.conv_transpose1d as used by ConvWeighted.__call__
 * istftnet.py:128-166 and nn.Conv1d noise_convs :745-758).
 *   GEMM: acc[b][m][n] = sum_{t<taps} sum_{c<C_in} x[b][m*stride + t*dil - pad_left][c] * W[n][t*Cw + c]
 *         (rows outside [0,L_in) read as zero: per-utterance zero padding)
 *   epilogue: ph = n / C_out, c = n % C_out, j = m*up + ph - up_off, out_row = j + left_pad;
 *         v = (0 <= j < L_true) ? acc + bias[c] : 0;  v += res[b][out_row >> res_shift][c] + res2[b][out_row][c];
 *         v *= scale; out[b][out_row][c] = v; stats[b][c] += (v, v*v);
 *         out_act[b][out_row][c] = leaky(v, act_slope) (optional, fp32 or bf16)
 * A plain conv is up=1, up_off=0, left_pad=0, L_true=L_out=M.  A ConvTranspose1d with kernel 2*up is the
 * polyphase form taps=2, dil=1, pad_left=1, N = up*C_out (weights from kk_pack_conv_weight mode 1). */
typedef struct kk_conv_desc {
  int B, L_in, C_in, x_pitch;      /* input tensor (B, L_in, x_pitch)                                  */
  int taps, dil, pad_left, stride; /* stride > 1 only in the fp32 kernel                               */
  int Cw;                          /* channel pitch inside one tap of W (>= C_in)                      */
  int M, N;                        /* GEMM rows per utterance, GEMM columns                            */
  int C_out, up, up_off, L_true, left_pad, L_out;
  const float* bias;
  const float* res; int res_pitch, res_shift, L_res;
  const float* res2; int res2_pitch;
  float scale;
  float* out; int out_pitch;
  double* stats;
  void* out_act; int act_pitch, act_is_bf16; float act_slope;
  float* out_add; int add_pitch;   /* optional: out_add (B, L_out, add_pitch) += result (the MRF sum over resblocks,
                                      istftnet.py:795-800); kk_conv1d_adain_bf16_tc only (TMA reduce-add)          */
  int res_is_bf16;                 /* `res` points at bf16 (B, L_res, res_pitch) data: the bf16 residual stream of
                                      AdaINResBlock1 (x = x + xt, istftnet.py:394) and of the generator's noise branch
                                      (x = x + x_source, :788); kk_conv1d_adain_bf16_tc, and the tcgen05 kernels when
                                      their TMA epilogue applies (N % 32 == 0, aligned rows, no res2, res_shift 0) */
  const int* lens; int len_mul, len_add; /* ragged batch (see top): utterance b has L_true_b = min(L_true, len_mul *
                                      lens[b] + len_add) true output rows (and L_out_b = L_true_b + L_out - L_true
                                      output rows); rows beyond are excluded from `stats`, written as zero to
                                      `out_act`, and - kk_conv1d_adain_bf16_tc, where L_in = M = L_true - masked on the
                                      input side too (zero AFTER the activation, statistics divisor L_true_b)      */
} kk_conv_desc;

/* fp32 CUDA-core kernel (strict-fp32 parity path; also the tiny / strided layers). x, w fp32. */
int kk_conv1d_f32(const float* x, const float* w, const kk_conv_desc* d, void* stream);
/* tcgen05/TMEM/TMA kernel: x, w bf16 (w from kk_pack_conv_weight(out_is_bf16=1), Cw multiple of 64,
 * x_pitch multiple of 8), fp32 accumulate in tensor memory.  block_n in {0(auto),32,64,128,256}.
 * Returns KK_ERR_UNSUPPORTED on a non-sm_100 device. */
int kk_conv1d_bf16_tc(const void* x, const void* w, const kk_conv_desc* d, int block_n, void* stream);

/* Same GEMM with fp32 operands on the tensor cores (kind::tf32): the strict-fp32 mode.  terms = 3: error-compensated
 * 3xTF32 - every operand is split into hi (its 19 leading bits) + lo (x - hi) and lo*hi + hi*lo + hi*hi accumulate in
 * fp32, which restores fp32-level accuracy (the reference arithmetic is fp32, istftnet.py:128-166); terms = 1: plain
 * TF32 (hi*hi only; 10-bit mantissa operands).  x: fp32 (B, L_in, x_pitch), x_pitch multiple of 4.  w_hi / w_lo:
 * kk_pack_conv_weight(out_is_bf16=0, Cp multiple of 32) split by kk_split_tf32.  The activation tile is split on the
 * fly in shared memory.  Epilogue identical to kk_conv1d_bf16_tc. */
int kk_split_tf32(const float* x, float* hi, float* lo, long long n, void* stream);
int kk_conv1d_tf32_tc(const float* x, const float* w_hi, const float* w_lo, const kk_conv_desc* d, int terms,
                      void* stream);

/* ---- fused AdaIN + activation + convolution (AdaINResBlock1's `conv(snake(adain(x, s)))` pairs,
 * istftnet.py:377-396, and AdainResBlk1d's `conv(leaky(norm(x, s)))`, :874-894) on the tensor cores.
 * The raw (pre-normalisation) tensor x is read ONCE per output tile (tile + dilation halo); the instance-norm
 * affine and the activation are applied in registers while the bf16 A operand is written to shared memory
 * (128B-swizzled), and every tap reads that tile through a row-shifted UMMA descriptor - the activated tensor
 * never exists in HBM.  Epilogue identical to kk_conv1d_bf16_tc (bias, residuals, scale, stats, out / out_act).
 *   A[b][l][c] = act(a_c * x[b][l][c] + b_c),  a_c = (1+gamma_c) / sqrt(var_c + 1e-5),  b_c = beta_c - a_c mean_c
 *   (rows outside [0, L_in) are zero AFTER the activation: the conv's own zero padding)
 * Restrictions: stride 1, up 1, (C_in, N) in {(128,128), (256,256), (128, <=32: `out` only)}, (taps-1)*dil <= 56,
 * no res2 (use out_add).  Outputs leave by TMA in 16-byte units: pad columns of `out` between C_out and the next
 * multiple of 4 are overwritten with zeros. */
typedef struct kk_adain_in {
  const void* x; int x_is_bf16; int x_pitch; /* raw input (B, L_in, x_pitch), fp32 or bf16                   */
  const double* stats;                       /* (B, C_in, 2) sum / sum of squares over time; NULL: no norm  */
  const float* h; int h_stride;              /* gamma = h[b*h_stride + c], beta = h[b*h_stride + C_in + c]  */
  const float* alpha;                        /* Snake alpha (C_in), KK_ACT_SNAKE* only                      */
  int act; float slope;                      /* KK_ACT_*; Snake always uses the hardware sine here          */
} kk_adain_in;
int kk_conv1d_adain_bf16_tc(const kk_adain_in* a, const void* w, const kk_conv_desc* d, void* stream);

/* ---- harmonic source (SineGen / SourceModuleHnNSF istftnet.py:531-680, interpolate.py:82-106,
 * nn.Upsample :718-720) fused with MLXSTFT.transform (:463-495, utils.stft n_fft=20 hop=5):
 * f0 (B, 2T) -> har_source (B, 600T) -> har (B, 120T+1, har_pitch) = [|X| (11) , angle(X) (11)].
 * noise (B, 600T, 9) N(0,1) or NULL (then a Philox counter RNG is used, seeded with *seed_dev when seed_dev is a
 * device pointer - so that a captured CUDA graph draws fresh noise on every replay, as mx.random.normal does on
 * every call (istftnet.py:620) - else with `seed`).
 * phase_ws: scratch (B, 9, 2T) floats.  lin_w (9), lin_b (1): m_source.l_linear.
 * lens (frames per utterance, or NULL): utterance b has T2_b = 2*lens[b] f0 samples (the interpolation's edge clamp and
 * index wrap follow T2_b); har_source samples at or beyond T2_b*upsample are written as zero. */
int kk_hnsf_source(const float* f0, const float* noise, unsigned long long seed, const unsigned long long* seed_dev,
                   const float* lin_w, const float* lin_b, float* phase_ws, float* har_source, int B, int T2,
                   int upsample, const int* lens, void* stream);
/* kk_hnsf_source without caller-provided scratch.  NO counterpart in the reference: added
 * for callers that cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The phase scratch is a static device array of KK_HNSF_NOALLOC_MAX_FLOATS floats (4 MiB, held by every
 * GPU that loads the library until exit); B*9*T2 beyond that is rejected with KK_ERR_INVALID_ARG.  The call allocates
 * nothing.  Every call shares the array (it keeps the last call's contents; nothing in it needs clearing): calls that
 * overlap on the GPU race - serialise them on one stream.  Otherwise identical to kk_hnsf_source. */
#define KK_HNSF_NOALLOC_MAX_FLOATS (1 << 20)
int kk_hnsf_source_noalloc(const float* f0, const float* noise, unsigned long long seed,
                           const unsigned long long* seed_dev, const float* lin_w, const float* lin_b,
                           float* har_source, int B, int T2, int upsample, const int* lens, void* stream);
/* lens (frames, or NULL): utterance b has S_b = len_mul*lens[b] samples (the reflect pad mirrors at S_b) and S_b/5 + 1
 * frames; frames beyond are written as zero. */
int kk_stft_n20(const float* har_source, float* har, int har_pitch, int B, int S, const int* lens, int len_mul,
                void* stream);
/* Same transform, additionally (or only, har == NULL) written as bf16 into (B, bf_rows, bf_pitch) with
 * bf_rows >= 120T+1: the tensor-core noise_convs (istftnet.py:778) read it as (B, bf_rows/6, 6*bf_pitch) rows
 * of six frames (stride-6 conv) or as is (1x1 conv).  Rows >= 120T+1 are not touched (caller keeps them zero). */
int kk_stft_n20_bf16(const float* har_source, float* har, int har_pitch, void* har_bf16, int bf_pitch, int bf_rows,
                     int B, int S, const int* lens, int len_mul, void* stream);

/* ---- iSTFT head (Generator tail istftnet.py:804-806, MLXSTFT.inverse :497-523, utils.istft :104-158):
 * y (B, F, y_pitch) conv_post output -> spec=exp(y[:11]), phase=sin(y[11:]) -> irfft(20) * periodic hann,
 * overlap-add hop 5, divide by window sum, trim 10 -> audio (B, 5*(F-1)).
 * lens (frames, or NULL): utterance b has F_b = len_mul*lens[b] + len_add spectrogram frames; samples beyond 5*(F_b-1)
 * are written as zero. */
int kk_istft_head_n20(const float* y, int y_pitch, float* audio, int B, int F, const int* lens, int len_mul, int len_add,
                      void* stream);

/* ---- bidirectional LSTM layer, recurrent half (mlx_audio/tts/models/kokoro/modules.py:93-286 `LSTM`, gate order
 * i, f, g, o; the `shared` layer of ProsodyPredictor.F0Ntrain :355-358).
 * xw  (B, L, 2, 4H): x Wx^T + b_ih + b_hh of every step, forward direction first (one GEMM, e.g. kk_conv1d_tf32_tc as
 *                    a 1x1 convolution with the two directions' weights stacked);
 * whT (2, H, 4H):    recurrent weights, transposed (whT[d][k][c] = Wh_d[c][k]);
 * out (B, L, 2H):    h of the forward direction in [0, H), of the backward direction in [H, 2H).
 * lens (B ints, device, or NULL): valid steps per sequence - the backward direction starts at each sequence's own last
 * step, rows past a sequence's end are zeros.  H in {64, 128, 192, 256}.  Thread-block clusters of 8 CTAs; h of a step
 * travels between them through distributed shared memory (no scratch). */
int kk_lstm_bidir_f32(const float* xw, const float* whT, float* out, int B, int L, int H, const int* lens,
                      void* stream);

/* ---- general STFT / iSTFT / log-mel (mlx_audio/utils.py:52-158,164-237), batched over signals.
 * window: n_fft floats (already zero-extended).  pad_mode: 0 none(center=False) 1 reflect 2 constant.
 * spec out: (B, frames, n_fft/2+1, 2) float (re, im).  n_fft must factor into {2,3,5} and be <= 4096.
 * Ragged batches: lens (device, B ints; NULL = every signal has n samples) = valid samples per signal.  Signal b is
 * transformed exactly as if it were given alone with lens[b] samples (its own reflect padding and last frame); its
 * frames past 1 + (lens[b] + pad - n_fft) / hop are written as zeros (all of them if it is too short to transform). */
int kk_stft_f32(const float* x, const float* window, float* spec, int B, int n, int n_fft, int hop,
                int pad_mode, int frames, const int* lens, void* stream);
/* power: 1 -> |X|, 2 -> |X|^2.  mel: fb (n_mels, n_fft/2+1).  log_mode: 0 none, 3 ln(. + floor) [parakeet], 1 ln(max(.,floor)),
 * 2 log10(max(.,floor)).  out (B, frames, n_mels).  Spectrogram is never materialised. */
int kk_logmel_f32(const float* x, const float* window, const float* fb, float* out, int B, int n, int n_fft,
                  int hop, int pad_mode, int frames, int n_mels, int power, int log_mode, float floor_,
                  const int* lens, void* stream);
/* spec (B, frames, n_fft/2+1, 2) [frames-major, i.e. the layout kk_stft_f32 writes] -> signal.
 * out (B, out_len): out[i] = ola[i + trim] / wsum (where wsum != 0), ola length (frames-1)*hop + n_fft.
 * Ragged batches: lens (device, B ints; NULL = all frames) = valid frames per signal; signal b is synthesised from its
 * first lens[b] frames only and ends at (lens[b]-1)*hop + n_fft - trim - tail samples (tail = what the caller cuts
 * from the end of a full-length signal: n_fft/2 with centre trimming, 0 with an explicit length); the rest is zeros. */
int kk_istft_f32(const float* spec, const float* window, float* out, int B, int frames, int n_fft, int hop,
                 int trim, int out_len, const int* lens, int tail, void* stream);
/* Vocos ISTFTHead polar stage (mlx_audio/codec/models/vocos/vocos.py:126-133): x (rows, x_pitch) holding
 * [log-magnitude (bins) | phase (bins)] per frame -> spec (rows, bins, 2) = min(exp(m), clip) * (cos p, sin p),
 * frame-major like kk_stft_f32's output, ready for kk_istft_f32. */
int kk_polar_to_spec(const float* x, int x_pitch, float* spec, long long rows, int bins, float clip, void* stream);
/* Whisper log-mel dynamic-range step (mlx_audio/stt/models/whisper/audio.py:80-81), in place, per signal of n_per
 * values: m = max(x); x = (max(x, m - range) + add) * scale   [reference: range 8, add 4, scale 1/4]. */
int kk_clamp_below_max_affine(float* x, int B, long long n_per, float range, float add, float scale, void* stream);
/* Parakeet feature normalisation (mlx_audio/stt/models/parakeet/audio.py:56-63), in place on x (B, frames, n_feat):
 * per_feature = 1: (x - mean_f) / (std_f + eps) with mean / population std over the frames of each feature;
 * per_feature = 0: one mean / std per signal. */
int kk_normalize_features(float* x, int B, int frames, int n_feat, int per_feature, float eps, void* stream);
/* Snake activations of the codec generators (mlx_audio/codec/models/bigvgan/activation.py:5-51 Snake / SnakeBeta,
 * codec/models/descript/nn/layers.py:116-129 snake / Snake1d): y = x + 1/(b_c + eps) * sin(a_c x)^2 over an
 * (outer, C, inner) tensor (inner = 1: channels-last); beta may be NULL (b = a); logscale: a, b = exp(parameter).
 * out may alias x. */
int kk_snake_f32(const float* x, float* out, const float* alpha, const float* beta, long long outer, int C,
                 long long inner, int logscale, float eps, void* stream);
/* ---- SNAC generator pieces (mlx_audio/codec/models/snac/layers.py), channels-last fp32:
 * depthwise WNConv1d, groups = C, 'same' padding ((K-1)*dil = 2*pad) (:18-61, ResidualUnit :205-214): w (C, K),
 * out = bias + sum_t w[c][t] x[l + t*dil - pad][c] (+ res);  NoiseBlock (:217-228): out = x + noise * h with noise
 * (B, C) - constant over time, the reference draws (B, 1, T) on a channels-last tensor, i.e. one value per channel;
 * Tanh head (:199). */
int kk_dwconv1d_f32(const float* x, const float* w, const float* bias, const float* res, float* out, int B, int L, int C,
                    int K, int dil, int pad, void* stream);
int kk_noise_mix_f32(const float* x, const float* h, const float* noise, float* out, int B, long long L, int C,
                     void* stream);
int kk_tanh_f32(const float* x, float* out, long long n, void* stream);
/* ---- BigVGAN anti-aliased activation (mlx_audio/codec/models/bigvgan/resample.py:152-177 `Activation1d` with its
 * default ratios 2 / 2 and 12-tap kaiser-sinc filters; activation.py:5-51), fused, channels-last fp32 (B, L, C):
 * edge-pad + depthwise transposed conv x2 (UpSample1d :104-134) -> v = u + sin^2(alpha u) / (beta + eps) (beta = NULL:
 * Snake, beta := alpha; logscale: exp() of both) -> edge-pad + depthwise stride-2 low-pass (DownSample1d :137-149).
 * up_filter / down_filter: HOST pointers to the 12 taps (`upsample.filter`, `downsample.lowpass.filter`). */
/* out = min(max(x, lo), hi): BigVGAN's head when use_tanh_at_final is off (bigvgan.py:125-126) */
int kk_clip_f32(const float* x, float* out, float lo, float hi, long long n, void* stream);
/* out = a x + b y (y may be NULL): the sum over a stage's resblocks and its division by their number (bigvgan.py:113-117) */
int kk_axpby_f32(const float* x, const float* y, float* out, float a, float b, long long n, void* stream);
int kk_aa_act_f32(const float* x, float* out, const float* alpha, const float* beta, const float* up_filter,
                  const float* down_filter, int B, int L, int C, int logscale, float eps, void* stream);
/* Weight-norm fold with the epsilon and operation order as parameters: g_first = 1, eps = 0 is the DAC form
 * w = (g * v) / ||v|| of mlx_audio/codec/models/descript/nn/layers.py:7-12,57 (norm over (K, I) per output channel). */
int kk_fold_weight_norm_ex(const float* v, const float* g, float* w, int O, int KI, float eps, int g_first,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KOKORO_B200_H */
