/*
 * ditsep_b200 -- C ABI of the B200 (sm_100a) kernels behind the DiffSep
 * predictor-corrector separation sampler.
 *
 * The reference (eduardburlacu/DiTSep) has no FFI for this path: its boundary is a
 * Python object API.  This library is what the Python mirror
 * of that API (ditsep_b200/{sdes,models,diffsep.py}) binds with ctypes; every
 * entry point cites the reference code it replaces.  Conventions:
 *   - plain pointers and sizes only; all pointers are DEVICE pointers unless named
 *     host_*; buffers are caller-allocated; nothing is freed or retained
 *     (except ditsep_conv_plan objects, which own their tensor maps);
 *   - every launch goes to the cudaStream_t passed as `stream` (void*);
 *   - return value 0 = success, non-zero = error (ditsep_last_error() has the text);
 *     no C++ exception crosses the ABI;
 *   - activations are NHWC ("pixels-major"): [B][H][W][ld] with `ld` >= C the pixel
 *     stride in elements, which lets a tensor live inside a wider concat buffer;
 *   - `dtype` of activations: DITSEP_F32 (0) or DITSEP_BF16 (1).
 * There is no CPU fallback anywhere in this library.
 */
#ifndef DITSEP_B200_H
#define DITSEP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DITSEP_F32 0
#define DITSEP_BF16 1

#define DITSEP_SDE_MIX 0  /* sdes.sdes.MixSDE   (src/sdes/sdes.py:182-351) */
#define DITSEP_SDE_OUVE 1 /* sdes.sdes.OUVESDE  (src/sdes/sdes.py:595-698) */

/* SDE + sampler scalars.  p0 = d_lambda (mix) or theta (ouve). */
typedef struct {
  int32_t sde;     /* DITSEP_SDE_* */
  int32_t n_src;   /* sources per mixture (rows that are averaged by the mixing matrix) */
  float p0;        /* d_lambda | theta */
  float sigma_min;
  float sigma_max;
  int32_t N;       /* discretisation steps; dt = 1/N always (reference quirk, sdes.py:104) */
  float snr;       /* corrector snr */
  const float* elem_scale; /* nullable, DEVICE pointer (B, 1, L): per-mixture, per-sample multiplier of the diffusion
                              (PriorMixSDE's energy envelope, sdes.py:453-472): score terms scale with its square, noise
                              terms with it.  ABI v3. */
} ditsep_sde_params;

int ditsep_abi_version(void);
const char* ditsep_last_error(void);
/* Number of kernel launches issued through this library since load (bench "gpu_launches"). */
int64_t ditsep_launch_count(void);
/* Adds n to that count: the host side calls it when it replays a captured CUDA graph of n library launches
 * (and with -n after the capture pass, which issues no work). */
void ditsep_note_launches(int64_t n);

/* ---- fused SDE / predictor / corrector updates (fp32 state) ----------------------
 * x, score, x_mean: (B, n_src, L) contiguous; y: (B, 1, L); t: (B,).
 * noise: (B, n_src, L) tape draw, or NULL -> in-register Philox4x32-10 normals keyed by
 * (seed, draw): element e of the state uses counter (e/4, draw) lane e%4.
 * x is updated in place; x_mean (nullable) receives the noise-free update.
 * lengths (nullable, (B,) int32): ragged batch -- row b is a zero-padded signal of lengths[b] <= L samples; x and
 * x_mean are written as exactly 0 past it, so each mixture's STFT frames (and with them its score, GroupNorm
 * statistics included) equal those of a batch-1 run of that mixture (the reference evaluates at batch 1,
 * src/evaluate_mp.py:289,311).
 */
/* AnnealedLangevinDynamics2.update_fn (mix, src/sdes/correctors.py:102-121) or
 * AnnealedLangevinDynamics.update_fn (ouve, correctors.py:70-84), one step. */
int ditsep_pc_corrector(const ditsep_sde_params* host_p, float* x, const float* score, const float* noise,
                        uint64_t seed, uint64_t draw, const float* t, float* x_mean, int B, int64_t L,
                        const int* lengths, void* stream);
/* ReverseDiffusionPredictor.update_fn (src/sdes/predictors.py:60-66) with
 * SDE.discretize / RSDE.discretize (src/sdes/sdes.py:94-108,165-173). */
int ditsep_pc_predictor(const ditsep_sde_params* host_p, float* x, const float* score, const float* noise,
                        uint64_t seed, uint64_t draw, const float* t, const float* y, float* x_mean, int B,
                        int64_t L, const int* lengths, void* stream);
/* MixSDE.prior_sampling (sdes.py:336-348, n_src must be 2) / OUVESDE.prior_sampling (:687-695). */
int ditsep_pc_prior(const ditsep_sde_params* host_p, float* x, const float* noise, uint64_t seed, uint64_t draw,
                    const float* y, int B, int64_t L, const int* lengths, void* stream);
/* Standard normals from the same Philox stream as above (tests / statistics). */
/* LangevinCorrector.update_fn, one step (correctors.py:43-55): x_mean = x + eps score, x = x_mean + sqrt(2 eps) noise with
 * ONE step size eps for the whole batch, read from DEVICE memory (*step): it depends on the batch means of |score| and
 * |noise| (the host side forms it with two reductions), which is why the noise must be given explicitly. */
int ditsep_pc_langevin(int n_src, float* x, const float* score, const float* noise, const float* step, float* x_mean,
                       int B, int64_t L, const int* lengths, void* stream);
/* PriorMixSDE._std_sigma_mix (sdes.py:480-492): out (B, 1, L) = 0.5 sqrt(max(box_{avg_len}(mix^2), 1e-4)). */
int ditsep_energy_envelope(const float* mix, float* out, int B, int64_t L, int avg_len, void* stream);
/* Training / validation side (forward values only; no backward kernels are provided).
 * ditsep_pc_perturb: the marginal sample of DiffSepModel.sample_prior's default branch (src/diffsep.py:113-121,176-181):
 * x holds the clean sources x0 (B, n_src, L) on entry and x_t = mean(x0, t) + L(t) z on return, with
 * MixSDE.marginal_prob / mult_std (src/sdes/sdes.py:288-330) or OUVESDE's (:662-685) in closed form; x_mean (optional)
 * receives the mean; z from `noise` or Philox(seed, draw).  PriorMixSDE (elem_scale set) is rejected. */
int ditsep_pc_perturb(const ditsep_sde_params* host_p, float* x, const float* noise, uint64_t seed, uint64_t draw,
                      const float* t, const float* y, float* x_mean, int B, int64_t L, const int* lengths, void* stream);
/* ditsep_score_loss: loss[b] = mean over (source, sample) of (L(t_b) pred + z)^2, i.e. compute_score_loss
 * (src/diffsep.py:344-358: sde.mult_std, MSELoss(L_score, -z), mean over the last two axes).  z from `noise`, or
 * regenerated from the Philox stream ditsep_pc_perturb consumed (same seed / draw).  workspace: DEVICE scratch of at
 * least 256 * B doubles (per-block partial sums, reduced in a fixed order: bit-reproducible). */
int ditsep_score_loss(const ditsep_sde_params* host_p, const float* pred, const float* noise, uint64_t seed, uint64_t draw,
                      const float* t, float* loss, double* workspace, int B, int64_t L, void* stream);
/* The same reduction with the per-mixture standard deviation given by the caller: loss[b] = mean (sigma[b] pred + z)^2,
 * DiffSepOU.compute_score_loss (src/diffsep.py:836-841) for SDEs whose closed forms stay on the host (SBVESDE). */
int ditsep_score_loss_sigma(const float* pred, const float* noise, uint64_t seed, uint64_t draw, const float* sigma,
                            float* loss, double* workspace, int B, int n_src, int64_t L, void* stream);
/* One step of the Schroedinger-bridge samplers (get_sb_sampler, src/sdes/__init__.py:284-389: sde_sampler :288-333,
 * ode_sampler :335-382) with DiffSepOU.forward's EDM preconditioning (src/diffsep.py:813-828, coefficients :1165-1202)
 * folded into the same pass.  x, f: (B, n_src, L) f32 state and raw network output (f may be NULL: est = 0);
 * y: (B, 1, L) mixture; coef: DEVICE (B, 8) f32 rows [w_prev, w_est, w_y, w_z, c_skip, c_out, f_div, c_in_next]:
 *     est = c_skip x + c_out (f / f_div)          (c_skip = 0, c_out = f_div = 1 for an un-preconditioned model)
 *     x'  = w_prev x + w_est est + (use_noise == 0: w_y y | 1: w_z z | 2: w_y y + w_z z)   z: `noise` or Philox(seed, draw)
 *           (2 = the marginal sample mean(x0, y, t) + sigma z of DiffSepOU.sample_prior, src/diffsep.py:800-811)
 *     x_in = c_in_next x',  y_in = c_in_next y    (the scaled inputs of the NEXT network evaluation)
 * evaluated in the reference's order with individually rounded products.  Outputs x_out (may alias x), x_in, y_in,
 * est_out are each optional (at least one); rows are kept at 0 past lengths[b] when lengths is given. */
int ditsep_sb_step(const float* x, const float* f, const float* y, const float* noise, int use_noise, uint64_t seed,
                   uint64_t draw, const float* coef, float* x_out, float* x_in, float* y_in, float* est_out, int B,
                   int n_src, int64_t L, const int* lengths, void* stream);
int ditsep_philox_normal(float* out, int64_t n, uint64_t seed, uint64_t draw, void* stream);
/* The uniform -> normal transform of that stream on caller-supplied 32-bit words (n % 4 == 0; word 4q, 4q+2 give
 * the radii, 4q+1, 4q+3 the angles): lets a test feed the endpoint uniforms (0, 0xFFFFFFFF) and check for NaN. */
int ditsep_box_muller(const uint32_t* bits, float* out, int64_t n, void* stream);

/* ---- STFT + compression / expansion + iSTFT --------------------------------------
 * ScoreModelNCSNpp.pre_process (src/models/diffsep/score_models.py:107-116):
 * cat(xt, mix) -> right pad n_fft-hop -> STFT(n_fft=510, hop=128, periodic Hann, centre
 * zero pad, onesided) -> |X|^e e^{j arg X} * factor -> channels [Re c0..cn, Im c0..cn]
 * -> frames zero-padded to W; optional NCSNpp "2x-1" (ncsnpp.py:348-350) folded in.
 * exponent == 0 selects transform="log" instead (score_models.py:49-51,66-68): log1p(|X|) e^{j arg X} * |factor|, and
 * (exp(|S / factor|) - 1) e^{j arg S} on the way back; exponent == 1 with factor 1 is transform="none".
 * xt: (B, n_src, T) f32, mix: (B, 1, T) f32, window: (510,) f32.
 * feat: NHWC [B][256][W][2(n_src+1)] of `dtype`.  n_frames = 1 + (T+382)/128 <= W. */
int ditsep_stft_compress(const float* xt, const float* mix, const float* window, void* feat, int dtype, int B,
                         int n_src, int64_t T, int W, float exponent, float factor, int affine_2x_minus_1,
                         void* stream);
/* The same with the mixture channel's finished (Re, Im) kept across calls: mix_cache is a caller-owned DEVICE buffer of
 * B * W * 256 * 2 floats; mix_mode 1 transforms everything and fills it, mix_mode 2 reads the mixture channel from it
 * and transforms only the n_src source channels (the mixture does not change between the score evaluations of one
 * sampler run: score_models.py:126-131 re-transforms cat(xt, mix) every call), mix_mode 0 = ditsep_stft_compress. */
int ditsep_stft_compress_cached(const float* xt, const float* mix, const float* window, void* feat, int dtype, int B,
                                int n_src, int64_t T, int W, float exponent, float factor, int affine_2x_minus_1,
                                float* mix_cache, int mix_mode, void* stream);
/* ScoreModelNCSNpp.post_process (score_models.py:118-124): net output NHWC
 * [B][256][W][2 n_src] f32 (channel k pairs with k+n_src) -> /factor -> |S|^{1/e} e^{j arg S}
 * -> iSTFT (torch.istft centre=True) -> first T samples.  out: (B, n_src, T) f32. */
int ditsep_expand_istft(const float* spec, const float* window, float* out, int B, int n_src, int64_t T, int W,
                        float exponent, float factor, void* stream);

/* ---- GroupNorm (+SiLU) -------------------------------------------------------------
 * nn.GroupNorm(G, C, eps=1e-6) + SiLU (layerspp.py:269-271,283-285,297,318).
 *
 * Statistics travel as ORDER-INDEPENDENT FIXED-POINT SUMS: sums[B][C][4] int64 =
 * {sum int, sum frac, sumsq int, sumsq frac}, value = int + frac * 2^-32.  Producers (the convolution epilogue,
 * ditsep_gn_sums) split each fp32 partial exactly and ADD both parts with integer atomics; integer addition
 * commutes, so the totals -- and every bf16 rounding downstream -- are bit-reproducible and independent of the
 * batch size and of CTA scheduling.  A sums buffer must be ZERO before the first producer of a step adds to it. */
/* Explicit statistics pass over x: NHWC [B][HW][ld] (for tensors whose producer left no sums). */
int ditsep_gn_sums(const void* x, int dtype, int B, int64_t HW, int C, int ld, long long* sums, void* stream);

/* One GroupNorm consumer: per-(b,c) scale/shift [B][2][C] from the sums of up to 4 producers laid end to end along
 * the channel axis (torch.cat([h, skip], dim=1) feeding GroupNorm_0, ncsnpp.py:413 + layerspp.py:297). */
typedef struct {
  const long long* src_sums[4];
  int32_t src_C[4];
  int32_t n_src;
  int32_t C;      /* = sum of src_C */
  int32_t G;      /* groups, C % G == 0, G <= 64 */
  float eps;
  int64_t HW;     /* pixels per image */
  const float* gamma;
  const float* beta;
  float* scale_shift;
} ditsep_gn_job;
/* scale/shift of one job (a tiny launch: B CTAs reading C * 32 bytes each). */
int ditsep_gn_finalize(const ditsep_gn_job* host_job, int B, void* stream);
/* Allocation-free GroupNorm statistics: scale/shift [B][2][C] of nn.GroupNorm(G, C, eps) over x (NHWC [B][HW][ld]) =
 * ditsep_gn_sums + ditsep_gn_finalize, bit-identical to that pair, with the fixed-point sums held in a static
 * device array inside the library instead of a caller-allocated buffer.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation).  The call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine): three kernel launches on `stream`.  Limit: B * C <= DITSEP_GN_STATIC_MAX_BC
 * (16384 (b, c) pairs = 512 KiB of device memory, held by every GPU that loads the library until exit); larger
 * sizes are rejected with an error, use the two-call form with your own buffer.  All calls share the one array: it
 * is cleared at the head of every call, and two calls whose kernels overlap on the GPU (different streams without
 * an ordering between them) race -- serialise them. */
#define DITSEP_GN_STATIC_MAX_BC 16384
int ditsep_gn_stats_static(const void* x, int dtype, int B, int64_t HW, int C, int ld, int G, const float* gamma,
                           const float* beta, float eps, float* scale_shift, void* stream);
/* GroupNorm statistics + normalise (+ SiLU) of SMALL images in one launch (one CTA per image; y may alias x):
 * act(GroupNorm(x)) of ResnetBlockBigGANpp.forward (layerspp.py:297,318) at the 16x16 .. 4x4 levels, where the
 * three-launch sequence gn_partial_sums -> gn_finalize -> gn_apply is pure launch latency.  HW <= 4096, C <= 256. */
int ditsep_gn_small(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, int G, const float* gamma,
                    const float* beta, float eps, int silu, void* y, int ld_y, void* stream);

/* ditsep_gn_finalize + ditsep_gn_apply in one launch: every CTA forms the scale / shift of its image from the job's sums
 * itself (host_job->scale_shift is not used and may be NULL).  C = host_job->C. */
int ditsep_gn_apply_job(const void* x, int dtype, int B, int64_t HW, const ditsep_gn_job* host_job, int ld_x, int silu,
                        void* y, int ld_y, void* stream);
/* y = act(x*scale + shift), act = SiLU if silu else identity.  y: NHWC with stride ld_y. */
int ditsep_gn_apply(const void* x, int dtype, int B, int64_t HW, int C, int ld_x, const float* scale_shift,
                    int silu, void* y, int ld_y, void* stream);

/* ---- FIR resampling ([1,3,3,1] separable, factor 2) ----------------------------------
 * upsample_2d / downsample_2d (ncsnpp_utils/up_or_down_sampling.py:206-273) =
 * upfirdn2d (op/upfirdn2d_kernel.cu:107-207 / op/upfirdn2d.py:159-200).
 * x: NHWC [B][H][W][ld_x], C channels.  out_x (nullable) = FIR(x);
 * out_h (nullable) = FIR(SiLU(x*scale+shift)) with scale_shift [B][2][C] (the res-block
 * applies the resampler to both, layerspp.py:299-312).  up=1: output 2H x 2W; up=0: H/2 x W/2.
 * acc (nullable, same layout as out_x): out_x += acc (progressive pyramid sum, ncsnpp.py:436-441). */
int ditsep_fir_resample(const void* x, int dtype, int B, int H, int W, int C, int ld_x, int up,
                        const float* scale_shift, void* out_x, int ld_ox, void* out_h, int ld_oh, void* stream);

/* ---- convolution (3x3 pad 1 / 1x1, stride 1), NHWC ------------------------------------
 * nn.Conv2d via ddpm_conv3x3 / ddpm_conv1x1 (ncsnpp_utils/layers.py:112-156), NIN
 * (layers.py:678-689) with the res-block epilogue fused (layerspp.py:314-328):
 *   out[b,p,co] = out_scale * ( sum_{tap,ci} in[b,p+tap,ci] * w[co,tap,ci] + bias[co]
 *                               + bias_b[b,co] + residual[b,p,co] )
 * CUDA-core fp32 path: any Cin/Cout; weights w: [Cout][taps][Cin] f32. */
int ditsep_conv_direct(const void* in, int dtype, int B, int H, int W, int Cin, int ld_in, const float* w,
                       const float* bias, const float* bias_b, int ld_bias_b, const void* residual, int ld_res,
                       float out_scale, int Cout, int ksize, void* out, int out_dtype, int ld_out, void* stream);

/* Widens a narrow fp32 NHWC tensor x [npix][C] (the 2(n+1)-channel STFT features feeding the network's first
 * ddpm_conv3x3, ncsnpp.py:352-356) to a bf16 tensor [npix][ld] the tcgen05 path accepts, keeping fp32 accuracy:
 * channels [0,C) = hi = bf16(x), [C,2C) = lo = bf16(x - hi), [2C,3C) = hi; channels [3C, 8*ceil(3C/8)) are written as 0 and
 * channels beyond that are NOT touched (zero the buffer once; C in {3,4,6,8}).  Pair it with a filter whose
 * input channels are [wh | wh | wl] (w = wh + wl in bf16).  Needs 3*C <= ld, ld % 8 == 0. */
int ditsep_split_hilo(const float* x, int64_t npix, int C, void* out_bf16, int ld, void* stream);

/* tcgen05 implicit-GEMM path (bf16 operands, fp32 TMEM accumulators, TMA-fed).
 * Requirements: Cin % 64 == 0, Cout padded to a multiple of 16 inside `w` (and to a multiple of
 * 64 when > 128), `in` 16-byte aligned with ld_in % 8 == 0.  w: [CoutPad][taps][Cin] bf16.
 * A plan holds the encoded TMA descriptors for one (tensor, shape) binding.  Images of at least
 * 8 x 16 pixels run the tile kernel (one TMA box per 64-channel block brings an 8 x 16 pixel tile
 * plus its halo; the taps are shifted views of it; persistent CTAs over contiguous tile ranges,
 * filter resident in shared memory when it fits); smaller images one TMA box per tap.
 * `stats` (nullable, tile kernel only: ditsep_conv_tc_has_stats): the epilogue also ADDS the per-image,
 * per-channel {sum, sum of squares} of the values it stores (the GroupNorm statistics of the output tensor,
 * layerspp.py:297,318) to the fixed-point sums [B][Cout][4] described above, so the consumer's GroupNorm needs
 * no extra pass over the tensor. */
typedef struct ditsep_conv_plan ditsep_conv_plan;
int ditsep_conv_tc_plan_create(ditsep_conv_plan** plan, const void* in, int B, int H, int W, int Cin, int ld_in,
                               const void* w_bf16, int Cout, int ksize);

/* Same, with a FUSED SHORTCUT: the res-block's skip path (ResnetBlockBigGANpp.forward, layerspp.py:322-327:
 * `x = Conv_2(x)` when channels change, then `(x + h) / sqrt(2)`) is accumulated in the same TMEM accumulator as one
 * extra 1x1 tap over a second input x_sc [B][H][W][ld_x] with Cx channels, instead of a separate 1x1 launch plus a
 * residual read in the epilogue.  w: [CoutPad][taps*Cin + Cx] bf16 -- the shortcut filter (Conv_2's, or the identity
 * for a plain residual) behind the taps along K; pass bias = b_conv + b_shortcut to ditsep_conv_tc_run.
 * Needs the tile kernel with the whole filter resident in shared memory; returns a nonzero status (2) otherwise
 * (the caller then keeps the unfused sequence). */
int ditsep_conv_tc_plan_create_sc(ditsep_conv_plan** plan, const void* in, int B, int H, int W, int Cin, int ld_in,
                                  const void* w_bf16, int Cout, int ksize, const void* x_sc, int Cx, int ld_x);
void ditsep_conv_tc_plan_destroy(ditsep_conv_plan* plan);
int ditsep_conv_tc_has_stats(const ditsep_conv_plan* plan);
int ditsep_conv_tc_run(const ditsep_conv_plan* plan, const float* bias, const float* bias_b, int ld_bias_b,
                       const void* residual, int res_dtype, int ld_res, float out_scale, void* out, int out_dtype,
                       int ld_out, long long* stats, void* stream);

/* Debug aid (scripts/trace_conv.py): while buf != NULL, the tile kernel's first 4 CTAs write clock64() stamps of their
 * first 96 tiles to buf ([4][96][16] int64: 0 accumulator free, 1 first A tile landed, 2 MMAs issued, 3 accumulator
 * full seen by the epilogue, 4 TMEM read + handed back, 5 tile stored, 6 TMA issued).  NULL disarms it. */
void ditsep_debug_conv_trace(void* buf);

/* ---- attention (single head, all H*W tokens) ------------------------------------------
 * AttnBlockpp.forward core (layerspp.py:88-92): o = softmax(q k^T / sqrt(C)) v per image.
 * q, k, v, o: NHWC [B][S][ld] with C channels.  bf16 with C = 128 / 256, S >= 64 and 16-byte aligned operands runs on
 * the tensor cores (tcgen05: Q K^T and P V as UMMA tiles, V as an MN-major operand; csrc/attention_tc.cuh); fp32, C = 64
 * and shorter sequences run the CUDA-core flash kernel.  C in {64, 128, 256}. */
int ditsep_attention(const void* q, const void* k, const void* v, int dtype, int B, int S, int C, int ld_qkv,
                     void* o, int ld_o, void* stream);

/* ---- Oobleck VAE ------------------------------------------------------
 * 1-D convolutions of LatentDiffSep.encode / decode (src/diffsep_latent.py:107-128; OobleckEncoder / OobleckDecoder,
 * stable_audio_tools/models/autoencoders.py:59-85,228-357) on the tensor cores.  x: [B][T_in][Cin] bf16, DENSE rows
 * (ld = Cin), signals bs_in elements apart; w_bf16: [Cout][taps * Cin] (tap-major); tap k reads input sample
 * t * s_view + k * dil - pad (zero outside [0, T_in)), for t in [0, T_out).  The three forms of the network:
 *   dilated / plain convolution:   s_view 1, taps 7 / 3 / 1, dil d, pad d * (taps - 1) / 2, T_out = T_in
 *   strided (kernel 2s, stride s): s_view s, taps 2s, dil 1, pad s / 2, T_out = T_in / s
 *   transposed (kernel 2s, stride s, padding s / 2) in polyphase form: s_view 1, taps 2, dil -1, pad 0, T_out = T_in + 1,
 *     Cout = s * C with w[f * C + co][m * Cin + ci] = W[ci][co][f + s * m]; the buffer's rows (j, f), flattened, are the
 *     output delayed by s / 2 samples: the consumer reads it from row s / 2 on (bs = (T_in + 1) * s * C).
 * Epilogue: + bias[Cout] (nullable) + residual (nullable, bf16, rows ld_res / signals bs_res apart); `out` (nullable)
 * receives the value (bf16, or fp32 when out_f32), `out_act` (nullable) ELU of it in bf16 -- the next convolution's
 * input, so the network's ELUs cost no pass of their own.  Rows ld_out, signals bs_out elements apart. */
int ditsep_conv1d_tc(const void* x, int B, int T_in, int Cin, int64_t bs_in, const void* w_bf16, int Cout, int taps,
                     int dil, int pad, int s_view, int T_out, const float* bias, const void* residual, int ld_res,
                     int64_t bs_res, void* out, void* out_act, int out_f32, int ld_out, int64_t bs_out, void* stream);
/* OobleckEncoder's first layer: x (B, T) fp32 -> conv7 (1 -> C channels, w [C][7], bias [C]) -> out, out_act = ELU(out),
 * both [B][T][C] bf16. */
int ditsep_conv1d_in(const float* x, const float* w, const float* bias, void* out, void* out_act, int B, int T, int C,
                     void* stream);
/* OobleckDecoder's last layer: x = ELU'd activations [B][T][C] bf16 (rows ld, signals bs apart) -> conv7 (C -> 1, w [7][C],
 * no bias) -> tanh (when final_tanh) -> out (B, T) fp32. */
int ditsep_conv1d_out(const void* x, int ld, int64_t bs, const float* w, float* out, int B, int T, int C, int final_tanh,
                      void* stream);
/* VAEBottleneck.encode (bottleneck.py:57-65): pre [B][T][2L] fp32 = (mean | scale), noise (B, L, T) fp32 ->
 * z (B, L, T) = noise * (softplus(scale) + 1e-4) + mean. */
int ditsep_vae_sample(const float* pre, const float* noise, float* z, int B, int T, int L, void* stream);

/* ---- time embedding ---------------------------------------------------------------------
 * ncsnpp.py:325-346 + layerspp.py:44-46: act_temb = SiLU(Linear2(SiLU(Linear1(
 * [sin, cos](2 pi W log t))))) -- the SiLU every res-block applies before Dense_0
 * (layerspp.py:317) is folded in.  W: (nf,), w1: (4nf, 2nf), w2: (4nf, 4nf).  out: (B, 4nf). */
int ditsep_temb_mlp(const float* t, const float* gfp_w, const float* w1, const float* b1, const float* w2,
                    const float* b2, float* act_temb, int B, int nf, void* stream);
/* All res-blocks' Dense_0 at once: out[b, r] = dense_w[r, :] . act_temb[b, :] + dense_b[r],
 * dense_w: (R, 4nf) rows of all blocks concatenated.  out: (B, R). */
int ditsep_temb_dense(const float* act_temb, const float* dense_w, const float* dense_b, float* out, int B, int K,
                      int R, void* stream);

/* ---- small elementwise helpers ------------------------------------------------------------
 * output layer (ncsnpp.py:473-477): out[b,p,:] = Wout (x[b,p,:] / t[b]) + bout, 1x1 conv
 * Cin -> Cout on the progressive pyramid, fp32 out.  x NHWC of `dtype`. */
int ditsep_output_layer(const void* x, int dtype, int B, int64_t HW, int Cin, int ld_x, const float* t,
                        int scale_by_sigma, const float* w, const float* bias, int Cout, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DITSEP_B200_H */
