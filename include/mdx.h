/*
 * mdx.h — C ABI of libmdx_b200.so, the sm_100a kernel library behind the minddiffusion denoising hot path
 * (UNetModel forward inside the PLMS / DPM-Solver++ / DDIM sampler loops).
 *
 * The reference (mindspore-lab/minddiffusion) has NO native/FFI layer: every op on the path is a MindSpore
 * primitive called from Python.  The entry points below are therefore the operator-level
 * boundary a maintainer would bind with ctypes in place of those primitives; each one cites the reference call
 * site it replaces (paths relative to vision/stablediffusionv2/ unless noted).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes.  Every pointer is a DEVICE pointer unless its name ends in _host.
 *  - The caller owns every buffer (inputs, outputs, scratch); the library allocates nothing.
 *  - `stream` is a cudaStream_t passed as void*.  All work is enqueued on it; nothing synchronises.
 *  - Return value: 0 = ok, negative = error (-22 invalid argument, -5 CUDA failure, -95 unsupported).
 *    mdx_last_error() returns a thread-local message for the last failure.
 *  - Activations are NHWC ("pixel-major"): a tensor [n, H, W, C] is a matrix of n*H*W pixel rows whose row
 *    stride (in elements) is passed explicitly (ld*), so channel slices of wider buffers are valid operands.
 *  - dtype codes: MDX_F32 = 0 (fp32 check mode), MDX_BF16 = 1.
 */
#ifndef MDX_H_
#define MDX_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MDX_F32 0
#define MDX_BF16 1

const char* mdx_last_error(void);
int mdx_version(void);
/* Number of kernels this library has launched in the calling process (all threads). bench.py reports it. */
int64_t mdx_launch_count(void);

/* ---------------------------------------------------------------------------------------------------------
 * Tensor-core implicit GEMM (tcgen05.mma, TMEM accumulators, TMA-fed, persistent).
 *   out[p, n] = epilogue( sum_seg sum_tap sum_c A_seg[pixel(p)+tap][c] * B[n][koff(seg,tap,c)] )
 * Replaces nn.Conv2d 3x3/1x1 (ldm/modules/diffusionmodules/util.py:38-52, openaimodel.py:151-174,205),
 * nn.Dense (util.py:31-35; ldm/modules/attention.py:44,108-114,63-67) and the ResBlock residual add
 * (openaimodel.py:205) / GEGLU product (attention.py:47-51) / emb add (openaimodel.py:199-200) as epilogues.
 *
 *  A: 1..3 K-segments. Segment s is a bf16 NHWC view [n_img, H, W, seg_c[s]] with pixel stride lda[s];
 *     seg_taps[s] is 9 (3x3, pad 1, stride 1) or 1 (1x1 / dense); 4 only with up_phase (below). seg_c[s] % 64 == 0.
 *     A dense matrix [M, K] is passed as n_img = 1, H = 1, W = M.
 *  B: bf16 [N, Ktot] row-major (K contiguous), Ktot = sum_s taps_s * seg_c[s], K order (segment, tap=(r,s), c).
 *  epilogue, in this order: + bias[n] (fp32, optional) ; + rowbias[img(p) * ld_rowbias + n] (fp32, optional)
 *     ; geglu (optional): N is 2*I laid out per BLOCK_N tile as [value half | gate half] (see pack_geglu in
 *     minddiffusion_b200/weights.py), out[p, j] = value * gelu_tanh(gate), out has I columns
 *     ; + residual[p * ldr + n] (bf16 or fp32, optional) ; store bf16 (out_dtype MDX_BF16) or fp32 (MDX_F32).
 *  block_n: 0 = choose automatically; otherwise one of 64, 128, 160, 192, 256 (geglu packing must match).
 * --------------------------------------------------------------------------------------------------------- */
typedef struct mdx_gemm_args {
  int32_t nseg;
  const void* a[3];
  int64_t lda[3];
  int32_t seg_c[3];
  int32_t seg_taps[3];
  int32_t n_img, H, W;
  const void* b;
  int32_t N;
  void* out;
  int64_t ldc;
  int32_t out_dtype;
  const float* bias;
  const float* rowbias;
  int64_t ld_rowbias;
  const void* residual;
  int64_t ldr;
  int32_t residual_dtype; /* MDX_BF16 or MDX_F32 (the fp32 residual trunk, see DESIGN.md "precision") */
  int32_t geglu;
  int32_t block_n;
  int32_t act; /* applied after bias/rowbias, before the residual: 0 none, 1 GELU (tanh form, nn.GELU() of
                  encoders/text_encoder.py:67-69), 2 QuickGELU x*sigmoid(1.702x) (wukong text_encoder.py:67-74),
                  3 tanh (SRGAN output layer, Taichu-GLIDE model/srgan.py:103) */
  int32_t up_phase; /* 0 = off.  1..4: this launch computes phase (py, px) = ((up_phase-1) >> 1, (up_phase-1) & 1) of
                  Upsample = nearest x2 followed by conv3x3 (openaimodel.py:53-63) WITHOUT materialising the
                  upsampled image: output pixel (2y+py, 2x+px) only sees the 2x2 source window at offset
                  (py-1, px-1), so segment 0 has seg_taps = 4 and B holds the phase's folded weights
                  [N, (a, b, c)], a/b = window row/column (see fold_upsample_weights in unet.py).  out, residual
                  and rowbias are addressed in the upsampled [n_img, 2H, 2W] image; H, W stay the source size. */
  float* chan_stats; /* optional [n_img, slots, 2, N] fp32 (16-byte aligned, need not be initialised), slots =
                  mdx_gemm_tc_stat_slots(n_img, H, W, up_phase != 0): for every 32 output rows (one slot) the epilogue
                  stores the per-channel sum (plane 0) and sum of squares (plane 1) of the values it wrote — the
                  GroupNorm statistics of the tensor's consumer (mdx_groupnorm_pre), so that no separate pass
                  re-reads the activation.  No atomics: each slot is written once and the consumer adds the slots in
                  a fixed order.  The four up_phase launches of one tensor fill disjoint slots of one array.
                  Needs images of >= 128 pixels (mdx_gemm_tc_stat_slots returns 0 otherwise). */
  int32_t stride; /* 0 / 1: as above.  2: the single 3x3 segment is a stride-2, pad-1 convolution (Downsample,
                  openaimodel.py:80-88): H, W stay the INPUT image, out / residual / rowbias / chan_stats are addressed in
                  the [n_img, (H-1)/2+1, (W-1)/2+1] output image (size chan_stats with those dimensions).  The taps are
                  strided TMA boxes over the input; no patch matrix is built. */
} mdx_gemm_args;
/* Launch form: calls with convolution segments (seg_taps > 1) run as thread-block clusters of two CTAs — as a CTA
   pair (tcgen05 cta_group::2: two consecutive 128-row tiles, one M = 256 MMA over both SMs, the weight block halved
   between them) when the number of 128-row tiles is even, else as two CTAs sharing each activation block by TMA
   multicast — only when every operand is ordinary device memory of the CURRENT device (checked with
   cudaPointerGetAttributes; anything else, e.g. a mapping of another GPU's memory, takes the plain launch).  Results
   are bit-identical in all forms, except that a CTA-pair launch of ONE plain 3x3 segment (stride 1, tiles inside one
   image, BLOCK_N 160) loads each input row block once per dx and accumulates the taps dx-outermost: the same sum in
   another fp32 order (<= 2e-5 relative).  Environment: MDX_GEMM_HALO=0 disables that, MDX_GEMM_CG2=0 the CTA-pair
   form, MDX_GEMM_MC=0 every cluster form. */
int mdx_gemm_tc(const mdx_gemm_args* args, void* stream);
/* Statistics slots per image for mdx_gemm_args.chan_stats (0: images too small, use the two-pass GroupNorm). */
int mdx_gemm_tc_stat_slots(int32_t n_img, int32_t H, int32_t W, int32_t up);
/* BLOCK_N the library would pick for a GEMM with N output columns (needed to pack GEGLU weights). */
int mdx_gemm_tc_pick_block_n(int32_t N, int32_t geglu);

/* ---------------------------------------------------------------------------------------------------------
 * CUDA-core implicit-GEMM convolution / dense layer (any Cin/Cout, stride, padding, optional nearest-2x
 * upsample of the input).  fp32 check mode for every conv/dense of the path, and in bf16 mode the layers
 * tensor cores do not fit: conv_in (Cin=4, openaimodel.py:350-353), conv_out (Cout=4, openaimodel.py:521-526),
 * the time-embedding MLP (openaimodel.py:340-345) and ResBlock emb_layers (openaimodel.py:153-160).
 *  x: [n, H, W, Cin] (pixel stride ldx), dtype in_dtype; w: [Cout, kh*kw*Cin] same dtype as x, K order (r,s,c).
 *  upsample != 0: x is read through ops.ResizeNearestNeighbor x2 (openaimodel.py:53-60): logical input is
 *     [n, 2H, 2W, Cin] with src = dst >> 1.
 *  out: [n, OH, OW, Cout] (pixel stride ldo), dtype out_dtype; bias/rowbias fp32 optional; residual has the
 *     out dtype.  pre_silu != 0 applies SiLU to x on load (emb_layers: SiLU -> Dense).
 * --------------------------------------------------------------------------------------------------------- */
typedef struct mdx_conv_args {
  const void* x;
  int64_t ldx;
  int32_t in_dtype;
  int32_t n, H, W, Cin;
  const void* w;
  int32_t Cout, kh, kw, stride, pad, upsample, pre_silu;
  void* out;
  int64_t ldo;
  int32_t out_dtype;
  const float* bias;
  const float* rowbias;
  int64_t ld_rowbias;
  const void* residual;
  int64_t ldr;
  int32_t act; /* as mdx_gemm_args.act */
} mdx_conv_args;
int mdx_conv_direct(const mdx_conv_args* args, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * GroupNorm (+ optional SiLU) over NHWC, optionally over the channel-concatenation of two tensors
 * (the skip concat ops.Concat(axis=1) of openaimodel.py:566-571 is never materialised separately).
 * Replaces GroupNorm32(32, C) fp32 + nn.SiLU (util.py:87-108; openaimodel.py:134-137,162-163,521-523) and
 * Normalize (attention.py:83-84).  Statistics are biased (1/N) variance, accumulated in fp64.
 *  x1: [n, HW, C1] (ld1), x2: [n, HW, C2] (ld2) or NULL with C2 = 0 (both in_dtype).  gamma/beta: fp32 [C1+C2].
 *  out: [n, HW, C1+C2] with pixel stride ldo (out_dtype).  Supported dtype pairs: fp32->fp32 (check mode),
 *     bf16->bf16, fp32->bf16 (fp32 residual trunk normalised into a bf16 GEMM operand).
 *  raw_copy: optional [n, HW, C1+C2] (ldraw, out_dtype): the un-normalised concat(x1, x2), written in the same
 *     pass — the bf16 operand of the ResBlock's 1x1 skip_connection (openaimodel.py:167-174).  NULL to skip.
 *  stats_scratch: caller-provided, >= n*groups*2 doubles; overwritten.
 * --------------------------------------------------------------------------------------------------------- */
int mdx_groupnorm(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2, int32_t n,
                  int32_t HW, int32_t groups, float eps, const float* gamma, const float* beta, int32_t silu,
                  void* out, int64_t ldo, void* raw_copy, int64_t ldraw, int32_t in_dtype, int32_t out_dtype,
                  double* stats_scratch, void* stream);

/* mdx_groupnorm_pre: mdx_groupnorm (and, with scale_shift != NULL, mdx_groupnorm_film) without the statistics pass.
 *  chan1 / chan2: [n, slots1, 2, C1] / [n, slots2, 2, C2] fp32 partial sums written by the producers of x1 / x2
 *  (mdx_gemm_args.chan_stats).  A small kernel (one warp per image and group, fixed order) folds them into the
 *  per-group sums — any grouping of concat(x1, x2), so the skip concat of openaimodel.py:566-571 works even where
 *  a group straddles the two tensors — then the apply pass runs. */
int mdx_groupnorm_pre(const void* x1, int64_t ld1, int32_t C1, const float* chan1, int32_t slots1, const void* x2,
                      int64_t ld2, int32_t C2, const float* chan2, int32_t slots2, int32_t n, int32_t HW,
                      int32_t groups, float eps, const float* gamma, const float* beta, const float* scale_shift,
                      int64_t ld_ss, int32_t silu, void* out, int64_t ldo, void* raw_copy, int64_t ldraw,
                      int32_t in_dtype, int32_t out_dtype, double* stats_scratch, void* stream);
/* GroupNorm without caller scratch.  (No counterpart in the reference: MindSpore hands its kernels a workspace;
 * this exists for callers of THIS library that cannot provide one.)
 * Same operands and dtype pairs as mdx_groupnorm, minus stats_scratch: the per-group sums live in a fixed
 * 256 KiB device array inside the library (present on every GPU that loads it, for the life of the process).
 * Meant for call sites that cannot allocate device memory when they reach the norm — e.g. a framework custom-op
 * body that receives only its input/output addresses and a stream, or a thread that must stay out of the
 * allocator because the host framework's memory pool is frozen.
 *  - limit: n * groups <= 16384, otherwise -22 with a message pointing back at mdx_groupnorm;
 *  - the first call on each device resolves the array's device address (cudaGetSymbolAddress, with that device
 *    current); issue that first call before beginning any stream capture;
 *  - never calls cudaMalloc/cudaFree or any allocating routine;
 *  - every call on a device shares the array: each call clears what it uses on `stream` before accumulating,
 *    so consecutive calls on one stream are fine, but calls that may overlap on the GPU race with each other. */
int mdx_groupnorm_noalloc(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2,
                          int32_t n, int32_t HW, int32_t groups, float eps, const float* gamma, const float* beta,
                          int32_t silu, void* out, int64_t ldo, void* raw_copy, int64_t ldraw, int32_t in_dtype,
                          int32_t out_dtype, void* stream);

/* LayerNorm over the last dim of [rows, C] (ldx/ldo row strides). Replaces nn.LayerNorm([dim], epsilon=1e-5)
 * (attention.py:176-178).  Biased variance, fp32 statistics, one warp per row.  dtype pairs as mdx_groupnorm. */
int mdx_layernorm(const void* x, int64_t ldx, int32_t rows, int32_t C, float eps, const float* gamma,
                  const float* beta, void* out, int64_t ldo, int32_t in_dtype, int32_t out_dtype, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Multi-head attention  o = softmax(q k^T * scale) v  without materialising q k^T
 * (CrossAttention.construct, attention.py:117-166; heads are interleaved in the channel dim, so the
 * rearange_in/out transposes at attention.py:124-133,154-163 become strides).
 *  q: [batch, Nq, heads*d] (row stride ldq), k/v: [batch, Nk, heads*d] (ldk/ldv), o: [batch, Nq, heads*d] (ldo).
 *  Batch strides are Nq*ldq etc. (tensors are dense over batch).  dtype MDX_BF16, d <= 64, not causal: the
 *  tcgen05 kernels of attention_tc.cu (S / P / O in tensor memory, TMA-fed; a two-CTA-per-SM single-tile variant
 *  when Nk <= 128; d < 64 is zero-padded to 64 by the TMA box); larger head dims and causal: flash kernel on
 *  mma.sync tensor cores; all with fp32 softmax.
 *  MDX_F32: CUDA-core check-mode kernel.  d <= 160, d % 8 == 0.
 *  causal != 0: key j is visible to query i only if j <= i (the text tower's triu(-inf, 1) mask,
 *  encoders/text_encoder.py:131-134); uses the mma.sync kernel for every head dim.
 * --------------------------------------------------------------------------------------------------------- */
int mdx_attention(const void* q, int64_t ldq, const void* k, int64_t ldk, const void* v, int64_t ldv, void* o,
                  int64_t ldo, int32_t batch, int32_t heads, int32_t Nq, int32_t Nk, int32_t d, float scale,
                  int32_t dtype, int32_t causal, void* stream);

/* Row softmax out[r, :] = softmax(x[r, :] * scale) (fp32 math).  Used by the VAE decoder's single-head attention over
 * 512 channels (diffusionmodules/model.py:190-204), whose head dim exceeds the flash kernels': the scores are
 * materialised by mdx_gemm_tc, normalised here, and multiplied with V by mdx_gemm_tc.  in: fp32 or bf16, out: bf16/fp32. */
int mdx_softmax_rows(const void* x, int64_t ldx, int64_t rows, int32_t cols, float scale, void* out, int64_t ldo,
                     int32_t in_dtype, int32_t out_dtype, void* stream);

/* Token embedding + positional embedding (encoders/text_encoder.py:143-148): out[i, :] = table[ids[i], :] + pos[i % L, :].
 * table: fp32 [vocab, width], pos: fp32 [L, width], ids: int32 [n] (n = batch * L), out: [n, width] (dtype). */
int mdx_embedding(const float* table, int32_t vocab, int32_t width, const float* pos, int32_t L, const int32_t* ids,
                  int32_t n, void* out, int32_t dtype, void* stream);

/* GEGLU product for check mode: out[m, j] = x[m, j] * gelu_tanh(x[m, I + j]) (attention.py:47-51). */
int mdx_geglu(const void* x, int64_t ldx, int32_t rows, int32_t I, void* out, int64_t ldo, int32_t dtype,
              void* stream);

/* Sinusoidal timestep embedding [cos | sin], fp32 (util.py:111-131). t: fp32 [n] (may be fractional). */
int mdx_timestep_embedding(const float* t, int32_t n, int32_t dim, float max_period, float* out, void* stream);

/* Layout / dtype movers.
 *  mdx_nchw_to_nhwc: x fp32 NCHW [n, C, H, W] -> out [n*repeat, H, W, C] (dtype), batch index taken mod n:
 *     the CFG ops.concat((x, x)) of plms.py:192 without a copy of its own.
 *  mdx_nhwc_to_nchw: x [n, H, W, C] (dtype) -> fp32 NCHW.
 *  mdx_upsample_nearest2x: ops.ResizeNearestNeighbor (openaimodel.py:53-60) on NHWC (fp32->bf16 converts).
 *  mdx_concat_channels: ops.Concat(axis=1) (openaimodel.py:567) on NHWC rows.
 *  mdx_im2col3x3: gathers 3x3 patches (stride, pad 1) into [n*OH*OW, 9*C] rows, K order (r,s,c) — the patch-matrix
 *     form of a strided conv (the UNet's Downsample now runs as mdx_gemm_args.stride = 2; kept for callers that
 *     want the matrix).
 *  mdx_cast: elementwise fp32 <-> bf16 conversion of count elements. */
int mdx_nchw_to_nhwc(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t repeat, void* out,
                     int32_t dtype, void* stream);
/* same, output rows padded with zeros to Cpad >= C channels (conv_in, Cin = 4, is fed to mdx_gemm_tc as 64) */
int mdx_nchw_to_nhwc_pad(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t repeat, int32_t Cpad,
                         void* out, int32_t dtype, void* stream);
int mdx_nhwc_to_nchw(const void* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t dtype, float* out,
                     void* stream);
/* same, reading the first C channels of rows with pixel stride ldx >= C */
int mdx_nhwc_to_nchw_ld(const void* x, int64_t ldx, int32_t n, int32_t C, int32_t H, int32_t W, int32_t dtype,
                        float* out, void* stream);
int mdx_upsample_nearest2x(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, void* out, int32_t in_dtype,
                           int32_t out_dtype, void* stream);
int mdx_concat_channels(const void* x1, int64_t ld1, int32_t C1, const void* x2, int64_t ld2, int32_t C2,
                        int64_t rows, void* out, int64_t ldo, int32_t dtype, void* stream);
int mdx_im2col3x3(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, int32_t stride, void* out,
                  int32_t in_dtype, int32_t out_dtype, void* stream);
int mdx_cast(const void* x, int32_t in_dtype, void* out, int32_t out_dtype, int64_t count, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Sampler arithmetic: ONE elementwise kernel per step (classifier-free-guidance combine + latent update).
 *
 * mdx_sampler_plms_step — PLMSSampler.p_sample_plms (ldm/models/diffusion/plms.py:182-248) and, with
 *   w = {1,0,0,0}, the DDIM update (get_x_prev_and_pred_x0, plms.py:210-228).
 *    eps_model: UNet output, NHWC [cfg ? 2B : B, h, w, C] with pixel stride ld_eps (dtype model_dtype); rows [0,B) are the
 *       unconditional half, [B,2B) the conditional half (concat order of plms.py:192-194).
 *    e_t    = cfg ? e_u + scale*(e_c - e_u) : eps            (plms.py:195-197)
 *    e'     = w[0]*e_t + w[1]*old1 + w[2]*old2 + w[3]*old3    (plms.py:231-244; old* fp32 NCHW or NULL)
 *    pred_x0 = (x - sqrt(1-a_t) e') / sqrt(a_t)
 *    x_prev  = sqrt(a_prev) pred_x0 + sqrt(1 - a_prev - sigma^2) e' + sigma*noise   (noise NULL => 0)
 *    coef_host[4] = {a_t, a_prev, sigma_t, sqrt_one_minus_a_t}; x, x_prev, pred_x0, e_t_out fp32 NCHW [B,C,h,w];
 *    pred_x0 / e_t_out may be NULL.  x_prev may alias x.
 *
 * mdx_sampler_dpmpp2m_step — DPM-Solver++ multistep (dpm_solver/dpm_solver.py:256-329 model_wrapper,
 *   366-373 data_prediction_fn, 488-520 first update, 742-777 second update).
 *    out    = cfg-combined model output as above; model_type 0: noise = out; 1 ("v", dpm_solver.py:281-284):
 *             noise = alpha_s*out + sigma_s*x
 *    x0_cur = (x - sigma_s*noise) / alpha_s
 *    x_next = c[2]*x + c[3]*x0_cur + c[4]*(x0_cur - x0_prev)      (x0_prev NULL => c[4] ignored)
 *    coef_host[5] = {alpha_s, sigma_s, c_x, c_0, c_1}.  x0_cur is written (history for the next step).
 * --------------------------------------------------------------------------------------------------------- */
int mdx_sampler_plms_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale, const float* x,
                          const float* old1, const float* old2, const float* old3, const float* w_host,
                          const float* coef_host, const float* noise, float* x_prev, float* pred_x0,
                          float* e_t_out, int32_t B, int32_t C, int32_t h, int32_t w, void* stream);
int mdx_sampler_dpmpp2m_step(const void* eps_model, int64_t ld_eps, int32_t model_dtype, int32_t cfg, float scale,
                             int32_t model_type, const float* x, const float* x0_prev, const float* coef_host,
                             float* x_next, float* x0_cur, int32_t B, int32_t C, int32_t h, int32_t w,
                             void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Struct form of the two step kernels above, with three additions:
 *  (1) per-step scalars either from the host (coef_host, one row) or from a DEVICE table coef_table[rows][MDX_STEP_COEFS]
 *      indexed by the device counter *step_dev — a step's launches are then identical from step to step, so the whole
 *      step (UNet + this kernel + mdx_step_advance) is captured once in a CUDA graph and replayed (the host loop of
 *      plms.py:146-177 / dpm_solver.py:1073-1094 only replays);
 *  (2) model_type 1 ("v") for the PLMS / DDIM update too: e_t = row[9] * out + row[10] * x_model, the conversion
 *      model_wrapper applies (dpm_solver.py:281-284) with (row[9], row[10]) = (sqrt(abar_t), sqrt(1 - abar_t)) of the
 *      timestep the model was evaluated at and x_model the latent it was evaluated on (NULL = x);
 *  (3) shift_hist: the PLMS eps history is advanced inside the kernel (hist3 <- hist2 <- hist1 <- e_t, plms.py:168-170)
 *      so the buffers keep their addresses across steps; unused history must be zero-filled, not uninitialised.
 * Rows (MDX_STEP_COEFS floats, unused entries 0):
 *   PLMS/DDIM  {w0, w1, w2, w3, sqrt(a_t), sqrt(a_prev), sqrt(1 - a_prev - sigma^2), sigma, sqrt(1 - a_t), v_a, v_s,
 *               use_noise} — fill with mdx_sampler_plms_coef_row (host; the arithmetic mdx_sampler_plms_step uses)
 *   DPM++(2M)  {alpha_s, sigma_s, c_x, c_0, c_1, use_prev}   (use_prev 0 = first-order update: hist[0] is not read)
 * hist[0..2]: PLMS old eps 1..3 (may be NULL without shift_hist); DPM: hist[0] = x0_prev, pred_x0 = x0_cur (may alias).
 * x_out may alias x.
 * --------------------------------------------------------------------------------------------------------- */
#define MDX_STEP_COEFS 16
typedef struct mdx_step_args {
  const void* model_out;
  int64_t ld_out;
  int32_t model_dtype;
  int32_t cfg;
  float scale;
  int32_t model_type;
  const float* x;
  const float* x_model;
  float* hist[3];
  const float* noise;
  float* x_out;
  float* pred_x0;
  float* e_t_out;
  int32_t shift_hist;
  int32_t B, C, h, w;
  const float* coef_host;
  const float* coef_table;
  const int32_t* step_dev;
} mdx_step_args;
int mdx_sampler_plms_step_ex(const mdx_step_args* args, void* stream);
int mdx_sampler_dpmpp2m_step_ex(const mdx_step_args* args, void* stream);
/* Host helper: w4 = AB weights, coef4 = {a_t, a_prev, sigma_t, sqrt_one_minus_a_t} (as mdx_sampler_plms_step),
 * v2 = {v_a, v_s} or NULL (eps model) -> row[MDX_STEP_COEFS]. */
int mdx_sampler_plms_coef_row(const float* w4_host, const float* coef4_host, const float* v2_host, int32_t use_noise,
                              float* row_host);
/* *step_dev += delta (one thread; the last launch of a captured step). */
int mdx_step_advance(int32_t* step_dev, int32_t delta, void* stream);
/* out[j][0..width) = table[*idx_dev + offset][0..width) for j < n_out (table row stride ld): the timestep vector of
 * the step (plms.py:148 `ts = full((b,), step)`) read from a device table inside a captured step. */
int mdx_gather_row(const float* table, int64_t ld, const int32_t* idx_dev, int32_t offset, float* out, int32_t n_out,
                   int32_t width, void* stream);
/* Inpainting blend of the sampler loop (wukong-huahua plms.py:153-156 with q_sample, ddpm.py:180-182):
 *   img = (sqrt_ac * x0 + sqrt_1m_ac * noise) * mask + (1 - mask) * img, in place.
 * x0, noise, img: fp32 NCHW [B, C, h, w]; mask: fp32 [B, mask_channels, h, w], mask_channels 1 (broadcast) or C. */
int mdx_sampler_mask_blend(const float* x0, const float* noise, const float* mask, int32_t mask_channels, float sqrt_ac,
                           float sqrt_1m_ac, float* img, int32_t B, int32_t C, int32_t h, int32_t w, void* stream);

/* Struct sizes as this library was compiled — a binding compares them with its own declaration before the first call
 * (a short struct would make the library read fields from whatever follows it in memory). */
int32_t mdx_sizeof_gemm_args(void);
int32_t mdx_sizeof_conv_args(void);
int32_t mdx_sizeof_step_args(void);
int32_t mdx_offsetof_gemm_args_chan_stats(void);

/* ---------------------------------------------------------------------------------------------------------
 * Taichu-GLIDE family (vision/Taichu-GLIDE/model/glide_text2im): the operators its UNet / sampler need beyond
 * the latent-diffusion set above.
 *
 * mdx_groupnorm_film — ResBlock with use_scale_shift_norm (model/unet.py:189-194):
 *      out = GroupNorm(x) * (1 + scale) + shift, optional SiLU (out_layers_1).  scale_shift: fp32 rows
 *      [n][scale(C) | shift(C)] with row stride ld_ss (the emb_layers output, split at unet.py:190).  Other
 *      arguments as mdx_groupnorm (single source, no raw copy).
 * mdx_avgpool2x2 — Downsample without conv, nn.AvgPool2d(2, 2) (unet.py:69-73; resblock_updown h_upd/x_upd) on NHWC.
 * mdx_copy_rows — dst[b][r][:] = src[b][r][:] for b < n_batch, r < rows, width_bytes per row; all strides in
 *      bytes, multiples of 16.  batch_stride_src_bytes = 0 broadcasts one block of rows.  Builds the
 *      [encoder K/V | spatial K/V] sequences of QKVAttention (unet.py:262-266, ops.Concat(axis=-1)) and
 *      CombinePrompt's broadcast of the per-step random prompt (guider.py:36-41).
 * mdx_embedding_masked — Text2ImUNet.get_text_emb (text2im_model.py:79-84):
 *      out[i] = mask[i] ? table[ids[i]] + pos[i % L] : pad[i % L]   (mask NULL => all ones)
 * mdx_resize_bilinear — ops.ResizeBilinear(size, align_corners=False) (text2im_model.py:166,191) on fp32 NCHW;
 *      quantize_u8 != 0 first applies PMeanVariance.superResPreprocess (gaussian_diffusion.py:224-230):
 *      round((v + 1) * 127.5) / 127.5 - 1 to every source pixel.
 * mdx_nchw2_to_nhwc_pad — mdx_nchw_to_nhwc_pad over ops.Concat(axis=1)((x1, x2)) (text2im_model.py:192; the
 *      'hybrid' / 'concat' DiffusionWrapper of wukong-huahua ddpm.py:363-372).  x1 has n images, read `repeat` times;
 *      x2 has n2 = n (same) or n2 = n*repeat images (one per output image: the [uncond | cond] c_concat).
 * mdx_sampler_glide_step — Guider + PMeanVariance + PSample / DDimSample in one pass
 *      (guider.py:60-70, gaussian_diffusion.py:158-181, :49-63, :82-103).
 *      model_out: NHWC [cfg ? 2P : P, h, w, >= 2C]: channels [0,C) eps, [C,2C) variance fraction v; rows [0,P) are
 *      the prompt half, [P,2P) the random-prompt half (guider.py:63: cond first — the opposite of plms.py).
 *      eps   = cfg ? e_u + scale*(e_c - e_u) : e_c
 *      x0    = clip(c[0]*x - c[1]*eps, -1, 1)                    (c[0] = sqrt(1/abar), c[1] = sqrt(1/abar - 1))
 *      mode 0 (ancestral): logvar = f*c[5] + (1-f)*c[4], f = (v+1)/2 read from the prompt half;
 *                          x_prev = c[2]*x0 + c[3]*x + c[9] * exp(0.5*logvar) * noise
 *      mode 1 (DDIM):      e' = (c[0]*x - x0)/c[1]; sigma = c[8]*sqrt((1-c[7])/(1-c[6]))*sqrt((1-c[6])/c[7]);
 *                          x_prev = sqrt(1-c[7]-sigma^2)*e' + sqrt(c[7])*x0 + c[9]*sigma*noise
 *      coef_host[10] = {sqrt_recip_abar, sqrt_recipm1_abar, post_coef1, post_coef2, min_log, max_log, abar,
 *                       abar_prev, eta, t != 0}.  x, noise, x_prev, pred_xstart: fp32 NCHW [P, C, h, w].
 * --------------------------------------------------------------------------------------------------------- */
int mdx_groupnorm_film(const void* x, int64_t ldx, int32_t C, int32_t n, int32_t HW, int32_t groups, float eps,
                       const float* gamma, const float* beta, const float* scale_shift, int64_t ld_ss, int32_t silu,
                       void* out, int64_t ldo, int32_t in_dtype, int32_t out_dtype, double* stats_scratch,
                       void* stream);
int mdx_avgpool2x2(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, void* out, int32_t in_dtype,
                   int32_t out_dtype, void* stream);
int mdx_copy_rows(const void* src, int64_t ld_src_bytes, int64_t batch_stride_src_bytes, void* dst,
                  int64_t ld_dst_bytes, int64_t batch_stride_dst_bytes, int32_t n_batch, int32_t rows,
                  int64_t width_bytes, void* stream);
int mdx_embedding_masked(const float* table, int32_t vocab, int32_t width, const float* pos, const float* pad,
                         int32_t L, const int32_t* ids, const int32_t* mask, int32_t n, void* out, int32_t dtype,
                         void* stream);
int mdx_resize_bilinear(const float* x, int32_t n, int32_t C, int32_t H, int32_t W, int32_t OH, int32_t OW,
                        int32_t quantize_u8, float* out, void* stream);
int mdx_nchw2_to_nhwc_pad(const float* x1, int32_t C1, const float* x2, int32_t C2, int32_t n, int32_t n2, int32_t H,
                          int32_t W, int32_t repeat, int32_t Cpad, void* out, int32_t dtype, void* stream);
int mdx_sampler_glide_step(const void* model_out, int64_t ld_out, int32_t model_dtype, int32_t cfg, float scale,
                           int32_t mode, const float* x, const float* coef_host, const float* noise, float* x_prev,
                           float* pred_xstart, int32_t P, int32_t C, int32_t h, int32_t w, void* stream);

/* SRGAN x4 stage of the GLIDE cascade (model/srgan.py).
 * mdx_prelu — nn.PReLU(C) on NHWC rows: out = prelu(x; slope[c]) (+ residual, added after the activation; slope
 *      NULL = plain add: `conv1 + conv2` of Generator.construct, srgan.py:100, and ResidualBlock's `out + x`, :54).
 * mdx_depth_to_space2 — ops.DepthToSpace(2) (srgan.py:62) on NHWC fused with the PReLU that follows it:
 *      out[n, 2h+i, 2w+j, c] = prelu(x[n, h, w, (2i+j)*C + c]); x has 4C channels, slope may be NULL. */
int mdx_prelu(const void* x, int64_t ldx, int64_t rows, int32_t C, const float* slope, const void* residual,
              int64_t ldr, void* out, int64_t ldo, int32_t in_dtype, int32_t out_dtype, void* stream);
int mdx_depth_to_space2(const void* x, int32_t n, int32_t H, int32_t W, int32_t C, const float* slope, void* out,
                        int32_t in_dtype, int32_t out_dtype, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MDX_H_ */
