/* b200gan.h — C ABI of libb200gan.so: the sm_100a kernels behind the SNDCGAN training step.
 *
 * The reference (jonathan-schilling/imageGeneration) has NO native/FFI interface: its hot path is Python that
 * composes Keras layers (sndcgan/SNDCGAN.py:25-128 builders, :241-269 train_step) and every device op is
 * dispatched by TensorFlow. The drop-in boundary is therefore the Python API (imagegeneration_b200/sndcgan.py
 * mirrors it name for name); this header is the C boundary *underneath* that API — one entry point per TF op
 * family that train_step executes. Each declaration cites the reference call site whose TF op it replaces.
 *
 * Conventions: plain pointers + sizes only (no torch types); every pointer is a DEVICE pointer unless stated;
 * all calls are asynchronous on `stream`; nothing is allocated or freed by the library; return 0 on success,
 * <0 = error code below (message via b2g_last_error()); never throws.
 */
#ifndef B200GAN_H_
#define B200GAN_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* b2g_stream_t; /* cudaStream_t */

enum {
  B2G_OK = 0,
  B2G_E_INVALID_ARG = -1,
  B2G_E_UNSUPPORTED_SHAPE = -2,
  B2G_E_CUDA = -3,
  B2G_E_DRIVER = -4
};

int b2g_version(void);
/* SM budget of the persistent kernels (process-wide, set from the one host thread that launches): every launcher sizes
 * its grid for min(device SMs, n) SMs; n = 0 restores the whole device.  Data-parallel runs leave a few SMs to the
 * NCCL all-reduce kernel that overlaps the discriminator steps — a persistent grid of one CTA per SM otherwise waits
 * for the SMs the collective holds and runs as two waves.  b2g_num_sms returns the current budget. */
int b2g_set_sm_limit(int n);
int b2g_num_sms(void);
const char* b2g_last_error(void); /* thread-local, valid until the next failing call on this thread */

/* ---- implicit-GEMM convolutions on tcgen05 tensor cores (TMA-fed, TMEM accumulators) -------------------
 * Activations are NHWC bf16. Weights are *packed* bf16 [phases][Cout][taps*Cin] (K-major rows), produced by
 * b2g_pack_conv_weight from the fp32 master weights kept in the reference's Keras layouts.
 *
 * kind:
 *   B2G_CONV_K3S1   3x3 stride-1 SAME conv (9 taps).           Hout=Hin,   Wout=Win
 *       replaces Conv2D k3 s1 fwd (SNDCGAN.py:87-90,101-104,115-118) and, with flipped/transposed packing,
 *       its Conv2DBackpropInput (tape.gradient at :248,:256,:266).
 *   B2G_CONV_K4S2   4x4 stride-2 SAME conv (16 taps).          Hout=Hin/2, Wout=Win/2
 *       replaces Conv2D k4 s2 fwd (SNDCGAN.py:80-83,94-97,108-111) and the input-gradient of
 *       Conv2DTranspose k4 s2 (generator backward, :248).
 *   B2G_CONVT_K4S2  4x4 stride-2 SAME transposed conv as 4 sub-pixel phases of 2x2 taps. Hout=2Hin, Wout=2Win
 *       replaces Conv2DTranspose k4 s2 fwd (SNDCGAN.py:36-39,43-46,50-53) and the input-gradient of
 *       Conv2D k4 s2 (discriminator backward).
 *   B2G_CONV_K1S1   1x1 conv (1 tap, no shift) = plain GEMM over pixels. Used for the 3-channel edge layers
 *       after b2g_im2col27 has expanded [N,H,W,3] into 27(+1) bf16 channels: D.conv1 forward
 *       (SNDCGAN.py:73-76) and the input-gradient of G.convT4 (:57-61).
 *   B2G_CONV_K3S1_C3  3x3 stride-1 SAME conv over a 3-CHANNEL fp32 tensor: `in` is fp32 [N,H,W,3], Cin = 3, and the
 *       27(+1 ones)-wide im2col patch of every pixel is built in shared memory by the kernel itself (no im2col27
 *       tensor in HBM).  wpacked is the K1S1 packing of the [27][Cout] matrix with ck_pad = 64; Cout = 64.
 *       Same layers as K1S1: D.conv1 forward (SNDCGAN.py:73-76) and the input-gradient of G.convT4 (:57-61).
 */
enum { B2G_CONV_K3S1 = 0, B2G_CONV_K4S2 = 1, B2G_CONVT_K4S2 = 2, B2G_CONV_K1S1 = 3, B2G_CONV_K3S1_C3 = 4 };
enum { B2G_ACT_NONE = 0, B2G_ACT_LEAKY = 1, B2G_ACT_RELU = 2, B2G_ACT_TANH = 3 };

typedef struct b2g_conv_args {
  int kind;
  int N, Hin, Win, Cin; /* input activation, NHWC */
  int Cout;             /* GEMM-N: channels of the produced tensor */
  const void* in;       /* bf16 [N,Hin,Win,Cin] */
  const void* wpacked;  /* bf16 [phases][Cout][taps*Cin] */
  void* out;            /* bf16 [N,Hout,Wout,Cout] */
  /* fused epilogue, applied in this order to the fp32 accumulator */
  const float* bias; /* per-Cout fp32, or NULL                  (BiasAdd / BatchNorm beta, SNDCGAN.py:40,73) */
  int act;           /* B2G_ACT_*                                (LeakyReLU :77, ReLU :41)                   */
  float alpha;       /* LeakyReLU slope                                                                      */
  int dropout_thresh8;         /* 0 = off; element dropped when its hash byte < thresh8 (p = thresh8/256);
                                  survivors scaled by 256/(256-thresh8)            (Dropout, SNDCGAN.py:78) */
  unsigned long long dropout_seed;
  /* backward use: multiply by the activation derivative recovered from the saved forward output `yprev`
     (same shape as out): dact=LEAKY: y>0 -> dscale, y<0 -> alpha*dscale, y==0 -> dzero (dropped element);
     dact=RELU: y>0 -> 1 else 0; dact=TANH: (1-y^2). */
  const void* yprev; /* same dtype/shape as out, or NULL */
  int dact;
  float dact_alpha, dact_scale, dact_zero;
  /* narrow fp32 outputs (G.convT4 forward + tanh, D.conv1 input-gradient): out_f32 = 1 stores fp32 and
     cout_store (> 0, phases == 1 only) is the number of REAL output channels: wpacked then has cout_store
     rows, `out`/`yprev` are [N,Hout,Wout,cout_store] and Cout is only the MMA N tile (32). */
  int out_f32;
  int cout_store;
  /* tf32 = 1: the fp32 / tf32 path of north_star: `in`, `out`, `yprev` are fp32 NHWC (Cin % 32 == 0), wpacked is the
     fp32 packing produced by b2g_pack_conv_weight_f32, the MMAs are tcgen05 kind::tf32 (fp32 accumulate). */
  int tf32;
  /* optional fp32 [Cout] (bf16 outputs, no bias, dact != NONE — a backward epilogue): += the per-channel sums of
     the stored outputs over all pixels, i.e. the bias / BatchNorm-beta gradient of the layer whose output
     gradient this call produces (replaces a separate reduction pass over `out`). */
  float* colsum;
  /* optional device pointer to the 64-bit dropout seed (overrides dropout_seed when non-NULL): lets a captured CUDA
     graph of a whole training step be replayed with fresh masks — b2g_step_prologue writes the seeds of a step. */
  const unsigned long long* dropout_seed_dev;
  /* optional fp32 [Cout], only together with colsum (then also allowed on a FORWARD call: dact == NONE, no bias, no
     dropout): += the per-channel sums of squares of the stored outputs.  colsum + colsumsq are the batch statistics of
     a training-mode BatchNormalization behind this conv (wasserstein_gan/WGAN.py:60-92,116-128) — no separate
     reduction pass over the activation (feed them to b2g_bn_finalize as copy 0 of its [8][2C] sums). */
  float* colsumsq;
} b2g_conv_args;

int b2g_conv_igemm(const b2g_conv_args* a, b2g_stream_t stream);

/* Repack an fp32 master conv weight (Keras layout [kh][kw][d2][d3]) into the bf16 K-major layout a
 * b2g_conv_igemm role needs, folding a scale (1/sigma of spectral norm, or gamma/sqrt(var+eps) of the
 * inference-mode BatchNorm that follows the conv, SNDCGAN.py:40) so the scaled fp32 weight is never
 * materialised.
 *   n_is_d3 = 1: GEMM-N indexes d3 and GEMM-K indexes d2 (Conv2D fwd with HWIO; ConvT dgrad with HWOI)
 *   n_is_d3 = 0: GEMM-N indexes d2 and GEMM-K indexes d3 (ConvT fwd with HWOI; Conv2D dgrad with HWIO)
 *   flip    = 1: spatially flipped taps (dgrad of a stride-1 conv)
 *   scale_scalar (device ptr to one float, may be NULL) multiplies everything (use 1/sigma);
 *   scale_d2 / scale_d3 (device ptr to per-d2 / per-d3 floats, may be NULL).
 *   ck_pad (0 = none): pad the per-tap GEMM-K extent with zeros up to ck_pad (K1S1 edge layers: 27 -> 64).
 */
int b2g_pack_conv_weight(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad, const float* w,
                         const float* scale_scalar, const float* scale_d2, const float* scale_d3, void* wpacked,
                         b2g_stream_t stream);

/* Same packing with fp32 output (tf32 path; the tensor core reads the top 19 bits of every fp32 operand). */
int b2g_pack_conv_weight_f32(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad,
                             const float* w, const float* scale_scalar, const float* scale_d2, const float* scale_d3,
                             void* wpacked, b2g_stream_t stream);

/* Several bf16 packings in ONE launch (a train step re-packs every conv weight of a model after each optimizer
 * update; see sndcgan/SNDCGAN.py:250,:258,:268 apply_gradients).  `jobs` is a HOST array of 1..24 entries whose
 * fields mean what the arguments of b2g_pack_conv_weight mean. */
typedef struct b2g_pack_job {
  int kind, kh, kw, D2, D3, n_is_d3, flip, ck_pad;
  const float* w;
  const float* scale_scalar;
  const float* scale_d2;
  const float* scale_d3;
  void* wpacked;
  int f32; /* 0: bf16 packing; 1: fp32 packing rounded to tf32 (what b2g_pack_conv_weight_f32 writes) */
} b2g_pack_job;
int b2g_pack_conv_weight_multi(const b2g_pack_job* jobs, int n_jobs, b2g_stream_t stream);

/* fp32 [N,H,W,3] -> bf16 [N,H,W,64]: channel tap*3+c (tap = ky*3+kx) holds in[y+ky-1, x+kx-1, c] (zero outside),
 * channel 27 holds 1.0 (its wgrad row is the bias gradient), channels 28..63 are zero. */
int b2g_im2col27(const float* in, void* out, int N, int H, int W, b2g_stream_t stream);

/* ---- conv / convT weight gradient on tensor cores ---------------------------------------------------------
 *   dw[tap][a][b] += sum_{n,y,x} shifted[n, tap(y,x), a] * unshifted[n, y, x, b]      (fp32 atomics; the
 *   caller zero-fills dw first).  `kind` is the tap plan of the shifted operand:
 *     Conv2D k3 s1  [3,3,Cin,Cout]:  kind=K3S1, shifted = layer input x   (Ca=Cin),  unshifted = dY (Cb=Cout)
 *     Conv2D k4 s2  [4,4,Cin,Cout]:  kind=K4S2, shifted = layer input x   (Ca=Cin),  unshifted = dY (Cb=Cout)
 *     Conv2DTranspose k4 s2 [4,4,Cout,Cin]: kind=K4S2, shifted = dY (Ca=Cout), unshifted = layer input (Cb=Cin)
 *     3-channel edge layers [3,3,3,64]:     kind=K1S1, shifted = im2col27(3-channel tensor), unshifted = 64-ch tensor
 *       or, without the im2col27 tensor:    kind=K3S1_C3, shifted = the fp32 [N,Hs,Ws,3] tensor itself (Ca=3),
 *       unshifted = bf16 [N,Hs,Ws,64] (Cb=64), dw = [27][64] (row tap*3+c), dbias = [64] or NULL — the patch is
 *       built in shared memory inside the kernel
 *   Hs, Ws = spatial size of the shifted operand; the unshifted one is (Hs,Ws) for K3S1, (Hs/2,Ws/2) for K4S2.
 *   Replaces Conv2DBackpropFilter behind tape.gradient (sndcgan/SNDCGAN.py:248,:256,:266). */
typedef struct b2g_wgrad_args {
  int kind;
  int N, Hs, Ws;
  int Ca, Cb;
  const void* shifted;   /* bf16 [N,Hs,Ws,Ca] */
  const void* unshifted; /* bf16 [N,Hg,Wg,Cb] */
  float* dw;             /* fp32 [taps][Ca][Cb], accumulated into */
  /* K1S1 edge layers (shifted = b2g_im2col27 output, Ca = 64): only rows a < ca_valid are real weights
     (dw is [ca_valid][Cb]); row a == ca_valid (the all-ones channel) is added to dbias[Cb] when non-NULL.
     ca_valid = 0 means Ca. */
  int ca_valid;
  float* dbias;
  /* 1: shifted / unshifted are fp32 and the products run as tcgen05 kind::tf32 (K3S1 / K4S2 only) — the
     fp32/tf32 per-layer parity path, see b2g_conv_args.tf32. */
  int tf32;
} b2g_wgrad_args;

int b2g_conv_wgrad(const b2g_wgrad_args* a, b2g_stream_t stream);

/* ---- 3-channel edge layers (direct convolutions; HBM / CUDA-core bound) ---------------------------------
 * w is the fp32 Keras kernel [3][3][3][64] of either D.conv1 (Conv2D HWIO, SNDCGAN.py:73-76) or G.convT4
 * (Conv2DTranspose HWOI with 3 filters, :57-61); wscale (device scalar or NULL) multiplies it (1/sigma).
 *   c3_to_c64 : out bf16 [N,H,W,64] = act(conv3x3(in fp32 [N,H,W,3]) + bias) [dropout] [* relu'(yprev)]
 *               -> D.conv1 forward (act=LEAKY, dropout) and G.convT4 input-gradient (dact=RELU, yprev=a3).
 *   c64_to_c3 : out fp32 [N,H,W,3] = act(flipped conv3x3(in bf16 [N,H,W,64])) [* tanh'(yprev fp32)]
 *               -> G.convT4 forward (act=TANH) and D.conv1 input-gradient (dact=TANH, yprev=g_output).
 *   wgrad     : dw[27][64] += sum_pix t3[pix+tap][c3] * t64[pix][c64]; dbias[64] += sum_pix t64 (or NULL)
 *               -> D.conv1 (t3 = images, t64 = dY) and G.convT4 (t3 = dY fp32, t64 = layer input a3). */
int b2g_conv3x3_c3_to_c64(const float* in, const float* w, const float* wscale, const float* bias, void* out, int N,
                          int H, int W, int act, float alpha, int dropout_thresh8, unsigned long long dropout_seed,
                          const void* yprev, int dact, b2g_stream_t stream);
int b2g_conv3x3_c64_to_c3(const void* in, const float* w, const float* wscale, float* out, int N, int H, int W,
                          int act, const float* yprev, int dact, b2g_stream_t stream);
int b2g_wgrad3x3_c3_c64(const float* t3, const void* t64, float* dw, float* dbias, int N, int H, int W,
                        b2g_stream_t stream);

/* ---- generator stem: Dense(128->F, no bias) + inference BatchNorm + ReLU (SNDCGAN.py:31-34) ------------
 * fwd: out bf16 [B,F] = relu(gamma*inv_std * (z W) + beta).   bwd: from dq bf16 [B,F] (= dL/d(BN output),
 * ReLU' already applied) computes dW [128,F], dgamma [F], dbeta [F] (overwritten). */
int b2g_dense_g_fwd(const float* z, const float* W, const float* gamma, const float* beta, float inv_std, void* out,
                    int B, int F, b2g_stream_t stream);
int b2g_dense_g_bwd(const float* z, const float* W, const float* gamma, float inv_std, const void* dq, float* dW,
                    float* dgamma, float* dbeta, int B, int F, b2g_stream_t stream);
/* Same forward with LeakyReLU(alpha) instead of ReLU, 0 <= alpha < 1 (WGAN generator stem, Dense + LeakyReLU(0.2),
 * wasserstein_gan/WGAN.py:113-114: pass gamma = 1, beta = 0, inv_std = 1). */
int b2g_dense_g_fwd_leaky(const float* z, const float* W, const float* gamma, const float* beta, float inv_std,
                          float alpha, void* out, int B, int F, b2g_stream_t stream);

/* ---- WGAN path: training-mode BatchNormalization + RMSprop with weight clipping --------
 * Activations are bf16 NHWC viewed as [P][C] (P = B*H*W; C a power of two in 8..2048).  Replaces the
 * FusedBatchNormV3(is_training=True) / FusedBatchNormGradV3 ops behind the BatchNormalization layers of
 * wasserstein_gan/WGAN.py:62-92 (critic) and :118-128 (generator) under train_on_batch (:298,:303,:317).
 * bn_stats:    sums = fp32 [8][2C] partial sums (zeroed inside; 8 copies spread the same-address reds):
 *              sum_q sums[q][c] = sum_p x[p][c], sum_q sums[q][C+c] = sum_p x[p][c]^2.
 * bn_finalize: stat[c] = mean, stat[C+c] = 1/sqrt(biased var + eps); when moving_mean/var are given they move with
 *              `momentum` (Keras 0.99), the variance Bessel-corrected (P/(P-1)) as on Keras' fused path.
 *              `bias` (or NULL): the conv in front ran WITHOUT its bias (BatchNorm cancels it exactly) and its fused
 *              colsum / colsumsq epilogue produced `sums`; the moving mean then tracks mean + bias.
 * bn_inference_stat: stat from the moving statistics (generator.predict, WGAN.py:224); `bias` as above (subtracted).
 * bn_apply_act: y = leaky_alpha(gamma (x - mean) invstd + beta), bf16 [P][C].
 * bn_bwd:      g = dL/d(BN output) (activation derivative already applied), x = the BN input of the forward:
 *              dx = gamma invstd (g - mean(g) - xhat mean(g xhat)); dgamma[c] = sum g xhat, dbeta[c] = sum g
 *              (either may be NULL); red = fp32 [8][2C] scratch.  Two launches.
 * rmsprop:     Keras RMSprop (rho, momentum 0, not centred): rms = rho rms + (1-rho) g^2; w -= lr g/(sqrt(rms)+eps)
 *              (WGAN.py:99,:150); g is multiplied by gscale first; the first clip_n elements of the arena are
 *              clamped to [-clip, clip] afterwards (ClipConstraint, WGAN.py:33-44 — store the clipped kernels first). */
int b2g_bn_stats(const void* x, float* sums, long P, int C, b2g_stream_t stream);
int b2g_bn_finalize(const float* sums, long P, int C, float eps, float momentum, float* stat, float* moving_mean,
                    float* moving_var, const float* bias, b2g_stream_t stream);
int b2g_bn_inference_stat(const float* moving_mean, const float* moving_var, int C, float eps, float* stat,
                          const float* bias, b2g_stream_t stream);
int b2g_bn_apply_act(const void* x, const float* stat, const float* gamma, const float* beta, float alpha, void* y,
                     long P, int C, b2g_stream_t stream);
int b2g_bn_bwd(const void* g, const void* x, const float* stat, const float* gamma, float* red, void* dx,
               float* dgamma, float* dbeta, long P, int C, b2g_stream_t stream);
int b2g_rmsprop(float* w, const float* g, float* rms, long n, float lr, float rho, float eps, float gscale,
                long clip_n, float clip, b2g_stream_t stream);

/* ---- evaluation path: AveragePooling2D(8, 8) of the discriminator's conv7 activation ----
 * x bf16 [N,H,W,C] -> out fp32 [N,H/8,W/8,C] (stride 8, 'valid': trailing rows / columns are dropped) — the layer the
 * reference appends to the popped discriminator for its FID-like score (sndcgan/generator_evaluation.py:134-138). */
int b2g_avgpool8_bf16(const void* x, float* out, int N, int H, int W, int C, b2g_stream_t stream);

/* ---- discriminator head: Flatten + Dense(1) (SNDCGAN.py:122-123) fused with the loss (:170) ------------
 * fwd_loss: logits[b] = wscale * <x[b,:], w> + bias; loss_kind 0 = BCE-with-logits vs `label`,
 *           1 = D hinge, 2 = G hinge, 3 = Wasserstein label*x (wasserstein_gan/WGAN.py:48-49); loss_out[0] = inv_batch*sum loss, loss_out[1] = inv_batch*sum sigmoid(x);
 *           dlogit[b] = inv_batch * dloss/dlogit.  workspace: >= 64*B floats.
 * bwd:      dx bf16 [B,F] = dlogit[b]*wscale*w[j]*act'(x[b,j]) (LeakyReLU/Dropout derivative of conv7 from its
 *           saved output x); dw[j] += sum_b dlogit[b]*x[b,j]; db += sum_b dlogit[b]  (dw/db may be NULL). */
int b2g_dhead_fwd_loss(const void* x, const float* w, const float* wscale, const float* bias, int B, int F,
                       int loss_kind, float label, float inv_batch, float* workspace, int workspace_floats,
                       float* logits, float* dlogit, float* loss_out, b2g_stream_t stream);
int b2g_dhead_bwd(const void* x, const float* w, const float* wscale, const float* dlogit, int B, int F, float alpha,
                  float dscale, float dzero, void* dx, float* dw, float* db, b2g_stream_t stream);
/* b2g_dhead_fwd_loss without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and results as b2g_dhead_fwd_loss (bf16 x) minus `workspace`:
 * the partial sums live in a 256 KiB static device array inside the library (held by every GPU that loads it until the
 * process exits).  Limit: B * min(64, ceil(4 * SMs / B)) <= 65536 floats, i.e. any batch up to 1024; larger batches
 * return B2G_E_UNSUPPORTED_SHAPE.  Calls no cudaMalloc / cudaFree / allocating routine.  All calls share the array:
 * calls that overlap on the GPU (different streams) race — serialise them. */
int b2g_dhead_fwd_loss_static(const void* x, const float* w, const float* wscale, const float* bias, int B, int F,
                              int loss_kind, float label, float inv_batch, float* logits, float* dlogit,
                              float* loss_out, b2g_stream_t stream);

/* ---- parameter-side kernels -------------------------------------------------------------------------------
 * adam: Keras-2 Adam on a flat fp32 arena (all variables of one network): one launch per apply_gradients
 *       (SNDCGAN.py:249,:257,:267).  lr_t = lr*sqrt(1-b2^t)/(1-b1^t) is computed by the caller; g is
 *       multiplied by gscale first (1/world_size after a sum all-reduce).
 * colsum_bf16: out[c] += sum_p x[p][c] (BiasAddGrad / BatchNorm-beta gradient).
 * bn_fold_finalize: Conv2DTranspose kernel W[taps][Co][Ci] whose forward used W*gamma[co]*inv_std:
 *       dgamma[co] = inv_std*sum W*G, then G *= gamma[co]*inv_std in place.
 * spectral_norm_multi: one power iteration for up to 8 matrices W_l [K_l][C_l] in ONE launch reading each W once
 *       (row blocks balanced by bytes; the last block of a layer finalises sigma and u);
 *       scal_l = float[8] {scratch, sigma, 1/sigma, |W u|, ticket (uint), -, -, -}; t_raw_l and scal_l must be
 *       zero on first use (the kernel re-zeroes its scratch).
 * sn_grad_finalize: G <- G/sigma - (<G,W>/sigma^2) v u^T in place (gradient through W/sigma). */
int b2g_adam(float* w, const float* g, float* m, float* v, long n, float lr_t, float beta1, float beta2, float eps,
             float gscale, b2g_stream_t stream);
/* Adam with lr_t read from device memory (one float) — the graph-replay form of b2g_adam. */
int b2g_adam_dev(float* w, const float* g, float* m, float* v, long n, const float* lr_t_dev, float beta1,
                 float beta2, float eps, float gscale, b2g_stream_t stream);
/* Data-parallel Adam over NVSwitch multicast (NVLS) memory: the gradient all-reduce, the update and the weight broadcast
 * of optimizer.apply_gradients (sndcgan/SNDCGAN.py:257,:267) under data parallelism (north_star item 4) in ONE pass.
 * w_mc / g_mc: MULTICAST addresses of the weight / gradient arenas (every rank's replica bound to one multicast object,
 * e.g. torch.distributed._symmetric_memory); w_local, m, v: this rank's own buffers.  The calling rank owns elements
 * [begin, end) (multiples of 4): it reads the sum of all replicas' gradients (multimem.ld_reduce), applies Keras Adam
 * with gscale (1/world) and its local m / v — optimizer state is sharded by slice — and stores the new weights into
 * every replica (multimem.st).  lr_t_dev non-NULL: the rate is read from device memory (graph replay).  The caller
 * must put a cross-rank barrier before the call (all gradients final) and after it (all weights written).
 * No counterpart in the reference (single-GPU TensorFlow). */
int b2g_adam_nvls(void* w_mc, const void* g_mc, const float* w_local, float* m, float* v, long begin, long end,
                  float lr_t, const float* lr_t_dev, float beta1, float beta2, float eps, float gscale,
                  b2g_stream_t stream);
/* First node of a captured training step (sndcgan/SNDCGAN.py:241-269 replayed as one CUDA graph): reads the step
 * index s = *step_ctr, writes the 21 dropout seeds of the step (3 discriminator passes x 7 layers:
 * splitmix64(seed_base ^ ((3 s + pass) * 8 + layer))) and the three bias-corrected Adam rates
 * lr_t = lr * sqrt(1 - b2^t) / (1 - b1^t) for t = s + 1 (generator) and t = 2 s + 1, 2 s + 2 (discriminator), then
 * stores s + 1. */
int b2g_step_prologue(long long* step_ctr, unsigned long long seed_base, unsigned long long* seeds21, float* lr3,
                      float lr_g, float lr_d, float beta1, float beta2, b2g_stream_t stream);
int b2g_colsum_bf16(const void* x, float* out, long P, int C, b2g_stream_t stream);
int b2g_bn_fold_finalize(const float* W, float* G, const float* gamma, float inv_std, float* dgamma, int taps, int Co,
                         int Ci, b2g_stream_t stream);
int b2g_spectral_norm_multi(int n_layers, const float* const* W, float* const* u, float* const* v_raw,
                            float* const* t_raw, float* const* scal, const int* K, const int* C,
                            b2g_stream_t stream);
int b2g_sn_grad_finalize(const float* W, float* G, const float* u, const float* v_raw, const float* scal,
                         float* dot_scratch, int K, int C, b2g_stream_t stream);
/* Spectral norm made free (north_star item 2): the whole spectral-norm bookkeeping of one apply_gradients rides on
 * the optimizer's own pass over the discriminator arena.
 * sn_dot_multi: dots[l] += <G_l, W_l> for the n_layers matrices at element offsets off[l] (K[l] x C[l], row-major)
 *       inside the arenas w / g — one launch.  dots must be zero on first use; b2g_adam_sn re-zeroes it.
 * adam_sn: Keras Adam over the arena [0, n) where, for each listed matrix, the gradient is first mapped through
 *       d(W/sigma)/dW (g <- g/sigma - (dots[l]/sigma^2) v u^T, with u, v_raw, scal as the forward pass used them),
 *       and the NEXT power iteration (v_raw = W' u, t = W'^T v_raw, sigma, u) is computed from the updated rows
 *       while they are in registers; same state layout as b2g_spectral_norm_multi.  lr_t_dev (device float) takes
 *       precedence over lr_t when non-NULL (CUDA-graph replay).  Elements outside the matrices (biases) get plain
 *       Adam.  Matrices must be listed in ascending offset order. */
int b2g_sn_dot_multi(const float* w, const float* g, int n_layers, const long* off, const int* K, const int* C,
                     float* dots, b2g_stream_t stream);
int b2g_adam_sn(float* w, const float* g, float* m, float* v, long n, int n_layers, const long* off, const int* K,
                const int* C, float* const* u, float* const* v_raw, float* const* t_raw, float* const* scal,
                float* dots, float lr_t, const float* lr_t_dev, float beta1, float beta2, float eps, float gscale,
                b2g_stream_t stream);

/* ---- fp32 variants for the fp32/tf32 parity path (SNDCGANEngine(dtype="tf32"), north_star: per-layer <= 1e-3) ----
 * The 64-channel side of the 3-channel edge layers, the discriminator head's input and the generator stem's output /
 * incoming gradient are fp32 tensors here; arithmetic is exact fp32 on CUDA cores for the edge layers and the head,
 * kind::tf32 (forward) / bf16 hi+lo split products (backward, `scratch` = 2*B*F bf16 elements) for the stem. */
int b2g_conv3x3_c3_to_c64_f32(const float* in, const float* w, const float* wscale, const float* bias, float* out, int N,
                              int H, int W, int act, float alpha, int dropout_thresh8, unsigned long long dropout_seed,
                              const float* yprev, int dact, b2g_stream_t stream);
int b2g_conv3x3_c64_to_c3_f32(const float* in, const float* w, const float* wscale, float* out, int N, int H, int W,
                              int act, const float* yprev, int dact, b2g_stream_t stream);
int b2g_wgrad3x3_c3_c64_f32(const float* t3, const float* t64, float* dw, float* dbias, int N, int H, int W,
                            b2g_stream_t stream);
int b2g_dhead_fwd_loss_f32(const float* x, const float* w, const float* wscale, const float* bias, int B, int F,
                           int loss_kind, float label, float inv_batch, float* workspace, int workspace_floats,
                           float* logits, float* dlogit, float* loss_out, b2g_stream_t stream);
int b2g_dhead_bwd_f32(const float* x, const float* w, const float* wscale, const float* dlogit, int B, int F, float alpha,
                      float dscale, float dzero, float* dx, float* dw, float* db, b2g_stream_t stream);
int b2g_dense_g_fwd_f32(const float* z, const float* W, const float* gamma, const float* beta, float inv_std, float* out,
                        int B, int F, b2g_stream_t stream);
int b2g_dense_g_bwd_f32(const float* z, const float* W, const float* gamma, float inv_std, const float* dq, void* scratch,
                        float* dW, float* dgamma, float* dbeta, int B, int F, b2g_stream_t stream);

/* ---- small kernels either side of the step (csrc/b2g_misc.cu) -------------------------------------------------
 * dact: out = g * act'(y) elementwise from the saved forward output y (LeakyReLU/Dropout: y > 0 -> scale, y < 0 ->
 *       alpha*scale, y == 0 -> zero; ReLU; tanh: 1 - y^2).  f32 = 0: bf16 tensors (n % 8 == 0); 1: fp32.  The stand-alone
 *       form of the derivative epilogues, for the torch.library autograd formulas (train_step uses the fused ones).
 * bce_logits: tf.keras.losses.BinaryCrossentropy(from_logits=True) (sndcgan/SNDCGAN.py:170): loss_out[0] =
 *       mean_b[max(x,0) - x*y + log1p(exp(-|x|))]; dlogit (may be NULL) = (sigmoid(x) - y)/B.
 * gather_u8_to_f32: out[b] = images[idx[b]] / 127.5 - 1 from a device-resident uint8 cache (get_dataset's Rescaling +
 *       cache + shuffle, SNDCGAN.py:141-143); per_image = bytes of one image, a multiple of 16.
 * colsum_f32: out[c] += sum_p x[p][c] (bias gradient on the fp32/tf32 path). */
int b2g_dact(const void* g, const void* y, void* out, long n, int f32, int kind, float alpha, float scale, float zero,
             b2g_stream_t stream);
int b2g_bce_logits(const float* logits, const float* labels, int B, float* loss_out, float* dlogit,
                   b2g_stream_t stream);
int b2g_gather_u8_to_f32(const void* images_u8, const long* idx, float* out, int B, long per_image,
                         b2g_stream_t stream);
int b2g_colsum_f32(const float* x, float* out, long P, int C, b2g_stream_t stream);

/* ---- CycleGAN path (cyclegan/CycleGAN.py) — csrc/b2g_cyclegan.cu ----------------------------
 * Convolutions of arbitrary geometry run as gather + GEMM: b2g_im2col writes the bf16 matrix [Mpad][Kpad] (row = output
 * pixel (n, oy, ox), column = (ky, kx, c); padding rows / columns zero) that b2g_conv_igemm (kind K1S1) multiplies with
 * a b2g_pack_gemm_weight matrix and b2g_conv_wgrad (kind K1S1) reduces into the Keras-layout weight gradient.
 *   im2col mode 0: in[n, oy*stride + ky - pad_t, ox*stride + kx - pad_l, c]  (Conv2D forward :98,:131,:142 — `reflect`
 *                  folds ReflectionPadding2D :37-57 into the gather —, Conv2DTranspose input gradient);
 *          mode 1: in[n, (oy + pad_t - ky)/stride, (ox + pad_l - kx)/stride, c] where the divisions are exact
 *                  (Conv2DTranspose forward :152, Conv2D input gradient).   `in`: bf16 or fp32 NHWC, channel stride ldc.
 *   pack_gemm_weight: out[n][tap*Ck + c] bf16 [Npad][Kpad] from fp32 w[tap][D2][D3]; (n, c) = (d3, d2) if n_is_d3 else (d2, d3).
 *   reflect_fold: gradient of ReflectionPadding2D: dx[N,H,W,C] from dpad[N,H+2p,W+2p,(ldc)].
 *   rownorm_fwd/bwd: tensorflow_addons InstanceNormalization(axis=1) on NHWC as the reference uses it (:65,69,99,132,143,
 *                  153) = statistics per image ROW (n, h) over W x C, gamma/beta [H], with the following activation
 *                  (none / ReLU / LeakyReLU / tanh) and, when `res` is given, the ResBlock's add + ReLU in front (:89-91);
 *                  a/da: bf16 with channel stride ldc, y/dy: bf16 or fp32 [.., C]; stat: 2 floats per row; dgamma/dbeta +=.
 *   l1_loss: loss_out[0] += scale * mean|a - b|, da = scale * sign(a - b) / n   (calc_cycle_loss / identity_loss :199-207). */
int b2g_im2col(const void* in, int in_f32, int N, int Hin, int Win, int C, int ldc, int kh, int kw, int mode, int stride,
               int pad_t, int pad_l, int reflect, int Hout, int Wout, void* out, long Mpad, int Kpad, b2g_stream_t stream);
int b2g_pack_gemm_weight(const float* w, int taps, int D2, int D3, int n_is_d3, void* out, int Npad, int Kpad,
                         b2g_stream_t stream);
int b2g_reflect_fold(const void* dpad, void* dx, int N, int H, int W, int C, int ldc, int pad, b2g_stream_t stream);
int b2g_rownorm_fwd(const void* a, const void* res, const float* gamma, const float* beta, void* y, float* stat, int N,
                    int H, int W, int C, int ldc, int act, float alpha, float eps, int y_f32, b2g_stream_t stream);
int b2g_rownorm_bwd(const void* dy, const void* y, const void* a, const void* res, const float* gamma, const float* stat,
                    void* da, float* dgamma, float* dbeta, int N, int H, int W, int C, int ldc, int act, float alpha,
                    int y_f32, b2g_stream_t stream);
int b2g_l1_loss(const float* a, const float* b, long n, float scale, float* loss_out, float* da, b2g_stream_t stream);

/* ---- host-only introspection (no CUDA calls; used by the CPU test-suite) ---------------------------------
 * conv_plan_query: tiling + tap tables of a conv role -> out[0..15] = {phases, taps, Hg, Wg, bw, bh, bn, tiles_w,
 *   tiles_h, tiles_n, m_tiles, Hout, Wout, out_stride, n_input_maps, 0}; out[16+i] = input map of tap entry i,
 *   out[32+i] = dy, out[48+i] = dx (entry = phase*taps + tap); out[64+p] / out[68+p] = output phase offsets.
 * pack_source_index: flat index into the fp32 Keras kernel that packed element i is taken from (-1 = padding).
 * conv_rows_plan_query: the row-slab schedule (csrc/b2g_conv_rows.cuh) b2g_conv_igemm uses for layers >= 128 pixels
 *   wide with Cout <= 128, for phase group pg: out[0..7] = {handled, R, P, PG, n_acc, row_blocks, n_slabs, n_steps};
 *   out[8+3i..] = slab i {input map, dy, channel chunk}; out[80+6j..] = step j {k-block, local phase, slab of row 0,
 *   dx+1, accumulator of row 0, slabs released after the step}.  out_len >= 464.
 */
int b2g_conv_rows_plan_query(int kind, int N, int Hin, int Win, int Cin, int block_n, int pg, int* out, int out_len);
/* Same, for the schedule with MERGED steps (transposed conv with two sub-pixel phases per group: the dx = 0 shift of an
 * input row feeds both phases with one MMA of twice the tile width); ring = slab ring size (0: default).  8 ints per step
 * from out[8 + 72]: {bcol, phase_l, slab0, dxi, acc0, rel, wide, bcol2}; out needs 8 + 72 + 8 * 64 ints. */
int b2g_conv_rows_plan_query_merged(int kind, int N, int Hin, int Win, int Cin, int block_n, int pg, int ring, int* out,
                                    int out_len);
int b2g_conv_plan_query(int kind, int N, int Hin, int Win, int tile_pixels, int* out, int out_len);
long b2g_pack_source_index(int kind, int kh, int kw, int D2, int D3, int n_is_d3, int flip, int ck_pad, long i);

#ifdef __cplusplus
}
#endif
#endif /* B200GAN_H_ */
