/* libsrb200 -- C ABI of the B200-native ESRGAN hot path.
 *
 * The reference (lukas-blecher/super-resolution) is pure Python and reaches its
 * numeric backend through torch.nn calls; this library is what replaces that
 * backend for the hot path.  Each entry point names the reference call site(s)
 * it stands in for.  All entry points:
 *   - are extern "C", take plain pointers / ints / floats (no torch types);
 *   - take DEVICE pointers owned by the caller (the library never allocates or
 *     frees device memory) and an explicit cudaStream_t passed as void*;
 *   - are asynchronous, re-entrant and CUDA-graph capturable; they never call
 *     cudaDeviceSynchronize;
 *   - return 0 on success, non-zero on error; srb_last_error() then returns a
 *     thread-local message that starts with "CUDA" (the Python shim raises
 *     RuntimeError with that text, which the reference's handlers at
 *     esrgan.py:829-836 and hyper_search.py:108-110 key on).
 *
 * Activation layout ("PNHWC"): bf16 [batch][H+1][W+1][C]; row H and column W
 * of every image are zero.  With that padding a 3x3/pad-1 convolution is nine
 * constant row-shifts of the flat [batch*(H+1)*(W+1), C] matrix, which is what
 * the implicit-GEMM kernels exploit.  Kernels only ever store to valid pixels,
 * so a buffer zero-filled once at allocation keeps its padding forever.
 */
#ifndef SRB200_H
#define SRB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRB_ABI_VERSION 3   /* 2: srb_conv3x3_args gained sign_out / mask_bits / colsum, srb_wgrad3x3_args gy_colsum (session 29)
                               3: srb_conv3x3_args gained res_f32 / y32 / res3 / chan_scale (round 2: fp32 residual stream of the RRDB
                                  trunk, num_final_layer_res skip, Dropout2d); srb_adam_step state is float[4] with the device-side lr */

const char* srb_last_error(void);
int srb_abi_version(void);
/* Reads and clears the device-side fault word (non-zero after a kernel trapped on a pipeline wait). */
int srb_read_fault(unsigned int* fault_out);

/* ---- 3x3 stride-1 pad-1 convolution, channels multiple of 64, PNHWC bf16 ----------------------
 * Replaces: nn.Conv2d in DenseResidualBlock b1..b5 (models.py:19,36-38) incl. torch.cat (38) and
 * out.mul(res_scale)+x (41,53); conv2 + torch.add (models.py:67,125-126); upsampling conv +
 * LeakyReLU + PixelShuffle (models.py:87-89); conv3.0 + LeakyReLU (models.py:97-98); and, with the
 * "dgrad" weight pack, the input-gradient of all of those (aten::convolution_backward).
 *
 *   v = (acc + bias) * chan_scale[image][n] * alpha + res1 * res1_scale + res2 * res2_scale + res3 * res3_scale
 *   act == SRB_ACT_LRELU      : v = v > 0 ? v : slope * v
 *   act == SRB_ACT_LRELU_MASK : v = v * (mask > 0 ? 1 : slope)          (LeakyReLU backward)
 *   store_mode == SRB_STORE_PLAIN  : y[pixel][y_coff + n]
 *   store_mode == SRB_STORE_SHUFFLE: cout = 4*Cy; packed output channel n = (2i+j)*Cy + c goes to
 *                                    y[(2y+i, 2x+j)][y_coff + c]; y has geometry (2H, 2W).
 * Alignment (SRB_IMPL_TCGEN05): x 16 bytes with x_coff a multiple of 64; y, res1, res2, res3 and y32 - pointer, row pitch
 * (cstride elements) and channel offset - 32 bytes each: the epilogue moves these per-row streams with 256-bit accesses.
 */
enum { SRB_ACT_NONE = 0, SRB_ACT_LRELU = 1, SRB_ACT_LRELU_MASK = 2 };
enum { SRB_STORE_PLAIN = 0, SRB_STORE_SHUFFLE = 1 };
enum { SRB_IMPL_TCGEN05 = 0, SRB_IMPL_NAIVE = 1 };

typedef struct srb_conv3x3_args {
  int batch, H, W;             /* un-padded geometry of x (and of y for SRB_STORE_PLAIN) */
  const void* x;               /* bf16 PNHWC */
  int x_cstride, x_coff, cin;  /* channels [x_coff, x_coff+cin) are read; cin % 64 == 0 */
  const void* w;               /* bf16 packed [9][cin/64][cout][64]  (srb_pack_conv3x3_weights) */
  const float* bias;           /* fp32 [cout] in packed channel order, or NULL */
  int cout;                    /* % 64 == 0 */
  void* y;                     /* bf16 PNHWC */
  int y_cstride, y_coff;
  int store_mode;
  float alpha;
  int act;
  float slope;
  const void* mask; int mask_cstride, mask_coff;
  const void* res1; int res1_cstride, res1_coff; float res1_scale;
  const void* res2; int res2_cstride, res2_coff; float res2_scale;
  int impl;                    /* SRB_IMPL_* */
  int flags;                   /* bit0: descriptor base-offset mode (bring-up switch); bits 1-15: timing-experiment
                                  switches; bits 16-21: kernel selection of the experimental builds and tile height
                                  (18 / 19: forbid / force 512-row tiles); bit 22 / 23: force / forbid the CTA-pair
                                  kernel (csrc/conv_chain2.cu: tcgen05.mma.cta_group::2, thread-block clusters of 2;
                                  the default for every eligible launch unless SRB_CONV_PAIR=0 is in the environment).
                                  Every kernel choice gives bit-identical results. */
  void* sign_out;              /* optional uint32 [pixels][sign_cstride]: word sign_coff + (channel / 32) of a pixel gets
                                  bit (channel % 32) = (stored output > 0).  Written next to a LeakyReLU output so that
                                  the input-gradient pass can read LeakyReLU' from 1 bit per element (mask_bits)
                                  instead of re-reading the bf16 activation. */
  int sign_cstride, sign_coff; /* sign_cstride > 0: row-major, word index = pixel * sign_cstride + word;
                                  sign_cstride < 0: plane-major uint32 [words][-sign_cstride pixels], word index =
                                  word * (-sign_cstride) + pixel (32 consecutive pixels of one word = one 128-byte access) */
  const void* mask_bits;       /* SRB_ACT_LRELU_MASK: if set, the mask comes from these sign words (same layouts, selected by
                                  the sign of maskb_cstride) and `mask` is ignored */
  int maskb_cstride, maskb_coff;
  float* colsum;               /* optional fp32 [64] (cout == 64 only): += sum over valid pixels of the stored outputs,
                                  per channel, with atomics - the bias gradient of the convolution whose pre-activation
                                  gradient this launch produces (torch: grad_bias of aten::convolution_backward).  The
                                  caller zeroes it; srb_wgrad3x3 consumes it through gy_colsum.  tcgen05 kernels only. */
  int res_f32;                 /* bit 0: res1 points at fp32 (strides / offsets in floats), bit 1: the same for res2.  The
                                  RRDB trunk keeps its residual stream x_{d+1} = x_d + 0.2 conv5(..) (models.py:41,53) in
                                  fp32 next to the bf16 copy the next convolution reads: 3 nb bf16 roundings of the stream
                                  are what pushes nb = 23 past the 2e-2 rel-L2 budget. */
  float* y32;                  /* optional fp32 copy of the stored output (plain store only): y32[pixel][y32_coff + n] */
  int y32_cstride, y32_coff;
  const void* res3; int res3_cstride, res3_coff; float res3_scale;   /* third bf16 residual: `out3 + out` of
                                  num_final_layer_res (models.py:128-130) and its gradient */
  const float* chan_scale;     /* optional fp32 [batch][cout]: per image and output channel factor on (acc + bias) - the
                                  Dropout2d mask (0 or 1/(1-p)) on out5 of a DenseResidualBlock (models.py:39-40).  Plain
                                  store only. */
} srb_conv3x3_args;

int srb_conv3x3(const srb_conv3x3_args* a, void* stream);

/* A chain of n (<= SRB_CONV_CHAIN_MAX) such convolutions over the same pixel grid in ONE persistent launch
 * (csrc/conv_chain.cu): layer l may read anything layers < l of the chain wrote - the five convolutions of a
 * DenseResidualBlock (models.py:34-41), the three dense blocks of an RRDB (models.py:52-53), or the input-gradient
 * chain of a dense block.  All layers share batch, H, W and cout, store plainly and use the tcgen05 kernels.  A tile
 * of layer l starts when the neighbouring tiles of layer l-1 are complete (per-tile counters in `flags`), so the
 * tensor pipe does not drain between layers.  `flags`: device workspace of srb_conv3x3_chain_flag_bytes(...) bytes,
 * ZERO on entry (the caller clears it; one workspace per chain launch in flight).  Results are identical to n
 * srb_conv3x3 calls in order. */
#define SRB_CONV_CHAIN_MAX 16
int srb_conv3x3_chain(const srb_conv3x3_args* layers, int n, void* flags, size_t flag_bytes, void* stream);
size_t srb_conv3x3_chain_flag_bytes(int batch, int H, int W, int n);

/* Weight packing, fp32 OIHW -> bf16 [9][cin/64][cout][64].
 *   mode 0 (forward):  wp[t][kb][n][k] = w[perm(n)][kb*64+k][t] * scale
 *   mode 1 (dgrad)  :  treats `w` as the forward weight [co][ci_total][3][3]; produces the pack of the
 *                      transposed/flipped filter for input channels [ci_off, ci_off+64):
 *                      wp[t][kb][n][k] = w[co_base... see DESIGN.md]
 * shuffle != 0 permutes output channels n = (2i+j)*C + c <- 4c + 2i + j (PixelShuffle(2) order).
 * Replaces nothing in the reference (layout glue); cites models.py:19 for the OIHW source layout. */
int srb_pack_conv3x3_fwd(const float* w_oihw, int cout, int cin, int shuffle, float scale, void* wp, void* stream);
/* dgrad pack of one forward conv restricted to input-channel slot [ci_off, ci_off+n_out) as OUTPUT channels:
 *   wp[t][kb][n][k] = scale * w[kb_off_co + kb*64 + k][ci_off + n][8 - t]
 * written at k-block offset kb_dst of a destination pack that has kb_total k-blocks (so several forward
 * convs can be concatenated along K: the dense-block backward of SURVEY.md section 7-2). */
int srb_pack_conv3x3_dgrad(const float* w_oihw, int cout, int cin, int ci_off, int n_out, int unshuffle,
                           float scale, void* wp, int kb_dst, int kb_total, void* stream);

/* Table-driven form of the two packs above: one launch for a whole model.  `items_dev` = device array of
 *   struct { const float* w; void* wp; int cout, cin, ci_off, n_out, flag, kind, kb_dst, kb_total; float scale; int pad; }
 * (kind 0: srb_pack_conv3x3_fwd with flag = shuffle; kind 1: srb_pack_conv3x3_dgrad with flag = unshuffle);
 * `block_prefix_dev[i]` = first 256-thread block of item i, `total_blocks` = sum of ceil(pairs_i / 256) with
 * pairs_i = cout * cin (kind 0) or cout * n_out (kind 1): one thread per channel pair handles its nine taps. */
int srb_pack_batch(const void* items_dev, const int* block_prefix_dev, int n_items, int total_blocks, void* stream);

/* ---- weight gradient of the same convolution -------------------------------------------------
 * Replaces aten::convolution_backward (filter part) for every 64-multiple conv of the generator.
 *   dw[co][ci][t] (+)= alpha * sum_pixels gy[p][co] * x[p + shift_t][ci];   dbias[co] (+)= alpha * sum gy
 */
typedef struct srb_wgrad3x3_args {
  int batch, H, W;
  const void* x;  int x_cstride, x_coff, cin;     /* bf16 PNHWC */
  const void* gy; int gy_cstride, gy_coff, cout;  /* bf16 PNHWC, gradient of the pre-activation */
  float* dw;       /* fp32 OIHW [cout][cin][3][3] */
  float* dbias;    /* fp32 [cout] or NULL */
  float alpha;
  int accumulate;  /* 0: overwrite, 1: add into dw/dbias */
  int unshuffle;   /* gy channel n = (2i+j)*C + c maps to dw row 4c+2i+j */
  void* workspace; size_t workspace_bytes;
  int impl;
  int flags;
  float* gy_colsum; /* optional fp32 [gy_cstride]: per-channel sums of gy over all pixels, accumulated by the launches
                       that wrote gy (srb_conv3x3_args.colsum).  When given, the bias gradient is taken from
                       gy_colsum[gy_coff .. gy_coff+cout) instead of a pass over gy, and those entries are reset to 0. */
} srb_wgrad3x3_args;
int srb_wgrad3x3(const srb_wgrad3x3_args* a, void* stream);
size_t srb_wgrad3x3_workspace_bytes(int batch, int H, int W, int cin, int cout);
/* Up to 8 weight gradients that share x, gy (different channel offsets), geometry, workspace and flags in ONE launch:
 * the five convolutions of a DenseResidualBlock (models.py:24-32) read the same growth buffer and the same gradient
 * buffer.  items[0].workspace must hold srb_wgrad3x3_group_workspace_bytes(...). */
int srb_wgrad3x3_group(const srb_wgrad3x3_args* items, int n, void* stream);
size_t srb_wgrad3x3_group_workspace_bytes(int batch, int H, int W, int n, const int* cin, const int* cout);
/* Allocation-free variant of srb_wgrad3x3.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and results as srb_wgrad3x3 on the tcgen05 kernel; a->workspace / workspace_bytes / impl
 * are ignored: the split-K scratch is a static device array inside the library, srb_wgrad3x3_static_max_bytes() =
 * 16 MiB, held by every GPU that loads the library until the process exits.  A problem whose
 * srb_wgrad3x3_workspace_bytes(...) exceeds that is rejected (returns 1, srb_last_error() names both sizes).  The call
 * allocates nothing (no cudaMalloc / cudaFree / allocating library routine); its first use on a device looks the
 * array's address up once (cudaGetSymbolAddress).  All calls share the array: calls that overlap on the GPU (two streams,
 * two host threads) race on it - serialise them on one stream. */
int srb_wgrad3x3_static(const srb_wgrad3x3_args* a, void* stream);
size_t srb_wgrad3x3_static_max_bytes(void);

/* ---- bandwidth-bound ends of the generator (csrc/gen_edge.cu) ------------------------------------
 * srb_conv_in_fwd : models.py:122-123  multiplier*x**power -> conv1; fp32 NCHW image -> PNHWC bf16 (two optional dsts)
 * srb_conv_in_bwd : conv1 weight/bias gradient from g (+ g2, the long-skip gradient of models.py:126)
 * srb_conv_out_fwd: models.py:99,131-135,114-118  conv3.2 / multiplier + output head; writes out, srs and z
 * srb_conv_out_bwd: adjoint of the above: head, conv3.2 dgrad (masked by LeakyReLU' of conv3.0's activation,
 *                   models.py:98) into a PNHWC bf16 buffer, conv3.2 weight/bias gradient
 * srb_unshuffle2  : adjoint of nn.PixelShuffle(2) (models.py:89) in the packed channel order (2i+j)*C + c      */
int srb_conv_in_fwd(const float* img, int B, int Cimg, int H, int W, float multiplier, float power, const float* w,
                    const float* bias, int F, void* y, int y_cstride, int y_coff, void* y2, int y2_cstride,
                    int y2_coff, void* stream);
int srb_conv_in_bwd(const float* img, int B, int Cimg, int H, int W, float multiplier, float power, const void* g,
                    int g_cstride, int g_coff, const void* g2, int g2_cstride, int g2_coff, int F, float* dw,
                    float* db, void* stream);
int srb_conv_out_fwd(const void* x, int x_cstride, int x_coff, int B, int H, int W, int F, const float* w,
                     const float* bias, int Cimg, float multiplier, float power, float thres, int training,
                     float* out, float* srs, float* z, void* stream);
int srb_conv_out_bwd(const float* z, const float* g_out, const float* g_srs, int B, int Cimg, int H, int W,
                     float multiplier, float power, float thres, int training, const float* w, int F, const void* act,
                     int act_cstride, int act_coff, float slope, float* gz_scratch, void* gy, int gy_cstride,
                     int gy_coff, float* dw, float* db, void* stream);
int srb_unshuffle2(const void* src, int s_cstride, int s_coff, int B, int H, int W, int C, void* dst, int d_cstride,
                   int d_coff, void* stream);
/* Dropout2d backward (models.py:29-31,39-40 under autograd): x = C-channel slot of a PNHWC bf16 buffer holding the
 * gradient w.r.t. a DenseResidualBlock's output; copy <- x (for the identity paths), x <- x * mask[image][channel]
 * in place (what the fifth convolution's weight / input gradients see).  mask: fp32 [B][C] (0 or 1/(1-p)). */
int srb_chan_mask_split(void* x, int x_cstride, int x_coff, void* copy, int c_cstride, int c_coff, const float* mask,
                        int B, int H, int W, int C, void* stream);

/* ---- discriminator stack and SumPool2d (csrc/disc.cu), fp32 NCHW ---------------------------------
 * srb_dconv_fwd/dgrad/wgrad: nn.Conv2d(k=3, padding=1, stride 1|2) of discriminator_block (models.py:140-146,168)
 *   and the two halves of its backward; composed by the host into first and second derivatives (the gradient
 *   penalty of esrgan.py:598-606 differentiates the input gradient again).
 * srb_lrelu_fwd/bwd: nn.LeakyReLU(0.2) (models.py:143,145).   srb_sumpool_fwd/bwd: SumPool2d (models.py:297-305). */
int srb_dconv_fwd(const float* x, const float* w, const float* bias, float* y, int B, int Cin, int H, int W, int Cout,
                  int stride, int act, float slope, void* stream);   /* act != 0: LeakyReLU(slope) fused */
int srb_set_dconv_impl(int impl);   /* 0: register-tiled kernels (default); 1: simple cross-check kernels */
int srb_dconv_dgrad(const float* gy, const float* w, float* gx, int B, int Cin, int H, int W, int Cout, int stride,
                    void* stream);
int srb_dconv_wgrad(const float* gy, const float* x, float* dw, float* db, int B, int Cin, int H, int W, int Cout,
                    int stride, void* stream);
int srb_lrelu_fwd(const float* x, float* y, long long n, float slope, void* stream);
int srb_lrelu_bwd(const float* g, const float* ref, float* gx, long long n, float slope, void* stream);
int srb_sumpool_fwd(const float* x, float* y, int BC, int H, int W, int k, void* stream);
int srb_sumpool_bwd(const float* gy, float* gx, int BC, int H, int W, int k, void* stream);

/* ---- discriminator stack on warp-level bf16 tensor-core MMA (csrc/disc_mma.cu) ---------------------
 * Activations bf16 NHWC [B][H][W][C] (C in {16,32,64}), images / predictions fp32 [B][H][W].  The hand-scheduled
 * discriminator engine (srb200/disc_engine.py) composes these into the forward of models.py:140-174 and into the
 * first and second derivatives the training step needs (esrgan.py:493-508, 569-616).
 * srb_dmma_conv : nn.Conv2d(k=3, p=1, stride 1|2) + bias + LeakyReLU(0.2) (models.py:142-145,168); with the
 *                 "dgrad" weight packs also the input gradient (stride 1: flipped conv; stride 2: transposed form,
 *                 `transposed` = 1, source = gradient map at half resolution).  act: 0 none, 1 LeakyReLU(slope),
 *                 2 multiply by LeakyReLU'(ref) (ref = stored activation with the geometry of the output).
 *                 Exactly one of in_bf16 / in_f32 (single channel) is given; out_f32 only for cout == 1.
 * srb_dmma_wgrad: filter/bias gradient, ACCUMULATED into dw [cout][cin][3][3] / db [cout] with atomics.
 * srb_dl1_*     : the Cin = 1 image end (first conv of models.py:142) in fp32 on CUDA cores: forward (optionally
 *                 with the mask epilogue), input gradient (fp32 image out), filter/bias gradient (accumulated).
 * srb_dpack     : fp32 OIHW parameters (offsets into one flat buffer) -> zero-padded bf16 [9][rows][k] packs;
 *                 kind 0 forward, 1 stride-1 input gradient (flipped, transposed), 2 stride-2 input gradient. */
int srb_dmma_conv(const void* in_bf16, const float* in_f32, int cin, const void* w_packed, const float* bias, int cout,
                  void* out_bf16, float* out_f32, const void* ref, int B, int Hs, int Ws, int Hd, int Wd, int stride,
                  int transposed, int act, float slope, void* stream);
int srb_dmma_wgrad(const void* gz_bf16, const float* gz_f32, int cout, const void* a_bf16, int cin, float* dw,
                   float* db, int B, int Hi, int Wi, int Ho, int Wo, int stride, void* stream);
int srb_dl1_fwd(const float* img, const float* w, const float* bias, void* out_bf16, const void* ref, int C, int B,
                int H, int W, int act, float slope, void* stream);
int srb_dl1_dgrad(const void* gz_bf16, const float* w, float* gx, int C, int B, int H, int W, void* stream);
int srb_dl1_wgrad(const void* gz_bf16, const float* img, float* dw, float* db, int C, int B, int H, int W,
                  void* stream);
int srb_dpack(const float* flat_params, void* dst_bf16, int n_items, const long long* src_off,
              const long long* dst_off, const int* cout, const int* cin, const int* kind, void* stream);

/* ---- losses and optimizer (csrc/optim_loss.cu) ---------------------------------------------------
 * srb_adam_step   : torch.optim.Adam (esrgan.py:299,305,555,624) over one flat fp32 buffer; `state` = device
 *                   float[4] {step, skip, lr, reserved}; the update is skipped unless *gate_value > gate_threshold
 *                   (the d_threshold test of esrgan.py:623 without a host sync); gate_value may be NULL.  lr < 0
 *                   means "use state[2]": the learning rate then lives on the device, so a captured CUDA graph
 *                   follows ReduceLROnPlateau (esrgan.py:308-310,666-667) without re-capture.
 * srb_pixel_losses: esrgan.py:471,487,490  L1(batch-mean image) and L1(SumPool(gen), lr) + gradient w.r.t. gen.
 * srb_ragan       : esrgan.py:507-508 (u=fake, v=real) and 578-588 (u=real, v=fake): loss, dL/du, dL/dv.
 * srb_gp_fwd/bwd  : esrgan.py:604-605 gradient-penalty reduction and its gradient.                           */
int srb_adam_step(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2, float eps,
                  float weight_decay, float grad_scale, float* state, const float* gate_value, float gate_threshold,
                  void* stream);
int srb_pixel_losses(const float* gen, const float* gt, const float* lr, int B, int H, int W, int k, float w_hr,
                     float w_lr, int accumulate, float* msign, float* dsign, float* acc, float* grad, void* stream);
/* The same with the low-resolution term taken on powered images (opt.scaling_power != 1, esrgan.py:472-473,490 for the
 * "pow" branch): loss_lr = L1(SumPool(gen)^lr_power, lr^lr_power), gradient through torch.pow's derivative. */
int srb_pixel_losses_pow(const float* gen, const float* gt, const float* lr, int B, int H, int W, int k, float w_hr,
                         float w_lr, float lr_power, int accumulate, float* msign, float* dsign, float* acc,
                         float* grad, void* stream);
int srb_ragan(const float* u, const float* v, int B, int n, float eps, float* loss, float* gu, float* gv,
              void* stream);
int srb_gp_fwd(const float* g, int B, int n, float coef, float* norms, float* loss, void* stream);
int srb_gp_bwd(const float* g, const float* norms, int B, int n, float coef, const float* gloss, float* gg,
               void* stream);

/* ---- optional generator loss terms (csrc/opt_losses.cu; lambda = 0 in options/default.json) --------------------
 * Each returns the loss as a device scalar and adds w * d loss / d generated into `grad` (accumulate != 0) or overwrites it.
 * srb_soft_hist / srb_kld_hist / srb_soft_hist_bwd: DiffableHistogram (models.py:308-342) of the entries > 0 and KLD_hist
 *     (utils.py:90-113) as used at esrgan.py:538-545: hist[j] = sum_{x>0} sigmoid(sigma (x - c_j + d_j/2)) -
 *     sigmoid(sigma (x - c_j - d_j/2)); kld: loss and dq = d loss / d (generated histogram); bwd: chain rule to pixels.
 * srb_nnz_loss : esrgan.py:530-533  mse(sum softgreater(gen, 0, sigma), count(gt > 0)) per image (utils.py:259-261).
 * srb_mask_loss: esrgan.py:534-537  L1(sigmoid(sigma gen), sigmoid(sigma gt)) (nnz_mask, utils.py:271-272).
 * srb_hit_loss : esrgan.py:546-550  mse of get_hitogram (utils.py:264-268) of generated and ground truth.          */
int srb_soft_hist(const float* x, long long n, const float* centers, const float* deltas, int nbins, float sigma,
                  float* hist, void* stream);
int srb_kld_hist(const float* q_hist, const float* p_hist, const float* binsizes, int nbins, float* loss, float* dq,
                 void* stream);
int srb_soft_hist_bwd(const float* x, long long n, const float* centers, const float* deltas, int nbins, float sigma,
                      const float* dq, float w, int accumulate, float* grad, void* stream);
int srb_nnz_loss(const float* gen, const float* gt, int B, int HW, float sigma, float w, int accumulate,
                 float* per_image, float* loss, float* grad, void* stream);
int srb_mask_loss(const float* gen, const float* gt, long long n, float sigma, float w, int accumulate, float* loss,
                  float* grad, void* stream);
int srb_hit_loss(const float* gen, const float* gt, int B, int H, int W, int factor, float thr, float sig, float w,
                 int accumulate, float* work, float* loss, float* grad, void* stream);

/* ---- sorted-Wasserstein / Sinkhorn generator loss term (csrc/sinkhorn.cu; lambda_wasser = 0 in options/default.json) ----
 * srb_sinkhorn_loss: esrgan.py:539-542 (per-image torch.sort of generated / ground truth, cut_smaller utils.py:241-255,
 *   SinkhornDistance(opt.sinkhorn_eps, 100, 'sum') utils.py:147-238; constructed at esrgan.py:321).
 *   gen, truth : fp32 [B][HW] images (HW <= 32768)
 *   eps, max_iter, thresh : SinkhornDistance's eps, max_iter (100) and stopping threshold (0.1, utils.py:194)
 *   cap1, cap2 : capacity (points per image) for the generated / true cloud, 12*cap1 + 16*cap2 <= 220 KB; a batch whose
 *                cut_smaller size exceeds it yields loss = NaN and info[3] = 1 (nothing is added to grad)
 *   loss       : device scalar, sum over the batch of the transport costs
 *   grad       : fp32 [B][HW] or NULL; w * d loss / d gen is ADDED into it (through all Sinkhorn iterations, as
 *                autograd does in the reference)
 *   info       : device int[4] or NULL: points of the two clouds, iterations run, capacity-overflow flag
 *   workspace  : srb_sinkhorn_workspace_bytes(...) bytes, 256-byte aligned; holds the duals of every iteration
 * No host synchronisation: max_iter forward and max_iter backward launches whose CTAs exit at once after convergence. */
size_t srb_sinkhorn_workspace_bytes(int B, int HW, int cap1, int cap2, int max_iter);
int srb_sinkhorn_loss(const float* gen, const float* truth, int B, int HW, float eps, int max_iter, float thresh,
                      int cap1, int cap2, float w, void* workspace, size_t workspace_bytes, float* loss, float* grad,
                      int* info, void* stream);

/* ---- data path (csrc/data_path.cu) ---------------------------------------------------
 * srb_extract_batch: datasets.py:136-145 (extract) + datasets.py:247-249 / models.py:297-305 (SumPool2d) for a batch.
 *   pairs : fp32 [B][N][2] = (flat index, energy) lists, zero padded (a zero energy ends a list)
 *   hr    : fp32 [B][1][etaBins][phiBins], hr[b,0,idx % etaBins, idx // etaBins] += energy in list order
 *   lr    : fp32 [B][1][etaBins/factor][phiBins/factor] = factor x factor sums of hr, or NULL
 *   n_bad : optional device counter, += entries whose index falls outside the image (the reference raises there) */
int srb_extract_batch(const float* pairs, int B, int N, int etaBins, int phiBins, int factor, float* hr, float* lr,
                      int* n_bad, void* stream);

/* ---- evaluation post-processing (csrc/eval_stats.cu) ----------------------------------
 * srb_image_stats: evaluation/eval.py:323-335 (collector modes max, min, mean, nnz, sum, meannnz) and eval.py:98-105
 * (get_nth_hardest, n = 1..nmax <= 8) for a batch in one pass.
 *   img : fp32 [B][HW];  out : fp32 [B][7 + nmax] = { max, min, mean, count(> thres), sum, mean of pixels > thres or 0,
 *   has-NaN flag (eval.py:319-321 skips such batches), 1st .. nmax-th largest pixel (duplicates counted) }, nmax <= 32
 *   (options/default.json's eval_modes reach E_30) */
int srb_image_stats(const float* img, int B, int HW, float thres, int nmax, float* out, void* stream);
/* calculate_metrics, evaluation/eval.py:477-485, one batch: out[b] = { mean |SumPool_factor(gen) - lr|, mean |gen - hr| }. */
int srb_l1_metrics(const float* gen, const float* hr, const float* lr, int B, int H, int W, int factor, float* out,
                   void* stream);
/* One image stack [B][H][W] added to running evaluation rasters (fp64, caller-owned, either may be NULL):
 *   raster[f][f]   += number of pixels > thr_raster per position inside the factor x factor cells  (SumRaster.add,
 *                     utils.py:290-293, evaluation mode 'hitogram')
 *   mean_acc[H][W] += energy ? (sum_b max(pixel, thr_mean)) * (count_b pixel > thr_mean) : count_b pixel > thr_mean
 *                     (MeanImage.add, utils.py:317-322, evaluation mode 'meanimg') */
int srb_eval_rasters(const float* img, int B, int H, int W, int factor, float thr_raster, float thr_mean, int energy,
                     double* raster, double* mean_acc, void* stream);

/* Timing experiment aid: clock64 stamps of CTA 0's pipeline when srb_conv3x3 ran with flags bit 4 (tools/exp_conv.py). */
int srb_igemm_read_stamps(long long* out16);

/* 1 when the library was built with SRB_EXPERIMENTAL (opt-in filter-row / 2+1-tap convolution kernels and the probe
 * below); the product build returns 0 and those kernels are absent. */
int srb_has_experimental(void);

/* Bring-up microbenchmark: sustained tcgen05.mma issue rate (csrc/probe.cu); SRB_EXPERIMENTAL builds only. */
int srb_probe_mma_rate(int cta_group, int N, int iters, int a_row_shift, int mn_major, int b_lbo, int n_ctas,
                       long long* cycles_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SRB200_H */
