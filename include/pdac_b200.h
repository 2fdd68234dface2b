/*
 * pdac_b200.h — C ABI of libpdac_b200.so: the B200 (sm_100a) kernels behind the drop-in nn.Module replacements
 * for moksu27/PDAC_pathological_image_segmentation's hot path.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer owned by the caller (PyTorch caching allocator); the library allocates
 *     nothing. Activations are NHWC bf16 "views" (PdacView: channel stride 1, element strides for W/H/N), so
 *     channel slices, parity sub-lattices (stride-2 convs, ConvTranspose k2s2) and concat halves are all views.
 *   - Every entry is stream-ordered on `stream` (a cudaStream_t passed as void*), never synchronises (except
 *     pdac_check_device_flag), and returns 0 or a negative PDAC_E_* code; pdac_last_error() gives the
 *     thread-local message.
 *   - No CPU fallback anywhere: host pointers or a missing GPU are errors.
 *
 * Reference interface each group replaces (paths relative to the reference repo; torchvision paths relative to
 * site-packages/torchvision/models):
 *   conv2d / conv-transpose       -> nn.Conv2d / nn.ConvTranspose2d built at models/resunet.py:12-21, run at :25-48;
 *                                    torchvision resnet.py:59-106 (BasicBlock), :197-204 (stem); smp decoders
 *                                    instantiated at train_worker.py:152-182
 *   batchnorm / relu / residual   -> torchvision resnet.py:82-101 (bn1/bn2/relu/+=identity), resnet.py:199
 *   maxpool                       -> torchvision resnet.py:200 (MaxPool2d 3x3 s2 p1), models/resunet.py:28
 *   resize                        -> F.interpolate at models/resunet.py:51-56 (bilinear, align_corners=False);
 *                                    smp UpsamplingBilinear2d (align_corners=True) and nearest x2
 *   dice / mask                   -> util.py:10-19 (DiceLoss), util.py:22-26 (dice_score),
 *                                    train_worker.py:24,72 and test.py:144 (sigmoid >= 0.5 mask)
 *   adam                          -> torch.optim.Adam(model.parameters(), lr) at train_worker.py:184, step at :80
 *   stitch                        -> visualize/predict_overlay.ipynb (missing blob) per configs/visualize_config.yaml:3-15
 *   prep                          -> dataset.py:24-30 + train_main.py:70-80 (Normalize / ColorJitter / OneOf flips, rot90)
 */
#ifndef PDAC_B200_H
#define PDAC_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PDAC_OK 0
#define PDAC_E_INVALID (-1)  /* bad argument / unsupported shape */
#define PDAC_E_CUDA (-2)     /* a CUDA runtime/driver call failed */
#define PDAC_E_TIMEOUT (-3)  /* a device-side pipeline barrier timed out (see pdac_check_device_flag) */
#define PDAC_E_NODEVICE (-4) /* no sm_100 device / driver entry point missing */

/* Reproducibility (the reference requests cudnn.deterministic, train_main.py:36-43): no result of this library depends
 * on the order in which thread blocks run. Cross-block reductions store one partial row per block into caller memory
 * and add the rows in index order; PDAC_STAT_ROWS is the row capacity those tables need (4 blocks x 148 SMs). The
 * split-K weight gradients and the bias gradient take an optional scratch buffer: with it they are reproducible,
 * without it they fall back to fp32 atomics (faster by a few per cent, last-bit nondeterministic). */
#define PDAC_STAT_ROWS 592

/* NHWC view: element (n,h,w,c) lives at ptr + n*sN + h*sH + w*sW + c (in elements of the view's dtype). */
typedef struct PdacView {
  void* ptr;
  int32_t N, H, W, C;
  int64_t sN, sH, sW;
} PdacView;

/* 2-D convolution geometry (square kernel stride/dilation, as every conv on the path has). */
typedef struct PdacConv {
  int32_t R, S;      /* kernel height, width */
  int32_t stride;    /* 1 or 2 */
  int32_t pad;       /* symmetric zero padding */
  int32_t dil;       /* dilation */
  int32_t relu;      /* fuse max(x,0) into the epilogue */
  int32_t wrows;     /* rows per tap in the packed weight matrix (>= rows used; lets a tap block be padded) */
  int32_t cout_valid; /* wgrad only: real channels of a channel-padded dY (0 = all of dy->C) */
} PdacConv;

int pdac_version(void);
const char* pdac_last_error(void);
/* Which tensor-core kernel this thread's last convolution call dispatched to: 1 conv_tc_kernel, 2 conv_halo_kernel,
 * 3 wgrad_tc_kernel, 4 wgrad_halo_kernel (0: none yet). Used by bench.py to attribute event times to kernels. */
int pdac_last_kernel_id(void);
/* Reads (and clears) the device-side pipeline error flag; synchronises `stream`. */
int pdac_check_device_flag(void* stream);
/* The same flag without synchronising: it lives in mapped pinned host memory, so the host sees what the kernels that
 * have FINISHED so far reported (0 = fine). The product path polls it once per step / batch (trainer.py, wsi.py,
 * module_base.py) and raises: a timed-out TMA / tcgen05 pipeline barrier never goes unnoticed into the optimizer. */
int pdac_peek_device_flag(void);

/* ---- tensor-core implicit-GEMM convolutions (conv_tc.cu) ----------------------------------------------------
 * Generalised "tap convolution": out[n,h,w,:] = act( sum_t W_t * cat_s(views[tap.src+s])[n, h+dh_t, w+dw_t, :]
 *                                                    + bias + addend ).
 * Every nn.Conv2d (any stride/dilation), its data gradient, ConvTranspose2d k2s2 and its data gradient are
 * instances (strided cases read/write parity sub-lattice views).
 * Packed weights: bf16 matrix [n_taps * wrows][Ktot], K contiguous (Ktot = sum of the n_cat source channels). */
#define PDAC_MAX_TAPS 16
typedef struct PdacTap {
  int32_t src;    /* index of the tap's first source view */
  int32_t dh, dw; /* pixel offset applied in that view's own lattice */
} PdacTap;
typedef struct PdacTapConv {
  int32_t n_taps;
  int32_t n_cat; /* 1 or 2 views concatenated along C per tap (torch.cat order) */
  int32_t wrows; /* rows per tap in the packed weight matrix */
  int32_t relu;
  PdacTap taps[PDAC_MAX_TAPS];
} PdacTapConv;
/* stats: optional fp32 [PDAC_STAT_ROWS][2][out->C], zeroed by the caller: every CTA (warpgroup) adds the per-channel
 * sum and sum of squares of the outputs it stored (bf16-rounded values) into its own row — BatchNorm batch statistics
 * as partial rows that pdac_bn_prepare adds in a fixed order. */
int pdac_tapconv_fwd(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed,
                     const float* bias, const PdacView* addend, const PdacView* out, float* stats, void* stream);
/* Same with a per-channel affine epilogue: out = act( acc * scale[c] + shift[c] + addend ) — eval-mode BatchNorm
 * (+ residual + ReLU) folded into the convolution (test.py:139-142 inference, torchvision resnet.py:92-103). */
int pdac_tapconv_fwd_affine(const PdacTapConv* tc, int n_views, const PdacView* views, const void* w_packed,
                            const float* scale, const float* shift, const PdacView* addend, const PdacView* out,
                            void* stream);
/* dw[t*tap_stride + a*a_stride + b*b_stride] += sum_pixels cat(views)[p + off_t, a] * dy[p, b]  (caller zeroes dw).
 * a runs over the concatenated source channels, b over the first b_valid channels of dy (0 = all; dy may be
 * channel-padded). scratch (fp32, 16-byte aligned, scratch_bytes >= n_taps * A * B * 4; 64 MB lets every layer of the
 * networks here keep its full split-K width): the split-K shares are stored there and added in a fixed order
 * (reproducible); scratch == NULL: fp32 atomics into dw. */
int pdac_tapconv_wgrad(const PdacTapConv* tc, int n_views, const PdacView* views, const PdacView* dy, float* dw,
                       int64_t tap_stride, int64_t a_stride, int64_t b_stride, int b_valid, float* scratch,
                       long long scratch_bytes, void* stream);

/* nn.Conv2d forward: out = [relu](conv(cat(src[0..n_src-1]), w) + bias + addend). w_packed: [R*S*wrows][Ctot] with
 * row (r*S+s)*wrows + cout. */
int pdac_conv2d_fwd(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed, const float* bias,
                    const PdacView* addend, const PdacView* out, float* stats, void* stream);
int pdac_conv2d_fwd_affine(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed, const float* scale,
                           const float* shift, const PdacView* addend, const PdacView* out, void* stream);
/* The same (without addend) stored nearest-neighbour x2 upsampled into out2x (N, 2Ho, 2Wo, Cout): inference form of an
 * smp decoder block's last Conv2dReLU in front of the next block's F.interpolate(x, scale_factor=2, mode="nearest").
 * *handled = 0 (and PDAC_OK): not eligible, nothing launched — run pdac_conv2d_fwd_affine + pdac_upsample_nearest2x_fwd. */
int pdac_conv2d_fwd_affine_up2(const PdacConv* cv, int n_src, const PdacView* src, const void* w_packed,
                               const float* scale, const float* shift, const PdacView* out2x, int* handled, void* stream);
/* dx = conv_transpose(dy, w) (+ addend): the data gradient of the conv described by cv. w_packed_t:
 * [R*S*wrows][Cout] with row (r*S+s)*wrows + cin (taps un-flipped: the library derives the offsets). For
 * stride 2 the four output parity classes are separate launches; classes no tap reaches get addend or zero. */
int pdac_conv2d_dgrad(const PdacConv* cv, const PdacView* dy, const void* w_packed_t, const PdacView* addend,
                      const PdacView* dx, void* stream);
/* The same data gradient when the conv's first dx_pooled->C input channels were F.interpolate(x, scale_factor=2,
 * mode="nearest") (smp UnetDecoder / DecoderBlock.forward: interpolate -> torch.cat([x, skip]) -> conv1): those channels'
 * gradient is returned at half resolution, summed over each 2x2 block (the interpolation's backward), in dx_pooled
 * (N, H/2, W/2, C0); the remaining channels (the skip connection's gradient) go to dx_rest (N, H, W, C1; NULL when there
 * is no skip). Stride 1, spatial-size-preserving convolutions only. Bit-identical to pdac_conv2d_dgrad followed by
 * pdac_upsample_nearest2x_bwd. *handled = 0 (and PDAC_OK): the geometry is not eligible for the fused kernel, nothing was
 * launched; run the two calls instead. */
int pdac_conv2d_dgrad_pool2(const PdacConv* cv, const PdacView* dy, const void* w_packed_t, const PdacView* dx_pooled,
                            const PdacView* dx_rest, int* handled, void* stream);
/* dw += sum_pixels dy[p,cout] * cat(src)[p*stride + off, cin]. dw_layout 0: fp32 [R*S][Cout][Ctot];
 * dw_layout 1: the nn.Conv2d state_dict layout [cout_valid][Ctot][R][S] (gradients land where autograd wants them). */
int pdac_conv2d_wgrad(const PdacConv* cv, int n_src, const PdacView* src, const PdacView* dy, float* dw,
                      int dw_layout, float* scratch, long long scratch_bytes, void* stream);
/* nn.ConvTranspose2d(k=2, s=2): out[2i+a,2j+b,:] = W[a,b] x[i,j,:] + bias. w_packed: [4*wrows][Cin], row
 * (a*2+b)*wrows + cout; w_packed_t: [4*wrows][Cout], row (a*2+b)*wrows + cin. */
int pdac_convt2x2_fwd(const PdacView* x, const void* w_packed, int wrows, const float* bias, int relu,
                      const PdacView* out, void* stream);
int pdac_convt2x2_dgrad(const PdacView* dy, const void* w_packed_t, int wrows, const PdacView* dx, void* stream);
/* dw_layout 0: [4][Cout][Cin]; 1: the nn.ConvTranspose2d state_dict layout [Cin][Cout][2][2]. */
int pdac_convt2x2_wgrad(const PdacView* x, const PdacView* dy, float* dw, int dw_layout, float* scratch,
                        long long scratch_bytes, void* stream);

/* ---- layout / packing (layout.cu) --------------------------------------------------------------------------- */
/* x: fp32 NCHW (N,C,H,W) -> out: NHWC bf16 view with out->C >= C (extra channels zeroed). */
int pdac_nchw_f32_to_nhwc_bf16(const float* x, int N, int C, int H, int W, const PdacView* out, void* stream);
/* first C_used channels of an NHWC bf16 view -> fp32 NCHW (N,C_used,H,W). */
int pdac_nhwc_bf16_to_nchw_f32(const PdacView* x, int C_used, float* out, void* stream);
/* Stem input: fp32 NCHW (N,3,H,W) -> space-to-depth bf16 (N,H/2,W/2,16), channel (p*2+q)*3+c = x[c][2i+p][2j+q],
 * channels 12..15 zero. The 7x7 s2 stem conv then is a 4x4 s1 conv over this tensor (pdac_pack_stem_weight). */
int pdac_stem_prep_f32(const float* x, int N, int H, int W, void* out, void* stream);
/* Same from uint8 HWC patches (N,H,W,3) with (x/255 - mean)/std fused (train_main.py:73) and a per-patch geometric
 * augmentation code aug[n] (0 none, 1 hflip, 2 vflip, 3 rot90; may be NULL); the uint8 label mask (N,H,W) gets the
 * same geometry (mask_in/mask_out may be NULL). mean/stdv are HOST pointers to 3 floats. */
int pdac_stem_prep_u8(const uint8_t* img, const uint8_t* mask_in, const int32_t* aug, int N, int H, int W,
                      const float* mean, const float* stdv, void* out, uint8_t* mask_out, void* stream);
/* Training augmentation fused into the stem input preparation (train_main.py:70-80: Normalize -> ColorJitter(0.3 x 4) ->
 * OneOf([HorizontalFlip, RandomRotate90, VerticalFlip], p=0.3)), drawn on the device from a counter-based hash of
 * (seed, *step, patch, draw): no host random numbers, so the step replays from a CUDA graph. Semantics: csrc/augment.cu
 * header (albumentations is absent and unpinned: PARITY UNPINNED; executable restatement oracle/augment_oracle.py). */
typedef struct PdacAugment {
  float brightness, contrast, saturation, hue; /* ColorJitter ranges: factors ~ U(1-r, 1+r), hue shift ~ U(-r, r) */
  float p_color;                               /* probability that ColorJitter is applied to a patch (library: 0.5) */
  float p_geom;                                /* probability of the OneOf group (reference: 0.3) */
  unsigned long long seed;
} PdacAugment;
/* step: optional DEVICE pointer to a step counter mixed into the seed. params: fp32 [N][8] written per patch:
 * [0] ColorJitter applied, [1] operation order 0..23, [2..4] brightness / contrast / saturation factors, [5] hue shift,
 * [6] geometry (0 none, 1 hflip, 2 vflip, 3/4/5 rot90 x 1/2/3 counter-clockwise), [7] the contrast step's gray mean. */
int pdac_stem_prep_u8_augment(const uint8_t* img, const uint8_t* mask_in, int N, int H, int W, const float* mean,
                              const float* stdv, const PdacAugment* cfg, const long long* step, float* params,
                              void* out, uint8_t* mask_out, void* stream);
/* state_dict weight fp32 [O][I][R][S] -> wp_f bf16 [R*S][wrows_f][kf] (row = o, col = i; forward) and/or
 * wp_t bf16 [R*S][wrows_t][kt] (row = i, col = o; data gradient). Padding rows/cols are zeroed. s2_pad >= 0 orders
 * wp_t's taps by stride-2 output parity class (see pdac_conv2d_dgrad), -1 keeps (r,s) order.
 * nn.ConvTranspose2d weights [Cin][Cout][2][2] pack with O=Cin, I=Cout: wp_f is then the dgrad matrix, wp_t the forward. */
int pdac_pack_conv_weight(const float* w, int O, int I, int R, int S, int wrows_f, int kf, void* wp_f, int wrows_t,
                          int kt, void* wp_t, int s2_pad, void* stream);
/* The same packing for n weight tensors in one launch per 24 tensors (descs is a HOST array): after an optimizer step
 * the whole network is re-packed at once. Fields as the arguments of pdac_pack_conv_weight. */
typedef struct PdacPackDesc {
  const float* w;
  void* wp_f;
  void* wp_t;
  int32_t O, I, R, S, wrows_f, kf, wrows_t, kt, s2_pad;
} PdacPackDesc;
int pdac_pack_conv_weights_batched(const PdacPackDesc* descs, int n, void* stream);
/* stem weight fp32 [O][3][7][7] -> bf16 [16][O][16] for the 4x4 space-to-depth form; and its gradient back. */
int pdac_pack_stem_weight(const float* w, int O, void* wp, void* stream);
int pdac_unpack_stem_wgrad(const float* dwp, int O, float* dw, void* stream);

/* ---- BatchNorm / ReLU / residual (bn.cu) -------------------------------------------------------------------- */
/* stats: fp32 [n_rows][2][C] partial sums / sums of squares over `count` values per channel in total (the conv
 * epilogues' rows; n_rows = PDAC_STAT_ROWS, or 1 for a plain [2][C] vector), added in a fixed order in double. Writes
 * scale_shift [2][C] (y = x*scale + shift), mean_invstd [2][C] (saved for backward) and, when training, updates
 * running_mean/var in place (momentum; unbiased variance) and increments the int64 num_batches_tracked scalar (may be
 * NULL). train=0 derives everything from the running stats. */
int pdac_bn_prepare(const float* stats, int n_rows, double count, const float* gamma, const float* beta, float eps,
                    float momentum, int train, float* running_mean, float* running_var,
                    long long* num_batches_tracked, float* scale_shift, float* mean_invstd, int C, void* stream);
/* y = [relu](x*scale[c] + shift[c] + residual); scale_shift / residual may be NULL. When y is 2H x 2W the result is
 * stored nearest-neighbour upsampled (each pixel four times): smp UnetDecoder's F.interpolate(scale_factor=2) fused into
 * the normalise pass of the layer that produces its input. */
int pdac_scale_shift_act(const PdacView* x, const float* scale_shift, const PdacView* residual, int relu,
                         const PdacView* y, void* stream);
/* BN backward pass 1: dz = dy * (y > 0); sums[c] = sum dz, sums[C+c] = sum dz*xhat. sums: fp32
 * [1 + PDAC_STAT_ROWS][2][C] (no zeroing needed): rows 1.. receive the blocks' partials, row 0 their fixed-order total,
 * which is what pdac_bn_bwd_apply reads. ReLU mask: from y; or, with y NULL and scale_shift (the forward's [2][C])
 * given, recomputed as x*scale + shift > 0 (y is then never read); with both NULL there is no ReLU. */
int pdac_bn_bwd_reduce(const PdacView* dy, const PdacView* y, const PdacView* x, const float* mean_invstd,
                       const float* scale_shift, float* sums, void* stream);
/* BN backward pass 2: dx = gamma*invstd*(dz - sums0/count - xhat*sums1/count) (train) or gamma*invstd*dz (eval);
 * optionally stores dz (the residual branch's gradient) and dgamma = sums1, dbeta = sums0. */
int pdac_bn_bwd_apply(const PdacView* dy, const PdacView* y, const PdacView* x, const float* mean_invstd,
                      const float* gamma, const float* scale_shift, const float* sums, double count, int train,
                      const PdacView* dx, const PdacView* dz_out, float* dgamma, float* dbeta, void* stream);
/* dz = dy * (y > 0) (y, dz may be NULL); dbias[c] += sum dz (may be NULL; caller zeroes). scratch: fp32
 * [PDAC_STAT_ROWS][C] for the reproducible fixed-order sum, or NULL (fp32 atomics). */
int pdac_relu_bwd_bias(const PdacView* dy, const PdacView* y, const PdacView* dz, float* dbias, float* scratch,
                       void* stream);
/* Allocation-free variant of pdac_relu_bwd_bias (reproducible mode). NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, serving loops,
 * harnesses that forbid it). Same arithmetic and result as pdac_relu_bwd_bias with caller scratch; the
 * [PDAC_STAT_ROWS][C] partial-row table lives in a static device array inside the library instead (1.2 MB per GPU that
 * loads the library, held until exit), and the call performs no cudaMalloc / cudaFree / allocating library call.
 * Limit: dy->C <= PDAC_NOALLOC_MAX_C, otherwise PDAC_E_INVALID with a message naming the limit. dy and dbias are
 * required (dbias[c] += sum, caller zeroes); y and dz may be NULL. Every call on a GPU shares the one array: calls that
 * overlap on the GPU (different streams, no ordering between them) race; calls ordered on one stream do not. */
#define PDAC_NOALLOC_MAX_C 512
int pdac_relu_bwd_bias_noalloc(const PdacView* dy, const PdacView* y, const PdacView* dz, float* dbias, void* stream);

/* ---- pooling / resampling (pool.cu) ------------------------------------------------------------------------- */
/* idx: uint8 [N][Ho][Wo][C] arg-max tap (r*3+s), saved for backward (may be NULL in inference). */
int pdac_maxpool3x3s2_fwd(const PdacView* x, const PdacView* y, uint8_t* idx, void* stream);
/* dx = scatter(dy by idx) (+ addend: the gradient another consumer of the pooled tensor's input has already left; may
 * alias dx). */
int pdac_maxpool3x3s2_bwd(const PdacView* dy, const uint8_t* idx, const PdacView* addend, const PdacView* dx,
                          void* stream);
int pdac_upsample_nearest2x_fwd(const PdacView* x, const PdacView* y, void* stream);
int pdac_upsample_nearest2x_bwd(const PdacView* dy, const PdacView* addend, const PdacView* dx, void* stream);
int pdac_bilinear_fwd(const PdacView* x, int align_corners, const PdacView* y, void* stream);
int pdac_bilinear_bwd(const PdacView* dy, int align_corners, const PdacView* dx, void* stream);
/* first C_used channels of NHWC bf16 x, resized to (Ho,Wo) -> fp32 NCHW logits; and the gradient back (all dx->C
 * channels are written, those >= C_used with zero). */
int pdac_bilinear_head_fwd(const PdacView* x, int C_used, int align_corners, float* out, int Ho, int Wo, void* stream);
int pdac_bilinear_head_bwd(const float* g, int C_used, int Ho, int Wo, int align_corners, const PdacView* dx,
                           void* stream);
int pdac_adaptive_avgpool_fwd(const PdacView* x, const PdacView* y, void* stream);
int pdac_adaptive_avgpool_bwd(const PdacView* dy, const PdacView* dx, void* stream);
int pdac_add(const PdacView* a, const PdacView* b, int relu, const PdacView* y, void* stream);

/* ---- GroupNorm / dropout / depthwise 3x3 (norm_dw.cu) ------------------------------------------------------
 * The layers smp.FPN / PSPNet / DeepLabV3Plus add to the ResNet18 path (train_worker.py:152-182):
 * nn.GroupNorm(32, 128) of Conv3x3GNReLU, nn.Dropout2d / nn.Dropout, SeparableConv2d's depthwise 3x3 (dilated). */
/* stats: fp32 [N][G][2] (sum, sum of squares per image and group). scratch (fp32, >= 1184 * G * 2 floats): the
 * blocks' partial rows are added in a fixed order and stats is overwritten (reproducible); scratch == NULL: fp32
 * atomics into stats, which the caller zeroes. */
int pdac_groupnorm_stats(const PdacView* x, int G, float* stats, float* scratch, void* stream);
/* y = [relu]((x - mean) * rstd * gamma + beta), mean / rstd from stats (biased variance, eps inside the sqrt). */
int pdac_groupnorm_fwd(const PdacView* x, int G, const float* stats, const float* gamma, const float* beta, float eps,
                       int relu, const PdacView* y, void* stream);
/* dz = dy * (y > 0) (y may be NULL); gsums fp32 [N][G][2] = (sum dz*gamma, sum dz*gamma*xhat);
 * dgamma[c] += sum dz*xhat, dbeta[c] += sum dz (may be NULL). scratch (fp32, >= 1184 * 2 * C floats): reproducible
 * fixed-order sums (gsums overwritten); NULL: fp32 atomics (gsums zeroed by the caller). */
int pdac_groupnorm_bwd_reduce(const PdacView* dy, const PdacView* y, const PdacView* x, int G, const float* stats,
                              const float* gamma, float eps, float* gsums, float* dgamma, float* dbeta, float* scratch,
                              void* stream);
/* dx = rstd * (dz*gamma - (gsums0 + xhat*gsums1) / (H*W*C/G)). */
int pdac_groupnorm_bwd_apply(const PdacView* dy, const PdacView* y, const PdacView* x, int G, const float* stats,
                             const float* gamma, float eps, const float* gsums, const PdacView* dx, void* stream);
/* y = x * keep / (1 - p) with keep a counter-based hash of (seed, image, channel) (per_channel = 1: nn.Dropout2d) or
 * of (seed, element) (nn.Dropout); the backward pass is the same call on the gradient. Not torch's Philox stream.
 * step: optional DEVICE pointer to a step counter that is mixed into the seed when the kernel runs (a fresh mask per
 * training step under CUDA-graph replay, where `seed` is baked into the graph). */
int pdac_dropout(const PdacView* x, const PdacView* y, float p, unsigned long long seed, const long long* step,
                 int per_channel, void* stream);
/* Depthwise 3x3, stride 1, padding = dilation. w: fp32 [C][9] (the state_dict (C,1,3,3) tensor). flip = 1 applies the
 * transposed taps (data gradient). */
int pdac_dwconv3x3(const PdacView* x, const float* w, int dil, int flip, const PdacView* y, void* stream);
/* dw[c][r*3+s] += sum dy[n,h,w,c] * x[n, h+(r-1)*dil, w+(s-1)*dil, c] (caller zeroes). scratch (fp32, >= 1184 * C * 9
 * floats): reproducible fixed-order sum over the blocks; NULL: fp32 atomics. */
int pdac_dwconv3x3_wgrad(const PdacView* x, const PdacView* dy, int dil, float* dw, float* scratch, void* stream);

/* ---- full-resolution segmentation head (head.cu) -----------------------------------------------------------
 * smp SegmentationHead of Unet: Conv2d(C, classes, 3, padding=1) + bias with C = 16 or 32 and 1 or 2 classes, as direct
 * HBM-bound kernels (a tensor-core GEMM with N = 1 would write and re-read 16-channel padded tensors).
 * weight: the fp32 state_dict tensor [classes][C][3][3], bias fp32 [classes] or NULL; logits / grad_logits: fp32 NCHW
 * (N, classes, H, W) — what the reference's callers see (train_worker.py:70-79). dw / dbias are accumulated into (state_dict
 * layouts); scratch: fp32, >= 592 * classes * (9*C + 1) floats (fixed-order partial rows: reproducible). */
int pdac_head3x3_supported(int channels, int classes); /* 1 / 0, no status code */
int pdac_head3x3_fwd(const PdacView* x, const float* weight, const float* bias, int classes, float* logits,
                     void* stream);
int pdac_head3x3_dgrad(const float* grad_logits, const float* weight, int classes, const PdacView* dx, void* stream);
int pdac_head3x3_wgrad(const PdacView* x, const float* grad_logits, int classes, float* dw, float* dbias, float* scratch,
                       long long scratch_bytes, void* stream);

/* ---- loss / metric / mask / stitch / optimiser (loss.cu) ---------------------------------------------------- */
/* tgt_dtype: 0 = fp32, 1 = uint8. sums: 3 * (1 + PDAC_STAT_ROWS) doubles: [0..2] = (sum p*t, sum p, sum t), kept for
 * backward; the rest takes the blocks' partial rows, added in a fixed order (no zeroing needed). loss: 1 float. */
int pdac_dice_fwd(const float* logits, const void* targets, int tgt_dtype, long long n, double* sums, float smooth,
                  float* loss, void* stream);
int pdac_dice_bwd(const float* logits, const void* targets, int tgt_dtype, long long n, const double* sums, float smooth,
                  const float* grad_out, float* grad_logits, void* stream);
/* The 1-class case of pdac_dice_bwd with grad_out = 1, written as the bf16 NHWC tensor [n][cpad] (class in channel 0, the
 * other cpad - 1 channels zero) that the head convolution's backward kernels read. */
int pdac_dice_bwd_nhwc(const float* logits, const void* targets, int tgt_dtype, long long n, const double* sums,
                       float smooth, void* grad_nhwc, int cpad, void* stream);
/* mask = ge(sigmoid(x), 0.5) as fp32 {0,1} and/or uint8. */
int pdac_sigmoid_mask(const float* logits, long long n, float* mask_f32, uint8_t* mask_u8, void* stream);
/* pred_kind: 0 = fp32 hard mask, 1 = fp32 logits (mask rule applied on the fly), 2 = uint8 mask. counts: 3 uint64
 * (intersection, pred, target). score = (2*I + smooth) / (pred + target + smooth). */
int pdac_dice_score(const void* pred, int pred_kind, const void* target, int tgt_dtype, long long n,
                    unsigned long long* counts, float smooth, float* score, void* stream);
int pdac_argmax_nchw(const float* logits, int N, int C, long long HW, uint8_t* out, void* stream);
/* tiles: uint8 [T][th][tw]; tile t lands at slide rows rows[t]*th - row0 .., columns cols[t]*tw ..; parts outside
 * the (slide_h x slide_w) band are dropped. */
int pdac_stitch_tiles(const uint8_t* tiles, const int32_t* rows, const int32_t* cols, int T, int th, int tw,
                      uint8_t* slide, long long slide_h, long long slide_w, long long row0, void* stream);
/* Tumour overlay (README.md:28-35; visualize/predict_overlay.ipynb is a missing blob, integer spec in
 * oracle/util_oracle.py overlay_blend): rgb / out uint8 HWC [npix][3], mask uint8 [npix]; where mask != 0
 * out = (rgb*(256-alpha_256) + color*alpha_256) >> 8, elsewhere out = rgb. out may alias rgb. */
int pdac_overlay_blend(const uint8_t* rgb, const uint8_t* mask, long long npix, int color_r, int color_g, int color_b,
                       int alpha_256, uint8_t* out, void* stream);
/* torch.optim.Adam semantics over one flat fp32 arena; grads are multiplied by grad_scale first. */
int pdac_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                   float beta1, float beta2, float eps, int step, float grad_scale, void* stream);
/* The same update with the step counter and the hyper-parameters in DEVICE memory, so that a whole train step replays
 * from one CUDA graph and a scheduler changes the learning rate by writing hyper[0] (ReduceLROnPlateau at
 * train_worker.py:240): *step += 1 first; hyper = float[12]: [0] lr, [1] beta1, [2] beta2, [3] eps, [4] grad_scale,
 * [7] 1 - beta1, [8] 1 - beta2 (the caller rounds them from double, as torch.optim.Adam does); [5..6] are written by
 * the library (lr / (1 - beta1^t), sqrt(1 - beta2^t), computed in double). */
int pdac_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, long long n, float* hyper,
                       long long* step, void* stream);

#ifdef __cplusplus
}
#endif
#endif
