/* dei2i.h - C ABI of libdei2i.so: the B200 (sm_100a) kernels behind the DefectGAN hot path.
 *
 * This is the drop-in boundary.  The reference (jason2714/de-i2i-gan) is pure PyTorch: its "FFI" for this path
 * is the set of ATen library calls made by its nn.Modules.  Each entry point below names the reference call
 * site(s) (path:line under /root/reference/defectGAN) whose device work it replaces.  Conventions:
 *   - plain pointers and sizes only; every pointer is DEVICE memory owned by the caller (PyTorch's allocator);
 *     the library never allocates, frees or retains device memory;
 *   - activations are NHWC ("channels_last") BF16 with the channel pitch a multiple of 8; parameters, gradients,
 *     statistics, losses and optimizer state are FP32;
 *   - all work is enqueued on `stream`; no host synchronisation, capturable in a CUDA graph;
 *   - return 0 on success, a negative dei2i_status otherwise; dei2i_last_error() gives the thread-local message;
 *   - no CPU fallback: on a machine without an sm_100 device every compute entry point fails with DEI2I_ERR_CUDA.
 */
#ifndef DEI2I_H
#define DEI2I_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* dei2i_stream; /* cudaStream_t */

enum dei2i_status { DEI2I_OK = 0, DEI2I_ERR_CUDA = -1, DEI2I_ERR_INVALID = -2 };
enum dei2i_act { DEI2I_ACT_NONE = 0, DEI2I_ACT_LRELU = 1 /* slope 0.2 */, DEI2I_ACT_RELU = 2, DEI2I_ACT_TANH = 3, DEI2I_ACT_SIGMOID = 4 };
/* upper bound of the block count of the block-partial reductions (scratch sizing, see the individual entry points) */
#define DEI2I_REDUCE_BLOCKS 2048

enum dei2i_pad { DEI2I_PAD_ZEROS = 0, DEI2I_PAD_REFLECT = 1 };

int dei2i_version(void);
const char* dei2i_last_error(void);
/* runtime tuning knobs (tile / pipeline selection); returns the previous value, -1 for an unknown key */
int dei2i_set_tuning(const char* key, int value);
/* number of kernels launched by this library since process start (bench.py's gpu_launches) */
long long dei2i_launch_count(void);
/* kernel variants enqueued by the calling thread's most recent dei2i_conv2d_{fwd,dgrad,wgrad} call, joined by '+'
 * (e.g. "conv_tma<128,4,2>", "wgrad_tma<1,4>"): lets parity tests assert which variant a geometry exercised */
const char* dei2i_last_kernel(void);

/* Geometry of one nn.Conv2d call (architecture.py:51-56,95-100,228-233,320-337; normalization.py:17-22). */
typedef struct dei2i_conv_geom {
  int N;        /* batch */
  int H, W;     /* logical input size seen by the convolution */
  int Hs, Ws;   /* stored input size; nearest-resized to (H, W) on the fly (F.interpolate(mode='nearest'),
                   normalization.py:29; nn.Upsample, architecture.py:49,203). Hs == H, Ws == W when no resize */
  int Cin;      /* stored input channels (pitch), multiple of 8 (real channels first, rest zero) */
  int Cout;     /* real output channels */
  int Cout_ld;  /* output channel pitch, multiple of 8, >= Cout */
  int kh, kw, stride, pad;
  int pad_mode; /* dei2i_pad */
} dei2i_conv_geom;

typedef struct dei2i_conv_epilogue {
  const float* scale;        /* device scalar multiplying the accumulator first (1/sigma of spectral norm, so the packed
                                weights stay those of W_orig and need no re-packing per forward) or NULL */
  const float* bias;         /* [Cout] or NULL (SPADE convs, normalization.py:17-22) */
  const float* noise;        /* [N*Ho*Wo] N(0,1) samples or NULL: NoiseInjection, architecture.py:385-389 */
  const float* noise_weight; /* device scalar; required when noise != NULL */
  const void* residual;      /* bf16 [N*Ho*Wo, Cout_ld] added last (NormResBlock, architecture.py:350) or NULL */
  int act;                   /* dei2i_act, applied after bias and noise (ConvBlock, architecture.py:106-107) */
  int out_fp32;              /* 0: y is bf16, 1: y is fp32 (same pitch) */
  /* The two fields below need the TMA / tcgen05 kernel (dei2i_conv2d_fwd_path(g) == 1) and a bf16 output: */
  float* stats_partial;      /* NULL, or fp32 [ceil(N*Ho*Wo / 128)][2][Cout_ld]: per 128-pixel tile the per-channel (mean, M2) of
                                the stored output - the producer-side half of the BatchNorm / InstanceNorm statistics
                                (generator.py:71,113,124; normalization.py:14), finished by dei2i_stats_finalize */
  int out_pad;               /* > 0: y is [N, Ho+2p, Wo+2p, Cout_ld]; the convolution writes the interior and the same call mirrors
                                the frame (the reflect padding of the NEXT convolution, architecture.py:95-100) */
} dei2i_conv_epilogue;

/* Packed K-major BF16 weight matrices consumed by the implicit-GEMM kernels.
 *   w        fp32 [Cout, Cin_real, kh, kw]  (nn.Conv2d.weight / weight_orig)
 *   inv_sigma device scalar or NULL: multiplies every element (spectral norm W/sigma, architecture.py:109-112)
 *   w_fwd    bf16 [Cout, Kf],  Kf = round_up(kh*kw*Cin, 64),      k = (r*kw+s)*Cin + ci
 *   w_dgrad  bf16 [Cin,  Kd],  Kd = round_up(kh*kw*Cout_ld, 64),  k = (r*kw+s)*Cout_ld + co   (or NULL)      */
int dei2i_pack_weight(const float* w, const float* inv_sigma, int Cout, int Cin_real, int kh, int kw, int Cin,
                      int Cout_ld, void* w_fwd, void* w_dgrad, dei2i_stream stream);

/* dei2i_pack_weight for many weights in one launch per 32 items (a DefectGAN network has ~50 convolutions whose ~8 us packing
 * launches are pure latency; same call sites).  items is a HOST array.  w2 != NULL packs the row-wise concatenation [w ; w2]
 * (rows_w rows from w, the rest from w2: SPADE's mlp_gamma | mlp_beta and the two generator heads, normalization.py:21-22,
 * generator.py:224-241) without materialising it.  Results are bit-identical to dei2i_pack_weight without inv_sigma. */
typedef struct dei2i_pack_item {
  const float* w;
  const float* w2; /* or NULL */
  void* w_fwd;
  void* w_dgrad;   /* or NULL */
  int rows_w;      /* rows taken from w (= Cout when w2 is NULL) */
  int Cout, Cin_real, kh, kw, Cin, Cout_ld;
} dei2i_pack_item;
int dei2i_pack_weight_multi(const dei2i_pack_item* items, int count, dei2i_stream stream);

/* y = act(scale*conv(x) + bias + noise_weight*noise) + residual: the nn.Conv2d forward of ConvBlock / DeConvBlock / NormConvBlock /
 * NormResBlock / SPADE (architecture.py:51-56,95-107,228-244,320-350; normalization.py:17-22,30-32) with the followers the modules apply
 * right after it (spectral-norm 1/sigma :109-112, NoiseInjection :385-389, activation :106-107, residual add :350) */
int dei2i_conv2d_fwd(const void* x, const void* w_fwd, void* y, const dei2i_conv_geom* g,
                     const dei2i_conv_epilogue* e, dei2i_stream stream);
/* which forward kernel family dei2i_conv2d_fwd picks for g: 1 = TMA-im2col tcgen05 (supports stats_partial / out_pad), 0 = other */
int dei2i_conv2d_fwd_path(const dei2i_conv_geom* g);
/* mirrors the frame of an [N, H+2p, W+2p, C] bf16 tensor whose interior is written (F.pad(mode='reflect') in place: the padding
 * nn.Conv2d(padding_mode='reflect') applies to its input, architecture.py:95-100) */
int dei2i_reflect_border(void* y, int N, int H, int W, int C, int pad, dei2i_stream stream);

/* Gradient w.r.t. the convolution input (autograd of the nn.Conv2d calls listed at dei2i_conv2d_fwd, reached from loss.backward(),
 * trainers/defectgan_trainer.py:160,177: conv_transpose of dy with W).
 * dy: bf16 [N,Ho,Wo,Cout_ld].  pad_mode zeros:   dx is [N,H,W,Cin].
 *                              pad_mode reflect: dx is over the PADDED domain [N,H+2p,W+2p,Cin]; fold it with
 *                              dei2i_reflect_fold (the backward of F.pad(mode='reflect')).
 * The output pitch is Cin; out_fp32 selects fp32 / bf16; scale (device scalar or NULL) multiplies the result.
 * Resized inputs (Hs != H) are not supported here. */
int dei2i_conv2d_dgrad(const void* dy, const void* w_dgrad, void* dx, const dei2i_conv_geom* g, const float* scale,
                       int out_fp32, dei2i_stream stream);
/* dx[n,h,w,c] = sum of dxp over all padded positions that reflect onto (h,w); bf16 in / bf16 out: autograd of the reflect padding in
 * front of every padding_mode='reflect' convolution (architecture.py:95-100) */
int dei2i_reflect_fold(const void* dxp, void* dx, int N, int H, int W, int C, int pad, dei2i_stream stream);
/* The same fold done in place: mirror images are added into the interior positions of the padded tensor itself, the frame keeps its
 * values (only the ~3 % of the pixels that have a mirror image are touched).  The interior of dxp is then the folded gradient, read
 * with the padded pitch by consumers that accept it (dei2i_bn_bwd with dy_pad < 0). */
int dei2i_reflect_fold_inplace(void* dxp, int N, int H, int W, int C, int pad, dei2i_stream stream);

/* F.pad(x, (pad,)*4, mode='reflect') (what nn.Conv2d(padding_mode='reflect') does before its convolution,
 * architecture.py:95-100): x bf16 [N,H,W,C] -> xp bf16 [N,H+2p,W+2p,C].  The TMA-fed convolution kernels take zero / no
 * padding only (TMA fills out-of-bounds with zeros, it cannot mirror), so a reflect convolution is this followed by a
 * pad = 0 convolution; its backward is dei2i_conv2d_dgrad over the padded domain followed by dei2i_reflect_fold. */
int dei2i_reflect_pad(const void* x, void* xp, int N, int H, int W, int C, int pad, dei2i_stream stream);

/* Gradient w.r.t. the weight (autograd of the same nn.Conv2d calls; the spectral-norm term is the backward of
 * torch.nn.utils.spectral_norm, architecture.py:68-71,109-112), split over pixels, deterministic two-pass reduction.
 *   dw  fp32 [Cout, Cin_real, kh, kw]; written (beta = 0) or accumulated (beta = 1).
 *   workspace: dei2i_conv2d_wgrad_workspace(g) bytes of device scratch.
 *   sn_u/sn_v/sn_w/sn_inv_sigma: when non-NULL apply the spectral-norm backward
 *        dW_orig = (dW - <dW, W/sigma> u v^T) * inv_sigma            (torch.nn.utils.spectral_norm autograd). */
size_t dei2i_conv2d_wgrad_workspace(const dei2i_conv_geom* g, int Cin_real);
int dei2i_conv2d_wgrad(const void* x, const void* dy, float* dw, int Cin_real, float beta, const dei2i_conv_geom* g,
                       const float* sn_w_orig, const float* sn_u, const float* sn_v, const float* sn_inv_sigma,
                       void* workspace, size_t workspace_bytes, dei2i_stream stream);

/* One power iteration of torch.nn.utils.spectral_norm (old hook API, architecture.py:7,68-71,109-112):
 *   v <- normalize(W^T u); u <- normalize(W v); sigma = u^T W v;  W viewed as [rows, cols] fp32.
 *   do_power_iter = 0 (eval mode) only recomputes sigma.  Writes inv_sigma[0] = 1/sigma, inv_sigma[1] = sigma.
 *   scratch: (rows + 17 * cols) floats. */
int dei2i_spectral_norm(const float* w, float* u, float* v, int rows, int cols, int do_power_iter, float eps,
                        float* inv_sigma, float* scratch, dei2i_stream stream);

/* The same power iteration for many weights in four launches (a DefectGAN network has ~45 spectral-normalised convolutions;
 * architecture.py:68-71,109-112 wrap every one).  Arithmetic and summation order are those of dei2i_spectral_norm, so the
 * results are bit-identical.  items is a HOST array.  u_snap / v_snap (optional, device) receive copies of the u / v the
 * returned sigma belongs to - what the spectral-norm backward of dei2i_conv2d_wgrad needs after later calls have moved u, v on.
 * scratch: dei2i_spectral_norm_multi_scratch(items, count) floats of device memory. */
typedef struct dei2i_sn_item {
  const float* w;
  float* u;
  float* v;
  float* inv_sigma; /* [2]: 1/sigma, sigma */
  float* u_snap;
  float* v_snap;
  int rows, cols;
} dei2i_sn_item;
size_t dei2i_spectral_norm_multi_scratch(const dei2i_sn_item* items, int count);
int dei2i_spectral_norm_multi(const dei2i_sn_item* items, int count, int do_power_iter, float eps, float* scratch,
                              size_t scratch_floats, dei2i_stream stream);

/* ---- WGAN / StarGAN-v2 resampling (bf16 NHWC, C % 8 == 0) ----
 * nn.Upsample(scale_factor=up, 'nearest') fused with the zero padding of the following nn.Conv2d(padding='same'), which is
 * asymmetric for even kernels (WGanGenerator, generator.py:15-37: 4x4 'same' -> pad 1 left/top, 2 right/bottom):
 * out [N, up*H + pt + pb, up*W + pl + pr, C].  _bwd is the adjoint (sums each up x up window of the interior). */
int dei2i_upsample_pad_fwd(const void* x, void* out, int N, int H, int W, int C, int up, int pad_top, int pad_left,
                           int pad_bottom, int pad_right, dei2i_stream stream);
int dei2i_upsample_pad_bwd(const void* dout, void* dx, int N, int H, int W, int C, int up, int pad_top, int pad_left,
                           int pad_bottom, int pad_right, dei2i_stream stream);
/* nn.MaxPool2d(k, stride, pad) (WGanDiscriminator, discriminator.py:23-25).  idx: uint8 [N,Ho,Wo,C] window offset of the first
 * maximum in scan order (the element torch routes the gradient to); _bwd gathers. */
int dei2i_maxpool_fwd(const void* x, void* out, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                      dei2i_stream stream);
int dei2i_maxpool_bwd(const void* dout, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                      dei2i_stream stream);
/* F.avg_pool2d(x, k) * scale (stargan-v2/core/model.py:53,60; scale folds the residual 1/sqrt(2), model.py:70) and adjoint. */
int dei2i_avgpool_fwd(const void* x, void* out, int N, int H, int W, int C, int k, float scale, dei2i_stream stream);
int dei2i_avgpool_bwd(const void* dout, void* dx, int N, int H, int W, int C, int k, float scale, dei2i_stream stream);
/* nn.AdaptiveAvgPool2d(1) (discriminator.py:35): out fp32 [N,C]; _bwd: dx bf16 [N,HW,C] = dout / HW. */
int dei2i_gap_fwd(const void* x, float* out, int N, int HW, int C, dei2i_stream stream);
int dei2i_gap_bwd(const float* dout, void* dx, int N, int HW, int C, dei2i_stream stream);

/* Row-packed image for few-channel, wide-filter convolutions (the generator stem: ConvBlock 3->64, 7x7, 'same' reflect
 * padding, generator.py:76-86 / architecture.py:95-100).  img: fp32 NCHW [N,C<=8,H,W] -> xp: bf16 [N, H+2*pad, W, 64] with
 * xp[n,hp,w,s*8+c] = img[n,c,refl(hp-pad),refl(w+s-pad)] (s < kw <= 8): the convolution becomes a kh x 1 'valid'
 * convolution with 64 input channels (weights re-laid accordingly by the caller).  _bwd is the adjoint (gradient w.r.t. img). */
int dei2i_rowpack_fwd(const float* img, void* xp, int N, int C, int H, int W, int kw, int pad, dei2i_stream stream);
int dei2i_rowpack_bwd(const void* dxp, float* dimg, int N, int C, int H, int W, int kw, int pad, dei2i_stream stream);

/* NCHW fp32 <-> NHWC bf16 (channel pitch C_ld >= C, pad channels zero) at the model edge: the images entering
 * DefectGanModel.forward (defectgan_model.py:49-52) and the images / logits leaving the networks (generator.py:268-275, discriminator.py:92-98). */
int dei2i_nchw_to_nhwc(const float* src, void* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream);
int dei2i_nhwc_to_nchw(const void* src, float* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream);

/* Per-channel (groups = 1) or per-(n,channel) (groups = N) mean and BIASED variance of a bf16 NHWC tensor:
 * nn.BatchNorm2d / nn.InstanceNorm2d statistics (generator.py:71,113,124; normalization.py:14).
 * stats: fp32 [groups, 2, C] ([g][0][c] = mean, [g][1][c] = variance), overwritten.  Deterministic: per-block (mean, M2) pairs
 * (shifted sums, no atomics) in `scratch` (>= dei2i_channel_stats_scratch(...) floats), merged in a fixed order (Chan et al.). */
size_t dei2i_channel_stats_scratch(int N, int HW, int C, int per_sample);
int dei2i_channel_stats(const void* x, int N, int HW, int C, int per_sample, float* stats, float* scratch, size_t scratch_floats,
                        dei2i_stream stream);
/* dei2i_channel_stats without a scratch argument, for callers that must not allocate (CUDA-graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.  Same arguments,
 * same kernels and bit-identical results; the per-block pairs go to a file-scope `static __device__` array of
 * DEI2I_STATS_STATIC_MAX_FLOATS floats (4 MiB, held by every GPU that loads the library until exit) whose device address is looked
 * up once per device with cudaGetSymbolAddress.  Calls no cudaMalloc / cudaFree / allocating routine.  Shapes that need more
 * scratch than that (dei2i_channel_stats_scratch(...) > the maximum) return DEI2I_ERR_INVALID with a message (dei2i_last_error) - use dei2i_channel_stats.
 * Every call on a device shares the one array (it keeps the last call's contents; nothing needs clearing, each call overwrites what
 * it reads): calls that overlap on the GPU - concurrent streams with no ordering between them - race. */
#define DEI2I_STATS_STATIC_MAX_FLOATS (1u << 20)
size_t dei2i_channel_stats_static_max_floats(void);
int dei2i_channel_stats_static(const void* x, int N, int HW, int C, int per_sample, float* stats, dei2i_stream stream);
/* Second half of the statistics when the producer emitted the per-tile pairs itself (dei2i_conv_epilogue.stats_partial, 128 rows
 * per tile): partial [groups * tiles_per_group][2][ld] (mean, M2), tile t of a group covering rows [t * rows_per_tile, ...) of the
 * group's rows_per_group rows -> stats [groups, 2, C] (mean, biased variance).  Fixed merge order. */
int dei2i_stats_finalize(const float* partial, int tiles_per_group, int rows_per_tile, long long rows_per_group, int groups, int C,
                         int ld, float* stats, dei2i_stream stream);

/* nn.BatchNorm2d (+ LeakyReLU(0.2) / residual add) forward, ConvBlock architecture.py:103-107, ResBlock :174-176.
 * training: uses batch statistics from `stats` (dei2i_channel_stats, groups=1), updates running stats
 * (momentum, unbiased var) and writes save_mean/save_rstd [C]; eval: uses running stats.
 * y = act(bn(x)) + residual */
int dei2i_bn_fwd(const void* x, void* y, const void* residual, const float* stats, const float* gamma,
                 const float* beta, float* running_mean, float* running_var, float* save_mean, float* save_rstd,
                 int N, int HW, int C, int training, float momentum, float eps, int act, dei2i_stream stream);
/* The same forward writing the reflect padding of the convolution that consumes it (ConvBlock -> ConvBlock chains of the
 * discriminator and of the generator's encoder, architecture.py:95-100 padding_mode='reflect'): y is
 * [N, H+2*out_pad, W+2*out_pad, C]; residual (optional) is unpadded like x.  The backward takes the gradient folded by
 * dei2i_reflect_fold and calls dei2i_bn_bwd. */
int dei2i_bn_fwd_pad(const void* x, void* y, const void* residual, const float* stats, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, float* save_mean, float* save_rstd, int N, int H, int W, int C,
                     int training, float momentum, float eps, int act, int out_pad, dei2i_stream stream);
/* backward of the above: dx, dbeta | dgamma (fp32 [2, C] as ONE buffer, dbeta first; dgamma = dbeta + C).  The activation mask is
 * recomputed from x (the forward output is not needed).  dy_pad > 0: dy is the gradient of the PADDED output of dei2i_bn_fwd_pad
 * ([N, H+2p, W+2p, C]) and is folded while it is read (no dei2i_reflect_fold pass); dy_pad < 0: the same padded layout with
 * p = -dy_pad whose border was already folded IN PLACE by dei2i_reflect_fold_inplace (interior read only).  Deterministic: per-block partial sums in
 * `scratch` (>= dei2i_bn_bwd_scratch(N, H*W, C) floats) finished in a fixed order. */
size_t dei2i_bn_bwd_scratch(int N, int HW, int C);
int dei2i_bn_bwd(const void* x, const void* dy, void* dx, const float* gamma, const float* beta, const float* save_mean,
                 const float* save_rstd, float* dgamma, float* dbeta, float* scratch, size_t scratch_floats, int N, int H, int W,
                 int C, int training, int act, int dy_pad, dei2i_stream stream);

/* SPADE modulation (normalization.py:24-37) + ReLU (architecture.py:244,347-348), optional 2x nearest up-sampling
 * of x (architecture.py:242): out = [relu](IN(x) * (1 + gamma) + beta), gamma/beta = first/second C columns of `gb`
 * (bf16 [N,H,W,2C], output of the fused gamma|beta convolution).  x is [N,H/up,W/up,C].
 * stats from dei2i_channel_stats(per_sample=1) on x; writes mean_rstd fp32 [N,2,C] for the backward. */
int dei2i_spade_fwd(const void* x, const void* gb, void* out, const float* stats, float* mean_rstd, int N, int H,
                    int W, int C, int up, int relu, float eps, dei2i_stream stream);
/* backward: dgb (bf16 [N,H,W,2C]) and dx (bf16 [N,H/up,W/up,C]); scratch >= dei2i_image_reduce_scratch(N, H*W, C) floats
 * (per-block partial sums, finished in a fixed order - no atomics). */
size_t dei2i_image_reduce_scratch(int N, int HW, int C);
int dei2i_spade_bwd(const void* x, const void* gb, const void* out, const void* dout, void* dgb, void* dx,
                    const float* mean_rstd, float* scratch, size_t scratch_floats, int N, int H, int W, int C, int up, int relu,
                    dei2i_stream stream);

/* SPADE modulation over a border-class map: the training path feeds SPADE a 1x1 label map (defectgan_model.py:185-205),
 * which normalization.py:29 nearest-resizes to a spatially constant map; mlp_shared -> mlp_gamma/mlp_beta (two stacked
 * zero-padded 3x3 convolutions, normalization.py:30-32) then take 5 distinct values per axis (rows 0, 1, interior, H-2,
 * H-1).  gb5: bf16 [N,5,5,2C] = the same convolutions evaluated on a constant 5x5 map; pixel (h,w) uses entry
 * (cls(h), cls(w)).  Exact restructuring of normalization.py:24-37, H, W >= 5.
 * backward: dgb5 fp32 [N,5,5,2C] (class-wise sums of the per-pixel gamma/beta gradients, written), dx bf16;
 * scratch >= dei2i_spade_cls_bwd_scratch(...) floats.  Deterministic: the interior class is reduced band by band, each frame
 * class by its own block, everything summed in a fixed order (no atomics); the backward takes out_pad in {0, 1}.
 * out_pad > 0: `out` (and `dout` of the backward) is [N, H+2*out_pad, W+2*out_pad, C] and already carries the reflect
 * padding of the convolution that consumes it (architecture.py:228-233,320-337 padding_mode='reflect'): the forward writes the
 * mirrored border itself, the backward folds the padded gradient while reading it.
 * The backward supports up = 1 or 2 (what architecture.py uses) and does not read `out`: the ReLU mask is re-evaluated from x
 * with the forward's own expression (`out` stays in the signature; it may be NULL). */
int dei2i_spade_cls_fwd(const void* x, const void* gb5, void* out, const float* stats, float* mean_rstd, int N, int H,
                        int W, int C, int up, int relu, float eps, int out_pad, dei2i_stream stream);
size_t dei2i_spade_cls_bwd_scratch(int N, int H, int W, int C, int up);
int dei2i_spade_cls_bwd(const void* x, const void* gb5, const void* out, const void* dout, float* dgb5, void* dx,
                        const float* mean_rstd, float* scratch, size_t scratch_floats, int N, int H, int W, int C, int up, int relu,
                        int out_pad, dei2i_stream stream);

/* ---- StarGAN-v2 normalisation (stargan-v2/core/model.py) ----
 * out = lrelu_slope(IN(x) * scale[n,c] + shift[n,c]): AdaIN with scale = 1 + gamma(s), shift = beta(s) (model.py:73-84), or
 * InstanceNorm2d(affine=True) with scale = weight[c], shift = bias[c] broadcast over n (model.py:42-43), fused with the
 * LeakyReLU(0.2) that always follows (model.py:56-63,104-112; slope = 1: none).  stats: dei2i_channel_stats(per_sample=1);
 * scale / shift fp32 [N,C]; mean_rstd fp32 [N,2,C] (written).
 * backward: dx bf16 and dshift_dscale fp32 [N,2,C] = (sum g, sum g*xhat), g = dout * lrelu'(out). */
int dei2i_adain_act_fwd(const void* x, void* out, const float* stats, const float* scale, const float* shift,
                        float* mean_rstd, int N, int HW, int C, float slope, float eps, dei2i_stream stream);
int dei2i_adain_act_bwd(const void* x, const void* out, const void* dout, void* dx, const float* scale,
                        const float* mean_rstd, float* dshift_dscale, float* scratch, size_t scratch_floats, int N, int HW, int C,
                        float slope, dei2i_stream stream);  /* scratch >= dei2i_image_reduce_scratch(N, HW, C) floats */
/* y = act(x) over n bf16 elements (pre-activation of the StarGAN-v2 ResBlk, model.py:56,62); n % 8 == 0. */
int dei2i_act_fwd(const void* x, void* y, long long n, int act, dei2i_stream stream);
/* out = sa * a + sb * b (b may be NULL) over n bf16 elements: residual merge (x + r) / sqrt(2), model.py:66-68,118-121. */
int dei2i_axpby(const void* a, const void* b, void* out, long long n, float sa, float sb, dei2i_stream stream);

/* act backward for fused conv epilogues (autograd of the LeakyReLU / ReLU / Tanh / Sigmoid that follow a convolution,
 * architecture.py:106-107, generator.py:224-241, discriminator.py:60-75): dx = dy * act'(y) computed from the saved OUTPUT y (bf16);
 * bias_grad (fp32 [C], may be NULL) receives the sum over pixels of dx (per-block partials in `scratch`, >= DEI2I_REDUCE_BLOCKS * C
 * floats, finished in a fixed order).  pad > 0: y and dy are the consumer-PADDED tensors [N, H+2p, W+2p, C] written by
 * dei2i_conv2d_fwd with out_pad: y is read at the interior, dy folded on read; dx is dense [N, H, W, C]; M = N * H * W. */
int dei2i_act_bwd(const void* y, const void* dy, void* dx, float* bias_grad, float* scratch, long long M, int C, int act, int H,
                  int W, int pad, dei2i_stream stream);

/* NoiseInjection weight gradient: dw[0] = beta * dw[0] + sum_{m,c} dy[m,c] * noise[m]  (architecture.py:389);
 * scratch >= DEI2I_REDUCE_BLOCKS floats. */
int dei2i_noise_wgrad(const void* dy, const float* noise, float* dw, float* scratch, long long M, int C, int C_real, float beta,
                      dei2i_stream stream);

/* Generator heads + composition (generator.py:268-270): hd = fp32 [M,8] raw head convolution output
 * (cols 0..2 foreground logits, col 3 distribution logit); x, out: fp32 NCHW images; prob: fp32 [N,1,H,W].
 * out = x*(1-p) + tanh(fg)*p,  p = sigmoid(hd[:,3]).  out_nhwc (bf16, pitch 8, may be NULL) gets a copy of out. */
int dei2i_compose_fwd(const float* hd, const float* x, float* out, float* prob, void* out_nhwc, int N, int H, int W,
                      int cycle_gan, dei2i_stream stream);
/* backward: dhd (bf16 [M,8]) and dx (fp32 NCHW, may be NULL) from dout (fp32 NCHW, may be NULL) and dprob. */
int dei2i_compose_bwd(const float* hd, const float* x, const float* dout, const float* dprob, void* dhd, float* dx,
                      int N, int H, int W, int cycle_gan, dei2i_stream stream);

/* Loss reductions (BaseModel._cal_loss, models/base_model.py:68-80).  Each writes loss[0] (mean) and, when grad is
 * non-NULL, grad = grad_scale * d(mean loss)/d(input) (fp32, same shape as input).
 * bce: binary_cross_entropy_with_logits(logits, targets) with targets either a constant (targets == NULL,
 * target_const) or an fp32 tensor;  l1: l1_loss(a, b) with b == NULL meaning zeros. */
int dei2i_bce_logits(const float* logits, const float* targets, float target_const, long long n, float* loss,
                     float* grad, float grad_scale, dei2i_stream stream);
int dei2i_l1(const float* a, const float* b, long long n, float* loss, float* grad_a, float* grad_b, float grad_scale,
             int accumulate_grad, float* scratch /* >= DEI2I_REDUCE_BLOCKS floats */, dei2i_stream stream);

/* mse: mse_loss(a, b) (b == NULL: zeros); same gradient conventions as dei2i_l1. */
int dei2i_mse(const float* a, const float* b, long long n, float* loss, float* grad_a, float* grad_b, float grad_scale,
              int accumulate_grad, float* scratch /* >= DEI2I_REDUCE_BLOCKS floats */, dei2i_stream stream);
/* cce: cross_entropy(logits [B,C], class-probability targets [B,C]) = mean_b(-sum_c t log_softmax(z)). */
int dei2i_ce_soft(const float* logits, const float* targets, int B, int C, float* loss, float* grad, float grad_scale,
                  dei2i_stream stream);

/* torch.optim.Adam step (trainers/base_trainer.py:75-89) over one flat fp32 parameter vector. step is 1-based. */
int dei2i_adam(float* p, const float* g, float* m, float* v, long long n, float lr, float beta1, float beta2,
               float eps, int step, dei2i_stream stream);

/* The same step with the per-step scalars in DEVICE memory: hyper[0] = lr / (1 - beta1^step),
 * hyper[1] = 1 / sqrt(1 - beta2^step), hyper[2] = 1 - lr * weight_decay for torch.optim.AdamW's decoupled decay (base_trainer.py:
 * 82-84, the MAE pre-training default) or 1 for Adam.  The launch arguments are then constant from step to step, so the optimizer step
 * can be part of a captured CUDA graph of the whole D-step / G-step (trainers/defectgan_trainer.py:138-180). */
int dei2i_adam_dev(float* p, const float* g, float* m, float* v, long long n, const float* hyper, float beta1, float beta2,
                   float eps, dei2i_stream stream);

/* torch.optim.RMSprop step (alpha 0.99, eps 1e-8, no momentum, not centered: trainers/base_trainer.py:76-77 with
 * options/wgan_options.py:49-57) and the WGAN weight clipping p.data.clamp_(lo, hi) (models/wgan_model.py:18-21). */
int dei2i_rmsprop(float* p, const float* g, float* square_avg, long long n, float lr, float alpha, float eps,
                  dei2i_stream stream);
/* the same step with the learning rate in DEVICE memory (hyper[0]): constant launch arguments, CUDA-graph capturable */
int dei2i_rmsprop_dev(float* p, const float* g, float* sq, long long n, const float* hyper, float alpha, float eps, dei2i_stream stream);
int dei2i_clamp(float* p, long long n, float lo, float hi, dei2i_stream stream);
/* p_ema = lerp(p, p_ema, beta): StarGAN-v2 moving_average (stargan-v2/core/solver.py:549-551). */
int dei2i_ema(const float* p, float* p_ema, long long n, float beta, dei2i_stream stream);

/* ---- DiffAugment (defectGAN/utils/diffaug.py:9-75) on FP32 NCHW images [N, C <= 4, H, W] - between the generator output and the
 * discriminator input (defectgan_model.py:200-203,266-271).  policy = OR of the flags below, applied in the reference's order
 * color (brightness, saturation, contrast) -> translation (ratio 1/8) -> cutout (ratio 1/2), fused into one pass (+ one per-image
 * reduction for the contrast mean).  params fp32 [N, 8], drawn by the caller with the reference's torch calls:
 * [0] brightness - 0.5, [1] 2 * saturation, [2] contrast + 0.5, [3] translation_x (rows), [4] translation_y (columns),
 * [5] cutout offset_x, [6] offset_y, [7] unused.  scratch >= dei2i_diffaug_scratch(N, H, W) floats.  bwd: dx = J^T dy. */
enum dei2i_aug { DEI2I_AUG_COLOR = 1, DEI2I_AUG_TRANSLATION = 2, DEI2I_AUG_CUTOUT = 4 };
size_t dei2i_diffaug_scratch(int N, int H, int W);
int dei2i_diffaug_fwd(const float* x, float* y, const float* params, float* scratch, int N, int C, int H, int W, int policy,
                      dei2i_stream stream);
int dei2i_diffaug_bwd(const float* dy, float* dx, const float* params, float* scratch, int N, int C, int H, int W, int policy,
                      dei2i_stream stream);

/* ---- MAE pre-training input (defectgan_model.py:361-383; MaskToken, architecture.py:392-418; generate_shifted_mask, utils/util.py:60-71)
 * masked = img * m + token * (1 - m) on FP32 NCHW images, m evaluated on the fly from the coarse Bernoulli grid `coarse` [N, gh, gw]
 * (nearest-upsampled by `patch`, cropped at (h_shift, w_shift)); `mask` (optional, [N,1,H,W]) receives m.  token: NULL (zero token) or a
 * broadcastable tensor of extents token_dims[4] (host int array), each 1 or the image's.  dei2i_mask_mean: the 'mean' token
 * [N, C] = mean_{h,w}(img * m) / mask_ratio.  dei2i_mask_token_bwd: gradient of a learnable token ([1,1|C,1,1] or [1,1|C,H,W]).
 * scratch >= dei2i_mask_token_scratch(N, C, H, W) floats. */
size_t dei2i_mask_token_scratch(int N, int C, int H, int W);
int dei2i_mask_token_fwd(const float* img, const float* coarse, const float* token, const int* token_dims, float* out, float* mask,
                         int N, int C, int H, int W, int patch, int h_shift, int w_shift, int gh, int gw, dei2i_stream stream);
int dei2i_mask_mean(const float* img, const float* coarse, float* token_mean, float* scratch, float mask_ratio, int N, int C, int H,
                    int W, int patch, int h_shift, int w_shift, int gh, int gw, dei2i_stream stream);
int dei2i_mask_token_bwd(const float* dy, const float* coarse, const int* token_dims, float* dtoken, float* scratch, int N, int C, int H,
                         int W, int patch, int h_shift, int w_shift, int gh, int gw, dei2i_stream stream);

/* ---- FP32-parity path (north-star: per-layer outputs and gradients within 1e-3 relative in FP32; the reference trains in FP32,
 * defectgan_model.py:54).  The same operators on FP32 NHWC activations [N, H, W, C_ld] with exact FP32 arithmetic (CUDA-core FMA,
 * fixed-order reductions): every function has the semantics of its BF16 sibling above, with `float*` activations.  Packed weights are
 * FP32 [Cout, Kf] / [Cin, Kd] (same k order).  dei2i_conv2d_fwd_f32 honours scale / bias / noise / act / residual (FP32) of the
 * epilogue struct (stats_partial, out_pad, out_fp32 unused).  Selected by tensor dtype in the Python wrappers (DEI2I_KERNELS=fp32). */
int dei2i_pack_weight_f32(const float* w, int Cout, int Cin_real, int kh, int kw, int Cin, int Cout_ld, float* w_fwd, float* w_dgrad,
                          dei2i_stream stream);
int dei2i_conv2d_fwd_f32(const float* x, const float* w_fwd, float* y, const dei2i_conv_geom* g, const dei2i_conv_epilogue* e,
                         dei2i_stream stream);
int dei2i_conv2d_dgrad_f32(const float* dy, const float* w_dgrad, float* dx, const dei2i_conv_geom* g, const float* scale,
                           dei2i_stream stream);
size_t dei2i_conv2d_wgrad_f32_workspace(const dei2i_conv_geom* g);
int dei2i_conv2d_wgrad_f32(const float* x, const float* dy, float* dw, int Cin_real, float beta, const dei2i_conv_geom* g,
                           const float* sn_w_orig, const float* sn_u, const float* sn_v, const float* sn_inv_sigma, void* workspace,
                           size_t workspace_bytes, dei2i_stream stream);
int dei2i_nchw_to_nhwc_f32(const float* src, float* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream);
int dei2i_nhwc_to_nchw_f32(const float* src, float* dst, int N, int C, int H, int W, int C_ld, dei2i_stream stream);
int dei2i_channel_stats_f32(const float* x, int N, int HW, int C, int per_sample, float* stats, dei2i_stream stream);
int dei2i_bn_fwd_f32(const float* x, float* y, const float* residual, const float* stats, const float* gamma, const float* beta,
                     float* running_mean, float* running_var, float* save_mean, float* save_rstd, int N, int HW, int C, int training,
                     float momentum, float eps, int act, dei2i_stream stream);
int dei2i_bn_bwd_f32(const float* x, const float* dy, float* dx, const float* gamma, const float* beta, const float* save_mean,
                     const float* save_rstd, float* dbeta_dgamma, int N, int HW, int C, int training, int act, dei2i_stream stream);
int dei2i_spade_cls_fwd_f32(const float* x, const float* gb5, float* out, const float* stats, float* mean_rstd, int N, int H, int W,
                            int C, int up, int relu, float eps, int out_pad, dei2i_stream stream);
int dei2i_spade_cls_bwd_f32(const float* x, const float* gb5, const float* dout, float* dgb5, float* dx, const float* mean_rstd,
                            float* red, int N, int H, int W, int C, int up, int relu, int out_pad, dei2i_stream stream);
int dei2i_spade_fwd_f32(const float* x, const float* gb, float* out, const float* stats, float* mean_rstd, int N, int H, int W, int C,
                        int up, int relu, float eps, dei2i_stream stream);
int dei2i_reflect_pad_f32(const float* x, float* xp, int N, int H, int W, int C, int pad, dei2i_stream stream);
int dei2i_reflect_fold_f32(const float* dxp, float* dx, int N, int H, int W, int C, int pad, dei2i_stream stream);
int dei2i_act_bwd_f32(const float* y, const float* dy, float* dx, float* bias_grad, long long M, int C, int act, dei2i_stream stream);
int dei2i_noise_wgrad_f32(const float* dy, const float* noise, float* dw, long long M, int C, int C_real, float beta,
                          dei2i_stream stream);
int dei2i_compose_bwd_f32(const float* hd, const float* x, const float* dout, const float* dprob, float* dhd, float* dx, int N, int H,
                          int W, int cycle_gan, dei2i_stream stream);

/* ---- training input pipeline on the device (train_defectgan.py:57-65): Resize(1.5 S) -> RandomResizedCrop(S) -> flips -> ColorJitter ->
 * ToTensor -> Normalize(0.5, 0.5) from decoded uint8 HWC images (concatenated in `src_u8`, image n at byte offset src_offsets[n] with
 * size src_hw[2n] x src_hw[2n+1], all in device memory) to the FP32 NCHW batch `out` [N, 3, S, S].  params fp32 [N, 12]: crop box in
 * source pixels (y0, x0, h, w), flip_h, flip_v, brightness / contrast / saturation factors, the order of the three jitter ops (op ids
 * 0 / 1 / 2), unused.  Antialiased bilinear (triangle) resampling as PIL / torchvision.  scratch >= dei2i_augment_scratch(N, S) floats. */
size_t dei2i_augment_scratch(int N, int S);
int dei2i_augment_train(const void* src_u8, const long long* src_offsets, const int* src_hw, const float* params, float* out,
                        float* scratch, int N, int S, int jitter, dei2i_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* DEI2I_H */
