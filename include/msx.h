/*
 * msx.h - C-ABI of libmsx.so: the sm_100a (B200) kernels behind the drop-in Python API of
 * medical_segmentation_pytorch_b200.
 *
 * The reference (AjayGaneshC/medical-segmentation-pytorch) is 100 % Python and has NO FFI of its
 * own: every FLOP goes through stock torch ops.  Each entry point below
 * therefore cites the reference Python call(s) whose device work it replaces; INTEGRATION.md
 * shows the ctypes stub a maintainer of the reference would add at that call site.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++ or torch types cross this line.
 *  - every function returns 0 (MSX_OK) or a negative msx_status; msx_last_error() gives the text
 *    (thread-local).  Unsupported shapes return MSX_ERR_UNSUPPORTED - there is no CPU fallback.
 *  - all pointers are DEVICE pointers unless named host_*; the CALLER owns every buffer,
 *    including workspaces (sizes from the msx_ws_bytes_* queries).  The library never allocates
 *    device memory in these calls.
 *  - all work is enqueued asynchronously on `stream` (a cudaStream_t passed as void*); no call
 *    synchronises the device.  Calls that share a workspace must be ordered on one stream.
 *  - activations are NHWC with a pixel stride `ld` (elements) >= round_up(C, 8); channels
 *    [C, round_up(C,8)) are padding and always hold zeros (every producer writes them as zero).
 *    `ld` > C lets a producer write into a channel slice of a wider buffer (zero-copy concat).
 *    Activation dtype is MSX_F32 or MSX_BF16; parameters, statistics, losses and all gradients of
 *    parameters are fp32; labels are int64 (the reference casts masks to torch.long,
 *    core/seg_trainer.py:36); logits at the loss/metric boundary are NCHW fp32 as in the reference.
 */
#ifndef MSX_H_
#define MSX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum msx_status {
    MSX_OK = 0,
    MSX_ERR_INVALID = -1,      /* bad argument (null pointer, misaligned ld, ...) */
    MSX_ERR_CUDA = -2,         /* a CUDA runtime / driver call failed              */
    MSX_ERR_UNSUPPORTED = -3,  /* shape / mode this build has no kernel for        */
    MSX_ERR_WORKSPACE = -4     /* caller's workspace is too small                  */
} msx_status;

enum { MSX_F32 = 0, MSX_BF16 = 1 };
enum { MSX_ENGINE_AUTO = 0, MSX_ENGINE_SIMT = 1, MSX_ENGINE_TCGEN05 = 2 };
enum { MSX_PACK_FPROP = 0, MSX_PACK_DGRAD = 1 };
enum { MSX_REDUCE_MEAN = 0, MSX_REDUCE_SUM = 1, MSX_REDUCE_NONE = 2 };
/* `relu` argument of the BatchNorm BACKWARD entry points is a bit set: bit 0 = a ReLU followed the BN; bit 1 = the
 * forward ran in eval mode (running statistics are constants: dx = gamma * invstd * g, no batch-statistics terms). */
enum { MSX_BN_RELU = 1, MSX_BN_EVAL = 2 };

const char* msx_last_error(void);
int msx_version(void);
/* which conv engine the last conv call on this thread used (MSX_ENGINE_SIMT / _TCGEN05). */
int msx_last_conv_engine(void);

/* Launch accounting and live per-kernel-class timing (measurement support for bench.py; no
 * counterpart in the reference).  msx_launch_count: kernels launched by this library so far.
 * With profiling enabled every launch is bracketed by CUDA events on its own stream and books its
 * ALGORITHMIC bytes / flops; msx_profile_collect synchronises, returns per-class totals
 * (arrays of msx_profile_num_classes() entries) and clears the records. */
long long msx_launch_count(void);
int msx_profile_enable(int on);
int msx_profile_num_classes(void);
const char* msx_profile_class_name(int cls);
int msx_profile_collect(double* host_ms, long long* host_launches, double* host_bytes, double* host_flops);

/* ------------------------------------------------------------------------------------------
 * Convolution geometry.  Describes the MODULE (nn.Conv2d or nn.ConvTranspose2d):
 *   x : [n, h, w, cin]  (pixel stride ldx)      y : [n, oh, ow, cout] (pixel stride ldy)
 * Conv2d:          oh = (h + 2*pad_h - dil*(kh-1) - 1)/stride + 1         weight [cout,cin,kh,kw]
 * ConvTranspose2d: oh = (h-1)*stride - 2*pad_h + dil*(kh-1) + out_pad + 1 weight [cin,cout,kh,kw]
 * Replaces nn.Conv2d / nn.ConvTranspose2d as built by models/modules.py:7-15, :73-85, :89-108.
 * ------------------------------------------------------------------------------------------ */
typedef struct msx_conv_desc {
    int32_t n, h, w, cin;
    int32_t oh, ow, cout;
    int32_t kh, kw, stride, dil, pad_h, pad_w;
    int32_t ldx, ldy;
    int32_t transposed;   /* 0 = Conv2d, 1 = ConvTranspose2d */
    int32_t dtype;        /* MSX_F32 | MSX_BF16 (activations and packed weights) */
    int32_t engine;       /* MSX_ENGINE_*: AUTO picks tcgen05 for bf16 shapes it supports */
} msx_conv_desc;

/* workspace / scratch sizes (bytes) */
size_t msx_conv_packed_bytes(const msx_conv_desc* d);           /* one packed copy of the weight */
size_t msx_ws_bytes_conv(const msx_conv_desc* d);               /* wgrad split-K partials + BN partials */
size_t msx_ws_bytes_bn(int32_t channels);                       /* BN / reduction partials */
size_t msx_ws_bytes_loss(void);
size_t msx_ws_bytes_ce(int64_t pixels);   /* msx_ws_bytes_loss() + room for one-byte label codes (see msx_ce_loss_fwd_bwd) */

/* Pack fp32 master weights (reference layout) into the [tap][N][K] operand of the given pass. */
int msx_conv_pack_weights(const msx_conv_desc* d, int mode, const float* w_master, void* packed, void* stream);

/* Pack plan: every conv weight of a model (both passes) repacked by ONE launch per step instead of one launch
 * per conv and pass.  No counterpart in the reference (its convs read the master weights directly); the
 * callers are models/unet.py:30-41 / models/ducknet.py:38-50 forward, once before the first layer.
 * An item names a master weight [cout,cin,kh,kw] (Conv2d) or [cin,cout,kh,kw] (ConvTranspose2d, transposed=1),
 * the pass (MSX_PACK_FPROP / MSX_PACK_DGRAD) and the destination (>= taps*round16(N)*round8(K) elements, N/K the
 * operand dims of that pass).  build copies a device table (synchronously: start-up work, not to be called
 * under stream capture); run launches the kernel.  Convs then take w_master = NULL and the pre-packed
 * operand as w_packed in msx_conv2d_fwd/bwd and msx_conv_bn_act_fwd/bwd. */
typedef struct msx_pack_item {
    const float* w_master;
    void* packed;
    int32_t cout, cin, kh, kw, transposed, mode;
} msx_pack_item;
size_t msx_conv_pack_plan_bytes(const msx_pack_item* host_items, int32_t count);
int msx_conv_pack_plan_build(const msx_pack_item* host_items, int32_t count, void* table, size_t table_bytes, int32_t* chunks_out);
int msx_conv_pack_plan_run(int dtype, const void* table, int32_t count, int32_t chunks, void* stream);

/* Plain convolution (no BN): ResidualBlock.upper_branch conv1x1, models/ducknet.py:93,104.
 * bias may be NULL.  fwd: y = op(x, W) (+bias).  bwd: dx (nullable), dw (fp32, reference layout),
 * dbias (nullable). w_packed: scratch of msx_conv_packed_bytes; with w_master == NULL it must already hold the
 * operand of the pass that reads it (fwd: MSX_PACK_FPROP, bwd: MSX_PACK_DGRAD), see msx_conv_pack_plan_*. */
int msx_conv2d_fwd(const msx_conv_desc* d, const void* x, const float* w_master, const float* bias,
                   void* w_packed, void* y, void* stream);
int msx_conv2d_bwd(const msx_conv_desc* d, const void* x, const float* w_master, const void* dy,
                   void* w_packed, void* dx, float* dw, float* dbias,
                   void* ws, size_t ws_bytes, void* stream);

/* ConvBNAct / DeConvBNAct forward (models/modules.py:73-85, :89-108):
 *   y_raw = op(x, W) (+bias);  training: batch statistics (biased var for normalisation, unbiased
 *   for running_var, eps, momentum) else running statistics;  z = relu(gamma*(y-mean)*invstd+beta).
 * save_mean / save_invstd (fp32[round_up(cout,8)]) are kept for backward.  relu: 1/0.
 * z is written with pixel stride ldz (zero-copy concat).
 * Inference (core/seg_trainer.py:97-146 validate, tools/test_speed.py): with training = 0 and y_raw = NULL no raw
 * conv output is produced; the eval-mode BN affine and the ReLU are applied in the conv epilogue from the fp32
 * accumulators (no backward can follow such a call: it needs y_raw). */
int msx_conv_bn_act_fwd(const msx_conv_desc* d, const void* x, const float* w_master, const float* bias,
                        const float* gamma, const float* beta,
                        float* running_mean, float* running_var, int64_t* num_batches_tracked,
                        float eps, float momentum, int training, int relu,
                        void* w_packed, void* y_raw, void* z, int32_t ldz,
                        float* save_mean, float* save_invstd,
                        void* ws, size_t ws_bytes, void* stream);

/* Backward of the above: from dz produces dx (nullable), dw, dbias (nullable), dgamma, dbeta.
 * dy_scratch: same shape/stride as y_raw.  The ReLU mask is recomputed from y_raw.  `relu` is a MSX_BN_* bit set
 * (MSX_BN_EVAL when the forward used running statistics - frozen-BN fine-tuning, input saliency). */
int msx_conv_bn_act_bwd(const msx_conv_desc* d, const void* x, const float* w_master,
                        const void* y_raw, const void* dz, int32_t lddz,
                        const float* gamma, const float* beta,
                        const float* save_mean, const float* save_invstd, int relu,
                        void* w_packed, void* dy_scratch, void* dx, float* dw, float* dbias,
                        float* dgamma, float* dbeta,
                        void* ws, size_t ws_bytes, void* stream);

/* Stand-alone (sum of inputs) -> BatchNorm -> ReLU:  DUCK.in_bn / out_bn and ResidualBlock.bn
 * (models/ducknet.py:99,107-108,118-121,137-139,151-152), the stage adds (models/ducknet.py:39-43)
 * and UpsampleBlock's nearest-2x + in-place add (models/ducknet.py:80-85).
 *   s = sum_i xs[i]   (if up2x_first, xs[0] is [n, h/2, w/2, C] and is nearest-upsampled 2x)
 *   z = relu(bn(s)).  With n_in == 1 and no upsample, `s` may be NULL (xs[0] is used directly).
 * host_xs / host_ldxs are HOST arrays of n_in (<= 8) device pointers / pixel strides. */
int msx_sum_bn_act_fwd(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs,
                       int up2x_first, int32_t n, int32_t h, int32_t w, int32_t channels,
                       void* s, int32_t lds,
                       const float* gamma, const float* beta,
                       float* running_mean, float* running_var, int64_t* num_batches_tracked,
                       float eps, float momentum, int training, int relu,
                       void* z, int32_t ldz, float* save_mean, float* save_invstd,
                       void* ws, size_t ws_bytes, void* stream);
/* ds = d(loss)/d(s) given dz; also dgamma, dbeta.  `relu`: MSX_BN_* bit set. */
int msx_bn_act_bwd(int dtype, const void* s, int32_t lds, const void* dz, int32_t lddz,
                   int64_t pixels, int32_t channels,
                   const float* gamma, const float* beta, const float* save_mean, const float* save_invstd,
                   int relu, void* ds, int32_t ldds, float* dgamma, float* dbeta,
                   void* ws, size_t ws_bytes, void* stream);
/* SyncBatchNorm (utils/parallel.py:37-38: nn.SyncBatchNorm.convert_sync_batchnorm, the reference's DEFAULT synBN=True;
 * torch/nn/modules/_functions.py SyncBatchNorm): the per-rank halves of a BatchNorm whose statistics span all ranks.  The
 * caller all-reduces (sum) the small `sums` buffer between the halves - double[2*cp + 1], cp = round_up(channels, 8):
 * per-channel sum, per-channel sum of squares, then this rank's sample count.
 *   forward : msx_bn_stats_sums -> all-reduce -> msx_bn_finalize_sums (mean / invstd over ALL ranks' samples, running
 *             statistics with the global unbiased variance, num_batches_tracked) -> msx_bn_apply
 *   backward: msx_bn_bwd_reduce_sums (dgamma / dbeta stay per-rank, as in torch; sums = [sum g, sum g*xhat]) -> all-reduce
 *             -> msx_bn_bwd_apply_sums (count_dev: the all-reduced forward buffer's last element)
 * ws: msx_ws_bytes_bn(channels). */
int msx_bn_stats_sums(int dtype, const void* x, int32_t ldx, int64_t pixels, int32_t channels, double* sums,
                      void* ws, size_t ws_bytes, void* stream);
int msx_bn_finalize_sums(const double* sums, int32_t channels, float eps, float momentum, float* save_mean, float* save_invstd,
                         float* running_mean, float* running_var, int64_t* num_batches_tracked, void* stream);
int msx_bn_apply(int dtype, const void* x, int32_t ldx, void* z, int32_t ldz, int64_t pixels, int32_t channels,
                 const float* mean, const float* invstd, const float* gamma, const float* beta, int relu, void* stream);
int msx_bn_bwd_reduce_sums(int dtype, const void* x, int32_t ldx, const void* dz, int32_t lddz, int64_t pixels, int32_t channels,
                           const float* gamma, const float* beta, const float* save_mean, const float* save_invstd, int relu,
                           float* dgamma, float* dbeta, double* sums, void* ws, size_t ws_bytes, void* stream);
int msx_bn_bwd_apply_sums(int dtype, const void* x, int32_t ldx, const void* dz, int32_t lddz, int64_t pixels, int32_t channels,
                          const float* gamma, const float* beta, const float* save_mean, const float* save_invstd, int relu,
                          const double* sums, const double* count_dev, void* dx, int32_t lddx,
                          void* ws, size_t ws_bytes, void* stream);

/* out = sum_i xs[i] (n_in <= 8): gradient accumulation where one activation feeds several
 * consumers (DUCK.in_bn output feeds nine convolutions, models/ducknet.py:144-149). */
int msx_sum_n(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs,
              int64_t pixels, int32_t channels, void* out, int32_t ldout, void* stream);
/* Same computation and results as msx_bn_act_bwd, for callers that cannot allocate device memory or
 * pass a workspace (engines with a fixed memory plan, deployments where cudaMalloc after start-up is
 * not allowed, plug-ins that only see their input/output buffers).  It has no counterpart in the
 * reference.  The reduction scratch is a fixed array inside the library image, so:
 *   - channels must be <= MSX_BN_STATIC_MAX_CHANNELS, otherwise MSX_ERR_UNSUPPORTED is returned;
 *   - each GPU that loads libmsx.so holds about 2.4 MB for it until the process exits;
 *   - the device the buffers live on must be the current device;
 *   - the first call on each device looks the scratch address up.  That call allocates nothing itself, but under
 *     CUDA's default lazy module loading it may be the moment the driver loads this library's device image and
 *     reserves the array's 2.4 MB - so make that first call at start-up, before any stream capture;
 *   - the address is cached per device for the life of the process: calls after cudaDeviceReset are unsupported;
 *   - every call uses the same scratch: two calls that can run at the same time on one GPU (different
 *     streams, or different host threads on different streams) race on it - keep them on one stream or
 *     order them yourself.  Nothing has to be cleared between calls.
 * The function performs no cudaMalloc / cudaFree and calls no allocating library routine (leave the
 * optional msx_profile_enable() measurement mode off: it creates CUDA events). */
#define MSX_BN_STATIC_MAX_CHANNELS 256
int msx_bn_act_bwd_static(int dtype, const void* s, int32_t lds, const void* dz, int32_t lddz,
                          int64_t pixels, int32_t channels,
                          const float* gamma, const float* beta, const float* save_mean, const float* save_invstd,
                          int relu, void* ds, int32_t ldds, float* dgamma, float* dbeta, void* stream);

/* Forward-only helpers of the distillation teacher (smp.Unet over a torchvision ResNet, un-vendored dependencies
 * segmentation-models-pytorch 0.3.2 / torchvision; call site models/__init__.py:43-58, core/seg_trainer.py:70-72):
 * out = [relu](sum_i xs[i]) with xs[0] optionally nearest-2x upsampled first.  n_in = 2, relu = 1 is a residual
 * block's `relu(out + identity)`; n_in = 1, up2x_first = 1 is the decoder's F.interpolate(scale_factor=2, 'nearest')
 * written straight into its slice of the concat buffer (ldout). */
int msx_sum_act(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs, int up2x_first,
                int32_t n, int32_t h, int32_t w, int32_t channels, int relu, void* out, int32_t ldout, void* stream);

/* backward of the nearest-2x upsample: da[n,y,x,:] = sum of the 2x2 block of ds. */
int msx_upsample2x_bwd(int dtype, const void* ds, int32_t ldds, int32_t n, int32_t h, int32_t w,
                       int32_t channels, void* da, int32_t ldda, void* stream);

/* nn.MaxPool2d(3, 2, 1), models/unet.py:49.  idx: int8 [n, oh, ow, ldidx] window position of the
 * (first) maximum, consumed by the backward. */
int msx_maxpool3s2_fwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                       void* z, int32_t ldz, int8_t* idx, int32_t ldidx, void* stream);
int msx_maxpool3s2_bwd(int dtype, const void* dz, int32_t lddz, const int8_t* idx, int32_t ldidx,
                       int32_t n, int32_t h, int32_t w, int32_t channels, void* dx, int32_t lddx, void* stream);

/* Layout boundary: images arrive NCHW fp32 (core/seg_trainer.py:35). */
int msx_nchw_to_nhwc(int dtype, const float* src, int32_t n, int32_t c, int32_t h, int32_t w,
                     void* dst, int32_t ld, void* stream);
int msx_nhwc_to_nchw(int dtype, const void* src, int32_t ld, int32_t n, int32_t c, int32_t h, int32_t w,
                     float* dst, void* stream);

/* seg_head = conv1x1(base_channel, num_class), bias-free (models/unet.py:27,41; models/ducknet.py:35,50).
 * x NHWC -> logits NCHW fp32 [n, num_class, h, w]; w_master [num_class, channels] fp32. */
int msx_head_fwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                 const float* w_master, int32_t num_class, float* logits, void* stream);
int msx_head_bwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                 const float* w_master, int32_t num_class, const float* dlogits,
                 void* dx, int32_t lddx, float* dw, void* ws, size_t ws_bytes, void* stream);

/* nn.CrossEntropyLoss(ignore_index, reduction, weight) forward AND gradient in one pass
 * (core/loss.py:29-31; called at core/seg_trainer.py:63).  logits [n,c,h*w] fp32 NCHW, labels int64.
 * loss_out: 1 float (MEAN / SUM).  dlogits (nullable) = d loss / d logits for upstream gradient 1.
 * MEAN divides by the summed class weight of non-ignored pixels (NaN if none, like PyTorch).
 * Workspace: msx_ws_bytes_loss() is enough; with msx_ws_bytes_ce(n*hw) (and c <= 255) the weight pre-pass leaves
 * one-byte label codes there and the main pass does not read the int64 labels a second time. */
int msx_ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                        const float* class_weights, int64_t ignore_index, int reduction,
                        float* loss_out, float* dlogits, void* ws, size_t ws_bytes, void* stream);
/* per-pixel CE (reduction='none'): loss_pix [n*hw]; used by OHEM (core/loss.py:11,15). */
int msx_ce_loss_none_fwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                         const float* class_weights, int64_t ignore_index, float* loss_pix, void* stream);
/* dlogits = coef[pix] * w[t] * (softmax - onehot) with per-pixel upstream coefficient. */
int msx_ce_loss_none_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                         const float* class_weights, int64_t ignore_index, const float* coef_pix,
                         float* dlogits, void* stream);
/* OhemCELoss.forward (core/loss.py:13-20) without the two host syncs: threshold branch and
 * top-k(n_min) fallback are both decided on the device.  loss_pix: scratch float[n*hw]. */
int msx_ohem_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                          float thresh_prob, int64_t ignore_index, float* loss_out, float* dlogits,
                          float* loss_pix, void* ws, size_t ws_bytes, void* stream);

/* kd_loss_fn (core/loss.py:42-50).  kind 0: kl_div with temperature T, F.kl_div default 'mean'
 * (over all n*c*h*w elements) times T^2;  kind 1: mse.  dstudent nullable. */
int msx_kd_loss_fwd_bwd(const float* student, const float* teacher, int32_t n, int32_t c, int64_t hw,
                        int kind, float temperature, float* loss_out, float* dstudent,
                        void* ws, size_t ws_bytes, void* stream);

/* sigmoid + Dice + BCE (north_star extension; absent from the reference):
 * loss = mean(softplus(z) - z*t) + 1 - (2*sum(p*t)+eps)/(sum(p)+sum(t)+eps).  logits [pixels] fp32,
 * targets int64 {0,1} (labels_are_float = 0) or float (1). */
int msx_dice_bce_loss_fwd_bwd(const float* logits, const void* targets, int labels_are_float, int64_t pixels,
                              float eps, float* loss_out, float* dlogits,
                              void* ws, size_t ws_bytes, void* stream);

/* dst[i] *= *scale_dev (device scalar): applies a non-unit upstream gradient to a saved dlogits. */
int msx_scale_by_device_scalar(float* data, int64_t count, const float* scale_dev, void* stream);

/* torchmetrics JaccardIndex / Dice update (utils/metrics.py:6-9, core/seg_trainer.py:119):
 * argmax over classes (first max wins, NaN is maximal) and confmat[target][pred] += 1, int64, on
 * device, no host sync.  ignore_index < 0: nothing ignored.  A label outside [0,c) that is not
 * ignored sets *bad_label_flag to 1 (checked at compute()). */
int msx_confmat_update(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                       int64_t ignore_index, int64_t* confmat, int32_t* bad_label_flag, void* stream);

/* The evaluation sweep without logits in HBM (core/seg_trainer.py:114-119: preds = ema_model.ema(images) followed by
 * metric.update(preds, masks); utils/metrics.py:4-12): seg_head 1x1 conv + argmax + confusion-matrix count in one pass
 * over the NHWC features x [pixels, channels].  Counts are identical to msx_head_fwd followed by msx_confmat_update
 * (same fp32 arithmetic in the same order, same tie / NaN rule).  labels int64 [pixels] (= the [n, h, w] mask). */
int msx_head_confmat_update(int dtype, const void* x, int32_t ldx, int64_t pixels, int32_t channels, const float* w_master,
                            int32_t num_class, const int64_t* labels, int64_t ignore_index, int64_t* confmat,
                            int32_t* bad_label_flag, void* stream);

/* Multi-tensor optimizer / EMA steps over one flat fp32 parameter arena
 * (utils/optimizer.py:8-16 -> torch.optim.Adam/AdamW/SGD; utils/model_ema.py:29-40).
 * grad_scale multiplies the gradient first (1/world_size after the DDP all-reduce).
 * step_dev: device int64 step counter, incremented by the Adam kernel.
 * hyper_dev (nullable): device float[4] = {lr, beta1 | momentum, beta2, weight_decay} that OVERRIDES the scalar
 * arguments.  The reference steps OneCycleLR every iteration (utils/scheduler.py:12-19, core/seg_trainer.py:85),
 * which also cycles momentum / beta1; scalars baked into a captured CUDA graph could not follow it, a device
 * array refreshed by one small copy per step can. */
int msx_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t count,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled_wd,
                  float grad_scale, int64_t* step_dev, const float* hyper_dev, void* stream);
int msx_sgd_step(float* param, const float* grad, float* momentum_buf, int64_t count,
                 float lr, float momentum, float weight_decay, float grad_scale, int first_step,
                 const float* hyper_dev, void* stream);
/* ModelEmaV2.update (utils/model_ema.py:29-40): e = decay * e + (1 - decay) * m as two rounded products and one
 * rounded sum in fp32 (what the reference's three torch kernels compute; no fused multiply-add), or a plain copy when
 * decay == 0.  The _i64 form is the same rule applied to the int64 num_batches_tracked counters, which the reference
 * also runs through the float expression and truncates back to int64 with copy_ (utils/model_ema.py:31-35). */
int msx_ema_update(float* ema, const float* model, int64_t count, double decay, void* stream);
int msx_ema_update_i64(int64_t* ema, const int64_t* model, int64_t count, double decay, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MSX_H_ */
