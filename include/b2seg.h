/*
 * b2seg.h — C ABI of libb2seg.so, the sm_100a (B200) kernels behind the
 * segmentation training step of SwagMag1213/image-segmentation-project.
 *
 * The reference has no FFI of its own: its hot path sits behind a Python module
 * API.  Each entry point below names the reference call site(s) it replaces
 * (file:line under the reference checkout).  The Python host layer
 * (image-segmentation-project_b200/b2seg/) binds these with ctypes; the stub a
 * reference maintainer would add is shown in INTEGRATION.md.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no torch / C++ types.
 *   - every function returns 0 (B2SEG_OK) or a negative B2SEG_E* code; the
 *     message is available from b2seg_last_error() (thread-local).
 *   - tensor arguments are raw DEVICE pointers into caller-owned memory unless
 *     a parameter is documented as host memory.  The library never allocates
 *     tensors, never synchronises the device implicitly and never takes
 *     ownership.  Scratch comes from caller-provided workspaces.
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued there,
 *     so whole steps are CUDA-graph capturable.
 *   - there is no CPU fallback: without a Blackwell GPU every compute entry
 *     point fails with B2SEG_ECUDA.
 */
#ifndef B2SEG_H_
#define B2SEG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2SEG_ABI_VERSION 3

#define B2SEG_OK 0
#define B2SEG_EINVAL (-1)  /* bad argument / unsupported shape            */
#define B2SEG_ECUDA (-2)   /* CUDA runtime or driver error                 */
#define B2SEG_ENOMEM (-3)  /* workspace too small                          */

int b2seg_abi_version(void);
const char* b2seg_last_error(void);
/* Fills name[len] with the device name, *sms with the SM count, *cc with
 * major*10+minor.  Fails with B2SEG_ECUDA when no CUDA device is usable. */
int b2seg_device_info(char* name, int len, int* sms, int* cc);

/* ======================================================================== *
 *  1. Per-pixel segmentation losses + metrics                               *
 *     replaces losses.py:13-171, 322-342 (Dice, BCE, WeightedBCE,          *
 *     BalancedBCE, Focal, Tversky, FocalTversky, SensitivitySpecificity,   *
 *     LogCoshDice, Combo, TripleCombo), utils.py:26-56 (dice_loss,         *
 *     combo_loss) and utils.py:120-151 (calculate_metrics).                *
 * ======================================================================== */

/* Every closed-form loss of the reference is a weighted sum of these terms;
 * a term with weight 0 costs nothing.  Field meanings follow the reference
 * constructors (losses.py:345-403 for the factory defaults). */
typedef struct b2seg_loss_spec {
  float w_dice;          /* DiceLoss weight;            losses.py:19-28   */
  float w_logcosh_dice;  /* log(cosh(dice_loss)) weight; losses.py:156-158 */
  float dice_smooth;
  float w_tversky;       /* (1 - Tversky)^tv_gamma;     losses.py:97-122  */
  float tv_alpha, tv_beta, tv_smooth, tv_gamma;
  float w_sens_spec;     /* losses.py:132-147 */
  float ss_w, ss_eps;
  float w_bce;           /* losses.py:36-65 */
  int32_t bce_mode;      /* 0 plain mean, 1 weighted (beta), 2 balanced, 3 per-pixel weight map */
  float bce_beta;
  float w_focal;         /* losses.py:76-86 (reduction='mean') */
  float focal_alpha, focal_gamma;
  /* ExponentialLogarithmicLoss, losses.py:174-194: w_explog_dice (-log(1 - dice_loss))^el_gamma_dice
   * + w_explog_cross mean((-log(clamp(1 - bce_i, 1e-7)))^el_gamma_cross).  The per-pixel term shares
   * the focal accumulator: w_focal and w_explog_cross are exclusive. */
  float w_explog_dice, w_explog_cross, el_gamma_dice, el_gamma_cross;
} b2seg_loss_spec;

/* Layout of the device-side result block written by the forward kernels
 * (double[B2SEG_STATS_LEN]).  Sums are over all n pixels of the batch. */
enum {
  B2SEG_S_P = 0,       /* sum sigmoid(x)                          */
  B2SEG_S_T = 1,       /* sum t                                   */
  B2SEG_S_PT = 2,      /* sum sigmoid(x)*t                        */
  B2SEG_S_BCE = 3,     /* sum bce_i (times weight map in mode 3)  */
  B2SEG_S_BCE_T = 4,   /* sum bce_i*t                             */
  B2SEG_S_FOCAL = 5,   /* sum (1-pt)^gamma * bce_i, or sum of the exp-log pixel term */
  B2SEG_S_PB = 6,      /* sum [sigmoid(x) > 0.5]                  */
  B2SEG_S_PB_T = 7,    /* sum [sigmoid(x) > 0.5]*t                */
  B2SEG_S_N = 8,       /* n                                       */
  B2SEG_S_LOSS = 9,    /* loss value                              */
  B2SEG_S_TP = 10, B2SEG_S_FP = 11, B2SEG_S_FN = 12, B2SEG_S_TN = 13,
  B2SEG_STATS_LEN = 16
};
#define B2SEG_LOSS_COEF_LEN 8

/* Scratch for the two-level reduction (must be zero-filled ONCE by the caller
 * before first use; the kernels leave it zeroed). */
size_t b2seg_seg_reduce_workspace_bytes(void);

/* Forward: one pass over logits+targets (8 B/px algorithmic).  Writes
 * stats[16] (double), loss_out[1] (float) and coef[8] (float; consumed by
 * b2seg_seg_loss_bwd without a host round trip).  weight_map may be NULL
 * unless spec->bce_mode == 3. */
int b2seg_seg_loss_fwd(const float* logits, const float* target, const float* weight_map,
                       int64_t n, const b2seg_loss_spec* spec /*host*/, void* workspace,
                       double* stats, float* loss_out, float* coef, void* stream);

/* Backward: grad_logits[i] = grad_out * dL/dlogits[i] (12 B/px algorithmic).
 * grad_out is a DEVICE scalar (NULL means 1). */
int b2seg_seg_loss_bwd(const float* logits, const float* target, const float* weight_map,
                       int64_t n, const float* coef, const float* grad_out, float* grad_logits,
                       void* stream);

/* Second half of b2seg_seg_loss_fwd as its own launch: loss, gradient coefficients and the
 * confusion block from the eight sums in stats[0..7] over n_total pixels.  Data-parallel replicas
 * all-reduce stats[0..8] (sums and pixel count) after their local forward and call this with
 * n_total = 0 (= take the count from stats[B2SEG_S_N]; no host round trip), so
 * Dice / Tversky / balanced-BCE terms see the whole batch exactly as the single-process
 * reference does (losses.py:21-27, 99-107). */
int b2seg_seg_loss_finish(const b2seg_loss_spec* spec, int64_t n_total, double* stats /*in/out [16]*/,
                          float* loss_out, float* coef /*[8]*/, void* stream);

/* calculate_metrics(pred, target) of utils.py:120-151: `probs` are
 * probabilities; fills stats[S_PB, S_PB_T, S_T, S_N, S_TP..S_TN]. */
int b2seg_seg_metrics_from_probs(const float* probs, const float* target, int64_t n,
                                 void* workspace, double* stats, void* stream);

/* Allocation-free variant of b2seg_seg_metrics_from_probs (same kernel, same grid, same result).
 * NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The reduction scratch is a file-scope static __device__ array inside the library
 * (75 792 B, held by every GPU that loads the library until exit); the call performs no
 * cudaMalloc / cudaFree and calls no allocating library routine.  The caller still owns probs,
 * target and stats[16] (device).  n > B2SEG_NOALLOC_MAX_PIXELS returns B2SEG_EINVAL.
 * All calls on a device share the one array: the entry point clears the arrival ticket on
 * `stream` before each launch, and two calls that OVERLAP on the GPU (different streams with no
 * ordering between them) race on it -- serialise them, or use the workspace variant above.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 * that device current); make that first call outside stream capture. */
#define B2SEG_NOALLOC_MAX_PIXELS (1LL << 28) /* 268 435 456 px = 4 096 masks of 256 x 256 */
int b2seg_seg_metrics_from_probs_noalloc(const float* probs, const float* target, int64_t n,
                                         double* stats, void* stream);

/* Weight maps of the boundary-style losses: they depend on the target only, are built on the
 * device and go into b2seg_seg_loss_fwd/_bwd as `weight_map` (bce_mode 3).
 *
 * BoundaryLoss (losses.py:283-301): weight = 1 + gain * conv2d(dilate(t) - erode(t), gauss) with
 * max_pool2d(theta0, stride 1, pad theta0/2) as dilation / erosion (window clipped at the image
 * edge) and a theta x theta kernel `gauss` (HOST, row-major; BoundaryLoss._gaussian_kernel,
 * losses.py:309-318) applied with zero padding.  theta0, theta odd, <= 15.  One launch, 8 B/px. */
int b2seg_loss_boundary_weight_map(const float* target, float* weight, int planes, int h, int w,
                                   int theta0, int theta, const float* gauss /*host*/, float gain,
                                   void* stream);

/* scipy.ndimage.distance_transform_edt on `planes` float32 planes (non-zero = foreground): exact
 * Euclidean distance to the nearest zero pixel; a plane without zeros gets scipy's distances to
 * (row -1, column 0).  mode 0: out = float32(distance)  (HausdorffDistanceLoss._compute_dtm,
 * losses.py:262-273).  mode 1: out = 1 + alpha * float32(d / (max_plane(d) + 1e-7))
 * (DistanceMapPenalizedCrossEntropy, losses.py:203-232).  Bit-identical to the scipy/numpy path. */
size_t b2seg_loss_edt_workspace_bytes(int planes, int h, int w);
int b2seg_loss_edt_f32(const float* src, float* out, int planes, int h, int w, int mode,
                       float alpha, void* workspace, void* stream);

/* ======================================================================== *
 *  2. Geometric augmentation gather                                         *
 *     replaces the cv2 calls albumentations bottoms out in at               *
 *     dataset.py:148-154,184 and forward_selection_integration.py:73-105.   *
 * ======================================================================== */
enum {
  B2SEG_AUG_HFLIP = 0,     /* cv2.flip(img, 1)            */
  B2SEG_AUG_VFLIP = 1,     /* cv2.flip(img, 0)            */
  B2SEG_AUG_ROT90 = 2,     /* np.rot90(img, k), k in arg  */
  B2SEG_AUG_TRANSPOSE = 3, /* img.T                       */
};
enum { B2SEG_INTERP_NEAREST = 0, B2SEG_INTERP_LINEAR = 1 };
enum { B2SEG_BORDER_CONSTANT = 0, B2SEG_BORDER_REFLECT = 1, B2SEG_BORDER_REFLECT101 = 2 };

/* Exact index permutation of `planes` uint8 planes of size h x w (image and
 * mask stacked).  Output plane size is (w x h) for odd rot90 k and transpose. */
int b2seg_aug_permute_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w, int mode,
                         int k, void* stream);

/* cv2.warpAffine(src, M, (w,h), flags=interp|WARP_INVERSE_MAP-free form):
 * `minv` (host, 6 doubles) is the already-inverted 2x3 matrix mapping
 * destination (x,y) to source coordinates, exactly what cv2 computes from M.
 * Bilinear uses cv2's fixed-point scheme (5-bit coordinate fractions, 15-bit
 * weights) so uint8 results match cv2 bit-for-bit away from rounding ties. */
int b2seg_aug_warp_affine_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                             const double* minv /*host[6]*/, int interp, int border,
                             int border_value, void* stream);

/* cv2.remap(src, map_x, map_y, interp, border): float32 absolute maps (device). */
int b2seg_aug_remap_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w,
                       const float* map_x, const float* map_y, int interp, int border,
                       int border_value, void* stream);

/* One entry of the device table the batched warps read: the dst->src matrix of the item (already
 * inverted, as cv2.warpAffine does; unused by remap), and per PLANE of the item (image, mask, ...)
 * the interpolation and the constant-border value; one border rule per item. */
typedef struct b2seg_warp_item {
  double minv[6];
  int32_t interp[4];
  int32_t cval[4];
  int32_t border;
  int32_t pad;
} b2seg_warp_item;

/* All augmented copies of a batch in ONE launch: planes are [n_items][ppi][h][w] (ppi <= 4, e.g.
 * image + mask), `items` a DEVICE table of n_items entries.  Arithmetic per plane is exactly that
 * of b2seg_aug_warp_affine_u8 / b2seg_aug_remap_u8 (replaces the per-image A.Affine /
 * A.GridDistortion calls of dataset.py:184 and forward_selection_integration.py:295).
 * remap: item i reads map_x + i*map_stride (map_stride 0 = one shared pair of maps). */
int b2seg_aug_warp_affine_batch_u8(const uint8_t* src, uint8_t* dst, int n_items, int ppi, int h, int w,
                                   const b2seg_warp_item* items /*device*/, void* stream);
int b2seg_aug_remap_batch_u8(const uint8_t* src, uint8_t* dst, int n_items, int ppi, int h, int w,
                             const float* map_x, const float* map_y, int64_t map_stride,
                             const b2seg_warp_item* items /*device*/, void* stream);

/* Non-geometric augmentations on uint8 planes (forward_selection_integration.py:107-178).
 * lut:        dst[i] = lut[src[i]] (InvertImg, Solarize, RandomGamma); lut = 256 device bytes.
 * hist:       256-bin histograms of n_images images of n_per_image bytes (also CLAHE / percentiles).
 * affine_lut: dst = uint8(clip(v*factor + mean(src)*mean_coef)) with the mean taken from `hist`
 *             (ColorJitter brightness: mean_coef 0, hist may be NULL; contrast: mean_coef 1-factor).
 * fill_rects: in place, rects = device int32 [n][4] = (y1, x1, y2, x2) (CoarseDropout, GridDropout).
 * gauss_noise: dst = uint8(clip(src + mean + sigma*N(0,1))), Philox-4x32-10 keyed by `seed`.
 * filter2d:   cv2.filter2D(src, -1, k x k float32 kernel), BORDER_REFLECT_101 (AdvancedBlur).
 * pad:        cv2.copyMakeBorder (CropAndPad).
 * uniform_f32 / blur1d_f32: U(-1,1) noise and one pass of scipy.ndimage.gaussian_filter
 *             (mode='reflect'); out = gain*blur(in) (+ x if add_coord 1, + y if 2): ElasticTransform maps. */
int b2seg_aug_lut_u8(const uint8_t* src, uint8_t* dst, int64_t n, const uint8_t* lut, void* stream);
int b2seg_aug_hist_u8(const uint8_t* src, int64_t n_per_image, int n_images, uint32_t* hist, void* stream);
int b2seg_aug_affine_lut_u8(const uint8_t* src, uint8_t* dst, int64_t n, const uint32_t* hist, double factor,
                            double mean_coef, void* stream);
int b2seg_aug_fill_rects_u8(uint8_t* plane, int h, int w, const int32_t* rects, int n_rects, int value,
                            void* stream);
int b2seg_aug_gauss_noise_u8(const uint8_t* src, uint8_t* dst, int64_t n, float mean, float sigma, uint64_t seed,
                             void* stream);
int b2seg_aug_filter2d_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w, const float* kernel,
                          int k, void* stream);
int b2seg_aug_pad_u8(const uint8_t* src, uint8_t* dst, int planes, int h, int w, int top, int bottom, int left,
                     int right, int border, int border_value, void* stream);
int b2seg_aug_uniform_f32(float* dst, int64_t n, uint64_t seed, void* stream);
int b2seg_aug_blur1d_f32(const float* src, float* dst, int planes, int h, int w, const float* weights,
                         int radius, int along_x, float gain, int add_coord, void* stream);

/* ======================================================================== *
 *  2b. Optimizer step (replaces torch.optim.Adam.step, train.py:49; the    *
 *      optimizer the reference builds at train.py:331-335 and               *
 *      cross_validation.py:93-97).                                          *
 * ======================================================================== */
/* One chunk of one parameter: n <= 2^31-1 consecutive fp32 elements of the parameter, its
 * gradient and its two moment buffers.  The table lives in DEVICE memory. */
typedef struct b2seg_adam_chunk {
  float* p;
  const float* g;
  float* m; /* exp_avg */
  float* v; /* exp_avg_sq */
  int32_t n;
  int32_t pad;
} b2seg_adam_chunk;

/* Adam (amsgrad = maximize = false) with L2 weight decay on every chunk, one launch of `nchunks`
 * CTAs.  `step` is the 1-based step count t the bias corrections 1 - beta^t use. */
int b2seg_adam_step(const b2seg_adam_chunk* chunks /*device*/, int nchunks, float lr, float beta1,
                    float beta2, float eps, float weight_decay, double step, void* stream);

/* ======================================================================== *
 *  3. Planned ops: created once per (pointers, shapes), replayed per step.  *
 *     TMA descriptors are encoded at creation, so run() is launch-only and  *
 *     CUDA-graph capturable.                                                 *
 * ======================================================================== */
typedef struct b2seg_op b2seg_op;   /* opaque */
typedef struct b2seg_seq b2seg_seq; /* opaque ordered list of ops */

int b2seg_op_run(b2seg_op* op, void* stream);
int b2seg_op_destroy(b2seg_op* op);
int b2seg_op_launches(const b2seg_op* op);
int b2seg_seq_create(b2seg_seq** out);
int b2seg_seq_append(b2seg_seq* seq, b2seg_op* op); /* does not take ownership */
/* lane 0 = the dependent chain (same as b2seg_seq_append).  lane 1 = an op whose outputs no later
 * op of the sequence reads and whose inputs no later op overwrites (weight gradients,
 * convolution_backward's grad_weight half at train.py:48): in the CUDA-graph replay it runs on a
 * side branch forked after the op appended before it and joined at the end of the sequence;
 * lane-1 ops keep their relative order.  Direct (non-graph) replays ignore lanes. */
int b2seg_seq_append_lane(b2seg_seq* seq, b2seg_op* op, int lane);
/* The side branch joins the chain before the next appended op (instead of at the end): for a
 * lane-1 op whose output IS read later, e.g. the weight repack that only the first tensor-core
 * convolution needs while the stem runs. */
int b2seg_seq_join(b2seg_seq* seq);
/* Replays the ops in order.  Sequences of >= 4 ops are captured into a CUDA graph on first use
 * (one graph launch per replay; B2SEG_GRAPH=0 in the environment keeps direct launches). */
int b2seg_seq_run(b2seg_seq* seq, void* stream);
int b2seg_seq_is_graph(const b2seg_seq* seq); /* 1 once the sequence replays as a CUDA graph */
int b2seg_seq_launches(const b2seg_seq* seq);
/* Replays the sequence with a CUDA event pair around every op, synchronises the stream and
 * writes each op's elapsed milliseconds to ms_per_op[0..n_ops) (profiling aid for bench.py's
 * per-kernel roofline; not for the timed path). */
int b2seg_seq_run_timed(b2seg_seq* seq, void* stream, float* ms_per_op /*host*/, int n_ops);
int b2seg_seq_destroy(b2seg_seq* seq);

/* NHWC bf16 activation view; strides in ELEMENTS, channel stride is 1.  A
 * channel slice of a wider buffer (fused skip-concat) or a stride-2 phase of a
 * feature map is just another view. */
typedef struct b2seg_tensor4 {
  void* ptr;
  int32_t n, h, w, c;
  int64_t sn, sh, sw;
} b2seg_tensor4;

/* One filter tap of a tap-gather GEMM: which source view it reads, the pixel
 * offset applied to the output pixel coordinate, and which weight slab. */
typedef struct b2seg_tap {
  int8_t src, dh, dw, wtap;
} b2seg_tap;

#define B2SEG_MAX_TAPS 9
#define B2SEG_MAX_SUB 4

/* Tap-gather implicit GEMM on tcgen05/TMEM, fed by TMA:
 *   out[pix, col] (+)= sum_taps sum_c src[tap.src][pix + (tap.dh,tap.dw), c] * W[tap.wtap][col][c]
 * replaces torch conv2d / conv_transpose2d and their data gradients:
 *   advanced_models.py:76,90-100,158-160,199,202 (Conv2d 3x3/1x1, ConvTranspose2d 2x2 s2),
 *   torchvision resnet BasicBlock/Bottleneck convs (advanced_models.py:81-87,111-117),
 *   AttentionGate 1x1 convs (advanced_models.py:11,16) and, with repacked weights,
 *   the dgrad half of convolution_backward triggered at train.py:48.
 * Stride-2 convs read four phase views; ConvTranspose writes through the
 * pixel-shuffle addressing (cq, oqh, oqw) straight into a channel slice of the
 * decoder's concat buffer (advanced_models.py:288-333 torch.cat fused away). */
typedef struct b2seg_tapgemm_desc {
  b2seg_tensor4 src[4];
  int32_t nsrc;
  int32_t nsub;                 /* independent sub-problems (e.g. 4 dgrad phases) */
  struct {
    int32_t ntaps;
    b2seg_tap taps[B2SEG_MAX_TAPS];
    int64_t out_offset;         /* elements added to `out` for this sub-problem */
  } sub[B2SEG_MAX_SUB];
  const void* weights;          /* bf16 [wtaps][ngemm][cin] */
  int32_t wtaps, ngemm, cin;
  int32_t on, oh, ow;           /* iteration (output pixel) grid */
  void* out;                    /* bf16 */
  int64_t osn, osh, osw;        /* output strides, elements */
  int32_t cq;                   /* columns per pixel-shuffle group (power of 2); 0 = none */
  int64_t oqh, oqw;             /* extra offset for group q: (q>>1)*oqh + (q&1)*oqw */
  const float* bias;            /* fp32 [cq ? cq : ngemm] or NULL */
  float* stats;                 /* fp32 [stat_rows][2][ngemm] partial (sum, sumsq) or NULL */
  int32_t accumulate;           /* out += result instead of out = result */
  int32_t stats_fixed;          /* 1: `stats` is int64 [2][ngemm] instead -- every CTA adds its partial sums as
                                   fixed-point integers (sum x 2^20, sumsq x 2^16) with 64-bit atomics: exact,
                                   order-independent, one row; the caller zeroes it before the launch and the
                                   BatchNorm consumer reads it with rows = -1 (B2SEG_EW_BN_APPLY_FUSED) */
} b2seg_tapgemm_desc;
#define B2SEG_STATS_FIX_SUM 1048576.0f   /* 2^20 */
#define B2SEG_STATS_FIX_SQ 65536.0f      /* 2^16 */
#define B2SEG_BWD_FIX_SCALE 1099511627776.0 /* 2^40: BatchNorm-backward sums (sum dz, sum dz*y) as fixed point */

int b2seg_tapgemm_create(const b2seg_tapgemm_desc* desc /*host*/, b2seg_op** out);
/* number of partial-statistics rows a tapgemm op with this iteration grid writes */
int b2seg_tapgemm_stat_rows(int on, int oh, int ow);

/* Weight-gradient GEMM on tcgen05 (both operands MN-major straight from the
 * NHWC tiles TMA lands), split-K over pixel tiles with a deterministic
 * two-phase reduction:
 *   G[tap][i][j] = sum_pix a_src[tap.src][pix + (tap.dh,tap.dw), i] * b_src[pix, j]
 *   grad[tap.out_base + i*row_stride + j*col_stride] (+)= G[tap][i][j]   (fp32)
 * replaces the wgrad half of convolution_backward for Conv2d (a = layer input,
 * b = dY) and ConvTranspose2d (a = dY phases, b = layer input) reached from
 * loss.backward() at train.py:48.
 * A 3x3 stride-1 problem (one source, the nine offsets -1..1) whose b side has at most 64
 * channels is planned on a halo-reuse kernel that reads all nine taps out of one input brick
 * per pixel tile; the result is the same, and b2seg_wgrad_workspace_bytes() already accounts
 * for whichever kernel the planner picks. */
typedef struct b2seg_wgrad_desc {
  b2seg_tensor4 a_src[4];
  int32_t na_src;
  b2seg_tensor4 b_src;
  int32_t ntaps;
  struct {
    int8_t src, dh, dw, pad;
    int64_t out_base;           /* fp32 elements into `grad` */
  } taps[B2SEG_MAX_TAPS];
  int32_t pn, ph, pw;           /* pixel grid reduced over (K dimension) */
  float* grad;                  /* fp32 destination (e.g. an OIHW parameter gradient) */
  int64_t row_stride, col_stride;
  int32_t accumulate;           /* grad += instead of grad = */
  void* workspace;              /* fp32 split-K partials, >= b2seg_wgrad_workspace_bytes() */
  size_t workspace_bytes;
} b2seg_wgrad_desc;

size_t b2seg_wgrad_workspace_bytes(int a_channels, int b_channels, int ntaps, int pn, int ph,
                                   int pw);
int b2seg_wgrad_create(const b2seg_wgrad_desc* desc /*host*/, b2seg_op** out);

/* ======================================================================== *
 *  3b. Input pipeline of CellSegmentationDataset.__getitem__ (dataset.py:30-66):   *
 *      cv2.resize (INTER_AREA / INTER_NEAREST / INTER_LINEAR), np.percentile(2, 98) clip,       *
 *      CLAHE(2.0, 8 x 8), min-max -- batched uint8 planes in, fp32 out, bit-exact vs cv2.       *
 *      Coefficient tables are device arrays built by the caller once per size pair              *
 *      (resize.cpp's own table formulas; b2seg/preproc.py, oracle/preproc_cv.py).               *
 * ======================================================================== */
/* nearest: xofs[dw], yofs[dh] source indices; dst u8 and/or dst_binary_f32 = (v > 0) (the mask path, dataset.py:47,58) */
int b2seg_resize_nearest_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, float* dst_binary_f32,
                            int dh, int dw, const int32_t* xofs, const int32_t* yofs, void* stream);
/* linear: xofs[dw] + int16 xcoef[dw][2] (x 2048), same for y (HResizeLinear / VResizeLinear fixed point) */
int b2seg_resize_linear_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh, int dw,
                           const int32_t* xofs, const int16_t* xcoef, const int32_t* yofs, const int16_t* ycoef,
                           void* stream);
/* area, integer ratios (ResizeAreaFast) */
int b2seg_resize_areafast_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh, int dw,
                             void* stream);
/* area, fractional ratios (ResizeArea): CSR tables xstart[dw+1] -> (xsi, xalpha), ystart[dh+1] -> (ysi, ybeta) */
int b2seg_resize_area_u8(const uint8_t* src, int planes, int sh, int sw, uint8_t* dst, int dh, int dw,
                         const int32_t* xstart, const int32_t* xsi, const float* xalpha, const int32_t* ystart,
                         const int32_t* ysi, const float* ybeta, void* stream);
/* lut[i][v] = uint8(clip(v, *np.percentile(image_i, [q_lo, q_hi]))) from 256-bin histograms (b2seg_aug_hist_u8);
 * `percentiles` (double [n][2]) optional */
int b2seg_pre_percentile_lut(const uint32_t* hist, int n_images, int64_t pixels_per_image, double q_lo, double q_hi,
                             uint8_t* lut, double* percentiles, void* stream);
size_t b2seg_pre_clahe_workspace_bytes(int n_images, int tiles_x, int tiles_y);
/* clipped = clip_lut[src]; eq = cv2.createCLAHE(clip_limit, (tiles_x, tiles_y)).apply(clipped);
 * out = (eq - min) / (max - min + 1e-8) as fp32 (four launches) */
int b2seg_pre_clahe_normalize(const uint8_t* src, int n_images, int h, int w, const uint8_t* clip_lut,
                              double clip_limit, int tiles_x, int tiles_y, uint8_t* clipped, uint8_t* eq,
                              float* out, void* workspace, size_t workspace_bytes, void* stream);
/* normalize=False path (dataset.py:55): out = v / 255 */
int b2seg_pre_scale255_f32(const uint8_t* src, int n_images, int64_t pixels_per_image, float* out, void* stream);

/* ======================================================================== *
 *  4. HBM-bound layer kernels (BatchNorm, ReLU, pooling, stem, 1-channel    *
 *     convs, attention gate, squeeze-excite, weight packing) as planned ops.*
 *     Activations are NHWC bf16 seen as [P pixels][C] with a leading         *
 *     dimension `ld` in elements (channel slices of the concat buffers work  *
 *     in place).  One generic descriptor: `kind` selects the kernel; slot    *
 *     meanings are listed per kind (p = device pointers, i = integers,       *
 *     f = floats).  Per-channel reductions are two-phase (block partials ->  *
 *     fixed-order fp64 finalize), never atomics: runs are bit-reproducible.  *
 *     C must be a multiple of 8 (<= 2048); ld multiples of 8; 16-B pointers. *
 * ======================================================================== */
#define B2SEG_EW_MAX_PARTIAL_ROWS 592 /* 148 SMs x 4: most rows any block-partial buffer has */

enum {
  /* BatchNorm2d statistics -> per-channel affine (torch native_batch_norm; every
   * nn.BatchNorm2d: advanced_models.py:12,17,22,200,203 + torchvision blocks).
   * p0 partial f32 [rows][2][C] (sum, sumsq; may be folded in place; unused in eval)
   * p1 gamma|NULL  p2 beta|NULL  p3 running_mean  p4 running_var  p5 num_batches_tracked (int64)|NULL
   * p6 out bnp f32 [4][C] = scale, shift, mean, invstd
   * i0 rows  i1 C  i2 count (N*H*W)  i3 training    f0 eps  f1 momentum */
  B2SEG_EW_BN_FINALIZE = 1,
  /* out = act(y*scale+shift (+ res*rscale+rshift | + res)), act = ReLU if i5.
   * p0 y bf16  p1 bnp  p2 out bf16  p3 res bf16|NULL  p4 res bnp|NULL
   * i0 P  i1 C  i2 ld_y  i3 ld_out  i4 ld_res  i5 relu */
  B2SEG_EW_BN_APPLY = 2,
  /* BatchNorm backward phase 1: dz = (da1 [+ da2]) * [a > 0]; block partials of (dz, dz*y) -- the
   * RAW pre-BN y; BN_BWD_FINALIZE with i5 = 1 turns them into (sum dz, sum dz*yhat).
   * p0 da1 bf16  p1 da2 bf16|NULL  p2 a bf16|NULL (ReLU mask source)  p3 y bf16 (pre-BN)  p4 bnp
   * p5 out partial f32 [rows][2][C], rows = b2seg_ew_partial_rows(kind, P, C)
   * i0 P  i1 C  i2 ld_da1  i3 ld_da2  i4 ld_a  i5 ld_y
   * i9 self-mask: with p2 NULL, the ReLU mask is recomputed as [bf16(y*scale+shift) > 0] from y
   *    and bnp (the activation of a plain conv-BN-ReLU is never re-read) */
  B2SEG_EW_BN_BWD_REDUCE = 3,
  /* p0 partial  p1 gamma|NULL  p2 bnp  p3 dgamma f32|NULL  p4 dbeta f32|NULL  p5 out coef f32 [3][C]
   * p6 dbias f32|NULL (gradient of a conv bias feeding this BN: 0 with batch statistics)
   * i0 rows  i1 C  i2 count  i3 accumulate into dgamma/dbeta/dbias  i4 training
   * i5 the second accumulator of `partial` is sum dz*y (BN_BWD_REDUCE), not sum dz*yhat */
  B2SEG_EW_BN_BWD_FINALIZE = 4,
  /* phase 2: dy = c1*(dz - c2 - yhat*c3) -> bf16; optionally also emits dz (residual branch).
   * p0..p4 as BN_BWD_REDUCE  p5 coef  p6 dy bf16  p7 dz bf16|NULL
   * i0..i5 as BN_BWD_REDUCE  i6 ld_dy  i7 ld_dz  i8 dz mode (1 store, 2 accumulate)  i9 self-mask */
  B2SEG_EW_BN_BWD_APPLY = 5,
  /* MaxPool2d(3, stride 2, pad 1) (advanced_models.py:83,272) + argmax byte (kh*3+kw) per element.
   * p0 x bf16 [N,H,W,C]  p1 out bf16 [N,H/2,W/2,C]  p2 idx u8 [N,H/2,W/2,C]
   * i0 N i1 H i2 W i3 C i4 ld_x i5 ld_out */
  B2SEG_EW_MAXPOOL_FWD = 6,
  /* p0 dout bf16  p1 idx u8  p2 dx bf16 [N,H,W,C]   i0 N i1 H i2 W i3 C i4 ld_dout i5 ld_dx i6 accumulate */
  B2SEG_EW_MAXPOOL_BWD = 7,
  /* input_conv 7x7 s2 p3, 1 -> 64 channels (advanced_models.py:76,268) + BatchNorm block partials.
   * p0 x f32 [N,H,W]  p1 w f32 [64][49]  p2 y bf16 [N,H/2,W/2,64]  p3 partial f32 [rows][2][64]|NULL
   * i0 N i1 H i2 W ; rows = b2seg_stem_stat_rows(N,H,W) */
  B2SEG_EW_STEM_FWD = 8,
  /* p0 x f32  p1 dy bf16 [N,H/2,W/2,64]  p2 out partial f32 [rows][3136] (rows =
   * b2seg_stem_wgrad_rows; reduce with COLSUM_FINALIZE into dW [64][49])   i0 N i1 H i2 W */
  B2SEG_EW_STEM_WGRAD = 9,
  /* 1-output-channel 1x1 conv (conv_final advanced_models.py:160,350; AttentionGate.psi :21):
   * out[p] = sum_c v[p,c]*w[c] + b (f32), optional block partials (sum, sumsq) for BatchNorm2d(1).
   * p0 v bf16  p1 w f32 [C]  p2 b f32 [1]|NULL  p3 out f32 [P]  p4 partial f32 [rows][2]|NULL
   * i0 P  i1 C  i2 ld_v */
  B2SEG_EW_DOT1_FWD = 10,
  /* backward of DOT1: ds[p] = dscalar[p], or BatchNorm2d(1)-backward of it when p5 is given
   * (coef p5, pre-BN values p6, bnp p7);  dv[p,c] = ds*w[c];  partial [rows][C+1] = (dW[C], db).
   * p0 dscalar f32  p1 v bf16  p2 w f32  p3 dv bf16  p4 partial  p5 coef|NULL  p6 ypre f32|NULL
   * p7 bnp|NULL     i0 P  i1 C  i2 ld_v  i3 ld_dv */
  B2SEG_EW_DOT1_BWD = 11,
  /* out[k] (+)= sum_r partial[r*stride + k].  p0 partial  p1 out f32  i0 rows i1 K i2 accumulate
   * i3 row stride (elements) */
  B2SEG_EW_COLSUM_FINALIZE = 12,
  /* per-channel sum over pixels (ConvTranspose2d bias gradient).
   * p0 x bf16  p1 partial f32 [rows][C]   i0 P i1 C i2 ld */
  B2SEG_EW_COLSUM = 13,
  /* every fp32 parameter -> its bf16 GEMM operand layouts, one launch.
   * p0 device table of b2seg_pack_job  i0 number of jobs  i1 total 32x32xT tiles (T <= 9) */
  B2SEG_EW_PACK = 14,
  /* AttentionGate tail (advanced_models.py:34-40): psi = sigmoid(BN1(psi_pre)); out[p,c] = x[p,c]*psi[p].
   * p0 psi_pre f32 [P]  p1 bnp (1 channel)  p2 x bf16  p3 out bf16  p4 psi f32 [P] (saved)
   * i0 P i1 C i2 ld_x i3 ld_out */
  B2SEG_EW_GATE_APPLY = 15,
  /* gate backward: dpsi[p] = sum_c dout[p,c]*x[p,c]; dz = dpsi*psi*(1-psi) (f32 out);
   * dx[p,c] (+)= dout[p,c]*psi[p]; partial [rows][2] = (sum dz, sum dz*yhat) for BatchNorm2d(1).
   * p0 dout bf16  p1 x bf16  p2 psi f32  p3 psi_pre f32  p4 bnp (1 channel)  p5 dz f32  p6 dx bf16
   * p7 partial    i0 P i1 C i2 ld_dout i3 ld_x i4 ld_dx i5 accumulate into dx */
  B2SEG_EW_GATE_BWD = 16,
  /* ChannelAttention (advanced_models.py:43-61) squeeze: per (n,c) sum, max, first argmax over H*W.
   * p0 x bf16  p1 part f32 [N][rows][3][C]   i0 N i1 HW i2 C i3 ld ;
   * rows = b2seg_ew_partial_rows(kind, HW, C) */
  B2SEG_EW_SE_POOL = 17,
  /* excite: finish pooling, shared MLP C->Ch->C (no bias), sigmoid.
   * p0 part  p1 w1 f32 [Ch][C]  p2 w2 f32 [C][Ch]  p3 saved f32 [N][4][C] (avg, max, gate, argmax bits)
   * p4 hidden f32 [N][2][Ch]   i0 N i1 HW i2 C i3 Ch i4 rows */
  B2SEG_EW_SE_MLP = 18,
  /* out[p,c] = x[p,c]*gate[n,c].  p0 x bf16  p1 saved  p2 out bf16  i0 N i1 HW i2 C i3 ld_x i4 ld_out */
  B2SEG_EW_SE_SCALE = 19,
  /* backward 1: part[n][row][c] = sum_p dout*x.  p0 dout bf16  p1 x bf16  p2 part f32 [N][rows][C]
   * i0 N i1 HW i2 C i3 ld_dout i4 ld_x */
  B2SEG_EW_SE_BWD_REDUCE = 20,
  /* backward 2 (two launches: per-sample, then per weight element summed in sample order): MLP
   * backward -> dW1, dW2, dpool = (davg, dmax).  Uses the tail of `part` ([N][rows][3][C]) as scratch.
   * p0 part  p1 w1  p2 w2  p3 saved  p4 hidden  p5 dW1 f32  p6 dW2 f32  p7 dpool f32 [N][2][C]
   * i0 N i1 HW i2 C i3 Ch i4 rows i5 accumulate into dW1/dW2 */
  B2SEG_EW_SE_MLP_BWD = 21,
  /* backward 3: dx = dout*gate + davg/HW + dmax*[p == argmax].
   * p0 dout bf16  p1 saved  p2 dpool  p3 dx bf16   i0 N i1 HW i2 C i3 ld_dout i4 ld_dx */
  B2SEG_EW_SE_BWD_APPLY = 22,
  /* p0 dst, i0 bytes (multiple of 4), i1 32-bit pattern */
  B2SEG_EW_FILL = 23,
  /* Training-mode BatchNorm forward in ONE launch (BN_FINALIZE + BN_APPLY): every block folds the
   * conv epilogue's partial rows itself (fixed order, fp64) into scale / shift, block 0 also writes
   * bnp and updates the running statistics; then out = act(y*scale+shift (+ res | + BN_r(res))).
   * With p10 the residual carries its own training-mode BatchNorm (downsample branch, gate W_x).
   * p0 y bf16  p1 out bf16  p2 res bf16|NULL
   * p3 partial f32 [rows][2][C]  p4 gamma  p5 beta  p6 running_mean  p7 running_var
   * p8 num_batches_tracked|NULL  p9 out bnp f32 [4][C]
   * p10 res partial|NULL  p11..p16 = gamma, beta, running_mean, running_var, nbt, bnp of the residual BN
   * i0 P  i1 C  i2 ld_y  i3 ld_out  i4 ld_res  i5 relu  i6 rows  i7 count  i8 res rows
   * f0 eps  f1 momentum
   * rows < 0 (res rows < 0): p3 (p10) is the int64 [2][C] fixed-point accumulator that the conv epilogues added
   * to (b2seg_tapgemm_desc::stats_fixed; sum x 2^20, sum of squares x 2^16) -- nothing to fold, the blocks derive
   * scale / shift straight from it.  The caller zeroes the accumulators before the convolutions of every pass. */
  B2SEG_EW_BN_APPLY_FUSED = 24,
  /* BatchNorm backward in ONE persistent launch (BN_BWD_REDUCE + _FINALIZE + _APPLY): phase 1 block
   * partials of (sum dz, sum dz*y); device-wide barrier (all blocks are co-resident: one per SM);
   * every block folds the partial rows (fixed order, fp64) into the three per-channel constants,
   * block 0 writes dgamma / dbeta / dbias; phase 2 walks the block's pixels in reverse (they are
   * the most recently read, L2-hot) and writes dy (and dz).
   * p0 da1 bf16  p1 da2 bf16|NULL  p2 a bf16|NULL  p3 y bf16  p4 bnp  p5 scratch partial f32 [grid][2][C]
   * p6 dy bf16  p7 dz bf16|NULL  p8 gamma|NULL  p9 dgamma|NULL  p10 dbeta|NULL  p11 dbias|NULL
   * i0 P  i1 C  i2 ld_da1  i3 ld_da2  i4 ld_a  i5 ld_y  i6 ld_dy  i7 ld_dz  i8 dz mode (1 store,
   * 2 accumulate)  i9 self-mask  i10 training  i11 accumulate into dgamma/dbeta/dbias */
  B2SEG_EW_BN_BWD_FUSED = 25,
};

/* out[t][a][b] = bf16(in[a*sa + b*sb + t*st]); `start` = number of 32x32xT tiles
 * (ceil(A/32)*ceil(B/32) per job) of all earlier jobs. */
typedef struct b2seg_pack_job {
  const float* in;
  void* out;
  int32_t T, A, B, pad;
  int64_t sa, sb, st;
  int64_t start;
} b2seg_pack_job;

#define B2SEG_EW_NPTR 18
#define B2SEG_EW_NINT 12
typedef struct b2seg_ew_desc {
  int32_t kind;
  void* p[B2SEG_EW_NPTR];
  int64_t i[B2SEG_EW_NINT];
  float f[4];
} b2seg_ew_desc;

int b2seg_ew_create(const b2seg_ew_desc* desc /*host*/, b2seg_op** out);
/* rows of the block-partial buffer an op of `kind` writes for P pixels x C channels */
int b2seg_ew_partial_rows(int kind, int64_t P, int C);
int b2seg_stem_stat_rows(int n, int h, int w);
int b2seg_stem_wgrad_rows(int n, int h, int w);

#ifdef __cplusplus
}
#endif
#endif /* B2SEG_H_ */
