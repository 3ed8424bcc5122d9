/*
 * include/weaklabel_cuda.h -- C-ABI of libweaklabel_cuda.so, the B200 (sm_100a)
 * replacement for the EM-Adapt E-step host library of xtudbxk/em-adapt-tensorflow.
 *
 * Every entry point is extern "C" with plain pointers and sizes; no torch types.
 * Citations are relative to the reference tree (estep/weaklabel.cpp etc.).
 *
 * There is NO CPU fallback behind any of these: without a CUDA device every
 * compute entry point fails loudly (legacy e_step: message on stderr + abort();
 * the rest: non-zero status + weaklabel_last_error()).
 */
#ifndef WEAKLABEL_CUDA_H
#define WEAKLABEL_CUDA_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define WL_API __attribute__((visibility("default")))
#else
#define WL_API
#endif

#define WEAKLABEL_MAX_CLASSES 32     /* label sets are 32-bit masks                  */
#define WEAKLABEL_OK 0
#define WEAKLABEL_ERR_CUDA 1         /* a CUDA runtime call or kernel failed         */
#define WEAKLABEL_ERR_ARG 2          /* bad shape / stride / parameter               */
#define WEAKLABEL_ERR_EMPTY_SET 3    /* an image has no class < C at even coordinates
                                        (the reference runs off its array there,
                                        weaklabel.cpp:83 with present_label_num==0)  */
#define WEAKLABEL_ERR_STATE 4        /* plan used before weaklabel_plan_prepare      */

/* ---------------------------------------------------------------------------
 * 1. Drop-in legacy entry point.  Replaces estep/weaklabel.cpp:6-7 exactly:
 *    same symbol, same argument order and types, same in-place contract, and it
 *    returns `map`.  Bound by the reference's ctypes callers without change
 *    (estep.py:15,30; estep/estep.py:16,31; estep/estep_with_ctype.py:17,22).
 *
 *    map    float32, logical [B,H,W,C], HOST memory, mutated in place; element
 *           (b,i,c) at map + b*stride[0]/4 + i*stride[2]/4 + c (weaklabel.cpp:50)
 *    shape  {B,H,W,C}; stride = BYTE strides (weaklabel.cpp:13-16)
 *    label  int32 dense [B,H,W] ground-truth map; the image-level label set is
 *           its stride-2 scan, values >= C are void (weaklabel.cpp:19-31)
 *    Class order: std::random_shuffle on the process-global libc rand(), one
 *    image after the other (weaklabel.cpp:83) -- this library consumes the very
 *    same rand() stream, max(P-2,0) draws per iteration per image.
 *    Batch > 1 (which crashes in the reference, weaklabel.cpp:117-118) is
 *    defined as B consecutive single-image calls in index order.
 *    Errors cannot be reported through this signature: they print to stderr
 *    and abort().
 * ------------------------------------------------------------------------- */
WL_API float *e_step(float *map, const int *shape, const int *stride, const int *label,
              const bool suppress_others, const int num_iter, const float margin_others,
              float bg, float fg);

/* ---------------------------------------------------------------------------
 * 2. Status.
 * ------------------------------------------------------------------------- */
WL_API const char *weaklabel_last_error(void);   /* thread-local, "" when none            */
WL_API int weaklabel_version(void);              /* 10000*major + 100*minor + patch       */
WL_API int weaklabel_device_count(void);         /* 0 when no CUDA device is usable       */

/* ---------------------------------------------------------------------------
 * 3. Explicit class-order stream.  The reference's only hidden state is the
 *    glibc rand() stream behind std::random_shuffle (weaklabel.cpp:83); this
 *    object reproduces that generator bit-for-bit from a seed so that a run is
 *    reproducible and shardable (srand(1) == process start).
 * ------------------------------------------------------------------------- */
typedef struct weaklabel_rng weaklabel_rng;
WL_API weaklabel_rng *weaklabel_rng_create(unsigned int seed);
WL_API void weaklabel_rng_destroy(weaklabel_rng *rng);
WL_API int weaklabel_rng_next(weaklabel_rng *rng);              /* == rand()             */
WL_API void weaklabel_rng_skip(weaklabel_rng *rng, long long n_draws);
WL_API long long weaklabel_rng_draws(const weaklabel_rng *rng); /* draws since seeding   */
/* draws one image's E-step consumes: num_iter * max(P-2, 0) */
WL_API long long weaklabel_draws_for(int n_present, int num_iter);

/* ---------------------------------------------------------------------------
 * 4. Image-level label sets (weaklabel.cpp:19-31): stride-2 scan of a dense
 *    int32 [B,H,W] HOST ground-truth map into a 0/1 matrix [B][C].
 * ------------------------------------------------------------------------- */
WL_API int weaklabel_label_set_host(const int *label, int B, int H, int W, int C,
                             unsigned char *present /* [B][C] out */);

/* Class visiting order of B consecutive images on one stream (weaklabel.cpp:36-43, 83):
 * order [B][num_iter][C] = class id at (iteration, visit position), -1 padded.
 * Pure host code; rng == NULL uses the process-global libc rand().  This is the table
 * weaklabel_plan_prepare uploads; exposed so that image shards can be checked and
 * positioned (weaklabel_rng_skip) without a device. */
WL_API int weaklabel_draw_order_host(const unsigned char *present, int B, int C, int num_iter,
                              weaklabel_rng *rng, int *order);

/* ---------------------------------------------------------------------------
 * 5. Device-resident E-step.  A plan owns the scratch (compact class planes,
 *    histograms) for one shape on one device; prepare() fixes the label sets,
 *    the five scalars of weaklabel.cpp:7 and the class order; run() only
 *    launches kernels on `stream` (a cudaStream_t passed as void*), never
 *    synchronises, and takes raw DEVICE pointers.  prepare() does not block
 *    either: the tables are staged in pinned memory and uploaded with one
 *    asynchronous copy (see weaklabel_plan_prepare_async).
 *
 *    Algorithm per image (weaklabel.cpp:35-112): absent-class suppression,
 *    floored per-pixel max, num_iter x P exact order-statistic biases
 *    (std::nth_element's value), re-centring by (mean_max_0 - mean_max_1) with
 *    the reference's sequential float / double summation order, then
 *    argmax over all C classes, lowest index on ties (deeplab.py:123).
 * ------------------------------------------------------------------------- */
typedef struct weaklabel_plan weaklabel_plan;

WL_API int weaklabel_plan_create(weaklabel_plan **out, int B, int H, int W, int C, int device);
WL_API void weaklabel_plan_destroy(weaklabel_plan *plan);

/* present: [B][C] 0/1 (from weaklabel_label_set_host or the caller's image-level
 * labels).  rng == NULL draws the class order from the process-global libc
 * rand() exactly as the reference would; otherwise from the explicit stream. */
WL_API int weaklabel_plan_prepare(weaklabel_plan *plan, const unsigned char *present,
                           bool suppress_others, int num_iter, float margin_others,
                           float bg, float fg, weaklabel_rng *rng);

/* The same, with the table upload queued on `stream` (a cudaStream_t as void*): the host side (label-set
 * validation, class order from rand()) runs in the caller's thread, the tables go up with ONE
 * asynchronous copy out of a pinned double buffer.  Never blocks on the device unless it is two
 * uploads ahead of it.  A later run() on a different stream waits for the upload through an event. */
WL_API int weaklabel_plan_prepare_async(weaklabel_plan *plan, const unsigned char *present,
                                 bool suppress_others, int num_iter, float margin_others,
                                 float bg, float fg, weaklabel_rng *rng, void *stream);

/* Ground-truth shrink + label sets ON THE DEVICE (deeplab.py:110 resize_nearest_neighbor to the score
 * size, then weaklabel.cpp:19-31): d_gt is a dense [B][Hg][Wg] DEVICE map, uint8 (gt_is_int32 = 0) or
 * int32 (1).  d_masks[b] gets bit c set <=> class c < C seen at an even (row, column) of the shrunk map;
 * d_flags[0] bit 0 <=> a negative label was seen.  Asynchronous on `stream`. */
WL_API int weaklabel_label_set_device(const void *d_gt, int gt_is_int32, int B, int Hg, int Wg, int H, int W,
                               int C, unsigned int *d_masks, int *d_flags, void *stream);
/* prepare() from a DEVICE ground-truth map: label sets by weaklabel_label_set_device, B + 1 words come
 * back to the host (one small stream synchronisation: the class order is a function of every label-set
 * size and of rand()), then weaklabel_plan_prepare_async.  No host pass over the GT map. */
WL_API int weaklabel_plan_prepare_device_gt(weaklabel_plan *plan, const void *d_gt, int gt_is_int32, int Hg, int Wg,
                                     bool suppress_others, int num_iter, float margin_others,
                                     float bg, float fg, weaklabel_rng *rng, void *stream);

/* d_map: const float32 scores, element (b,i,c) at d_map + b*batch_stride +
 * i*pixel_stride + c (ELEMENT strides; contiguous NHWC: H*W*C and C).
 * d_labels: uint8 [B][H*W] out.  The score map is NOT modified. */
WL_API int weaklabel_estep_labels_device(weaklabel_plan *plan, const float *d_map,
                                  long long batch_stride, long long pixel_stride,
                                  unsigned char *d_labels, void *stream);

/* What run() produces.  with_sums = 1 (default): labels AND mean_max_0 / mean_max_1 / shift of every image
 * (weaklabel.cpp:69-80, 102-110) -- needed by weaklabel_estep_finalize_map_device and the probes.
 * with_sums = 0: labels only.  The re-centring shift is one constant per image added to every class
 * (weaklabel.cpp:108-112), so it can change the argmax only through float rounding of near-ties; the
 * kernel bounds its magnitude, labels every pixel that is safe against the bound, and forms the two
 * order-exact sums only for images in which some pixel is not (rare).  The labels are bit-identical in
 * both modes; mean0 / mean1 of the other images read NaN.  The training path (labels -> loss) uses 0. */
WL_API int weaklabel_plan_set_outputs(weaklabel_plan *plan, int with_sums);

/* The reference's in-place contract on DEVICE memory: d_map becomes the
 * bias-shifted, re-centred score map (what e_step leaves in `map`).  Must be
 * called after weaklabel_estep_labels_device on the same (unmodified) d_map;
 * d_labels of that call may be NULL if only the map is wanted. */
WL_API int weaklabel_estep_finalize_map_device(weaklabel_plan *plan, float *d_map,
                                        long long batch_stride, long long pixel_stride,
                                        void *stream);

/* Device-side status of every run() since it was last read (read-and-clear): synchronises the plan's
 * last stream and reads ONE int into *code (may be NULL).  0 = fine (returns WEAKLABEL_OK); 2xx / 3xx =
 * a selection found no consistent rank -- NaN / Inf among the scores, the labels of that image are not
 * to be trusted; 4xx = a bounded wait inside the kernel gave up (returns WEAKLABEL_ERR_CUDA). */
WL_API int weaklabel_plan_status(weaklabel_plan *plan, int *code);

/* Probes for parity tests (host copies; synchronises the plan's last stream):
 * order/bias [B][num_iter][C] indexed (iteration, visit position) with -1 / 0
 * padding, mean0 [B] float, mean1 [B] double.  Any pointer may be NULL. */
WL_API int weaklabel_plan_probes(weaklabel_plan *plan, int *order, float *bias,
                          float *mean0, double *mean1);

/* tuning knobs, mostly for tests: cluster size (0 = auto; 1..9 for the on-chip kernel, the generic
 * kernel rounds down to 1,2,4,8 or takes 16), tile size in pixels (0 = auto), number of resident
 * clusters (0 = auto). */
WL_API int weaklabel_plan_tune(weaklabel_plan *plan, int cluster_size, int tile_pixels,
                        int max_clusters);
/* Two kernels implement the E-step: the on-chip one (estep_onchip.cu: suppress_others with a
 * margin >= 0, contiguous pixels, image fits a cluster of <= 9 CTAs) and the generic one
 * (estep_kernel.cu: everything else).
 * path: 0 = choose automatically (default), 1 = force generic, 2 = force on-chip (run() fails if
 * it does not apply).  weaklabel_plan_last_path: which one the last run() launched. */
WL_API int weaklabel_plan_set_path(weaklabel_plan *plan, int path);
WL_API int weaklabel_plan_last_path(const weaklabel_plan *plan);
/* Per-image phase time stamps of the on-chip kernel (globaltimer ns): out[B][32], [0..7] = {claimed,
 * streamed-in, mean_max_0 digest done, iterations done, sums known, labels written, selections
 * performed, label-set size}, [8..31] finer stamps of the lead thread (tools/phase_report.py).
 * Reads the stamps of the last run into `out` (may be NULL), then enables (1) / disables (0)
 * stamping for the following runs. */
WL_API int weaklabel_plan_profile(weaklabel_plan *plan, int enable, unsigned long long *out);
/* kernels launched by this plan since creation (bench.py's gpu_launches) */
WL_API long long weaklabel_plan_launches(const weaklabel_plan *plan);

/* One-call convenience over a cached plan: host GT label map, device scores.
 * d_map_out may be NULL (labels only) or == a writable copy target of d_map's
 * layout that receives the shifted map.  rng as in weaklabel_plan_prepare. */
WL_API int e_step_labels_device(const float *d_map, const int *shape, const int *stride_bytes,
                         const int *h_label, bool suppress_others, int num_iter,
                         float margin_others, float bg, float fg, weaklabel_rng *rng,
                         unsigned char *d_labels, float *d_map_out, void *stream);

/* ---------------------------------------------------------------------------
 * 6. Fused softmax cross-entropy, forward + backward, ignore_label (255).
 *    Replaces deeplab.py:182 (tf.nn.sparse_softmax_cross_entropy_with_logits +
 *    reduce_mean) with the ignore semantics of network.py:77-89: pixels whose
 *    label == ignore_label or >= C contribute no loss, no gradient, and are
 *    left out of the mean's denominator.
 *    d_logits [n_pix][C] float32 (pre-E-step fc8 scores), d_labels [n_pix] uint8,
 *    d_loss_mean: 1 float out, d_valid: 1 int out (kept pixels),
 *    d_dlogits [n_pix][C] out = (softmax - onehot) / valid, or NULL (forward only).
 *    d_work: >= 16 bytes of device scratch (zeroed by the call).
 * ------------------------------------------------------------------------- */
WL_API int weaklabel_softmax_ce_device(const float *d_logits, const unsigned char *d_labels,
                                long long n_pix, int C, int ignore_label,
                                float *d_loss_mean, int *d_valid, float *d_dlogits,
                                void *d_work, void *stream);

/* 6b. The same loss without a scratch argument and without any allocation, for callers that
 *    cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 *    forbid it).  NO COUNTERPART IN THE REFERENCE.
 *    The 16 bytes of scratch live in a static device array of WEAKLABEL_CE_STATIC_SCRATCH_BYTES
 *    that every GPU loading the library holds until exit; the call clears it on `stream`.  All
 *    calls on a device share it: calls that overlap on the GPU (different streams) race -- keep
 *    them on one stream or order them with events.  The first call on each device resolves the
 *    array's device address (set-up; make it before a graph capture starts).  n_pix above
 *    WEAKLABEL_CE_STATIC_MAX_PIXELS is refused (WEAKLABEL_ERR_ARG): the kept-pixel counter in
 *    the scratch is 32 bits wide.  Arguments and outputs as in weaklabel_softmax_ce_device. */
#define WEAKLABEL_CE_STATIC_SCRATCH_BYTES 256
#define WEAKLABEL_CE_STATIC_MAX_PIXELS 2147483647LL
WL_API int weaklabel_softmax_ce_static(const float *d_logits, const unsigned char *d_labels,
                                long long n_pix, int C, int ignore_label,
                                float *d_loss_mean, int *d_valid, float *d_dlogits, void *stream);

/* ---------------------------------------------------------------------------
 * 7. E-step + loss as one call (deeplab.py:120-123 + 182): labels and the softmax
 *    cross-entropy that consumes them, queued back to back on `stream`.  d_map is
 *    the contiguous [B][H*W][C] fc8 score map (the logits); every pixel is kept
 *    (the E-step never emits ignore_label), so the loss is a single launch.
 *    Outputs as in weaklabel_softmax_ce_device; d_labels [B][H*W] uint8.
 * ------------------------------------------------------------------------- */
WL_API int weaklabel_estep_ce_device(weaklabel_plan *plan, const float *d_map, unsigned char *d_labels,
                              float *d_loss_mean, int *d_valid, float *d_dlogits, void *d_work,
                              void *stream);

/* ---------------------------------------------------------------------------
 * 7b. EM-Fixed labels -- the fixed-bias variant of the paper the reference's readme
 *    cites (readme.md:3; NOT implemented in the reference, so there is no parity
 *    target: the test restates it in NumPy).  label = the first maximum over the PRESENT
 *    classes (d_masks[b], as weaklabel_label_set_device writes them) of score + bias_bg
 *    (class 0) / bias_fg (others); an image without any present class gets 255; where
 *    d_strong (uint8 [B][H*W], may be NULL) holds a class < C -- a pixel-level annotation of
 *    a strongly labelled image -- that class is the label.  d_map contiguous [B][H*W][C].
 * ------------------------------------------------------------------------- */
WL_API int weaklabel_fixed_bias_labels_device(const float *d_map, int B, int H, int W, int C,
                                       const unsigned int *d_masks, float bias_bg, float bias_fg,
                                       const unsigned char *d_strong, unsigned char *d_labels, void *stream);

/* ---------------------------------------------------------------------------
 * 7c. Bilinear up-sampling of the fc8 scores (network.py:40: tf.image.resize_bilinear with
 *    the TF1 defaults align_corners = False, no half-pixel centres; TensorFlow is a
 *    third-party dependency of the reference, its published kernel is restated in
 *    csrc/upsample.cuh and oracle/oracle.py).  d_src [B][h][w][C] -> d_dst [B][H][W][C],
 *    both contiguous float32.
 *    weaklabel_estep_labels_upsampled_device runs the E-step (deeplab.py:120-123) on the
 *    up-sampled scores WITHOUT materialising them: the plan is created
 *    and prepared for the full-resolution shape, d_fc8 is the low-resolution tensor; the
 *    labels are bit-identical to up-sampling first and calling weaklabel_estep_labels_device.
 *    On-chip kernel only (suppress_others, margin >= 0); WEAKLABEL_ERR_STATE otherwise.
 * ------------------------------------------------------------------------- */
WL_API int weaklabel_upsample_bilinear_device(const float *d_src, int B, int h, int w, int C, int H, int W,
                                       float *d_dst, void *stream);
WL_API int weaklabel_estep_labels_upsampled_device(weaklabel_plan *plan, const float *d_fc8, int h, int w,
                                            unsigned char *d_labels, void *stream);

/* ---------------------------------------------------------------------------
 * 8. DLPack front door.  weaklabel_dltensor is the DLTensor of dlpack.h (dmlc/dlpack,
 *    third-party; restated field by field so that no foreign header is needed).  A
 *    DLManagedTensor begins with its DLTensor, so a pointer taken from a "dltensor"
 *    capsule (tensor.__dlpack__()) can be passed as is.  scores: float32 [B,H,W,C] in CUDA
 *    memory on the plan's device, class stride 1, dense rows (batch and pixel strides
 *    free); labels: uint8 [B,H,W], contiguous.
 * ------------------------------------------------------------------------- */
typedef struct {
    void *data;
    int32_t device_type;      /* DLDevice.device_type: 2 = kDLCUDA                 */
    int32_t device_id;
    int32_t ndim;
    uint8_t dtype_code;       /* DLDataType.code: 1 = uint, 2 = float               */
    uint8_t dtype_bits;
    uint16_t dtype_lanes;
    int64_t *shape;
    int64_t *strides;         /* in elements; NULL = compact row-major              */
    uint64_t byte_offset;
} weaklabel_dltensor;
WL_API int weaklabel_estep_labels_dlpack(weaklabel_plan *plan, const void *dl_scores, const void *dl_labels,
                                  void *stream);

#ifdef __cplusplus
}
#endif
#endif /* WEAKLABEL_CUDA_H */
