/*
 * ts2d_b200.h - C ABI of libts2d_b200.so: the B200-native TS2D inference hot path.
 *
 * This is the drop-in boundary.  The reference (risc-mi/totalsegmentator2D) is pure
 * Python; its hot path runs inside SimpleITK, nnunetv2 and torch.  A maintainer binds
 * this library with ctypes and each entry point replaces the
 * reference call cited next to it (paths relative to the reference checkout).
 *
 * Conventions
 *   - every pointer named *_dev is a DEVICE pointer owned by the caller (e.g. a torch
 *     tensor's data_ptr()); pointers named *_host are host pointers read synchronously
 *     during the call; nothing is allocated or freed by a call except opaque handles;
 *   - every launching call takes the cudaStream_t to launch on (as void*), returns
 *     0 on success or a negative ts2d_status; the message is in ts2d_last_error()
 *     (thread local);
 *   - no CPU fallback: a device that is not sm_100 yields TS2D_ERR_DEVICE;
 *   - 2-D images are planar float32 [C][H][W] (nnU-Net's (C,1,H,W)); volumes are
 *     [n0][n1][n2] with n2 fastest (sitk.GetArrayFromImage order [z][y][x]).
 */
#ifndef TS2D_B200_H
#define TS2D_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TS2D_ABI_VERSION 1

typedef enum {
    TS2D_OK = 0,
    TS2D_ERR_ARG = -1,      /* invalid argument */
    TS2D_ERR_CUDA = -2,     /* a CUDA runtime/driver call failed */
    TS2D_ERR_DEVICE = -3,   /* no sm_100 device / kernel image not loadable */
    TS2D_ERR_STATE = -4,    /* handle not finalised, weights missing, batch too large ... */
    TS2D_ERR_UNSUPPORTED = -5
} ts2d_status;

typedef enum { TS2D_U8 = 0, TS2D_I16 = 1, TS2D_U16 = 2, TS2D_I32 = 3, TS2D_F32 = 4, TS2D_F64 = 5 } ts2d_dtype;
typedef enum { TS2D_PROJ_MEAN = 0, TS2D_PROJ_MAX = 1, TS2D_PROJ_MIN = 2 } ts2d_proj_mode;
typedef enum { TS2D_NORM_NONE = 0, TS2D_NORM_ZSCORE = 1, TS2D_NORM_CT = 2, TS2D_NORM_RESCALE01 = 3, TS2D_NORM_RGB01 = 4 } ts2d_norm_scheme;
typedef enum { TS2D_PREC_F32 = 0, TS2D_PREC_BF16 = 1, TS2D_PREC_F16 = 2 } ts2d_precision;

typedef void *ts2d_stream;           /* cudaStream_t */
typedef struct ts2d_unet ts2d_unet;  /* opaque engine: plan + repacked weights + workspace + TMA descriptors */

const char *ts2d_last_error(void);
int ts2d_abi_version(void);
/* 0 if `device` is an sm_100 GPU whose kernels load; negative status otherwise */
int ts2d_device_check(int device);

/* ---- K1: projection ------------------------------------------------------------------
 * Replaces reorient_image + project + Cast + Compose:
 *   ts2d/core/util/image.py:32-43 (sitk.DICOMOrient), :46-101 (Max/MeanProjectionImageFilter,
 *   SetProjectionDimension), ts2d/tool.py:154-160,181-185.
 * One pass over the volume produces every requested channel.  The axis permutation and
 * flips of DICOMOrient are folded into the addressing: `reduce_dim` is the memory axis
 * (0,1,2 = n0,n1,n2) that is collapsed, `h_dim` the memory axis that becomes output rows
 * (the remaining axis becomes columns); rows/columns are reversed when flip_h/flip_w.
 * out_dev: float32 [n_modes][H][W].  max/min are exact; mean = (double)sum / N rounded
 * once to float32 (integer input: exact int64 sum, i.e. bit-identical to ITK).
 */
int ts2d_project(const void *vol_dev, int dtype, int64_t n0, int64_t n1, int64_t n2,
                 int reduce_dim, int h_dim, int flip_h, int flip_w,
                 const int *modes_host, int n_modes, float *out_dev, ts2d_stream stream);

/* ---- K2: nnU-Net preprocessing (DefaultPreprocessor.run_case_npy) ------------------------
 * Replaces preprocessor.run_case at ts2d/core/inference/prediction_worker.py:194-199
 * (upstream nnunetv2: crop_to_nonzero -> _normalize -> resampling_fn_data).
 */
/* bbox_dev: int32[4] = {h0, h1, w0, w1} (half open) of OR_c(img[c] != 0); full frame when empty */
int ts2d_bbox_nonzero(const float *img_dev, int C, int H, int W, int *bbox_dev, ts2d_stream stream);
/* stats_dev: double[C][4] = {sum, sum of squares, min, max} over the bbox */
int ts2d_channel_stats(const float *img_dev, int C, int H, int W, int h0, int h1, int w0, int w1,
                       double *stats_dev, ts2d_stream stream);
/* out_dev: float32 [C][h1-h0][w1-w0]; scheme_host int[C]; params_host float[C][4] =
 * {lower, upper, mean, std} (CTNormalization constants from plans.json); stats_dev from above */
int ts2d_normalize(const float *img_dev, int C, int H, int W, int h0, int h1, int w0, int w1,
                   const int *scheme_host, const float *params_host, const double *stats_dev,
                   float *out_dev, ts2d_stream stream);
/* skimage.transform.resize(order in {1,3}, mode='edge', anti_aliasing=False, clip=True) per channel.
 * scratch_dev: double[ C * (h+24) * (w+24) + 2*C ] for order 3 (may be NULL for order 1). */
int ts2d_resample2d(const float *in_dev, int C, int h, int w, float *out_dev, int h2, int w2, int order,
                    double *scratch_dev, ts2d_stream stream);

/* The order-3 resampler for callers that cannot allocate device memory.  NO COUNTERPART IN THE REFERENCE (its resampling
 * runs in scipy on the host and allocates freely); it exists for users of this library only.
 * Who needs it: code that is called back from inside another application's CUDA work and must not call cudaMalloc there
 * (the call serialises against the whole context and can stall every stream), and bindings that are handed just an input
 * and an output pointer by their framework and own no device memory pool from which a work area could come.
 * Same arithmetic and same results as ts2d_resample2d, but the double-precision work area is a fixed array built into
 * the library: each GPU that loads the library carries it (16 MiB + 512 B) until the process ends.  The call itself
 * allocates and frees nothing.
 *   limit   C*(h+24)*(w+24) + 2*C <= ts2d_resample2d_static_capacity() doubles (= 2^21 + 64; two channels of 1000 x 1000
 *           fit exactly) for order 3; anything larger returns TS2D_ERR_ARG with the sizes in ts2d_last_error().
 *           Order 1 needs no work area and has no limit.
 *   sharing every call on a GPU uses the same array: calls that may overlap on the device (two streams, two threads)
 *           race and produce garbage.  Keep them on one stream, or order them with events.  The array is left dirty
 *           between calls; each call re-initialises what it needs, so results do not depend on earlier calls.
 *   set-up  the first call on each GPU (with that GPU current, as for every entry point) asks the runtime for the array's
 *           device address; issue that first call outside of any stream capture. */
int ts2d_resample2d_static(const float *in_dev, int C, int h, int w, float *out_dev, int h2, int w2, int order,
                           ts2d_stream stream);
int64_t ts2d_resample2d_static_capacity(void);

/* ---- sliding-window tiling ---------------------------------------------------------------
 * Replaces pad_nd_image + slicers + torch.flip in nnUNetPredictor.predict_sliding_window_return_logits
 * (reached from prediction_worker.py:209).  Writes NHWC tiles [T*M][Ph][Pw][C] in `precision`;
 * variant m of tile t is image index t*M+m; mirror variants are enumerated as the upstream
 * axes_combinations: m=0 none, then (H), (W), (H,W) for mirror_axes (0,1).
 * mirror_mask: bit0 = axis H allowed, bit1 = axis W allowed. M = 1,2 or 4 accordingly.
 */
int ts2d_extract_tiles(const float *img_dev, int C, int h, int w, int pad_top, int pad_left, /* T <= 1024 */
                       int Ph, int Pw, const int *tile_yx_host, int T, int mirror_mask,
                       int precision, void *out_dev, ts2d_stream stream);

/* ---- K4-K7: the U-Net ------------------------------------------------------------------------
 * Replaces the network built by nnUNetPredictor.initialize_from_trained_model_folder
 * (ts2d/core/inference/nnu.py:149-166) = dynamic_network_architectures PlainConvUNet.
 */
#define TS2D_MAX_STAGES 12
typedef struct {
    int32_t n_stages;
    int32_t in_channels;
    int32_t num_heads;
    int32_t features[TS2D_MAX_STAGES];
    int32_t strides[TS2D_MAX_STAGES][2];
    int32_t kernel_sizes[TS2D_MAX_STAGES][2];
    int32_t n_conv_enc[TS2D_MAX_STAGES];
    int32_t n_conv_dec[TS2D_MAX_STAGES];   /* [d], d = 0 is the deepest decoder stage */
    float eps;                              /* InstanceNorm eps */
    float negative_slope;                   /* LeakyReLU */
    int32_t patch_h, patch_w;
    int32_t precision;                      /* ts2d_precision of activations / MMA operands */
    int32_t max_batch;                      /* images per forward (tiles x mirror variants x studies) */
    int32_t use_tensor_cores;               /* 0: CUDA-core reference kernels only (parity lock) */
} ts2d_unet_plan;

int ts2d_unet_create(const ts2d_unet_plan *plan_host, ts2d_unet **out);
/* one tensor of the upstream state dict (float32, host), by its upstream name, e.g.
 * "encoder.stages.0.0.convs.0.conv.weight", "decoder.transpconvs.2.bias", "decoder.seg_layers.6.weight".
 * Unknown / duplicate-alias names (all_modules.*, decoder.encoder.*) are accepted and ignored (returns 1). */
int ts2d_unet_set_weight(ts2d_unet *u, const char *name, const float *data_host, int64_t count);
/* checks that every parameter arrived, repacks to kernel layouts, allocates the workspace, encodes TMA maps */
int ts2d_unet_finalize(ts2d_unet *u);
size_t ts2d_unet_workspace_bytes(const ts2d_unet *u);
/* x_dev: NHWC tiles [batch][Ph][Pw][Cin] in plan precision. Leaves the last decoder features in the workspace. */
int ts2d_unet_forward(ts2d_unet *u, const void *x_dev, int batch, ts2d_stream stream);
/* per-tile logits of the last forward: float32 [batch][heads][Ph][Pw] (inspection / parity tests) */
int ts2d_unet_tile_logits(ts2d_unet *u, int batch, float *logits_dev, ts2d_stream stream);
/* runs the forward up to and including layer `name` ("enc{s}.{i}" / "dec{d}.{i}" / "up{d}") and writes its
 * normalised+activated output as float32 NCHW (layer-wise parity tests). */
int ts2d_unet_forward_debug(ts2d_unet *u, const void *x_dev, int batch, const char *layer, float *out_dev,
                            ts2d_stream stream);
/* ---- K8/K9: mirror-TTA average + Gaussian blend + 1x1 head + threshold + un-pad/un-crop + merge -------
 * Replaces the accumulate/normalise loop of predict_sliding_window_return_logits, the seg head,
 * export_prediction_from_logits (prediction_worker.py:215-221) and, through ch_off/ch_total, the
 * channel concatenation of combine_segmentations (ts2d/core/util/image.py:490-510).
 * The head is linear, so it is applied after blending the 32-channel decoder features.
 * Tiles t=0..T-1 at tile_yx (in padded coordinates), M mirror variants each (same order as extract).
 * (h, w): preprocessed image size; the padded image is (h+pad..)>=patch.
 * logits_dev (optional): float32 [heads][h][w].
 * mask_dev (optional): uint8 [H0][W0][ch_total]; this model writes channels [ch_off, ch_off+heads) of the
 *   pixels inside the crop bbox (h0.., w0..) - requires (h, w) == bbox size (no resampling). The caller
 *   zero-fills mask_dev once per study.
 * present_dev (optional): uint8 [ch_total]; element ch_off+c is set to 1 when label c occurs anywhere
 *   (reference ts2d/core/util/meta.py:15-33 get_labels(fetch=True), which decides the merged channel order).
 * WRITES OUTSIDE THE CROP BOX: when ts2d_unet_blend_covers_image(u) is 1 (fp16 / bf16 plans whose first stage has 8, 16, 32 or 64 features) the call stores this model's
 *   channels [ch_off, ch_off+heads) of EVERY pixel of the H0 x W0 mask - zeros outside the crop box - so the caller need not clear the
 *   mask, and a caller that pre-fills or composes those channels itself will find them overwritten.  Channels of other models are never
 *   touched.  When it returns 0 only the crop box is written and the caller zero-fills mask_dev once per study, as stated above.
 */
int ts2d_unet_blend_head(ts2d_unet *u, const int *tile_yx_host, int T, int M, int mirror_mask,
                         int pad_top, int pad_left, int h, int w, const float *gaussian_dev,
                         float *logits_dev, uint8_t *mask_dev, int H0, int W0, int h0, int w0,
                         int ch_off, int ch_total, uint8_t *present_dev, ts2d_stream stream);
/* 1 when ts2d_unet_blend_head writes the model's channel window of every pixel of the mask (zeros outside the crop box): the
 * caller need not clear the merged mask first (saves a 60 MB memset per study) */
int ts2d_unet_blend_covers_image(const ts2d_unet *u);
/* export when the plan resamples: logits float32 [heads][h][w] -> order-1 resize to (h2,w2) -> >0 -> mask */
int ts2d_threshold_logits(const float *logits_dev, int heads, int h, int w, int h2, int w2,
                          uint8_t *mask_dev, int H0, int W0, int h0, int w0, int ch_off, int ch_total,
                          ts2d_stream stream);
/* dst[pix][j] = src[pix][perm[j]] for pix < n_pix: the data-dependent channel order of combine_segmentations
 * (labels present in the image first, reference ts2d/core/util/meta.py:280-286) applied on the device.
 * perm_host: int[C]. */
int ts2d_permute_channels(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_pix, int C, const int *perm_host,
                          ts2d_stream stream);
/* the same with the reference's merged order computed on the device (no host round trip between the models and the merge):
 * per model window offsets_host[m]..offsets_host[m+1], labels with present_dev[c] != 0 first (ascending), then the absent ones.
 * perm_dev: int32[C], receives the permutation (the host reads it to name the channels). */
int ts2d_merge_present_order(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_pix, int C, const uint8_t *present_dev,
                             const int *offsets_host, int n_models, int *perm_dev, ts2d_stream stream);
/* Bit-packed transport of a 0/1 label mask (59.9 MB as uint8, 7.5 MB bit-packed; the reference has no
 * counterpart - its masks go through gzip-compressed temp files, ts2d/core/inference/prediction_worker.py:235-240).
 * pack: dst_dev[i] bit k = (src_dev[8i + k] != 0), ceil(n_bytes / 8) output bytes; src 16-byte aligned.
 * unpack (HOST memory, no CUDA call): out[j] = bit (j & 7) of packed[j / 8], with up to `threads` worker threads. */
int ts2d_pack_mask_bits(const uint8_t *src_dev, uint8_t *dst_dev, int64_t n_bytes, ts2d_stream stream);
int ts2d_unpack_mask_bits_host(const uint8_t *packed_host, uint8_t *out_host, int64_t n_bytes, int threads);
/* Upload buffers for the volume ingest (the reference decodes into pageable memory with sitk.ReadImage,
 * ts2d/core/util/image.py:225-238): page-locked host memory, optionally write-combined (written once by the CPU, read once by the
 * GPU over PCIe without cache snoops). Returns NULL on failure (ts2d_last_error). */
void *ts2d_host_alloc(size_t bytes, int write_combined);
int ts2d_host_free(void *ptr);
/* Upload of a volume held in ordinary (pageable) host memory - what sitk.ReadImage returns (ts2d/core/util/image.py:225-238) - at close to
 * page-locked speed: the library stages it through its own ring of page-locked slots (4 x 32 MiB per GPU, allocated at the first call on that
 * GPU and kept), `threads` CPU threads filling slot i+1 while slot i is in flight.  Returns once src_host has been read completely; the
 * device copy completes in stream order.  One staged upload runs at a time per process. */
int ts2d_upload_pageable(void *dst_dev, const void *src_host, size_t bytes, int threads, ts2d_stream stream);
/* re-allocates the workspace for a larger/smaller max_batch (weights are kept) */
int ts2d_unet_resize(ts2d_unet *u, int max_batch);
/* number of kernel launches this engine issued since creation (bench.py's gpu_launches) */
int64_t ts2d_unet_launch_count(const ts2d_unet *u);
/* number of conv / transposed-conv layers of a REDUCED-PRECISION tensor-core plan that run on the fp32 CUDA-core parity kernel
 * (conv_direct_kernel) because no tcgen05 kernel takes their shape (channels not a multiple of 32, kernels other than 3x3 / 1x1,
 * strides other than 1 / 2 per axis).  ts2d_unet_finalize logs each such layer to stderr once; 0 for every nnU-Net 2-D plan tested
 * (reference plans come from nnunetv2ml's planner: ts2d/core/inference/nnu.py:149-166). */
int ts2d_unet_direct_layer_count(const ts2d_unet *u);
/* debug aid (compute-sanitizer is not available on every GPU pool): with TS2D_ARENA_GUARD=<bytes> in the environment at
 * ts2d_unet_finalize time every workspace buffer is followed by a canary band; returns the number of damaged bands (0 = no kernel
 * stored outside its buffer), the number of bands being 0 when the variable is unset; -1 on a CUDA error. Synchronises the device. */
int ts2d_unet_check_guards(ts2d_unet *u);
int ts2d_unet_guard_count(const ts2d_unet *u);
/* conv/transposed-conv/head FLOPs (MAC = 2) of one forward at `batch` images */
double ts2d_unet_flops(const ts2d_unet *u, int batch);
/* device time (ms) the conv kernels of the last forward took; needs ts2d_unet_set_profiling(u, 1) */
int ts2d_unet_set_profiling(ts2d_unet *u, int enable);
double ts2d_unet_last_conv_ms(ts2d_unet *u);
/* same, summed over EVERY launch of the last forward (convs, transposed convs, InstanceNorm-apply and split-K finish passes):
 * the denominator of the honest U-Net roofline fraction (bench.py roofline.frac_unet) */
double ts2d_unet_last_forward_ms(ts2d_unet *u);
/* per-launch record of the last profiled forward: kind 0 = conv, 1 = transposed conv, 2 = InstanceNorm-apply;
 * flops / bytes are the ALGORITHMIC figures of that launch (MAC = 2; bytes = activations read once + written once) */
int ts2d_unet_profile_count(const ts2d_unet *u);
int ts2d_unet_profile_entry(ts2d_unet *u, int i, char *name_out, int name_cap, int *kind, double *ms,
                            double *flops, double *bytes);
int ts2d_unet_destroy(ts2d_unet *u);

/* global launch counter of the non-engine kernels (projection, preprocessing, tiling) */
int64_t ts2d_launch_count(void);

/* ---- PNG visuals (reference ts2d/core/util/image.py:383-481 create_visual / get_auto_window, ts2d/tool.py:272-284) ----
 * labels: uint8 mask [H][W][C] -> label = 1 + highest-index non-zero component -> nearest-neighbour uniform resampling
 * (output index j samples input index c_in + (j - c_out) * ratio, c = size / 2, default 0 outside) -> sitk.LabelToRGB with the
 * colour map palette_host[n_colors][3] (label L -> entry L % n_colors, background -> black).  rgb_dev: uint8 [H2][W2][3]. */
int ts2d_visual_labels(const uint8_t *mask_dev, int C, int H, int W, int H2, int W2, double ratio_y, double ratio_x,
                       const uint8_t *palette_host, int n_colors, uint8_t *rgb_dev, ts2d_stream stream);
/* intensity: float32 [H][W] -> (resample != 0: cubic B-spline uniform resampling, ITK mirror boundaries) -> window
 * [lower, upper] (auto_lower / auto_upper: minimum / maximum of the resampled image) -> IntensityWindowing to 0..255 ->
 * uint8 by truncation.  scratch_dev: H*W + 2 doubles; resampled_dev: H2*W2 floats; out_dev: uint8 [H2][W2]. */
int ts2d_visual_intensity(const float *img_dev, int H, int W, int H2, int W2, double ratio_y, double ratio_x, int resample,
                          double lower, double upper, int auto_lower, int auto_upper, double *scratch_dev,
                          float *resampled_dev, uint8_t *out_dev, ts2d_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* TS2D_B200_H */
