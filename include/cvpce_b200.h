/*
 * cvpce_b200.h -- C ABI of the B200-native (sm_100a) proposal -> product classification hot path.
 *
 * This is the drop-in boundary: every entry point replaces one Python/torch call site of the reference
 * (laitalaj/cvpce, paths relative to the reference root; "tv:" = torchvision, a third-party dependency of it).
 * Plain pointers and sizes only; no torch types.  All pointers are DEVICE pointers unless marked [host].
 * Every call is stream-ordered on `stream` (a cudaStream_t passed as void*), never synchronises the device,
 * never allocates or frees caller-visible memory and never falls back to the CPU.
 *
 * Return value: 0 (CVPCE_OK) or a cvpce_status code; cvpce_last_error() gives a thread-local message.
 * Empty inputs (0 boxes / crops / queries) are valid and return CVPCE_OK without launching.
 */
#ifndef CVPCE_B200_H
#define CVPCE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CVPCE_B200_VERSION 100

#if defined(__GNUC__)
#define CVPCE_API __attribute__((visibility("default")))
#else
#define CVPCE_API
#endif

typedef enum cvpce_status {
  CVPCE_OK = 0,
  CVPCE_ERR_INVALID_ARG = 1,   /* null pointer, negative size, misaligned pointer, unsupported enum value */
  CVPCE_ERR_UNSUPPORTED = 2,   /* shape outside what the kernels implement (message says which limit) */
  CVPCE_ERR_CUDA = 3,          /* a CUDA runtime / driver call failed (message carries cudaGetErrorString) */
  CVPCE_ERR_NO_DEVICE = 4,     /* no sm_100 device is current: there is NO CPU fallback */
  CVPCE_ERR_WORKSPACE = 5      /* workspace pointer null or smaller than *_workspace_bytes() */
} cvpce_status;

typedef enum cvpce_dtype { CVPCE_F32 = 0, CVPCE_BF16 = 1, CVPCE_F16 = 2 } cvpce_dtype; /* F16: crop output, max-pool, dense NHWC MAC */
typedef enum cvpce_layout { CVPCE_NCHW = 0, CVPCE_NHWC = 1, CVPCE_NHWC4 = 2 } cvpce_layout; /* NHWC4: crop output only (4th channel 0) */

/* Thread-local message of the last failing call on this thread ("" if none). */
CVPCE_API const char* cvpce_last_error(void);
/* CVPCE_B200_VERSION the library was built from. */
CVPCE_API int cvpce_version(void);
/* CVPCE_OK iff the current CUDA device is compute capability 10.x (B200); else CVPCE_ERR_NO_DEVICE. */
CVPCE_API int cvpce_device_check(void);
/* Frees lazily created module state (TMA descriptor cache). Safe to call more than once. */
CVPCE_API int cvpce_shutdown(void);
/* Development overrides of launch heuristics (A/B measurements, cross-checks of alternative kernels): process-wide named
 * integers, 0 = automatic.  Knobs: "crop_kernel", "crop_band_rows", "sim_split", "sim_plan", "sim_keys", "mac_cluster",
 * "mac_kernel", "mac_ctas_per_sm", "rmac_kernel", "rmac_ws", "rmac_stages" (csrc/common.cu lists their values).  Results never depend on them beyond fp32 summation order; the library reads no
 * environment variables.  Unknown names return CVPCE_ERR_INVALID_ARG. */
CVPCE_API int cvpce_tuning_set(const char* name, int value);
CVPCE_API int cvpce_tuning_get(const char* name);

/* ------------------------------------------------------------------------------------------------------------
 * a1 + a2  Batched greedy IoU-NMS.
 * Replaces: tv:models/detection/retinanet.py:559-561  box_ops.batched_nms(boxes, scores, labels, nms_thresh)
 *           followed by keep[:detections_per_img]  (reached from cvpce/models/proposals.py:176-177), i.e.
 *           tv:ops/boxes.py:20-48 torch.ops.torchvision.nms for the single-class GLN detector; and, optionally,
 *           the confidence filter res['boxes'][res['scores'] > thr] of cvpce/production.py:13-15.
 *
 * Semantics (bit-exact with the torchvision CPU op): candidates of one image are visited in (score descending,
 * index ascending) order (stable sort; NaN scores first); candidate j is suppressed iff an earlier KEPT candidate
 * i has  inter / (area_i + area_j - inter) > iou_threshold  evaluated in IEEE fp32 without contraction, with
 * area = (x2-x1)*(y2-y1) (no +1) and inter = max(0,.)*max(0,.); 0/0 is NaN and does not suppress.
 * `iou_threshold` is compared as fp32; pass the double threshold rounded toward -inf to reproduce the CPU op's
 * float-vs-double comparison (identical for 0.5, the only value cvpce uses).
 * `min_score`: candidates with score <= min_score are dropped (pass -INFINITY to disable); because kept boxes are
 * in descending score order this equals the reference's post-NMS filter.  `max_keep` <= 0 means unlimited.
 *
 * boxes   f32 [total,4] xyxy, 16-byte aligned; scores f32 [total]; seg_offsets i32 [n_images+1] (CSR, device).
 * max_n   [host] upper bound of candidates per image (<= CVPCE_NMS_MAX_N); an image exceeding it gets
 *         keep_counts[b] = -1.
 * keep    i64 [total]: for image b the kept LOCAL indices (0-based within the segment, score-descending) are
 *         written at keep[seg_offsets[b] ...]; keep_counts i32 [n_images].
 * kept_boxes  (nullable) f32 [total,4]: the kept boxes in the same positions, multiplied by box_scale
 *         (nullable f32 [n_images,2] = (x ratio, y ratio), the fp32 multiply of tv:models/detection/transform.py:306-319).
 * kept_scores (nullable) f32 [total].
 */
#define CVPCE_NMS_MAX_N 8192
CVPCE_API int cvpce_nms_batched(const float* boxes, const float* scores, const int32_t* seg_offsets, int n_images, int max_n,
                      float iou_threshold, float min_score, int max_keep,
                      int64_t* keep, int32_t* keep_counts,
                      const float* box_scale, float* kept_boxes, float* kept_scores, void* stream);

/* Same kernel over fixed-stride segments: image b owns rows [b*seg_stride, b*seg_stride + seg_counts[b]) of boxes / scores
 * (seg_counts i32 [n_images], device; seg_stride <= CVPCE_NMS_MAX_N) -- the layout cvpce_select_candidates writes, so the
 * detector post-processing (f1 -> a1 -> a2) is two launches with no host synchronisation in between.  keep / kept_boxes /
 * kept_scores use the same row positions. */
CVPCE_API int cvpce_nms_segments(const float* boxes, const float* scores, int seg_stride, const int32_t* seg_counts, int n_images,
                       float iou_threshold, float min_score, int max_keep,
                       int64_t* keep, int32_t* keep_counts,
                       const float* box_scale, float* kept_boxes, float* kept_scores, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * f1  Pre-NMS candidate selection of the RetinaNet / GLN head (the step that feeds a1).
 * Replaces: the per-image, per-FPN-level loop of tv:models/detection/retinanet.py:527-557
 *           (RetinaNet.postprocess_detections, inherited unchanged by cvpce/models/proposals.py:161 GaussianLayerNetwork):
 *           torch.sigmoid(logits).flatten() -> `> score_thresh` -> topk(min(topk_candidates, n)) -> anchor / label index
 *           -> BoxCoder.decode_single (tv:models/detection/_utils.py:183-224) -> clip_boxes_to_image -> torch.cat over
 *           levels; ~25 small launches and one host sync (torch.where) per level per image in the reference.
 *
 * levels [host] n_levels (<= CVPCE_SELECT_MAX_LEVELS) descriptors of the head outputs split per FPN level:
 *   logits     f32 [n_images, n_anchors, num_classes], image stride logits_stride (elements), inner dims contiguous;
 *   regression f32 [n_images, n_anchors, 4], image stride regression_stride, rows 16-byte aligned;
 *   anchors    f32 [n_anchors, 4] xyxy, image stride anchors_stride (0: one anchor set shared by all images).
 * image_sizes i32 [n_images,2] (device) = (height, width) of the resized images boxes are clipped to.
 * weights [host] nullable f32[4] BoxCoder weights (RetinaNet: 1,1,1,1); bbox_xform_clip = log(1000/16) in torchvision.
 * Candidates of a level are ordered (score descending, flat (anchor, class) index ascending) -- torch.topk leaves the
 * order of equal scores unspecified -- and levels are concatenated in order.  Scores are sigmoid values computed with a
 * fixed IEEE fp32 operation sequence (reproducible on the host; within a few ulp of ATen's), boxes follow
 * decode_single's fp32 operation order with the same exp.
 *
 * Outputs: image b's candidates are rows [b*seg_stride, b*seg_stride + seg_counts[b]) with seg_stride = n_levels*topk:
 *   boxes f32 [n_images*seg_stride, 4] (16-byte aligned), scores f32 [n_images*seg_stride],
 *   labels (nullable) i64 [n_images*seg_stride], seg_counts i32 [n_images],
 *   level_counts (nullable) i32 [n_images, n_levels].  1 <= topk <= CVPCE_SELECT_MAX_TOPK.
 */
typedef struct cvpce_fpn_level {
  const float* logits;
  const float* regression;
  const float* anchors;
  int64_t logits_stride, regression_stride, anchors_stride;
  int32_t n_anchors;
} cvpce_fpn_level;
#define CVPCE_SELECT_MAX_LEVELS 8
#define CVPCE_SELECT_MAX_TOPK 2048
CVPCE_API int cvpce_select_candidates(const cvpce_fpn_level* levels, int n_levels, int n_images, int num_classes,
                            float score_thresh, int topk, const int32_t* image_sizes, const float* weights,
                            float bbox_xform_clip, float* boxes, float* scores, int64_t* labels,
                            int32_t* seg_counts, int32_t* level_counts, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * a3 + a4 (+ a5, a6)  Batched ROI crop -> pad-to-square -> (antialiased) bilinear resize.
 * Replaces: the per-box python loop  image[:, y1:y2, x1:x2]  over boxes.to(long) at cvpce/production.py:20
 *           (same expression: classification_eval.py:24, detection_eval.py:29, production.py:111) and
 *           cvpce/datautils.py:234-239 resize_for_classification (canvas filled with 0.5, crop at top-left,
 *           ttf.resize to S x S); optionally fuses utils.scale_to_tanh (cvpce/utils.py:280-281, called at
 *           production.py:63) and the MACVGG input normalisation (cvpce/models/classification.py:41-44) as a
 *           per-channel affine y = x*a_c + b_c.
 *
 * image_table i64 [n_images,5] (device): {data pointer (f32, CHW, pixel stride 1), H, W, channel stride, row stride}
 *             strides in elements.
 * boxes f32 [K,4] xyxy in image pixels; coordinates are truncated toward zero (Tensor.to(long)) and then clamped to
 *             the image like a python slice end; box_image_idx i32 [K] (nullable: all boxes belong to image 0).
 * antialias   1: ATen _upsample_bilinear2d_aa (installed torchvision >= 0.17 default), 0: plain bilinear,
 *             align_corners=False (torchvision 0.9, the version the reference pins).
 * affine      [host] nullable f32[6] = {a_0,b_0,a_1,b_1,a_2,b_2}.
 * out         [K,3,S,S] (NCHW), [K,S,S,3] (NHWC) or [K,S,S,4] (NHWC4: a zero 4th channel, 16-byte pixels in fp32 -- the
 *             input layout on which cuDNN runs VGG16's first convolution without its channel-padding pass, with a
 *             weight zero-padded to 4 input channels; bit-identical convolution output) of out_dtype.
 * valid       (nullable) i32 [K]: 1, or 0 for an empty crop (x2<=x1 or y2<=y1 after truncation, where the reference
 *             raises); an empty crop's output is pad_value passed through the affine.
 */
CVPCE_API int cvpce_crop_resize(const int64_t* image_table, int n_images, const float* boxes, const int32_t* box_image_idx,
                      int K, int S, int antialias, float pad_value, const float* affine,
                      int out_dtype, int out_layout, void* out, int32_t* valid, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * a8 + a9  Region max-pool embedding: MAC (reference) or R-MAC (extension) + L2 normalisation.
 * Replaces: x.amax(dim=(-2,-1)) x2, torch.cat, desc / norm(desc).clamp(min=eps) of
 *           cvpce/models/classification.py:46-51 (MACVGG) and :79-85 (MACResNet).
 *
 * maps [host] n_maps (<= 4) feature maps of the same batch, each [B,C,H,W] with arbitrary element strides
 *      (NCHW-contiguous and channels-last are the fast cases), dtype f32 or bf16.
 *      Measured fractions of HBM bandwidth on B200 (two 512-channel maps, 128-256 images): MAC channels-last 0.75-0.87,
 *      MAC NCHW 0.74-0.80.  Region lists / per-region L2 (R-MAC, 3 levels) on dense maps of at least SMs/2 images run a
 *      persistent TMA-staged kernel (csrc/rmac_staged.cu: cp.async.bulk ring, one CTA per image): channels-last f32 0.64-0.84 (128-800 images)
 *      (the cluster kernel it replaces there: 0.48-0.51), channels-last bf16 0.47-0.56 (0.17).  R-MAC on NCHW maps is SUPPORTED
 *      BUT NOT FAST (0.27 f32 / 0.15 bf16 staged, 0.17 / 0.06 before): an extension mode on a legacy layout -- hand the maps
 *      over channels-last (what the fast trunk produces) when R-MAC throughput matters.  Smaller batches, strided maps, maps
 *      with rows that are not whole 16-byte units or more than 24 distinct region boundaries per axis take the cluster kernel.
 * regions [host] i32 [sum_m n_regions[m], 4] = (y0,x0,y1,x1), end exclusive, in feature-map cells;
 *      n_regions [host] i32 [n_maps]; both null => one whole-map region per map (MAC, the reference's mode).
 * per_region_l2 0: region maxima of a map are summed as they are (with one region: the plain MAC descriptor);
 *               1: each region vector is divided by max(||v||_2, eps) before the sum (R-MAC, Tolias et al. 2016).
 * final_l2      1: out row = desc / max(||desc||_2, eps) over the concatenation of all maps (classification.py:51).
 * out f32 [B, sum_m C_m] with row stride out_stride (elements).
 */
typedef struct cvpce_feature_map {
  const void* data;
  int32_t C, H, W;
  int64_t stride_n, stride_c, stride_h, stride_w;
} cvpce_feature_map;
#define CVPCE_POOL_MAX_MAPS 4
#define CVPCE_POOL_MAX_REGIONS 32
CVPCE_API int cvpce_region_pool_embed(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                            const int32_t* regions, const int32_t* n_regions,
                            int per_region_l2, int final_l2, float eps,
                            float* out, int64_t out_stride, void* stream);
/* Same operation with a caller-owned device workspace.  With a workspace, dense channels-last MAC batches (whole-map
 * regions, no per-region L2 -- the reference's mode behind the fast trunk) run the balanced streaming kernel: the batch is
 * cut into equal contiguous byte ranges over a grid that is a multiple of the SM count, instead of one cluster per image.
 * workspace [device] >= cvpce_region_pool_workspace_bytes(...) bytes, 256-byte aligned.  Its first
 *      max(65536, (4*B + 255) & ~255) bytes are per-image counters: they must be ZERO at the first launch and every
 *      launch leaves them zero, so a buffer zeroed once can be reused by all later launches (of any B <= 16384) ordered
 *      on the same stream -- not by concurrent launches on other streams.  The rest needs no initialisation.
 * workspace null, or a case the streaming kernel does not cover (cvpce_region_pool_workspace_bytes returns 0):
 *      identical to cvpce_region_pool_embed.  Results are the same in both forms (maxima exact; the L2 norm sums the
 *      squares in a fixed channel order in each form). */
CVPCE_API int64_t cvpce_region_pool_workspace_bytes(const cvpce_feature_map* maps, int n_maps, int dtype, int B);
CVPCE_API int cvpce_region_pool_embed_ws(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                               const int32_t* regions, const int32_t* n_regions,
                               int per_region_l2, int final_l2, float eps,
                               float* out, int64_t out_stride,
                               void* workspace, int64_t workspace_bytes, void* stream);
/* Allocation-free MAC pooling.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as cvpce_region_pool_embed(maps, ..., regions = NULL, n_regions = NULL, per_region_l2 = 0, ...) -- one whole-map
 * region per map, the reference's MAC mode -- through the balanced streaming kernel, whose scratch is a file-scope device
 * array of CVPCE_POOL_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  The call makes no
 * cudaMalloc / cudaFree and calls no allocating routine; the caller passes inputs and outputs only.
 * Limits: dense channels-last maps only (unit channel stride, stride_w == C, stride_h == W*C, 16-byte aligned data,
 *      C*elemsize/16 dividing 256, sum of C <= 4096), and cvpce_region_pool_workspace_bytes(maps, n_maps, dtype, B) <=
 *      CVPCE_POOL_STATIC_SCRATCH_BYTES (B = 256 images of two 512 x 16 x 16 f32 maps need 2.1 MB); anything else returns
 *      CVPCE_ERR_UNSUPPORTED with the limit in the message.
 * Sharing: every call on a device uses the SAME array.  Calls ordered on one stream (or otherwise serialised on the GPU)
 *      are safe -- the entry point clears the counters it needs on `stream` before the kernel; calls that overlap on the
 *      GPU (different streams, different host threads without ordering) RACE on it and give wrong results. */
#define CVPCE_POOL_STATIC_SCRATCH_BYTES (8 * 1024 * 1024)
CVPCE_API int cvpce_region_pool_embed_static(const cvpce_feature_map* maps, int n_maps, int dtype, int B,
                                   int final_l2, float eps, float* out, int64_t out_stride, void* stream);
/* [host] R-MAC region grid of Tolias et al. for an H x W map and `levels` scales; returns the region count and
 * writes up to max_regions (y0,x0,y1,x1) quadruples. */
CVPCE_API int cvpce_rmac_regions(int H, int W, int levels, int32_t* regions, int max_regions);
/* In-place row L2 normalisation x / max(||x||_2, eps); x f32 [B,D] with row stride `stride`. */
CVPCE_API int cvpce_l2norm_rows(float* x, int B, int D, int64_t stride, float eps, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * f3  Pairwise box IoU and the greedy IoU claim pass (detection evaluation; planogram expected-vs-detected matching).
 * Replaces: tv:ops/boxes.py box_iou called from cvpce/metrics.py:11-13 (iou_matrices) and cvpce/planograms.py:224;
 *           the Python loops over the sorted matrix at cvpce/metrics.py:15-34 (check_matches) and
 *           cvpce/planograms.py:214-232 (inside finalize_via_ransac, feeding the second-round re-classification of
 *           cvpce/production.py:101-115).
 *
 * cvpce_box_iou: iou f32 [n_a, n_b] row-major = inter / ((area_a + area_b) - inter), area = (x2-x1)*(y2-y1),
 * inter = max(w,0)*max(h,0), IEEE fp32 without contraction: bit-equal to torchvision's matrix (0/0 = NaN).
 * Boxes f32 [n,4] xyxy, 16-byte aligned.
 *
 * cvpce_iou_match: problem b (one image) owns rows [row_offsets[b], row_offsets[b+1]) and columns
 * [col_offsets[b], col_offsets[b+1]) (i32 CSR, device).  Rows are visited in index order; row i marks EVERY column j of
 * its problem that is not marked yet, has the row's label (when labels are given: i64, both arrays or neither) and has
 * !(iou_ij < iou_threshold) in fp32 (so a NaN IoU claims, as in the reference where NaN sorts first and never breaks the
 * walk); row_matched[i] = 1 iff it marked at least one, col_matched[j] = 1 iff some row marked it.  This is exactly what
 * the reference's loops compute: they `continue` past used columns and never stop after the first hit, so the descending
 * walk order inside a row is immaterial.  check_matches: rows = predictions in confidence order, columns = targets,
 * true_positive = row_matched, false_positive = 1 - row_matched.  finalize_via_ransac: rows = projected planogram boxes,
 * columns = detections, labels = the integer label ids, matched_expected = row_matched, matched_actual = col_matched.
 * max_cols [host]: upper bound of the columns of any one problem (<= CVPCE_MATCH_MAX_COLS); the IoU matrix is never
 * written to memory.  row_matched u8 [total rows], col_matched u8 [total cols].
 */
#define CVPCE_MATCH_MAX_COLS 16384
CVPCE_API int cvpce_box_iou(const float* boxes_a, int n_a, const float* boxes_b, int n_b, float* iou, void* stream);
CVPCE_API int cvpce_iou_match(const float* row_boxes, const int64_t* row_labels, const int32_t* row_offsets,
                    const float* col_boxes, const int64_t* col_labels, const int32_t* col_offsets, int n_problems,
                    int max_cols, float iou_threshold, uint8_t* row_matched, uint8_t* col_matched, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * a7 (driving the torch trunk)  2x2 / stride-2 max pooling of a channels-last feature map.
 * Replaces: nn.MaxPool2d(kernel_size=2, stride=2) between the convolutions of the VGG16 trunk (tv:models/vgg.py make_layers,
 *           cfg 'D' entries 'M'; reached from cvpce/models/classification.py:31-37 MACVGG.block1 / block2).  The convolutions
 *           themselves stay torch/cuDNN; at::max_pool_forward_nhwc was 28 % of the step's GPU time.
 * in  [B,H,W,C] dense NHWC (a torch channels_last tensor's memory), f32 or bf16, 16-byte aligned, C*elemsize % 16 == 0;
 * out [B,H/2,W/2,C] dense NHWC, floor mode, no padding.  Bit-exact max; a NaN in the window propagates (as at::max_pool2d).
 */
CVPCE_API int cvpce_maxpool2x2_nhwc(const void* in, int dtype, int B, int H, int W, int C, void* out, void* stream);

/* ------------------------------------------------------------------------------------------------------------
 * a10 + a11  Cosine similarity + per-row top-k.
 * Replaces: cvpce/models/classification.py:87-95  distance() = 1 - cosine_similarity and nearest_neighbors()
 *           (gather [Q,G,D] x2, reduce, full argsort, [:, :k]).
 * Result order: (cosine similarity descending == distance ascending, gallery index ascending on equal scores).
 *
 * cvpce_rows_prepare: one-time operand preparation (gallery at build_index time, queries per batch):
 *   x f32 [R,D] -> x_bf16 [R,D] = bf16(x * inv_norm) and inv_norm f32 [R] = 1/max(||x||_2, eps) (normalize=1),
 *   or a plain cast with inv_norm = 1 (normalize=0; rows already unit-norm, as MACVGG outputs are).
 */
CVPCE_API int cvpce_rows_prepare(const float* x, int64_t R, int D, int normalize, float eps,
                       void* x_bf16, float* inv_norm, void* stream);
/* Gallery-sharded search with the QUERIES partitioned over the ranks too (each rank received 1/W of them from its host):
 * the same preparation of this rank's R rows, with the results stored into this block's row slot of EVERY rank's query
 * buffers -- peer_f32[p] / peer_bf16[p] / peer_inv_norm[p] [host arrays of n_peers <= 16 device pointers, each array nullable]
 * address row 0 of this rank's block inside rank p's [Q,D] fp32, [Q,D] bf16 and [Q] inverse-norm buffers (peer memory mapped
 * over NVLink; the own buffers are among them).  The all-gather of the query block IS the kernel's stores; the caller
 * synchronises the ranks before the search reads the buffers.  D % 4 == 0; fp32 targets 16-byte, bf16 targets 8-byte aligned. */
CVPCE_API int cvpce_rows_prepare_scatter(const float* x, int64_t R, int D, int normalize, float eps,
                               float* const* peer_f32, void* const* peer_bf16, float* const* peer_inv_norm, int n_peers,
                               void* stream);
/* tcgen05 path.  q_bf16 [Q,D], g_bf16 [G,D] bf16 row-major, 16-byte aligned, D % 64 == 0, 64 <= D <= 4096, 1 <= k <= 32
 * (k <= 24 with the exact re-score: k + 8 <= 32 candidates per row are kept; larger k returns CVPCE_ERR_UNSUPPORTED).
 * The kernel computes S = Q x G^T tile by tile on the tensor cores (TMA-fed, fp32 accumulators in TMEM) and keeps a
 * per-row candidate list in registers; S never reaches HBM.  When q_f32/g_f32 are given (parity mode), the k' >= k + 8
 * candidates of each row are re-scored in exact fp32 (score = dot(q,g) * q_inv_norm * g_inv_norm; inv norms nullable = 1)
 * before the final selection, so indices match an fp32 reference; otherwise the bf16-operand scores are returned.
 * idx_base is added to the gallery row index (gallery shards).  out_scores f32 [Q,k], out_idx i64 [Q,k].
 * Rows with fewer than k gallery entries are padded with score -INFINITY and index -1. */
CVPCE_API size_t cvpce_sim_topk_workspace_bytes(int Q, int G, int D, int k);
CVPCE_API int cvpce_sim_topk(const void* q_bf16, const void* g_bf16, int Q, int G, int D, int k, int64_t idx_base,
                   const float* q_f32, const float* g_f32, const float* q_inv_norm, const float* g_inv_norm,
                   float* out_scores, int64_t* out_idx, void* workspace, size_t workspace_bytes, void* stream);
/* Gallery-sharded search (e: one process per GPU, each holding a shard): the same computation, but the final [Q,k] block of
 * this rank is stored by the selection kernel's epilogue DIRECTLY into every peer's gather buffer -- peer_scores[p] /
 * peer_idx[p] [host arrays of n_peers <= 16 device pointers] address this rank's [Q,k] slot inside rank p's [W,Q,k] buffers
 * (peer memory mapped over NVLink, e.g. torch symmetric memory / cudaIpc; the own buffer is one of the n_peers).  The
 * all-gather that follows the search in the reference formulation (SURVEY §8 e) thus happens row by row while other rows
 * are still being re-scored; the caller only needs a cross-rank barrier before cvpce_topk_merge reads the gather buffers. */
CVPCE_API int cvpce_sim_topk_scatter(const void* q_bf16, const void* g_bf16, int Q, int G, int D, int k, int64_t idx_base,
                           const float* q_f32, const float* g_f32, const float* q_inv_norm, const float* g_inv_norm,
                           float* const* peer_scores, int64_t* const* peer_idx, int n_peers,
                           void* workspace, size_t workspace_bytes, void* stream);
/* CUDA-core fp32 path (any D >= 1, k <= 64): exact fp32 dot products, same ordering contract; used for tiny shapes and
 * as the on-device cross-check of the tensor-core path.  normalize=1 divides by max(norm,eps) of both rows (distance()). */
CVPCE_API int cvpce_sim_topk_f32(const float* q, const float* g, int Q, int G, int D, int k, int64_t idx_base,
                       int normalize, float eps, float* out_scores, int64_t* out_idx, void* stream);
/* Merge W per-shard results [W,Q,k] (scores f32, global idx i64) into the overall top-k per row with the same
 * ordering contract (used after the NCCL all-gather of the gallery-sharded search). */
CVPCE_API int cvpce_topk_merge(const float* scores, const int64_t* idx, int W, int Q, int k,
                     float* out_scores, int64_t* out_idx, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CVPCE_B200_H */
