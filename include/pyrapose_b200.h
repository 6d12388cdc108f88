/* pyrapose_b200 — C-ABI of the B200 (sm_100a) implementation of PyraPose's data-parallel hot path.
 *
 * Conventions
 *  - Every function returns 0 on success or a negative ppb_status; ppb_last_error() then holds a
 *    message (thread-local). The Python host layer raises ValueError for PPB_ERR_INVALID and
 *    RuntimeError otherwise.
 *  - All buffers are caller-owned DEVICE pointers unless a parameter name ends in _host.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream). Calls are
 *    asynchronous with respect to the host unless stated otherwise.
 *  - No entry point has a CPU fallback: without a CUDA device they fail with PPB_ERR_CUDA.
 *
 * Each declaration cites the reference interface (file:line under sThalham/PyraPose) it replaces.
 */
#ifndef PYRAPOSE_B200_H_
#define PYRAPOSE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PPB_API __attribute__((visibility("default")))

typedef enum ppb_status {
  PPB_OK = 0,
  PPB_ERR_INVALID = -1,
  PPB_ERR_CUDA = -2,
  PPB_ERR_FAULT = -3,
  PPB_ERR_STATE = -4
} ppb_status;

/* ------------------------------------------------------------------ library */
PPB_API int ppb_version(void);
PPB_API const char* ppb_last_error(void);
/* Number of kernels this library has launched in the calling process (bench.py's gpu_launches). */
PPB_API long long ppb_launch_count(void);

/* ------------------------------------------------------------------ convolution (tcgen05 implicit GEMM)
 * Replaces keras.layers.Conv2D(+BatchNormalization(freeze)+Activation('relu')+Add) as used by
 * PyraPose/models/retinanet.py:29-46,73-88,118-126,192-212 and keras-resnet's bottleneck_2d
 * (PyraPose/models/resnet.py:87).
 *
 * x        : bf16 [batch, h+2, w+2, cin]   "padded-flat": one-pixel zero border around every image
 * weight   : bf16 [cout_pad, ksize*ksize*cin], K index = (r*ksize + s)*cin + c  (Keras HWIO -> O,HWI);
 *            cout_pad = cout rounded up to a multiple of 16, padding rows zero
 * bias     : f32 [cout] or NULL (frozen BatchNorm is folded into weight and bias by the caller)
 * residual : bf16 [batch, h+2, w+2, cout] or NULL, added before the ReLU
 * y        : bf16 [batch, h+2, w+2, cout], border written as zeros
 * ksize    : 1 or 3; stride 1; 'same' zero padding. cin % 64 == 0, cout % 16 == 0.
 */
PPB_API int ppb_conv2d_bf16(const void* x, const void* weight, const float* bias, const void* residual, void* y,
                            int batch, int h, int w, int cin, int cout, int ksize, int relu, void* stream);

/* Weight gradient of the same convolution (training step; the filter-gradient op TensorFlow runs under
 * model.fit_generator, PyraPose/bin/train.py:381-390):
 *   dw[co][(r*ksize+s)*cin + c] += sum over pixels of  g[b,y,x,co] * x[b,y+r-k/2,x+s-k/2,c]
 * x : bf16 [batch,h+2,w+2,cin] padded-flat (zero border), g : bf16 [batch,h+2,w+2,g_ld] padded-flat with ZERO border,
 * g_ld >= cout and a multiple of 8 (columns >= cout are ignored); dw : f32 [cout, ksize*ksize*cin], ACCUMULATED into
 * (the caller zeroes it). cin % 64 == 0.
 */
PPB_API int ppb_conv2d_wgrad_bf16(const void* x, const void* g, float* dw, int batch, int h, int w, int cin, int cout,
                                  int g_ld, int ksize, void* stream);

/* Data gradient of the same convolution (the input-gradient op of Conv2D under model.fit_generator): the tcgen05
 * implicit-GEMM kernel run over g with transposed weights and negated tap offsets,
 *   dx[b,y,x,ci] = sum_{r,s,co} g[b, y-(r-k/2), x-(s-k/2), co] * w[co][(r*k+s)*cin + ci]   (+ residual) (masked)
 * g        : bf16 [batch,h+2,w+2,g_ld] padded-flat with ZERO border; g_ld >= cout, multiple of 64, columns >= cout zero
 * weight_t : bf16 [cin, ksize*ksize*g_ld] with weight_t[ci][(r*k+s)*g_ld + co] = w[co][(r*k+s)*cin + ci]
 * residual : bf16 [batch,h+2,w+2,cin] or NULL, added (skip-connection gradient); may alias dx
 * relu_mask: bf16 [batch,h+2,w+2,cin] or NULL: the forward activation; dx is zeroed where it is <= 0 (ReLU backward)
 * dx       : bf16 [batch,h+2,w+2,cin], border written as zeros. cin % 64 == 0.
 */
PPB_API int ppb_conv2d_dgrad_bf16(const void* g, const void* weight_t, const void* residual, const void* relu_mask,
                                  void* dx, int batch, int h, int w, int cin, int cout, int g_ld, int ksize,
                                  void* stream);

/* ------------------------------------------------------------------ anchor <-> ground-truth IoU (float64, bit-exact)
 * ppb_compute_overlap_f64 replaces compute_overlap(boxes, query_boxes) -> overlaps
 * (PyraPose/utils/compute_overlap.pyx:13-53; call sites PyraPose/utils/anchors.py:310, utils/eval.py:201).
 *   boxes    : f64 [n,4] (boxes_shared=1: one set for every batch item) or [batch,n,4]
 *   query    : f64 [batch,k,4]
 *   overlaps : f64 [batch,n,k], fully written
 */
PPB_API int ppb_compute_overlap_f64(const double* boxes, const double* query, double* overlaps, long long n, int k,
                                    int batch, int boxes_shared, void* stream);

/* Fused compute_gt_annotations (PyraPose/utils/anchors.py:290-318) without materialising the matrix.
 *   query       : f64 [batch,k_max,4]; query_count: i32 [batch] valid boxes per image, or NULL (= k_max)
 *   argmax      : i64 [batch,n]   (numpy.argmax: first maximum)
 *   positive    : u8  [batch,n]   max >= positive_overlap
 *   ignore      : u8  [batch,n]   max >  negative_overlap && !positive
 *   max_overlap : f64 [batch,n] or NULL
 */
PPB_API int ppb_gt_assign_f64(const double* boxes, const double* query, const int* query_count, int n, int k_max,
                              int batch, int boxes_shared, double negative_overlap, double positive_overlap,
                              long long* argmax, unsigned char* positive, unsigned char* ignore,
                              double* max_overlap, void* stream);

/* Per-anchor part of anchor_targets_bbox (PyraPose/utils/anchors.py:72-287: states :138-145, one-hot labels :148,
 * box3D_transform of the matched object's projected corners :267 / :515-559, out-of-image ignore :279-285), from the
 * outputs of ppb_gt_assign_f64. All DEVICE pointers.
 *   anchors f64 [n,4]; argmax i64 / positive u8 / ignore u8 [batch,n]; labels i32 [batch,k_max];
 *   boxes16 f64 [batch,k_max,16] projected 3D-box corners per object; counts i32 [batch] objects per image;
 *   image_wh i32 [batch,2] (width, height; width < 0: image without a shape -> no out-of-image rule)
 *   out_regression f32 [batch,n,17], out_labels f32 [batch,n,num_classes+1]   (last column = anchor state)
 */
PPB_API int ppb_anchor_targets(const double* anchors, const long long* argmax, const unsigned char* positive,
                               const unsigned char* ignore, const int* labels, const double* boxes16,
                               const int* counts, const int* image_wh, int n, int k_max, int batch, int num_classes,
                               float* out_regression, float* out_labels, void* stream);
/* Mask targets of anchor_targets_bbox (anchors.py:152-164): masks u8 [batch,height,width] instance-id images, resized
 * to the P3 grid (mask_h x mask_w) with PIL's NEAREST rule; an object marks its cells when it covers more than one.
 *   mask_ids / labels i32 [batch,k_max]; cells i32 [batch,k_max] workspace; out_mask f32 [batch,mask_h*mask_w,C+1] */
PPB_API int ppb_mask_targets(const unsigned char* masks, int height, int width, const int* mask_ids, const int* labels,
                             const int* counts, int k_max, int batch, int mask_h, int mask_w, int num_classes,
                             int* cells, float* out_mask, void* stream);

/* ------------------------------------------------------------------ anchors
 * One pyramid level: out[(y*fw + x)*num_base + a] = base[a] + ((x|y)+0.5)*stride.
 * f64: PyraPose/utils/anchors.py:415-444 (shift), level loop :402-412.
 * f32: PyraPose/backend/common.py:93-116 as used by the Anchors layer (PyraPose/layers/_misc.py:24-93).
 * base: [num_base,4] DEVICE (the nine generate_anchors() boxes, computed by the host as anchors.py:447-478 does).
 */
PPB_API int ppb_anchors_shift_f64(const double* base, int num_base, int fh, int fw, double stride, double* out,
                                  void* stream);
PPB_API int ppb_anchors_shift_f32(const float* base, int num_base, int fh, int fw, float stride, float* out,
                                  void* stream);

/* ------------------------------------------------------------------ per-anchor decode (layer API)
 * RegressBoxes3D / DenormBoxes3D -> box3D_transform_inv / box3D_denormalization
 * (PyraPose/layers/_misc.py:165-256, PyraPose/backend/common.py:25-90).
 *   anchors : f32 [anchors_per_image,4] (anchors_batched=0) or [batch,anchors_per_image,4]
 *   deltas  : f32 [batch,anchors_per_image,16];  out: same shape
 *   hull    : optional f32 [batch,anchors_per_image,4] axis-aligned hull of the 8 decoded corners (or NULL)
 *   mean_host / std_host : 16 floats on the HOST, or NULL for the defaults 0 / 0.2
 */
PPB_API int ppb_box3d_transform_inv_f32(const float* anchors, const float* deltas, float* out, float* hull, int batch,
                                        long long anchors_per_image, int anchors_batched, const float* mean_host,
                                        const float* std_host, int denormalize, void* stream);
/* RegressBoxes -> bbox_transform_inv (PyraPose/layers/_misc.py:118-162, backend/common.py:6-22); 4 values. */
PPB_API int ppb_bbox_transform_inv_f32(const float* anchors, const float* deltas, float* out, int batch,
                                       long long anchors_per_image, int anchors_batched, const float* mean_host,
                                       const float* std_host, void* stream);
/* ClipBoxes (PyraPose/layers/_misc.py:259-280): x to [0,width], y to [0,height]. boxes f32 [total_boxes,4]. */
PPB_API int ppb_clip_boxes_f32(const float* boxes, float* out, long long total_boxes, float width, float height,
                               void* stream);
/* UpsampleLike (PyraPose/layers/_misc.py:96-115 -> backend.resize_images(..., method='nearest'),
 * backend/tf_backend.py:28-35): source f32 [batch, src_h, src_w, channels] -> out [batch, dst_h, dst_w, channels],
 * out[y, x] = source[min(floor((y + 0.5) src_h / dst_h), src_h - 1), ...]. (Inside the engine the same resize is
 * fused into the pyramid's add kernel; this is the stand-alone layer.) */
PPB_API int ppb_upsample_like_f32(const float* source, float* out, int batch, int src_h, int src_w, int dst_h,
                                  int dst_w, int channels, void* stream);

/* ------------------------------------------------------------------ FilterDetections
 * Replaces FilterDetections.call / filter_detections (PyraPose/layers/filter_detections.py:21-118, 153-191).
 *   boxes [batch,n,4], boxes3D [batch,n,16], classification [batch,n,classes]   f32 (boxes*: 16-byte aligned)
 *   out_boxes [batch,D,4], out_boxes3D [batch,D,16], out_scores [batch,D] f32, out_labels [batch,D] i32,
 *   out_indices [batch,D] i32 or NULL (selected anchor index, for gathering `other` tensors); all -1 padded.
 *   workspace: >= ppb_filter_workspace_bytes(...) bytes of device memory. D = max_detections <= 1024.
 */
PPB_API size_t ppb_filter_workspace_bytes(int batch, int n, int num_classes, int class_specific, int max_detections);
PPB_API int ppb_filter_detections(const float* boxes, const float* boxes3D, const float* classification, int batch,
                                  int n, int num_classes, int class_specific, int nms, float score_threshold,
                                  float nms_threshold, int max_detections, float* out_boxes, float* out_boxes3D,
                                  float* out_scores, int* out_labels, int* out_indices, void* workspace,
                                  size_t workspace_bytes, void* stream);
/* ppb_filter_detections with NO allocation anywhere: no workspace argument, no cudaMalloc / cudaFree / allocating
 * library call inside. No counterpart in the reference — added for callers that
 * cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The scratch is
 * a file-scope `static __device__` array of ppb_filter_static_capacity_bytes() = 16 MiB inside the library (held by
 * every GPU that loads it, until the process exits). Shapes whose ppb_filter_workspace_bytes(...) exceeds the capacity
 * are rejected with PPB_ERR_INVALID and a message naming both sizes (one 480x640 image — 56 700 anchors — fits with up
 * to 31 classes class-specific; batch 2 with 15; class-agnostic up to batch 17). Same inputs, outputs and results as
 * ppb_filter_detections. All calls on a device share the one array, which keeps the previous call's contents (no part
 * of it needs to start at zero): calls that overlap on the GPU — different streams without an ordering between them,
 * or different host threads — race; serialise them. */
PPB_API size_t ppb_filter_static_capacity_bytes(void);
PPB_API int ppb_filter_detections_static(const float* boxes, const float* boxes3D, const float* classification,
                                         int batch, int n, int num_classes, int class_specific, int nms,
                                         float score_threshold, float nms_threshold, int max_detections,
                                         float* out_boxes, float* out_boxes3D, float* out_scores, int* out_labels,
                                         int* out_indices, void* stream);
/* dst[b,d,:] = indices[b,d] >= 0 ? src[b, indices[b,d], :] : -1   (filter_detections.py:97,108 for `other`). */
PPB_API int ppb_gather_rows_f32(const float* src, const int* indices, float* dst, int batch, int n,
                                int max_detections, int row_elems, void* stream);

/* ------------------------------------------------------------------ evaluator vote filter
 * Replaces `np.where(scores[0,:,c] > thr)` + `boxes3D[0, idx, :]` (PyraPose/utils/linemod_eval.py:317-319, 427-430;
 * PyraPose_ROS_wrapper/scripts/pyrapose_node.py:584-586, 638, 648).
 * Step 1: mask u32 [batch,classes,ceil(n/32)] and counts i32 [batch,classes] (counts may be NULL).
 * Step 2: offsets i64 [batch*classes+1] = exclusive prefix sum of counts (caller computes it);
 *         indices i64 [total] ascending per (image,class); votes f32 [total,16]; confidence f32 [batch,classes]|NULL.
 */
PPB_API int ppb_score_mask(const float* scores, int batch, int n, int num_classes, float threshold, uint32_t* mask,
                           int* counts, void* stream);
PPB_API int ppb_vote_gather(const uint32_t* mask, const float* scores, const float* boxes3D, int batch, int n,
                            int num_classes, const long long* offsets, long long* indices, float* votes,
                            float* confidence, void* stream);

/* ------------------------------------------------------------------ pose from corner votes (the step after the path)
 * Replaces, for ALL (image, class) groups of a batch in one call, the evaluators' per-class
 *   cv2.solvePnPRansac(obj_points (k*8,1,3), est_points (k*8,1,2), K, iterationsCount=300, reprojectionError=5.0,
 *                      confidence=0.99, flags=cv2.SOLVEPNP_ITERATIVE)
 * (PyraPose/utils/linemod_eval.py:479-485; PyraPose_ROS_wrapper/scripts/pyrapose_node.py:652-659) on the votes that
 * ppb_vote_gather compacted. Hypotheses = POSIT on single votes (a vote holds all 8 box corners), scored against every
 * correspondence of the group (inlier <=> reprojection error <= reprojection_error px), best one refined on its
 * inliers by damped Gauss-Newton on the reprojection error. All pointers DEVICE.
 *   votes      f32 [total_votes,16]   (x0,y0,...,x7,y7 per vote; 16-byte aligned)
 *   offsets    i64 [groups+1]         votes of group g are rows offsets[g] .. offsets[g+1]
 *   obj_points f32 [groups,8,3]       the 3D box corners of the group's class, in the votes' corner order
 *   cam        f32 [groups,4]         fx, fy, cx, cy
 *   out_rt     f32 [groups,6]         Rodrigues rotation vector, translation (zeros when no pose)
 *   out_inliers i32 [groups]          inlier correspondences of the best hypothesis; 0 = no pose (retval False):
 *                                     fewer than min_inliers (cv2 needs >= 5 model points)
 *   out_rmse   f32 [groups] or NULL   RMS reprojection error over the inliers after refinement
 */
PPB_API size_t ppb_pnp_workspace_bytes(int groups, long long total_votes, int iterations);
PPB_API int ppb_pnp_ransac_votes(const float* votes, const long long* offsets, const float* obj_points,
                                 const float* cam, int groups, long long total_votes, int iterations,
                                 float reprojection_error, int min_inliers, float* out_rt, int* out_inliers,
                                 float* out_rmse, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ PFPN inference engine
 * Replaces `model.predict_on_batch(images)` for the models built by Backbone.retinanet(num_classes)
 * (PyraPose/models/resnet.py:79-110, PyraPose/models/retinanet.py:260-299) and models.convert_model
 * (PyraPose/models/__init__.py:74-76 -> retinanet_bbox, retinanet.py:302-335).
 *
 * The handle owns weights (bf16, BatchNorm folded), the activation arena and a CUDA graph of the
 * whole forward pass. Thread-compatible: one caller at a time per handle; calls cudaSetDevice on entry.
 */
typedef struct ppb_engine ppb_engine;

PPB_API int ppb_engine_create(ppb_engine** out, int device, int max_batch, int height, int width, int num_classes,
                              int num_anchors);
/* precision 0: the production path (bf16 operands, fp32 accumulation; feature maps / scores / corners within rel.
 * 1e-2 of the fp32 reference). precision 1: the "fp32 path" of BASELINE.json north_star (tolerance 1e-4): every
 * activation and weight is carried as three bf16 parts (24 significant bits) and every product is the six tcgen05
 * MMAs whose terms reach 2^-16 of it, summed in the fp32 accumulator; the stem, Add / UpsampleLike, MaxPool and the
 * sigmoid run in IEEE fp32. Inference only (ppb_engine_train_bind returns PPB_ERR_STATE); about 6x the tensor work and
 * 3x the activation bytes. Replaces running the reference's float32 Keras graph
 * (PyraPose/models/retinanet.py:260-335) where parity tighter than bf16 matters. */
PPB_API int ppb_engine_create_ex(ppb_engine** out, int device, int max_batch, int height, int width, int num_classes,
                                 int num_anchors, int precision);
PPB_API void ppb_engine_destroy(ppb_engine* engine);

/* Layer table (Keras layer names; keras-resnet 0.1.0 for the backbone). */
PPB_API int ppb_engine_num_layers(const ppb_engine* engine);
PPB_API int ppb_engine_layer_info(const ppb_engine* engine, int index, char* name, int name_cap, int* ksize,
                                  int* cin, int* cout, int* stride, int* has_bn, int* has_bias);
/* Weights in Keras layout, all HOST pointers: kernel f32 [k,k,cin,cout] (HWIO); bias f32 [cout] or NULL;
 * frozen BatchNorm gamma/beta/moving_mean/moving_variance f32 [cout] (NULL for layers without BN). */
PPB_API int ppb_engine_set_layer(ppb_engine* engine, const char* name, const float* kernel_hwio_host,
                                 const float* bias_host, const float* bn_gamma_host, const float* bn_beta_host,
                                 const float* bn_mean_host, const float* bn_var_host, float bn_eps);
/* Anchors layer parameters (PyraPose/layers/_misc.py:24-93): base anchors f32 [3, num_anchors, 4] already cast to
 * float32 as keras.backend.variable does, strides f32 [3]; RegressBoxes3D mean/std f32 [16] or NULL (0 / 0.2). */
PPB_API int ppb_engine_set_anchors(ppb_engine* engine, const float* base_anchors_host, const float* strides_host,
                                   const float* mean16_host, const float* std16_host);
/* Production arena: the backbone's bottleneck intermediates share a handful of buffers (every branch2a / branch2b
 * output one buffer each, block outputs ping-pong between two, one projection-shortcut and one subsample buffer)
 * instead of one buffer per layer; 6.4 -> 3.7 GB at batch 32, 480x640. Call before the first ppb_engine_bind.
 * Inference plans only (a training plan needs every activation and ignores it); ppb_engine_activation then returns
 * only tensors that still hold their value when the pass ends. No reference counterpart (TensorFlow's allocator). */
PPB_API int ppb_engine_set_arena_alias(ppb_engine* engine, int enable);

/* Input staging (the step before the path: read_image_bgr -> preprocess_image, PyraPose/utils/image.py:35-62;
 * call sites utils/linemod_eval.py:265-267, preprocessing/generator.py:320-336): with enable=1 the `images` pointer of
 * the next ppb_engine_bind / ppb_engine_train_bind is read as RAW uint8 BGR pixels [batch,H,W,3] and
 * preprocess_image(mode='caffe') — float32(pixel) - mean[c], mean3 = {103.939, 116.779, 123.68} when NULL — is fused
 * into the stem's loader. Bit-identical activations to the float32 path, a quarter of the host->device bytes. */
PPB_API int ppb_engine_set_input_u8(ppb_engine* engine, int enable, const float* mean3_host);
/* Binds caller-owned DEVICE buffers and captures the graph for `batch` images:
 *   images    f32 [batch,H,W,3]  BGR, ImageNet-mean-subtracted (PyraPose/utils/image.py:35-62)
 *   out_box3d f32 [batch,N,16]   decode=1: decoded corners ('boxes3D'); decode=0: raw regression ('3Dbox')
 *   out_cls   f32 [batch,N,C]    sigmoid scores ('cls')
 *   out_mask  f32 [batch,H/8*W/8,C] or NULL to skip the mask head
 *   out_hull  f32 [batch,N,4] or NULL: axis-aligned hull of the decoded corners (decode=1 only)
 * N = 9 * sum over levels P3..P5 of ceil(H/2^l)*ceil(W/2^l). */
PPB_API int ppb_engine_bind(ppb_engine* engine, int batch, const float* images, float* out_box3d, float* out_cls,
                            float* out_mask, float* out_hull, int decode, void* stream);
PPB_API int ppb_engine_run(ppb_engine* engine, void* stream);
/* The forward pass in three parts cut where an output is complete (0: through the classification head -> 'cls';
 * 1: regression head -> 'boxes3D' / '3Dbox'; 2: mask head -> 'mask'). Parts 0, 1, 2 in order on one stream equal
 * ppb_engine_run; model.predict_on_batch (PyraPose/utils/linemod_eval.py:303 returns all three arrays on the host)
 * uses the cuts to copy each output to the host while the remaining heads compute. */
PPB_API int ppb_engine_run_part(ppb_engine* engine, int part, void* stream);
PPB_API int ppb_engine_run_eager(ppb_engine* engine, void* stream);
PPB_API int ppb_engine_kernels_per_run(const ppb_engine* engine);
PPB_API size_t ppb_engine_device_bytes(const ppb_engine* engine);
PPB_API int ppb_engine_activation(ppb_engine* engine, const char* name, float* out_host, int* shape4_host);
/* Per-launch device times of one eager forward pass (CUDA events); kinds: 0 conv, 1 stem-s2d, 2 max-pool,
 * 3 parity planes, 4 add. Fills min(cap, kernels_per_run) entries of the HOST arrays. */
/* Name (layer name for convolutions) and algorithmic FLOPs of launch `index` of the bound plan. */
PPB_API int ppb_engine_op_info(const ppb_engine* engine, int index, char* name, int name_cap, double* flops);
/* Diagnostic (tools/trace_conv.py): cycle-stamped trace of CTA `cta` of convolution launch `index` of the bound plan —
 * every barrier hand-off of its TMA producer, MMA issuer, residual producer and epilogue groups. trace_dev: DEVICE
 * buffer of 8 * cap 8-byte entries; info_out_host: 12 ints (M tiles, N tiles, block_n, taps, cin blocks, stages, ring
 * slots, pair, slab3, bres, staged, epilogue groups) or NULL. No reference counterpart (TensorFlow's profiler). */
PPB_API int ppb_engine_trace_op(ppb_engine* engine, int index, int cta, void* trace_dev, int cap, int* info_out_host,
                                void* stream);
PPB_API int ppb_engine_profile(ppb_engine* engine, void* stream, float* ms_out_host, int* kind_out_host, int cap);

/* ------------------------------------------------------------------ input staging
 * preprocess_image('caffe') + resize_image + compute_inputs of ONE image in one kernel
 * (PyraPose/utils/image.py:35-62, 281-323; PyraPose/preprocessing/generator.py:320-336):
 *   src  DEVICE raw pixels [src_h,src_w,3], uint8 (src_is_u8) or float32, BGR
 *   dst  DEVICE f32 [pad_h,pad_w,3]: the image's slot of the zero-padded batch tensor; rows < dst_h / cols < dst_w
 *        receive the bilinearly resized (cv2.resize(img, None, fx, fy), OpenCV's float32 INTER_LINEAR arithmetic),
 *        mean-subtracted pixels, everything else zeros
 *   dst_h = cvRound(src_h * fy), dst_w = cvRound(src_w * fx) (the caller computes them, like cv::resize does)
 *   mean3_host: HOST float[3] subtracted per channel, or NULL for none. */
PPB_API int ppb_resize_preprocess_f32(const void* src, int src_is_u8, int src_h, int src_w, float* dst, int dst_h,
                                      int dst_w, int pad_h, int pad_w, double fx, double fy, const float* mean3_host,
                                      void* stream);

/* ------------------------------------------------------------------ training step (BASELINE.json config 4)
 * Replaces one batch of `training_model.fit_generator(...)` (PyraPose/bin/train.py:381-390) for the model compiled at
 * train.py:95-102: loss = {'3Dbox': orthogonal_l1(), 'cls': focal(), 'mask': focal()} (PyraPose/losses.py:22-68,
 * 321-408) + the l2(0.001) kernel regulariser of the regression sub-model (PyraPose/models/retinanet.py:108),
 * optimizer = keras.optimizers.Adam(lr, clipnorm) (Keras 2.3.1: clip by the GLOBAL gradient norm). Trainable set as in
 * PyraPose/models/resnet.py:87,100-103: BatchNorm frozen everywhere, conv1 and stage 2 frozen.
 *
 * ppb_engine_train_bind: binds caller-owned DEVICE buffers and captures forward + losses + backward as one CUDA graph.
 *   images f32 [batch,H,W,3]; y_reg f32 [batch,N,17]; y_cls f32 [batch,N,C+1]; y_mask f32 [batch,H/8*W/8,C+1]
 *   (the three arrays anchor_targets_bbox returns, PyraPose/utils/anchors.py:72-287; last column = anchor state)
 *   out_reg f32 [batch,N,16] raw '3Dbox'; out_cls f32 [batch,N,C]; out_mask f32 [batch,H/8*W/8,C]  (model outputs)
 * ppb_engine_train_forward_backward: one launch; afterwards the flat fp32 gradient buffer (ppb_engine_train_grads)
 *   holds dLoss/dW of every trainable tensor (w.r.t. the BatchNorm-folded kernels; the fold is undone in _apply).
 *   A data-parallel caller all-reduces (sums) that buffer over NCCL and passes grad_scale = 1/world_size to _apply.
 * ppb_engine_train_apply: regulariser gradient, global-norm clip, Adam moments and step, refresh of the bf16 forward
 *   and data-gradient weight matrices.
 * ppb_engine_train_losses: synchronises `stream`; out5 = {3Dbox, cls, mask, l2 regulariser, squared gradient norm}
 *   (the last two are valid after _apply), counts3 = positives of the three outputs.
 */
PPB_API int ppb_engine_train_bind(ppb_engine* engine, int batch, const float* images, const float* y_reg,
                                  const float* y_cls, const float* y_mask, float* out_reg, float* out_cls,
                                  float* out_mask, void* stream);
PPB_API int ppb_engine_train_forward_backward(ppb_engine* engine, int eager, void* stream);
/* Data-parallel step (keras.utils.multi_gpu_model, PyraPose/bin/train.py:82-86).
 * ppb_engine_train_set_bucket_bytes: call before ppb_engine_train_bind. With bucket_bytes > 0 the step is captured
 *   as ppb_engine_train_num_segments graphs, cut after the launch that completes each ~bucket_bytes slice of the flat
 *   gradient buffer (the backward pass fills it from the end — heads — to the beginning — res3). Segment i leaves
 *   [grad_offset, grad_offset + grad_count) final, so the caller starts the NCCL all-reduce of that slice on its
 *   communication stream and launches segment i + 1 at once (grad_count may be 0 for the trailing segment).
 * ppb_engine_train_count: the positives of the three outputs (the loss normalisers of losses.py:61-66, 395-397) into
 *   the device int[4] that ppb_engine_train_counts returns; they depend on the targets only. multi_gpu_model computes
 *   each loss once over the CONCATENATED replica outputs, i.e. normalises by the positives of the global batch: a
 *   data-parallel caller sums the counters over the replicas between _count and the first segment, then sums (not
 *   averages) the gradients. If _count was not called explicitly the first segment / _forward_backward runs it. */
PPB_API int ppb_engine_train_set_bucket_bytes(ppb_engine* engine, long long bucket_bytes);
PPB_API int ppb_engine_train_count(ppb_engine* engine, void* stream);
PPB_API int ppb_engine_train_counts(ppb_engine* engine, int** counts_dev);
PPB_API int ppb_engine_train_num_segments(const ppb_engine* engine);
PPB_API int ppb_engine_train_segment_info(const ppb_engine* engine, int segment, long long* grad_offset,
                                          long long* grad_count, int* n_launches);
PPB_API int ppb_engine_train_run_segment(ppb_engine* engine, int segment, void* stream);
PPB_API int ppb_engine_train_grads(ppb_engine* engine, float** grads, long long* count);
PPB_API int ppb_engine_train_apply(ppb_engine* engine, float lr, float beta1, float beta2, float epsilon,
                                   float clipnorm, float grad_scale, void* stream);
PPB_API int ppb_engine_train_losses(ppb_engine* engine, double* out5_host, int* counts3_host, void* stream);
/* The same read-back without a host synchronisation: the per-batch `logs` Keras hands to the callbacks of
 * fit_generator (PyraPose/bin/train.py:381-390) arrive one step late, so the GPU never waits for the host. Copies the
 * five loss doubles (3Dbox, cls, mask, l2, squared gradient norm) and optionally the three positive counts into
 * PINNED host memory with cudaMemcpyAsync on `stream`; valid once work queued on `stream` up to here has completed. */
PPB_API int ppb_engine_train_losses_async(ppb_engine* engine, double* out5_host, int* counts3_host, void* stream);
/* Offsets (elements) of a layer's kernel [cout][k*k*cin] and bias [cout] inside the flat buffers; -1 if absent. */
PPB_API int ppb_engine_train_param_info(const ppb_engine* engine, const char* layer, long long* kernel_offset,
                                        long long* bias_offset, int* trainable);
/* Copies a range of a flat buffer to the host (synchronises): which = 0 gradients, 1 master weights, 2 Adam m, 3 v. */
PPB_API int ppb_engine_train_read(ppb_engine* engine, int which, long long offset, long long count, float* out_host);
PPB_API int ppb_engine_train_num_ops(const ppb_engine* engine);
PPB_API int ppb_engine_train_op_info(const ppb_engine* engine, int index, char* name, int name_cap, double* flops);
/* Per-launch device times (ms) of one eager training step, in op order; synchronises `stream`. */
PPB_API int ppb_engine_train_profile(ppb_engine* engine, void* stream, float* ms_out_host, int cap);

/* ------------------------------------------------------------------ peer-to-peer gradient exchange
 * Data-parallel training step (the reference: keras.utils.multi_gpu_model + fit_generator, PyraPose/bin/train.py:82-102,
 * 381-390). One process per GPU on one node. Every replica owns a staging buffer with one slot
 * per peer; a finished gradient bucket is pushed into the peers' slots with asynchronous copies (copy engines over
 * NVLink: no SM is taken from the backward pass, unlike an overlapped NCCL all-reduce), and after the backward pass
 * ppb_p2p_sum_f32 adds the slices in rank order (bit-identical on every replica).
 *   ppb_p2p_alloc / _free      device memory that can be exported (plain cudaMalloc on the current device, zeroed)
 *   ppb_p2p_export             64-byte CUDA IPC handle of such a buffer, to be sent to the peers (any host channel)
 *   ppb_p2p_import / _release  map / unmap a peer's buffer from its handle (enables peer access lazily)
 *   ppb_p2p_copy               asynchronous copy between local and peer-mapped device memory on `stream`
 *   ppb_p2p_sum_f32            out[i] = srcs[0][i] + srcs[1][i] + ... (srcs_host: HOST array of n_srcs device pointers,
 *                              16-byte aligned; out may be one of them) */
PPB_API int ppb_p2p_alloc(size_t bytes, void** dev_ptr);
PPB_API int ppb_p2p_free(void* dev_ptr);
PPB_API int ppb_p2p_export(const void* dev_ptr, unsigned char* handle64);
PPB_API int ppb_p2p_import(const unsigned char* handle64, void** peer_ptr);
PPB_API int ppb_p2p_release(void* peer_ptr);
PPB_API int ppb_p2p_copy(void* dst, const void* src, size_t bytes, void* stream);
PPB_API int ppb_p2p_sum_f32(float* out, const float* const* srcs_host, int n_srcs, long long count, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PYRAPOSE_B200_H_ */
