/*
 * tcb200.h — C ABI of libtcb200.so: the B200 (sm_100a) pre/post-processing hot path of
 * niqbal996/triton_client, rebuilt from scratch.
 *
 * Conventions:
 *   - every entry point returns int: 0 = ok, <0 = one of TCB_ERR_*; the text of the last failure is
 *     tcb_last_error(ctx). Nothing here throws, aborts or calls exit() (the reference's own native
 *     code does: OpenPCDet pcdet/ops/iou3d_nms/src/iou3d_nms.cpp:14-38 — deliberately not mirrored).
 *   - all `*_dev` pointers are CUDA device pointers on the context's device (e.g. torch
 *     Tensor.data_ptr()); `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *   - inputs are borrowed and never written; outputs are caller-allocated.
 *   - after tcb_create no entry point allocates: scratch comes out of the context's workspace and a
 *     call that needs more than the workspace holds fails with TCB_ERR_WORKSPACE.
 *   - calls on one context are issued in order on the caller's stream; one frame/batch in flight per
 *     context (the reference keeps one frame in flight per client object, communicator/ros_inference.py:117-175).
 *
 * Paths quoted below are relative to the reference repository root; OPCD/ abbreviates
 * utils/export_models/OpenPCDet/.
 */
#ifndef TCB200_H
#define TCB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct tcb_ctx tcb_ctx;

enum {
    TCB_OK = 0,
    TCB_ERR_ARG = -1,       /* invalid argument */
    TCB_ERR_CUDA = -2,      /* a CUDA runtime call failed; see tcb_last_error */
    TCB_ERR_WORKSPACE = -3, /* context workspace too small for this call */
    TCB_ERR_ALLOC = -4,     /* allocation failed in tcb_create */
    TCB_ERR_LIMIT = -5      /* size beyond a compiled-in limit */
};

/* ---- context ------------------------------------------------------------------------------- */

/* Creates a context on `device` owning `workspace_bytes` of device scratch (one cudaMalloc, here and
 * nowhere else). The pre/post objects of the reference are built once per process
 * (communicator/ros_inference.py:39-41): that is where this is called. */
int tcb_create(int device, size_t workspace_bytes, tcb_ctx** out);
int tcb_destroy(tcb_ctx* ctx);
/* Last error text of `ctx` (or of the calling thread when ctx is NULL, e.g. after a failed tcb_create). */
const char* tcb_last_error(const tcb_ctx* ctx);
size_t tcb_workspace_bytes(const tcb_ctx* ctx);
const char* tcb_version(void);
/* Number of kernels launched through this context since creation (bench.py's gpu_launches). */
long long tcb_launch_count(const tcb_ctx* ctx);

/* ---- 2D preprocess -------------------------------------------------------- */

enum { TCB_RESIZE_STRETCH = 0, TCB_RESIZE_LETTERBOX = 1, TCB_RESIZE_PAD = 2 };

/*
 * Fused u8 HWC -> [resize] -> [channel swap] -> x/div - sub[c] -> f32 NCHW, n frames per launch.
 * Replaces, in one pass: cv2.resize(img,(W,H)) INTER_LINEAR (communicator/ros_inference.py:140,
 * bag_inference2d.py:110, evaluate_inference.py:305, utils/preprocess.py:170), the BGR->RGB flip
 * (ros_inference.py:131, utils/preprocess.py:171), Yolov5preprocess.image_adjust
 * (clients/preprocess/yolov5_preprocess.py:12-24; div=255), FCOSpreprocess.image_adjust
 * (clients/preprocess/detectron_preprocess.py:12-24; div=1), the scaling modes of
 * utils/preprocess.py:127-175 (INCEPTION div=127.5 sub=1; VGG div=1 sub=(123,117,104)) and
 * image_adjust_ros' bottom padding (utils/preprocess.py:177-183; TCB_RESIZE_PAD, pad 0).
 *
 * Resampling is bit-identical to OpenCV's 8-bit INTER_LINEAR (11-bit fixed-point taps, half-pixel
 * centres, 2x2 averaging when both scale factors are exactly 2).
 *   mode STRETCH   : src (h,w) resampled to the whole (out_h,out_w) output (what the reference does).
 *   mode LETTERBOX : YOLOv5 convention (not in the reference): r=min(out_h/h,out_w/w),
 *                    content round(w r) x round(h r) centred, rest filled with pad_value.
 *   mode PAD       : no resampling, src placed top-left, rest filled with pad_value.
 * src_dev: n frames, row pitch src_row_stride bytes, frame pitch src_img_stride bytes, 3 bytes/pixel.
 * channel_swap: 0 keep order, 1 reverse (BGR<->RGB).
 * inv_affine_host (may be NULL): float[4] = {gain_x, gain_y, pad_left, pad_top} so that
 *   x_src = (x_out - pad_left) / gain_x  — the inverse used to map boxes back
 *   (communicator/ros_inference.py:100-115 is the stretch case).
 */
int tcb_preprocess_image(tcb_ctx* ctx, const uint8_t* src_dev, int n, int h, int w,
                         long long src_row_stride, long long src_img_stride, int channel_swap,
                         int mode, int out_h, int out_w, float div, const float sub[3],
                         int pad_value, float* dst_dev, float* inv_affine_host, void* stream);

/*
 * Host boundary of the preprocess result: the reference's image_adjust returns a host float32
 * array that goes into raw_input_contents (clients/preprocess/yolov5_preprocess.py:12-24, communicator/
 * ros_inference.py:143-146), so the model input has to come back over PCIe — but its padding is a constant.
 * tcb_content_rect: where tcb_preprocess_image puts the resampled frame for (mode, h, w, out_h, out_w):
 *   rect = {left, top, width, height} in output pixels; everything outside it holds the pad value. Host only.
 * tcb_copy_rect_d2h: asynchronous device -> (pinned) host copy of that rectangle of every plane of a
 *   [planes, plane_h, plane_w] array, both sides in the same layout; the rest of the host array is not touched,
 *   so a host buffer that already holds the pad value receives the same bytes as a full download
 *   (configs[1] letterbox: 360 of 640 rows per plane).
 */
int tcb_content_rect(int mode, int h, int w, int out_h, int out_w, int rect[4]);
int tcb_copy_rect_d2h(tcb_ctx* ctx, const void* src_dev, void* dst_host, int planes, int plane_h, int plane_w, int elem_bytes,
                      int left, int top, int rect_w, int rect_h, void* stream);

/* ---- compressed camera frames ------------------------------------------------- */

/*
 * Baseline JPEG decode on the device, bit-identical to what the reference's compressed-image branch computes:
 *   cv_image = cv2.imdecode(buf, cv2.IMREAD_ANYCOLOR)   # BGR8      (communicator/ros_inference.py:123-130)
 *   cv_image = cv_image[..., ::-1]                                  (:131; here: bgr = 0 writes RGB directly)
 * i.e. libjpeg-turbo's default decode path (Huffman, dequantisation + islow integer IDCT, fancy chroma
 * upsampling, fixed-point YCbCr -> RGB). Scope: 8-bit baseline / extended-sequential Huffman streams with one
 * interleaved scan, 3 components (4:4:4, 4:2:2, 4:4:0, 4:2:0) or 1 (grey), with or without restart intervals;
 * anything else (progressive, arithmetic, 12-bit, CMYK, RGB-coded) is refused with TCB_ERR_ARG and a message —
 * there is no CPU fallback.
 * tcb_jpeg_info: frame size and component count of one stream (host only).
 * tcb_jpeg_decode: n streams of the SAME size and component count -> out_dev [n, out_h, out_w, out_channels] u8
 *   (channels 3: RGB, or BGR when bgr != 0; channels 1: grey). jpeg_host / jpeg_len: n host pointers and
 *   lengths. The marker parse runs on the calling thread (batches above 32 MB of streams: on a few host threads
 *   that end with the call); the entropy-coded bytes go through staging_host
 *   (PINNED host memory of at least tcb_jpeg_staging_bytes(sum of lengths, n) bytes, in use until `stream`
 *   has passed this call) to the device in one copy; everything else is kernels on `stream`, no host sync.
 *   Needs tcb_jpeg_workspace(sum of lengths, n, out_h, out_w) bytes of context workspace.
 */
int tcb_jpeg_info(const unsigned char* jpeg_host, size_t len, int* height, int* width, int* channels);
size_t tcb_jpeg_staging_bytes(size_t total_jpeg_bytes, int n);
size_t tcb_jpeg_workspace(size_t total_jpeg_bytes, int n, int height, int width);
int tcb_jpeg_decode(tcb_ctx* ctx, const unsigned char* const* jpeg_host, const size_t* jpeg_len, int n, void* staging_host,
                    size_t staging_bytes, int bgr, unsigned char* out_dev, int out_h, int out_w, int out_channels, void* stream);

/* ---- YOLOv5 postprocess ---------------------------------------------------- */

/*
 * Confidence filter + conf=obj*cls + best class + xywh2xyxy + class-aware greedy NMS + max_det, for a
 * whole batch, entirely on device. Replaces Yolov5postprocess.extract_boxes
 * (clients/postprocess/yolov5_postprocess.py:28-125; twin utils/postprocess.py:105-199) from the
 * reshape at :37 on; the per-element struct.unpack of :36 (base_postprocess.py:15-25) disappears
 * because the little-endian wire bytes are the f32 tensor.
 * pred_dev: [b, n, 5+nc] f32 rows (cx, cy, w, h, obj, cls...). Arithmetic is float64 exactly where the
 * reference's is (Python-float products, float64 torchvision.ops.nms on boxes offset by cls*4096),
 * boxes are rounded to f32 where the reference rounds them (:79). Order: score descending, ties by
 * lower row index (torchvision's stable sort on the row-ordered candidates).
 * classes_host: optional whitelist (n_classes ints, host memory; ids < 256), NULL/0 = all.
 * out_dev: [b, max_det, 6] f64 rows (x1,y1,x2,y2,conf,cls); count_dev: [b] rows valid per image;
 * xc_count_dev (optional): [b] number of rows with obj > conf_thres (the reference asserts any() :44).
 * max_det <= 1024. Needs tcb_yolo_nms_workspace(b, n) bytes of context workspace.
 */
size_t tcb_yolo_nms_workspace(int b, int n);
int tcb_yolo_nms(tcb_ctx* ctx, const float* pred_dev, int b, int n, int nc, double conf_thres,
                 double iou_thres, const int* classes_host, int n_classes, int agnostic, int max_det,
                 int max_nms, double* out_dev, int* count_dev, int* xc_count_dev, void* stream);

/* ---- YOLO raw-head decode, YOLOv4-style postprocess, axis-aligned helpers ------ */

/*
 * Decode one detection head [b, A*(5+nc), h, w] (NCHW, as the network emits it) into rows ordered
 * (anchor, y, x), written at row_offset of a [b, total_rows, ...] output so that several heads concatenate.
 * variant 0 — YOLOv4, tools/yolo_layer.py:148-288 (yolo_forward_dynamic; YoloLayer.forward :313-321 divides the
 *   anchors by the stride first; heads are concatenated by get_region_boxes, tools/torch_utils.py:48-64):
 *   boxes_dev [b,total_rows,4] = normalised xyxy (the reference's [b,N,1,4]), confs_dev [b,total_rows,nc] =
 *   sigmoid(cls)*sigmoid(obj). anchors_host = 2*A floats (w,h) already divided by the stride.
 * variant 1 — YOLOv5 Detect (ultralytics; not in the reference): pred_dev [b,total_rows,5+nc] =
 *   (xy = (2s-0.5+grid)*stride, wh = (2s)^2*anchor, obj, cls), anchors in pixels — the tensor tcb_yolo_nms takes.
 */
int tcb_yolo_head_decode(tcb_ctx* ctx, const float* head_dev, int b, int h, int w, int num_anchors, int num_classes,
                         const float* anchors_host, int variant, double scale_x_y, float stride, int row_offset,
                         int total_rows, float* boxes_dev, float* confs_dev, float* pred_dev, void* stream);

/*
 * numpy greedy NMS of Postprocess.nms_cpu (clients/postprocess/base_postprocess.py:72-106 = utils/postprocess.py:69-103
 * = tools/utils.py:61-95), float32: order by score descending (ties: lower index first — the reference's
 * argsort()[::-1] leaves ties unspecified), suppress unless over <= thresh, over = inter/(a_i+a_j-inter) or
 * inter/min(a_i,a_j) when min_mode. keep_dev[post_max] int32 indices, count_dev[2] = {kept, 1 if n > 4096
 * (only the 4096 best-scoring boxes are considered)}. post_max <= 1024.
 */
int tcb_nms_axis_aligned(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, int n, float thresh, int min_mode,
                         int post_max, int* keep_dev, int* count_dev, void* stream);

/*
 * YOLOv4 two-tensor postprocess: extract_boxes_onnx (utils/postprocess.py:265-322) / post_processing
 * (tools/utils.py:166-233): max and first arg-max over classes, keep max > conf_thresh, per-class nms_cpu(nms_thresh),
 * rows [x1,y1,x2,y2,conf,conf,cls] grouped by class ascending, score descending inside a class.
 * boxes_dev [b,n,4] (the reference's [b,n,1,4]), confs_dev [b,n,nc]; out_dev [b,post_max,7] f32, count_dev [2*b]:
 * count_dev[i] = rows kept for image i, count_dev[b+i] = 1 when more than 4096 rows of image i passed the confidence
 * cut (the reference is unbounded; only the first 4096 in output order reach the NMS, so the result is then NOT the
 * reference's and the caller must treat it as an error — the Python drop-ins raise). A count equal to post_max means
 * the output may be truncated. post_max <= 1024.
 */
size_t tcb_yolo4_post_workspace(int b, int n);
int tcb_yolo4_post(tcb_ctx* ctx, const float* boxes_dev, const float* confs_dev, int b, int n, int nc, float conf_thresh,
                   float nms_thresh, int post_max, float* out_dev, int* count_dev, void* stream);

/* Pairwise IoU of xyxy boxes, Postprocess.box_iou (clients/postprocess/base_postprocess.py:48-70): out[n,m] f32. */
int tcb_box_iou(tcb_ctx* ctx, const float* boxes_a_dev, int n, const float* boxes_b_dev, int m, float* out_dev, void* stream);

/*
 * detectron2 dense-detector inference for a batch of b images (RetinaNet variant 0, FCOS variant 1) — the part of the served
 * model that produces what FCOSpostprocess.extract_boxes unpacks (clients/postprocess/detectron_postprocess.py:26-37,
 * utils/postprocess.py:368-468; tensors per examples/RetinaNet_detectron/config.pbtxt:13-33). detectron2 is not
 * vendored by the reference: restated from its published algorithm, parity unpinned.
 * Per level l: logits [b, A*K, h_l, w_l], deltas [b, A*4, h_l, w_l] (FCOS: + centerness [b,1,h_l,w_l], score =
 * sqrt(sigmoid(cls)*sigmoid(ctr))); keep score > score_thresh, top-k by score, decode against the anchors
 * (location*stride + cell anchor; Box2BoxTransform with dw,dh clamped at log(1000/16), or linear ltrb*size for
 * FCOS); concatenate levels, clip to (img_h,img_w), torchvision batched_nms(nms_thresh), keep max_det.
 * The `*_dev` level arguments are HOST arrays of device pointers; h, w, strides host int arrays; cell_anchors_host
 * [num_levels*A*4] (x0,y0,x1,y1). Outputs: boxes [b,max_det,4], scores [b,max_det], classes int64 [b,max_det], count [b];
 * every image is processed on its own (detectron2 loops over the images of a batch).
 * topk <= 4096, num_levels*topk <= 16384 (every one of those rows can reach the NMS, as in detectron2); max_det <= 1024.
 */
size_t tcb_detectron_post_workspace(int b, int num_levels, const int* h, const int* w, int num_anchors, int num_classes, int topk);
int tcb_detectron_post(tcb_ctx* ctx, int b, int num_levels, const float* const* logits_dev, const float* const* deltas_dev,
                       const float* const* ctrness_dev, const int* h, const int* w, const int* strides, int num_anchors,
                       int num_classes, const float* cell_anchors_host, int variant, float score_thresh, int topk,
                       float nms_thresh, int max_det, int img_h, int img_w, float* out_boxes_dev, float* out_scores_dev,
                       long long* out_classes_dev, int* out_count_dev, void* stream);

/* ---- LiDAR voxelisation + pillar decoration --------------------------------- */

enum { TCB_VOX_CONTINUE = 0, TCB_VOX_BREAK = 1 };

/*
 * Hard voxelisation of a batch of b sweeps, every sweep bit-identical to the sequential first-seen scan of the
 * voxel generators the reference calls: spconv's VoxelGenerator via
 * OPCD/pcdet/datasets/processor/data_processor.py:15-60,115-143 (PointpillarPreprocess.filter_pc,
 * clients/preprocess/preprocess_3d.py:33-51) and det3d's VoxelGenerator via
 * clients/preprocess/voxelize.py:19-24,38-42. The batch is what the reference's layouts
 * carry — the leading column of voxel_coords [b,z,y,x] (preprocess_3d.py:50, voxelize.py:46-47) is the sweep's
 * index in the batch — and a sweep's result does not depend on what else is in the call.
 * pts_dev: [n, pt_stride] f32 rows of all sweeps back to back, the first nfeat (>= 3: x,y,z,...) columns are
 * kept as features; sweep_offsets_host: b+1 ascending row offsets (host ints, [0] = 0, [b] = n; ragged and
 * empty sweeps allowed; NULL when b == 1). b <= 64.
 * range = {x0,y0,z0,x1,y1,z1}, vsize = {sx,sy,sz}; grid = round((max-min)/size).
 * overflow_mode: what happens at the (max_vox+1)-th new cell of a sweep — TCB_VOX_CONTINUE drops that point and
 *   goes on (spconv), TCB_VOX_BREAK stops the scan of that sweep (SECOND / det3d lineage).
 * f64_math: 0 = cell index computed in float32 from float32 range/size (KITTI .bin, Triton FP32),
 *   1 = in double (the det3d path as the reference client runs it on float64 arrays).
 * range_mask_xy: 1 applies mask_points_by_range first (OPCD/pcdet/utils/common_utils.py:60-63: x,y inclusive,
 *   z not tested). z_offset / intensity_col fuse the caller-side conditioning of
 *   communicator/ros_inference3d.py:127-128 (z += offset; column intensity_col /= its per-sweep max;
 *   0.f and -1 switch them off).
 * Outputs (caller-allocated): voxels [b, max_vox, max_pts, nfeat] f32 (rows < M_s of sweep s written, zero
 * padded), coords [b, max_vox, 4] i32 as [s, z, y, x], num_pts [b, max_vox] i32, m_out_dev [b] i32 = M_s.
 * Needs tcb_voxelize_workspace(n, b, max_vox, grid cells) bytes.
 */
size_t tcb_voxelize_workspace(int n, int b, int max_vox, long long cells);
int tcb_voxelize(tcb_ctx* ctx, const float* pts_dev, int n, const int* sweep_offsets_host, int b, int pt_stride, int nfeat,
                 const double range[6], const double vsize[3], int max_pts, int max_vox, int overflow_mode, int f64_math,
                 int range_mask_xy, float z_offset, int intensity_col, float* voxels_dev, int* coords_dev,
                 int* num_pts_dev, int* m_out_dev, void* stream);

/*
 * Pillar feature decoration, PillarVFE.forward up to the PFN layers
 * (OPCD/pcdet/models/backbones_3d/vfe/pillar_vfe.py:76-81,86-92,94-118): per voxel mean of xyz, offsets to
 * the mean and to the pillar centre, concatenated after the raw features, padded slots zeroed:
 * [b, m_max, max_pts, nfeat] -> [b, m_max, max_pts, nfeat+6] — the layout tcb_voxelize leaves (m_max = its
 * max_vox). The voxel count of sweep s is read from m_dev[s] when m_dev is non-NULL (device-resident, straight
 * from tcb_voxelize, no host sync), else all m_max rows of every sweep are decorated.
 */
int tcb_pillar_decorate(tcb_ctx* ctx, const float* voxels_dev, const int* coords_dev, const int* num_pts_dev,
                        const int* m_dev, int b, int m_max, int max_pts, int nfeat, const double vsize[3],
                        const double range_min[3], float* out_dev, void* stream);

/*
 * Dynamic pillarisation, DynamicPillarVFE.forward up to the PFN layers
 * (OPCD/pcdet/models/backbones_3d/vfe/dynamic_pillar_vfe.py:90-142): in-grid mask on x,y, pillar key
 * b*gx*gy + cx*gy + cy, pillars numbered in sorted key order (torch.unique), per-pillar mean of xyz
 * (torch_scatter.scatter_mean), per-POINT features [point features, xyz - mean, xy - pillar centre, z - z centre].
 * pts_dev: [n, ncol] rows (batch_idx, x, y, z, extra features...). Outputs for the n_valid in-grid points, input
 * order kept: feats_dev [n, (ncol-1)+6], unq_inv_dev [n] int64 (pillar id of each point), voxel_coords_dev
 * [n, 4] int32 (b, 0, y, x) for the n_unique pillars; counts_dev [2] = {n_valid, n_unique}. Pillar means are
 * accumulated in float64 (order independent to well below one float32 ulp).
 */
size_t tcb_dynamic_pillar_workspace(int n, int batch, int gx, int gy);
int tcb_dynamic_pillar(tcb_ctx* ctx, const float* pts_dev, int n, int ncol, int batch, const double range[6],
                       const double vsize[3], float* feats_dev, long long* unq_inv_dev, int* voxel_coords_dev,
                       int* counts_dev, void* stream);

/* ---- 3D detection postprocess ------------------------------------------------ */

/* Rotated BEV IoU matrix out[n,m] of [x,y,z,dx,dy,dz,heading] boxes — boxes_iou_bev
 * (OPCD/pcdet/ops/iou3d_nms/iou3d_nms_utils.py:12-36; kernel src/iou3d_nms_kernel.cu:104-265, CPU twin
 * src/iou3d_cpu.cpp:232-252). Same float32 operation order as the reference. */
int tcb_rotated_iou(tcb_ctx* ctx, const float* boxes_a_dev, int n, const float* boxes_b_dev, int m,
                    float* out_dev, void* stream);

/* NMS_CONFIG.NMS_TYPE of the pcdet configs, looked up by name in iou3d_nms_utils (model_nms_utils.py:15):
 * nms_gpu = rotated BEV IoU (iou3d_nms_utils.py:84-99, nms_kernel src/iou3d_nms_kernel.cu:267-311);
 * nms_normal_gpu = axis-aligned BEV IoU, heading ignored (iou3d_nms_utils.py:102-116, iou_normal +
 * nms_normal_kernel src/iou3d_nms_kernel.cu:314-367). Both: sort descending, bit mask, host sweep
 * (src/iou3d_nms.cpp:90-136 / :139-187). */
#define TCB_NMS_GPU 0
#define TCB_NMS_NORMAL 1

/* class_agnostic_nms without the score cut (OPCD/pcdet/models/model_utils/model_nms_utils.py:13-19) =
 * topk(min(pre_max, n)) by score + the NMS nms_type names + [:post_max], for b independent box sets (the per-sample loop of
 * OPCD/pcdet/models/detectors/detector3d_template.py:196-199): boxes_dev [b,n,7], scores_dev [b,n].
 * Order: score descending, ties by lower index. keep_dev [b,post_max]: indices into the set, in kept order;
 * count_dev [b]. pre_max <= 4096, post_max <= 1024. */
int tcb_rotated_nms(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, int b, int n, float thresh,
                    int nms_type, int pre_max, int post_max, int* keep_dev, int* count_dev, void* stream);

/*
 * PointPillars anchor head post for a batch of b frames: analytic anchors (AnchorGenerator.generate_anchors,
 * OPCD/pcdet/models/dense_heads/target_assigner/anchor_generator.py:17-60, align_center False, one size per
 * class) + ResidualCoder.decode_torch (OPCD/pcdet/utils/box_coder_utils.py:45-77) + direction bins
 * (generate_predicted_boxes(batch_size, ...), OPCD/pcdet/models/dense_heads/anchor_head_template.py:225-272,
 * limit_period OPCD/pcdet/utils/common_utils.py:21-24) + post_processing (OPCD/pcdet/models/detectors/
 * detector3d_template.py:178-284: per sample :196-199 sigmoid, max over classes, label+1) +
 * class_agnostic_nms(score_thresh, pre_max, nms_thresh, post_max). A frame's result does not depend on the batch.
 * cls_dev [b,h,w,A*num_class], box_dev [b,h,w,A*7], dir_dev [b,h,w,A*num_dir_bins] (NULL = no direction
 * classifier), A = num_class*num_rot anchors per location ordered (class, rotation); anchor_sizes [num_class*3]
 * (dx,dy,dz), anchor_bottoms [num_class], anchor_rotations [num_class*num_rot] are host arrays.
 * Outputs: boxes [b,post_max,7], scores [b,post_max], labels [b,post_max] int64 (1-based), count [b].
 */
size_t tcb_pp_anchor_post_workspace(int b, int h, int w, int num_class, int num_rot);
int tcb_pp_anchor_post(tcb_ctx* ctx, const float* cls_dev, const float* box_dev, const float* dir_dev, int b, int h, int w,
                       int num_class, int num_rot, const float anchor_range[6], const float* anchor_sizes,
                       const float* anchor_bottoms, const float* anchor_rotations, float dir_offset,
                       float dir_limit_offset, int num_dir_bins, float score_thresh, float nms_thresh, int nms_type,
                       int pre_max, int post_max, float* out_boxes_dev, float* out_scores_dev, long long* out_labels_dev,
                       int* out_count_dev, void* stream);

/*
 * CenterPoint head post for a batch of b frames: CenterHead.generate_predicted_boxes(batch_size, pred_dicts)
 * (OPCD/pcdet/models/dense_heads/center_head.py:253-304) = per head sigmoid + exp, _topk and
 * decode_bbox_from_heatmap (OPCD/pcdet/models/model_utils/centernet_utils.py:136-216), centre-range and
 * score > thresh mask, class id mapping, class_agnostic_nms(pre_max, nms_thresh, post_max); per frame the heads
 * concatenated, labels + 1. Per-head tensors are NCHW: hm [b,c,h,w], center [b,2,h,w], center_z [b,1,h,w],
 * dim [b,3,h,w] (log), rot [b,2,h,w] (cos, sin), vel [b,2,h,w] (vel_dev or its entries may be NULL).
 * The `*_dev` arguments are HOST arrays of num_heads device pointers; head_num_class [num_heads] and
 * class_id_map (concatenated per head, 0-based) are host arrays. Outputs: boxes [b, heads*post_max, 9],
 * scores [b, heads*post_max], labels int64 (1-based), count [b]. K <= 4096.
 */
size_t tcb_centerpoint_post_workspace(int b, int num_heads, int pre_max, int post_max);
int tcb_centerpoint_post(tcb_ctx* ctx, int b, int num_heads, const float* const* hm_dev, const float* const* center_dev,
                         const float* const* center_z_dev, const float* const* dim_dev, const float* const* rot_dev,
                         const float* const* vel_dev, const int* head_num_class, const int* class_id_map, int h, int w,
                         int feature_map_stride, const float voxel_size_xy[2], const float range_min_xy[2],
                         const float limit_range[6], int k, float score_thresh, float nms_thresh, int nms_type, int pre_max,
                         int post_max, float* out_boxes_dev, float* out_scores_dev, long long* out_labels_dev, int* out_count_dev,
                         void* stream);

/* Per-class score cut of PointPillarPostprocess.remove_low_score_nu / get_annotations_indices
 * (clients/postprocess/detector_3d_postprocess.py:55-64,98-133): for each (class, threshold) group in the
 * given order, the ascending indices i with labels[i] == class and scores[i] >= threshold, concatenated.
 * labels int64, group tables are host arrays (<= 64 groups; copied through pinned slots of the context, the call does
 * not wait for the device); out_idx_dev holds up to n_groups*n indices. */
int tcb_class_score_filter(tcb_ctx* ctx, const long long* labels_dev, const float* scores_dev, int n, int n_groups,
                           const int* group_class_host, const float* group_thresh_host, int* out_idx_dev,
                           int* out_count_dev, void* stream);

/* ---- the step after the path: box rescale + overlay drawing ------------------ */

/* _scale_box_array (communicator/evaluate_inference.py:237-256) / _scale_boxes (communicator/ros_inference.py:
 * 100-115) on n rows [x1,y1,x2,y2,conf,cls] float64: x *= orig_w / input_size0 (sic: the reference divides the
 * width by input_size[0]), y = y * orig_h / input_size1; normalized != 0: x *= orig_w, y *= orig_h. Columns 4,5
 * are copied. math_mode picks the arithmetic the reference's expression runs in:
 *   0  float64 throughout (the array version)
 *   1  box rounded to float32, float64 arithmetic  (np.float32 scalar * Python number under numpy < 2)
 *   2  box rounded to float32, float32 arithmetic  (same expression under numpy >= 2) */
int tcb_scale_boxes(tcb_ctx* ctx, const double* det_dev, int n, int orig_h, int orig_w, int input_size0, int input_size1,
                    int normalized, int math_mode, double* out_dev, void* stream);

/* The per-detection drawing loop of RosInference._callback (communicator/ros_inference.py:148-169) and of
 * Yolov5postprocess.plot_boxes_cv2 (clients/postprocess/yolov5_postprocess.py:127-169), in place on b device
 * frames (h x w x 3 uint8, row_stride / frame_stride in bytes): each of the first count_dev[f] (NULL: max_det) rows of
 * det_dev[f] (max_det x 6 float64: x1, y1, x2, y2, conf, class) is scaled like _scale_boxes (math_mode 1 or 2; orig
 * size = h, w; normalized != 0: x * w, y * h as plot_boxes_cv2 does), truncated like Python's int(), and drawn as
 * cv2.rectangle(thickness 1 or 2) plus, when labels != 0, cv2.putText(text, (x1, int(y1 + text_dy)),
 * FONT_HERSHEY_SIMPLEX, font_scale, color, text_thickness) with text = '{:.2f} {}'.format(conf, name) (labels 1:
 * ros_inference.py:163-169, font_scale 0.5, text_thickness 2, text_dy -10) or the class name alone (labels 2:
 * plot_boxes_cv2, font_scale 1.2, text_thickness 1, text_dy 0). Detections are painted in order (a later one
 * overwrites an earlier one). class_names: host memory, num_classes strings back to back, each '\0'-terminated
 * (<= 128 classes, <= 2560 bytes; labels longer than 64 characters are cut). colors: host memory, 3 bytes in frame
 * channel order, or num_classes x 3 when per_class_color != 0 (plot_boxes_cv2's get_color table). Pixels are
 * identical to OpenCV's for every scale and thickness, labels cut by the frame edge included: the strokes of the
 * Hershey font are rasterised with OpenCV's own rules (clipLine, LineIterator, FillConvexPoly, Circle). */
size_t tcb_draw_detections_workspace(int b, int max_det);
int tcb_draw_detections(tcb_ctx* ctx, unsigned char* frames_dev, int b, int h, int w, long long row_stride,
                        long long frame_stride, const double* det_dev, const int* count_dev, int max_det,
                        int input_size0, int input_size1, int normalized, int math_mode, const char* class_names,
                        int num_classes, const unsigned char* colors, int per_class_color, int thickness, int labels,
                        double font_scale, int text_thickness, int text_dy, void* stream);

/* ---- ROS message decode/encode around the 3D path ----------------------------- */

/* np.array(list(sensor_msgs.point_cloud2.read_points(msg, field_names, skip_nans))) of
 * communicator/ros_inference3d.py:125, from the message bytes on the device. sensor_msgs is not part of the
 * reference tree; restated from its published algorithm: points in storage order (row v at v*row_step, point u at
 * +u*point_step), the selected fields unpacked in ascending byte offset (pass them in that order; PointField
 * datatypes 1 INT8, 2 UINT8, 3 INT16, 4 UINT16, 7 FLOAT32 — the ones float32 rows hold exactly), a point dropped
 * when skip_nans != 0 and any selected value is NaN. out_dev: [width*height, nfields] f32; rows [0, count) are the
 * kept points, the rest is filled with NaN (columns 0-2) / -inf (other columns) so that tcb_voxelize, which ignores
 * points with NaN coordinates and -inf in its column maximum, can take all width*height rows without a host sync.
 * field tables are host arrays. */
size_t tcb_pointcloud2_unpack_workspace(int width, int height);
int tcb_pointcloud2_unpack(tcb_ctx* ctx, const unsigned char* data_dev, int width, int height, int point_step,
                           long long row_step, int nfields, const int* field_offsets, const int* field_datatypes,
                           int is_bigendian, int skip_nans, float* out_dev, int* count_dev, void* stream);

/* The BoundingBoxArray fields RosInference3D publishes (communicator/ros_inference3d.py:150-177): the ascending
 * indices i with labels[i] == label_sel and scores[i] > score_thresh (the reference: 2 and 0.5), and per index 12
 * float64 values: position x, y, z - z_offset; orientation x, y, z, w of pyquaternion's Quaternion(axis=[0,0,1],
 * radians=yaw) = (0*s, 0*s, s, c) with s, c = sin, cos(yaw/2) in float64; dimensions box[4], box[3], box[5] (x and y
 * swapped as the reference does); score; label. boxes [n,7] f32, scores [n] f32, labels [n] int64. */
int tcb_boxes_to_jsk(tcb_ctx* ctx, const float* boxes_dev, const float* scores_dev, const long long* labels_dev, int n,
                     long long label_sel, float score_thresh, double z_offset, double* out_dev, int* out_idx_dev,
                     int* out_count_dev, void* stream);

/*
 * Evaluation math (communicator/evaluate_inference.py): calculate_metrics :400-446, ap_per_class :158-218, compute_ap
 * :131-156, with box_iou (clients/postprocess/base_postprocess.py:48-70) in float64 as the reference's float64 arrays make
 * it. A batch of frames per call, one CTA per frame. Rows are float64 [x1,y1,x2,y2,score,class] (gt_callback :363-374
 * after xywh2xyxy; _scale_box_array :237-256 of extract_boxes' rows); frame f owns rows [offsets[f], offsets[f+1]) —
 * the offset tables are HOST arrays of frames+1 ints. At most 2048 ground truths and 2048 predictions per frame.
 *
 * tcb_eval_match: IoU >= 0.5 with equal class, each detection to its highest-IoU label, each label to the first
 *   detection that chose it; correct_dev u8 [total predictions, 10] at IoU 0.5..0.95 (float32 comparison as in :425),
 *   matched_dev i32 [frames] = 0 where no pair passed (the reference records nothing for such a frame). Where the reference
 *   leaves an order to an unstable sort (equal IoUs at :419), the later label wins.
 * tcb_ap_per_class: per frame, tp u8 [n,nt] (nt <= 10), conf / pred_cls / target_cls float64 with an element stride
 *   (6 when they are columns of the rows above, 1 when dense). Outputs per frame with room for max_classes (<= 256)
 *   classes: num_classes i32 [frames] (the number of distinct target classes; if it exceeds max_classes the frame's
 *   other outputs are incomplete), classes i32, p, r, f1 float64 [frames,max_classes], ap float64 [frames,max_classes,nt].
 *   Equal confidences keep their input order.
 * tcb_calculate_metrics: both, ap_per_class on each matched frame's own rows as :428-436 does.
 */
size_t tcb_eval_workspace(int frames, int max_classes);
int tcb_eval_match(tcb_ctx* ctx, const double* gts_dev, const int* gt_offsets_host, const double* preds_dev,
                   const int* pred_offsets_host, int frames, unsigned char* correct_dev, int* matched_dev, void* stream);
int tcb_ap_per_class(tcb_ctx* ctx, const unsigned char* tp_dev, int nt, const double* conf_dev, const double* pred_cls_dev,
                     long long pred_stride, const int* pred_offsets_host, const double* target_cls_dev, long long target_stride,
                     const int* target_offsets_host, int frames, int max_classes, int* num_classes_dev, int* classes_dev,
                     double* p_dev, double* r_dev, double* f1_dev, double* ap_dev, void* stream);
int tcb_calculate_metrics(tcb_ctx* ctx, const double* gts_dev, const int* gt_offsets_host, const double* preds_dev,
                          const int* pred_offsets_host, int frames, int max_classes, unsigned char* correct_dev,
                          int* matched_dev, int* num_classes_dev, int* classes_dev, double* p_dev, double* r_dev,
                          double* f1_dev, double* ap_dev, void* stream);

/*
 * tcb_yolo_nms without a context and without any allocation. No counterpart in the reference: it is
 * here for callers that are not allowed to allocate GPU memory — a plugin inside a host process that
 * has already committed all of HBM to its own pools, or a per-frame callback that cannot afford the
 * device-wide synchronisation and unbounded latency of cudaMalloc/cudaFree (tcb_create performs one).
 * Same arguments, results and error codes as tcb_yolo_nms; errors are read with tcb_last_error(NULL).
 * Scratch is one fixed device array of tcb_yolo_nms_static_scratch_bytes() (about 8.4 MB) that every
 * GPU loading the library holds until process exit, so:
 *   - limits: b <= 64, b*n <= 131072, and b*pow2ceil(n) <= 262144 when n > 4096; beyond that the call
 *     returns TCB_ERR_LIMIT and nothing is launched;
 *   - all calls share the array: calls that overlap on the GPU race (keep them on one stream); the
 *     array keeps the previous call's contents and the call clears the counters it needs zeroed;
 *   - the first call on a device does a one-time set-up (device address of the array, kernel
 *     attribute); make it with that device current and before any stream capture begins.
 * The device that owns pred_dev/out_dev must be the current device.
 */
size_t tcb_yolo_nms_static_scratch_bytes(void);
int tcb_yolo_nms_static(const float* pred_dev, int b, int n, int nc, double conf_thres, double iou_thres,
                        const int* classes_host, int n_classes, int agnostic, int max_det, int max_nms,
                        double* out_dev, int* count_dev, int* xc_count_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TCB200_H */
