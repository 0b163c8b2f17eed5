/*
 * absolutetrack_b200 - C-ABI of the B200 (sm_100a) implementation of UmeTrack's per-frame
 * multi-view inference hot path (crop resample -> per-view encoder -> FTL/fusion/temporal ->
 * pose head -> FK/skinning).
 *
 * The reference (2InfinityN6eyond/AbsoluteTrack) is 100% Python and has no FFI layer: its
 * boundary for this path is the Python surface lib.tracker.tracker.HandTracker /
 * lib.models.umetrack_model.UmeTrackModel / lib.common.hand_skinning.skin_landmarks.  The
 * host mirror of that surface lives in absolutetrack_b200/lib/ and binds the entry points
 * below with ctypes (absolutetrack_b200/_cabi.py); INTEGRATION.md shows the stub a reference
 * maintainer would add.  Each entry point cites the reference code it replaces
 * (file:line relative to the reference root).
 *
 * Conventions: plain pointers and sizes only; every data pointer is a DEVICE pointer owned by
 * the caller unless marked HOST; every launch goes onto `stream` (a cudaStream_t passed as
 * void*); functions return 0 or a negative AtStatus, never throw, never allocate device
 * memory, never synchronise.  at_last_error_string() describes the last failure on the
 * calling thread.
 */
#ifndef ABSOLUTETRACK_B200_H_
#define ABSOLUTETRACK_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AT_VERSION 100 /* major*10000 + minor*100 + patch */

typedef enum AtStatus {
  AT_OK = 0,
  AT_ERR_INVALID_ARGUMENT = -1,
  AT_ERR_CUDA = -2,
  AT_ERR_WORKSPACE_TOO_SMALL = -3,
  AT_ERR_UNSUPPORTED = -4
} AtStatus;

typedef enum AtDType { AT_U8 = 0, AT_F32 = 1, AT_BF16 = 2 } AtDType;
typedef enum AtMapMode { AT_MAP_FP32 = 0, AT_MAP_FP64 = 1 } AtMapMode;
typedef enum AtPrecision { AT_PREC_FP32 = 0, AT_PREC_BF16 = 1, AT_PREC_FP16 = 2 } AtPrecision;
typedef enum AtHeadMode { AT_HEAD_KNOWN_SKELETON = 0, AT_HEAD_PREDICT_SCALE = 1 } AtHeadMode;

int at_version(void);
const char* at_last_error_string(void);
/* Number of CUDA kernels this library has launched in this process (all threads). */
int64_t at_launch_count(void);
/* 1 when the library was compiled with the tcgen05 encoder (always, for sm_100a builds). */
int at_has_tcgen05(void);

/* ------------------------------------------------------------------------------------------
 * (1) Perspective-crop resampler.  Replaces tracker._warp_image (lib/tracker/tracker.py:61-89):
 * per dst pixel: pinhole unproject (camera.py:320-329,69-75), rotate into the source camera
 * (camera.py:296-306), arctan projection (camera.py:80-85), Fisheye62 distortion
 * (camera.py:122-143), window coords (camera.py:308-312), z<0 -> (-1,-1) (tracker.py:78-80),
 * cast float32, then cv2.remap INTER_LINEAR / BORDER_CONSTANT 0 in OpenCV's fixed point
 * (tracker.py:87) and optionally the /255 -> float of tracker.py:332.
 * The host precomposes M = R_src^T R_dst and d = R_src^T (t_dst - t_src) in fp64.
 * ---------------------------------------------------------------------------------------- */
typedef struct AtCropParams {
  int32_t frame;   /* source frame index in the batch */
  int32_t view;    /* source view (camera) index inside the frame */
  int32_t src_w;   /* source image width  (636) */
  int32_t src_h;   /* source image height (480) */
  double M[9];     /* row-major 3x3: src_eye = M * dst_eye + d */
  double d[3];
  double f_dst[2]; /* crop (pinhole) focal lengths */
  double c_dst[2]; /* crop principal point */
  double f_src[2]; /* fisheye focal lengths */
  double c_src[2]; /* fisheye principal point */
  double k[8];     /* k1 k2 k3 k4 p1 p2 k5 k6 (camera.py:111-118 field order) */
  double reserved[2];
} AtCropParams;    /* 256 bytes */

/* src pixel (frame,view,y,x) = src[frame*frame_stride + view*view_stride + y*row_stride + x].
 * For the recording layout uint8[480,4,636] (video_pose_data.py:123-125): frame_stride =
 * 480*2544, view_stride = 636, row_stride = 2544.  out: [n_crops, dst_h, dst_w] of out_dtype
 * (AT_U8: grey levels; AT_F32 / AT_BF16: level/255). */
int at_crop_resample_fisheye62_u8(const uint8_t* src, int64_t frame_stride, int64_t view_stride,
                                  int64_t row_stride, const AtCropParams* params, int n_crops,
                                  int dst_w, int dst_h, void* out, int out_dtype, int map_mode,
                                  void* stream);
/* Same, plus an optional device counter.  In AT_MAP_FP32 mode each crop's source bounding box is
 * staged once through shared memory, so `src` may also be PINNED HOST memory (zero-copy ingest:
 * only the boxes cross PCIe).  *staged_bytes (device, may be NULL) is incremented by the bytes each
 * crop pulled from `src`. */
int at_crop_resample_fisheye62_u8_ex(const uint8_t* src, int64_t frame_stride, int64_t view_stride,
                                     int64_t row_stride, const AtCropParams* params, int n_crops,
                                     int dst_w, int dst_h, void* out, int out_dtype, int map_mode,
                                     unsigned long long* staged_bytes, void* stream);
/* Host->device ingest without copying whole frames (the PCIe side of tracker._warp_image's source
 * read, lib/tracker/tracker.py:61-89, for frames that live in pinned host memory).  For every crop
 * the source region its map touches is bounded from the crop's boundary pixels (fp32 map) and
 * copied from `src` into `dst`, a DEVICE buffer of the same layout ([n_frames] x frame_stride
 * bytes); at_crop_resample_fisheye62_u8 then runs on `dst` and never reads outside the boxes.
 * src, dst and frame_stride must be 16-byte aligned, n_frames * frame_stride < 4 GiB.  The launch
 * is small on purpose (max_ctas one-warp CTAs, 0 = one per SM; 32 registers, <= 4 KB shared memory)
 * so that it stays co-resident with the encoder kernels of the previous step on a side stream.
 * *pulled_bytes (device, may be NULL) is incremented by the bytes read from `src`. */
int at_crop_pull_source_boxes_u8(const uint8_t* src, uint8_t* dst, int64_t n_frames, int64_t frame_stride,
                                 int64_t view_stride, int64_t row_stride, const AtCropParams* params,
                                 int n_crops, int dst_w, int dst_h, int max_ctas,
                                 unsigned long long* pulled_bytes, void* stream);

/* Config-3 (torch_data) preprocessing on the device (the arithmetic of the loader).
 * Replaces data_transform._gen_crop_matrices (lib/batched_dataset/data_transform.py:145-209) with
 * crop.gen_crop_parameters_from_points (lib/common/crop.py:31-82) and _compute_resample_matrix
 * (data_transform.py:57-75) for all n_frames * n_views pinhole views of a sequence in one launch:
 *   orig_extrinsics [n_frames,n_views,4,4] f32 rigid world->eye, orig_intrinsics [..,3,3] f32 (fx, fy, cx, cy
 *   are read), crop_points [n_frames,n_pts,3] f32, mirror_image uint8 [n_frames] (NULL = none; 1 = right
 *   hand, data_transform.py:249), focal_multiplier 0.95 in the reference
 *   -> extrinsics_xf [..,4,4] (world->eye of the crop camera), new_intrinsics [..,3,3], resample_xf [..,4,4]
 *      (crop pixel -> original pixel: the `xfs` of at_resample_homography_f32), all f32.
 * *n_errors (device int, accumulated) counts the views for which crop.py:25-26 raises
 * ValueError("Unable to create crop camera"): a point closer than 1e-4 in front of the crop camera,
 * or a focal below 5 px. */
int at_gen_crop_matrices_f32(const float* orig_extrinsics, const float* orig_intrinsics, const float* crop_points,
                             const uint8_t* mirror_image, int64_t n_frames, int n_views, int n_pts, int crop_w,
                             int crop_h, float focal_multiplier, float* extrinsics_xf, float* new_intrinsics,
                             float* resample_xf, int* n_errors, void* stream);

/* The torch_data pinhole->pinhole resampler.  Replaces data_transform._resample_images_batched
 * (lib/batched_dataset/data_transform.py:78-144): (u,v,1) -> xf[0:3,0:3]*(u,v,1)+xf[0:3,3],
 * /z, mask 0<=x<W-1 & 0<=y<H-1, float bilinear, 0 outside.  images [n,H,W] f32, xfs [n,4,4] f32
 * row-major, out [n,oh,ow] f32, divided (IEEE float32 division) by out_divisor: 255 reproduces the
 * normalisation of data_transform.py:275, 1 leaves grey levels.  The reference's numpy promotes
 * the map to float64 (float32 matrix x int32 grid, :99); so does the kernel. */
int at_resample_homography_f32(const float* images, int n, int H, int W, const float* xfs, int oh,
                               int ow, float* out, float out_divisor, void* stream);

/* ------------------------------------------------------------------------------------------
 * (4) Hand forward kinematics + landmark skinning.  Replaces hand_skinning.skin_landmarks
 * (lib/common/hand_skinning.py:189-209: _joint_local_transform :35-53 with pytorch3d
 * so3_exp_map, _finger_fk :17-32, _hand_skinning_transform :100-127, _get_skinning_weights
 * :70-97, _lbs :56-67) and the right-hand mirror of perspective_crop.landmarks_from_hand_pose
 * (lib/tracker/perspective_crop.py:40-51).  float32 throughout, like the reference.
 * ---------------------------------------------------------------------------------------- */
typedef struct AtHandModel {
  const float* joint_rotation_axes;          /* [22,3] or [n,22,3] */
  const float* joint_rest_positions;         /* [22,3] or [n,22,3] */
  const float* landmark_rest_positions;      /* [21,3] or [n,21,3] */
  const float* landmark_rest_bone_weights;   /* [21,K] or [n,21,K] */
  const int32_t* landmark_rest_bone_indices; /* [21,K] or [n,21,K] */
  int32_t n_weights;                         /* K (3 in every reference hand model) */
  int32_t per_sample;                        /* 0: one model shared by all n; 1: one per sample */
} AtHandModel;

/* joint_angles [n,22], wrist_xforms [n,4,4] row-major, hand_idx [n] or NULL (1 = right hand:
 * column 0 of the wrist transform is negated first), landmarks out [n,21,3]. */
int at_fk_skin_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                   const AtHandModel* model, int64_t n, float* landmarks, void* stream);

/* The same with the unit change of the tracker folded in: HandTracker._track_frame turns the model's
 * wrist transforms from metres into millimetres (tracker.py:34,379: output_wrist_xforms[..., :3, 3] *=
 * M_TO_MM) before anything downstream uses them.  The translation column of every wrist transform is multiplied by translation_scale
 * before use, and the scaled transforms are written to wrist_out [n,4,4] when it is not NULL
 * (wrist_out must not overlap wrist_xforms).  translation_scale = 1, wrist_out = NULL is at_fk_skin_f32. */
int at_fk_skin_scaled_f32(const float* joint_angles, const float* wrist_xforms, const int64_t* hand_idx,
                          const AtHandModel* model, int64_t n, float translation_scale, float* wrist_out,
                          float* landmarks, void* stream);

/* Evaluation sums on the device.  Replaces the per-frame bookkeeping of
 * run_eval_known_skeleton.py:92-93 (mean keypoint error of a hand-frame) for a batch:
 *   sums[0] += sum over valid hands of mean_l |landmarks[h,l] - gt_landmarks[h,l]|   (input units, mm)
 *   sums[1] += sum over valid hands of mean_j |joint_angles[h,j] - gt_joint_angles[h,j]| in degrees, j < 20
 *              (skipped when both joint-angle pointers are NULL)
 *   sums[2] += number of valid hands
 * landmarks [n,21,3], joint_angles [n,22], valid [n] (NULL = all), sums: 3 float64 accumulators on the
 * device - the per-shard numbers that the only collective of the path all-reduces. */
int at_eval_accumulate_f32(const float* landmarks, const float* gt_landmarks, const float* joint_angles,
                           const float* gt_joint_angles, const uint8_t* valid, int64_t n, double* sums,
                           void* stream);

/* Evaluation table of one recording on the device.  Replaces
 * load_eval._compute_metrics (load_eval.py:26-44) and the per-threshold counts of
 * metric_utils._PCK_curve (lib/common/metric_utils.py:34-44) with sums that add up across
 * recordings and ranks (aggregate_metrics, load_eval.py:47-92, divides them at the end):
 *   sums[0]   += sum over valid hand-frames of e = mean_l |gt - tracked|            (keypoint_errors)
 *   sums[1]   += sum over valid triples (t,t+1,t+2) of mean_l |p[t] + p[t+2] - 2 p[t+1]|, p = tracked
 *   sums[2]   += the same of gt
 *   counts[0] += valid hand-frames   counts[1] += n_hands * n_frames   counts[2] += valid triples
 *   counts[3 + k] += #(valid e <= thresholds[k]),  k < n_thresholds <= 128
 * tracked_keypoints / gt_keypoints: float32 [n_hands, n_frames, 21, 3] (mm; the reference keeps
 * float32 values in float64 arrays), valid_tracking: uint8 [n_hands, n_frames], thresholds: float64
 * on the device.  Arithmetic is float64 in numpy's summation order: e is bit-identical to the
 * reference's, so the integer counts are exact. */
int at_eval_recording_f32(const float* tracked_keypoints, const float* gt_keypoints, const uint8_t* valid_tracking,
                          int64_t n_hands, int64_t n_frames, const double* thresholds, int n_thresholds,
                          double* sums, int64_t* counts, void* stream);

/* ------------------------------------------------------------------------------------------
 * (5) Crop-camera generation + model-input assembly on the device.  Replaces
 * HandTracker.gen_crop_cameras (lib/tracker/tracker.py:222-260) with everything below it -
 * perspective_crop.gen_crop_cameras_from_pose (:136-180), _get_crop_points_from_hand_pose
 * (:89-133: landmarks of the given pose, the neutral pose and the open pose under one wrist
 * transform, float32 FK), rank_hand_visibility_in_cameras (:54-86: Fisheye62 projection of the 21
 * landmarks into every camera, float64), crop.gen_crop_parameters_from_points (crop.py:31-82:
 * bounding-box centre, affine.make_look_at_matrix affine.py:47-76 with from_two_vectors :34-44 and
 * the camera roll, x mirror for right hands, gen_intrinsics_from_bounding_pts crop.py:15-28) - and
 * the per-crop half of HandTracker._make_inputs (tracker.py:315-368: intrinsics matrix,
 * world->crop-eye extrinsics in metres, sample_range, hand_idx, memory_idx, use_memory from the
 * tracking history of tracker.py:105,409-410) plus the AtCropParams record of every crop.
 * Camera math in float64 like the reference's numpy; rigid transforms are inverted as rigid
 * transforms (R^T, -R^T t) instead of LAPACK's general inverse (difference: fp64 round-off).
 * Output order is the reference's: poses in input order, per pose the selected cameras (by camera
 * index when sort_camera_index, else most visible landmarks first, stable), poses with fewer than
 * min_num_crops views dropped.
 * ---------------------------------------------------------------------------------------- */
typedef struct AtRigCamera { /* one physical Fisheye62 camera of the headset (constant per recording) */
  double f[2];
  double c[2];
  double k[8];       /* k1 k2 k3 k4 p1 p2 k5 k6 */
  double angle_deg;  /* camera_angles[cam]: physical roll of the camera (affine.py:73) */
  int32_t width, height;
} AtRigCamera; /* 112 bytes */

typedef struct AtCropCameraOpts {
  int32_t num_crop_points;            /* 21, 42 or 63 (HandTrackerOpts.num_crop_points, tracker.py:51) */
  int32_t min_required_vis_landmarks; /* 19 (tracker.py:56) */
  int32_t max_view_num;               /* 2 (tracker.py:37 MAX_VIEW_NUM); <= 4 */
  int32_t min_num_crops;              /* gen_crop_cameras(min_num_crops) (tracker.py:228,257) */
  int32_t sort_camera_index;          /* 1 (tracker.py:250) */
  int32_t mirror_right_hand;          /* 1 (tracker.py:252) */
  int32_t crop_w, crop_h;             /* model.getInputImageSizes(): 96 x 96 */
  double focal_multiplier;            /* hand_ratio_in_crop 0.8 (tracker.py:55,251) */
  double min_focal;                   /* 5 (crop.py:16) */
} AtCropCameraOpts;

typedef struct AtCropCameraInputs {
  const float* joint_angles;         /* [n_poses,22] */
  const float* wrist_xforms;         /* [n_poses,4,4] wrist->world (mm) as in SingleHandPose: right hands NOT mirrored */
  const int32_t* pose_frame;         /* [n_poses] frame (sequence slot) of each pose */
  const int32_t* pose_hand;          /* [n_poses] 0 = left, 1 = right */
  const uint8_t* pose_valid;         /* [n_poses] hand_confidence >= 0.5 (tracker.py:236), NULL = all valid */
  const float* neutral_joint_angles; /* [22] perspective_crop.neutral_joint_angles(hand_model) (:17-24) */
  const double* camera_to_world;     /* [n_frames,n_cams,4,4] row-major */
  const AtRigCamera* cameras;        /* [n_cams] (device) */
  int32_t n_poses, n_frames, n_cams; /* n_cams <= 8 */
  int32_t reserved;
} AtCropCameraInputs;

typedef struct AtCropCameraOutputs { /* capacity: n_poses samples, n_poses*max_view_num crops */
  AtCropParams* crop_params;    /* [c] frame = pose_frame, view = camera index */
  float* intrinsics;            /* [c,3,3] uv_to_window_matrix (camera.py:371-379) */
  float* extrinsics;            /* [c,4,4] world->crop-eye, translation in metres (tracker.py:343-345) */
  double* crop_camera_to_world; /* [c,4,4] of the PinholePlaneCameraModel (crop.py:81), mm; may be NULL */
  int32_t* crop_camera_index;   /* [c] source camera of each crop; may be NULL */
  int64_t* sample_range;        /* [n,2] */
  int64_t* hand_idx;            /* [n] */
  int64_t* memory_idx;          /* [n] = pose_frame*2 + hand */
  uint8_t* use_memory;          /* [n] tracking_history[memory_idx] before this call */
  int32_t* pose_index;          /* [n] input pose of each emitted sample; may be NULL */
  uint8_t* tracking_history;    /* [n_frames*2] in/out: slot of every input pose := "emitted" (tracker.py:409-410); may be NULL */
  int32_t* counts;              /* [4]: n samples, c crops, poses for which crop.py:25-26 would raise, 0 */
} AtCropCameraOutputs;

int64_t at_crop_cameras_workspace_bytes(int n_poses, int max_view_num);
int at_gen_crop_cameras(const AtCropCameraInputs* in, const AtHandModel* hand_model /* shared, mm */,
                        const AtCropCameraOpts* opts, const AtCropCameraOutputs* out, void* workspace,
                        int64_t workspace_bytes, void* stream);

/* at_gen_crop_cameras without a workspace argument and without any allocation.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot allocate device memory (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The scratch between the
 * three launches lives in a file-scope `static __device__` array of the library, sized for
 * AT_CROP_CAMERAS_STATIC_MAX_POSES poses x 4 views (about 4 MB; every GPU that loads the library holds it until the
 * process exits).  in->n_poses beyond that limit returns AT_ERR_WORKSPACE_TOO_SMALL with a message naming the
 * limit.  The call itself runs no cudaMalloc / cudaFree / allocating library routine; the first call on each
 * device looks the array's device address up (cudaGetSymbolAddress) - make one call per device before capturing
 * a graph.  Inputs, outputs and results are exactly at_gen_crop_cameras's.  All calls on a device SHARE the
 * array: it keeps the previous call's contents (nothing in it has to start at zero - every header and slot is
 * written before it is read), and two calls that overlap on the GPU (different streams, or graphs launched
 * concurrently) race on it - serialise them or give the others their own workspace through at_gen_crop_cameras. */
#define AT_CROP_CAMERAS_STATIC_MAX_POSES 2048
int at_gen_crop_cameras_static(const AtCropCameraInputs* in, const AtHandModel* hand_model /* shared, mm */,
                               const AtCropCameraOpts* opts, const AtCropCameraOutputs* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * (2)+(3) Network.  Replaces UmeTrackModel.regress_pose_use_skeleton / regress_pose_pred_skel_scale
 * (lib/models/umetrack_model.py:188-242) and everything below them: backbone
 * (model_utils.py:107-138, backbone_resnet.py:14-158), FTL (model_utils.py:57-104,166-192),
 * 2-view fusion (feature_extractor.py:61-141), SimpleConvRNN (temporal.py:51-139),
 * SkeletonEncoder (skeleton_encoder.py:43-53), PoseRegressor + decoders + Procrustes
 * (regressor.py:76-186, model_utils.py:17-54), wrist->world (umetrack_model.py:77-97).
 *
 * Weights: ONE float32 device blob holding every layer with BatchNorm (eval, eps 1e-5) already
 * folded into the preceding conv/linear.  Layer order and per-layer layout are fixed by
 * at_model_layer_table(); conv weights are stored [kh][kw][Cin][CoutP] with CoutP = Cout
 * rounded up to 4 (zero padded), followed by bias [CoutP].  The host packer is
 * absolutetrack_b200/lib/models/model_loader.py (it consumes the reference's 252-key
 * state dict unchanged, model_loader.py:84-87).
 * ---------------------------------------------------------------------------------------- */
typedef struct AtLayerInfo {
  char name[48];       /* e.g. "backbone.layer2.0.conv1" */
  int32_t cin, cout, ksize, stride;
  int64_t weight_offset; /* float index into the blob */
  int64_t bias_offset;
} AtLayerInfo;

/* Fills up to `capacity` entries; returns the number of layers (or a negative AtStatus). */
int at_model_layer_table(AtLayerInfo* table, int capacity);
/* Total float count of the blob. */
int64_t at_model_blob_floats(void);
/* Prepack the 16-bit operand blob used by the tcgen05 encoder (AT_PREC_BF16 / AT_PREC_FP16) from
 * the fp32 blob: per backbone conv, blocks [k-chunk][tap] of [kc/8][Cout padded to 16][8] elements
 * (the no-swizzle K-major UMMA canonical layout, so a block is one bulk-TMA copy).
 * tc_blob must hold at_model_bf16_blob_bytes() bytes. */
int64_t at_model_bf16_blob_bytes(void);
int at_model_prepack_tc(const float* blob, void* tc_blob, int precision, void* stream);
int at_model_prepack_bf16(const float* blob, void* bf16_blob, void* stream); /* = prepack_tc(AT_PREC_BF16) */

/* Per-view encoder: crops [c,96,96] (AT_U8 grey levels, or AT_F32 already /255) ->
 * features [c,6,6,72] float32 (NHWC; the reference's [c,72,6,6] transposed).
 * workspace must hold at_encoder_workspace_bytes(c, precision) bytes. */
int64_t at_encoder_workspace_bytes(int c, int precision);
int at_encoder_forward(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                       int c, float* features, void* workspace, int64_t workspace_bytes,
                       int precision, void* stream);
/* Same call for a workspace the caller keeps between calls.  The 16-bit encoder works on chunks of
 * at most at_encoder_chunk_crops() crops and relies on halo cells of its activation layout that it
 * zeroes ("prepares") per chunk size.  prepared_chunk = the size of the LAST chunk of the previous
 * call on this workspace (min(c, chunk) or c % chunk) when nothing else wrote the workspace since,
 * else 0; chunks of that size skip the preparation.  AT_PREC_FP32 ignores it. */
int at_encoder_chunk_crops(void);
int at_encoder_forward_prepared(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                                int c, float* features, void* workspace, int64_t workspace_bytes,
                                int precision, int prepared_chunk, void* stream);

/* Same call with a device status word (may be NULL).  AT_PREC_FP16 stores activations as fp16, whose
 * range ends at 65504 where the reference's fp32 activations (backbone_resnet.py:56-72: twelve residual
 * blocks without normalisation after the BN fold) have none: a value beyond it is stored as +-65504, and
 * bit 1 of *status (value AT_STATUS_FP16_OVERFLOW) is set by the layer that produced it - also for inf /
 * NaN - so that an overflow is reported instead of silently clamped.  *status is only ever OR-ed into;
 * it is the same word at_tail_forward sets to AT_STATUS_BAD_SAMPLE_RANGE (bit 0). */
#define AT_STATUS_BAD_SAMPLE_RANGE 1
#define AT_STATUS_FP16_OVERFLOW 2
int at_encoder_forward_status(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                              int c, float* features, void* workspace, int64_t workspace_bytes,
                              int precision, int prepared_chunk, int32_t* status, void* stream);

/* Same call, plus a release point for side-stream work: when release_stream is not NULL it is made to wait
 * (cudaStreamWaitEvent, asynchronous) until the encoder of this call is under way - the 16-bit path records the event on
 * `stream` right after layer1.0.conv1 of the last chunk has been launched, the fp32 path behind the whole encoder.
 * BatchedHandTracker releases the NEXT step's source-box pull (at_crop_pull_source_boxes_u8 on its ingest stream) there:
 * its one-warp CTAs then meet an SM array full of conv CTAs and spread over the registers those leave free, instead of
 * piling up on whichever SMs are idle at a kernel boundary and pushing one CTA of the 148-CTA persistent conv grids into
 * a second wave (measured: compute at half speed while the pull was in flight). */
int at_encoder_forward_release(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                               int c, float* features, void* workspace, int64_t workspace_bytes,
                               int precision, int prepared_chunk, int32_t* status, void* release_stream, void* stream);

/* Same call, plus a FRONT stream (may be NULL): the stream on which the caller produced `crops` (at_crop_resample_* of
 * this step).  The 16-bit path launches the stem of the first chunk there as well and makes `stream` wait for it before
 * layer1.0.conv1, so that the resampler and the stem - ordinary grids, 8 % of a step - run NEXT TO whatever `stream`
 * still has queued ahead of this call instead of behind it.  BatchedHandTracker queues the previous step's tail (18 small
 * latency-bound launches, 14 % of a step) on `stream` and the next step's resampler + stem on the front stream.  The caller
 * orders the front stream after the previous encoder call (the stem overwrites the level-0 planes) and after the
 * producer of the frames; a call that has to prepare the workspace, later chunks, and the fp32 path do not split:
 * there `stream` simply waits for the front stream.  No counterpart in the reference (tracker.py:315-368 runs serially). */
int at_encoder_forward_front(const float* blob, const void* bf16_blob, const void* crops, int crop_dtype,
                             int c, float* features, void* workspace, int64_t workspace_bytes,
                             int precision, int prepared_chunk, int32_t* status, void* front_stream,
                             void* release_stream, void* stream);

/* Test/debug entry: ONE backbone conv layer (layer_id from at_model_layer_table, 1..28) through the
 * tcgen05 path on float32 NHWC tensors: in [n,H,W,cin] -> out [n,H/stride,W/stride,cout], optional
 * residual res [n,H/stride,W/stride,cout], optional ReLU.  mt = 128-pixel sub-tiles per CTA tile (0 = default). */
int64_t at_debug_conv_tc_workspace_bytes(int n_img, int H, int W);
int at_debug_conv_tc(const float* blob, const void* tc_blob, int layer_id, const float* in, int n_img, int H, int W,
                     const float* res, int relu, int precision, int mt, float* out, void* workspace,
                     int64_t workspace_bytes, void* stream);

void at_debug_set_prof(long long* dev_ptr); /* debug: per-CTA cycle counters of conv_tc launches */
void at_debug_set_prof_seq(long long* dev_ptr); /* debug: same, launch k of the sequence writes at dev_ptr + k*148*16 */

/* Recurrent state of SimpleConvRNN (temporal.py:41-42,105-125), owned by the caller. */
typedef struct AtTemporalState {
  float* mem_features;    /* [capacity,6,6,18] NHWC */
  float* prev_extrinsics; /* [capacity,4,4] */
  int64_t capacity;
} AtTemporalState;

typedef struct AtTailInputs {
  const float* features;       /* [c,6,6,72] from at_encoder_forward */
  const float* intrinsics;     /* [c,3,3] */
  const float* extrinsics;     /* [c,4,4] world->crop-eye, metres */
  const int64_t* sample_range; /* [n,2] (start,end) into the c crops; end-start in {1,2} */
  const int64_t* memory_idx;   /* [n]; values must be distinct within a call: two samples sharing a slot race on it
                                * (the reference, temporal.py:77,134, is last-write-wins; trackers give every (sequence, hand) its own slot) */
  const uint8_t* use_memory;   /* [n] bool */
  const int64_t* hand_idx;     /* [n] */
  const float* skel_axes;      /* [1 or n,22,3]  (NULL for AT_HEAD_PREDICT_SCALE) */
  const float* skel_rest;      /* [1 or n,22,3]  metres */
  int32_t n;                   /* hand samples */
  int32_t c;                   /* crops */
  int32_t n_skel;              /* 1 or n */
  int32_t head_mode;           /* AtHeadMode */
} AtTailInputs;

typedef struct AtTailOutputs {
  float* joint_angles; /* [n,22] */
  float* wrist_xfs;    /* [n,4,4] wrist->world (metres), right hands mirrored */
  float* skel_scales;  /* [n] (AT_HEAD_PREDICT_SCALE only, else may be NULL) */
  float* sigmas;       /* [n,21] */
  int32_t* status;     /* [1] device flag: set non-zero on a bad sample_range / memory_idx */
} AtTailOutputs;

int64_t at_tail_workspace_bytes(int n);
int at_tail_forward(const float* blob, const AtTailInputs* in, AtTemporalState* state,
                    const AtTailOutputs* out, void* workspace, int64_t workspace_bytes,
                    void* stream);
/* Same arithmetic contract (fp32-grade: every product carries 22 significant bits), twice the
 * tensor throughput: the 11 convolutions of the tail run as error-compensated fp16 x 3 GEMMs
 * (x = hi + lo in fp16; hi*hi + lo*hi + hi*lo accumulated in fp32, mma.m16n8k16) instead of
 * TF32 x 3.  The four 3x3 convolutions of the pose head (2/3 of the tail's FLOPs) run the same
 * three-product scheme as tcgen05 shift-GEMMs over [hi | lo | hi] activation planes and
 * [hi | hi | lo] weight planes (fp32 accumulation in TMEM).  f16_blob = at_tail_f16_blob_bytes()
 * bytes filled once per weight blob by at_tail_prepack_f16 (pre-split, pre-transposed weights for
 * both kernels); NULL selects the TF32 x 3 kernels, i.e. at_tail_forward. */
int64_t at_tail_f16_blob_bytes(void);
int at_tail_prepack_f16(const float* blob, void* f16_blob, void* stream);
int at_tail_forward_f16(const float* blob, const void* f16_blob, const AtTailInputs* in,
                        AtTemporalState* state, const AtTailOutputs* out, void* workspace,
                        int64_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ABSOLUTETRACK_B200_H_ */
