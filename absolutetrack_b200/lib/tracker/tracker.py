"""HandTracker on the B200 (mirror of the reference's lib/tracker/tracker.py API).

Same names and signatures: ViewData / InputFrame / HandTrackerOpts (:40-58), _warp_image (:61-89),
HandTracker.{gen_crop_cameras :222-260, track_frame :262-289, track_frame_and_calibrate_scale
:291-313, reset_history :107}.  The perspective-crop warp runs in csrc/crop_resample.cu and the
network in csrc/{conv_*,tail}.cu; the only host work per frame is building the 256-byte
AtCropParams records and the small intrinsics/extrinsics arrays.
"""
import ctypes
import logging
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _cabi
from ..common import camera
from ..common.hand import HandModel, NUM_HANDS, scaled_hand_model
from ..models.regressor import RegressorOutput
from ..models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData
from .perspective_crop import gen_crop_cameras_from_pose
from .tracking_result import SingleHandPose, TrackingResult

logger = logging.getLogger(__name__)

MM_TO_M = 0.001
M_TO_MM = 1000.0
MIN_OBSERVED_LANDMARKS = 21
CONFIDENCE_THRESHOLD = 0.5
MAX_VIEW_NUM = 2

MAP_MODES = {"fp32": _cabi.AT_MAP_FP32, "fp64": _cabi.AT_MAP_FP64}


@dataclass
class ViewData:
    image: np.ndarray
    camera: camera.CameraModel
    camera_angle: float


@dataclass
class InputFrame:
    views: List[ViewData]


@dataclass
class HandTrackerOpts:
    num_crop_points: int = 63
    enable_memory: bool = True
    use_stored_pose_for_crop: bool = True
    hand_ratio_in_crop: float = 0.8
    min_required_vis_landmarks: int = 19
    # B200 extension: "fp64" reproduces the reference's float64 map bit for bit, "fp32" is the
    # HBM-roofline mode (<= 1.2e-4 px map error)
    map_mode: str = "fp64"
    # B200 extension: "device" builds the crop cameras with csrc/crop_cameras.cu (one launch set per
    # frame), "host" with the numpy mirror of perspective_crop.py / crop.py (FK still on the GPU)
    crop_cameras: str = "device"


def pack_crop_params(records) -> np.ndarray:
    """records: list of (frame, view, dict from camera.crop_kernel_params) -> uint8 [c,256]."""
    arr = (_cabi.AtCropParams * len(records))()
    for a, (frame, view, p) in zip(arr, records):
        a.frame, a.view, a.src_w, a.src_h = frame, view, p["src_w"], p["src_h"]
        a.M[:] = p["M"].tolist()
        a.d[:] = p["d"].tolist()
        a.f_dst[:] = p["f_dst"].tolist()
        a.c_dst[:] = p["c_dst"].tolist()
        a.f_src[:] = p["f_src"].tolist()
        a.c_src[:] = p["c_src"].tolist()
        a.k[:] = p["k"].tolist()
    buf = np.frombuffer(arr, dtype=np.uint8).reshape(len(records), _cabi.CROP_PARAMS_BYTES)
    return buf.copy()


def crop_resample_cuda(
    src: torch.Tensor,
    frame_stride: int,
    view_stride: int,
    row_stride: int,
    params: torch.Tensor,
    dst_wh=(96, 96),
    out_dtype: torch.dtype = torch.float32,
    map_mode: str = "fp64",
    out: Optional[torch.Tensor] = None,
    staged_bytes: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """Launch at_crop_resample_fisheye62_u8.  `src` uint8 CUDA tensor - or, in map_mode "fp32", a
    PINNED host tensor (zero-copy ingest: the kernel pulls each crop's source bounding box over
    PCIe once, through shared memory); `params` uint8 CUDA [c,256]."""
    lib = _cabi.lib()
    assert src.dtype == torch.uint8 and params.is_cuda and params.dtype == torch.uint8
    if not src.is_cuda and not (src.is_pinned() and map_mode == "fp32"):
        raise _cabi.AbsoluteTrackError("crop source must be a CUDA tensor (or a pinned host tensor with map_mode='fp32')")
    c = params.shape[0]
    W, H = int(dst_wh[0]), int(dst_wh[1])
    code = {torch.uint8: _cabi.AT_U8, torch.float32: _cabi.AT_F32, torch.bfloat16: _cabi.AT_BF16}[out_dtype]
    if out is None:
        out = torch.empty((c, H, W), dtype=out_dtype, device=params.device)
    with torch.cuda.device(params.device):
        rc = lib.at_crop_resample_fisheye62_u8_ex(
            src.data_ptr(), frame_stride, view_stride, row_stride, params.contiguous().data_ptr(), c, W, H,
            out.data_ptr(), code, MAP_MODES[map_mode], _cabi.ptr(staged_bytes), _cabi.stream_ptr(params.device),
        )
    _cabi.check(rc, "at_crop_resample_fisheye62_u8")
    return out


def crop_pull_source_boxes_cuda(
    src: torch.Tensor,
    dst: torch.Tensor,
    frame_stride: int,
    view_stride: int,
    row_stride: int,
    params: torch.Tensor,
    dst_wh=(96, 96),
    max_ctas: int = 0,
    pulled_bytes: Optional[torch.Tensor] = None,
) -> None:
    """Launch at_crop_pull_source_boxes_u8 on the current stream: copy, for every crop in `params`,
    the source box its map touches from the PINNED host frames `src` into the device buffer `dst`
    of the same shape (bytes outside the boxes are left as they are and never read by the
    resampler)."""
    lib = _cabi.lib()
    assert src.dtype == torch.uint8 and dst.dtype == torch.uint8 and dst.is_cuda and params.is_cuda
    if src.shape != dst.shape or not src.is_contiguous() or not dst.is_contiguous():
        raise _cabi.AbsoluteTrackError("pull source and destination must be contiguous and of one shape")
    if not (src.is_cuda or src.is_pinned()):
        raise _cabi.AbsoluteTrackError("the pull source must be pinned host memory (pageable memory is not GPU-readable)")
    n_frames = src.numel() // frame_stride
    max_frames = ((1 << 32) - 32) // frame_stride          # the kernel addresses a call's frames with 32-bit offsets
    if n_frames > max_frames:                               # rare (> 4 GiB of frames per step): one call per frame range
        frame_of = params.view(torch.int32)[:, 0]
        flat_src, flat_dst = src.reshape(-1), dst.reshape(-1)
        for f0 in range(0, n_frames, max_frames):
            f1 = min(n_frames, f0 + max_frames)
            sel = ((frame_of >= f0) & (frame_of < f1)).nonzero().squeeze(1)
            if sel.numel() == 0:
                continue
            sub = params[sel].contiguous()
            sub.view(torch.int32)[:, 0] -= f0
            crop_pull_source_boxes_cuda(flat_src[f0 * frame_stride: f1 * frame_stride], flat_dst[f0 * frame_stride: f1 * frame_stride],
                                        frame_stride, view_stride, row_stride, sub, dst_wh, max_ctas, pulled_bytes)
        return
    with torch.cuda.device(dst.device):
        rc = lib.at_crop_pull_source_boxes_u8(
            src.data_ptr(), dst.data_ptr(), src.numel() // frame_stride, frame_stride, view_stride, row_stride,
            params.contiguous().data_ptr(), params.shape[0], int(dst_wh[0]), int(dst_wh[1]), int(max_ctas),
            _cabi.ptr(pulled_bytes), _cabi.stream_ptr(dst.device),
        )
    _cabi.check(rc, "at_crop_pull_source_boxes_u8")


def _warp_image(
    src_camera: camera.CameraModel,
    dst_camera: camera.CameraModel,
    src_image: np.ndarray,
    interpolation: int = 1,  # cv2.INTER_LINEAR
    depth_check: bool = True,
    map_mode: str = "fp64",
) -> np.ndarray:
    """Reference tracker.py:61-89 on the GPU: uint8 [H,W] fisheye image -> uint8 crop."""
    if interpolation != 1 or not depth_check:
        raise NotImplementedError("the CUDA resampler implements INTER_LINEAR with depth_check=True")
    img = torch.from_numpy(np.ascontiguousarray(src_image)).cuda()
    H, W = img.shape
    params = torch.from_numpy(pack_crop_params([(0, 0, camera.crop_kernel_params(src_camera, dst_camera))])).cuda()
    out = crop_resample_cuda(
        img, 0, 0, W, params, (dst_camera.width, dst_camera.height), torch.uint8, map_mode
    )
    return out[0].cpu().numpy()


class HandTracker:
    def __init__(self, model, opts: HandTrackerOpts) -> None:
        if not torch.cuda.is_available():
            raise _cabi.AbsoluteTrackError("absolutetrack_b200.HandTracker needs a CUDA device (no CPU fallback)")
        self._device = torch.device("cuda", torch.cuda.current_device())
        logger.info(f"Using device: {self._device}")
        self._model = model
        self._model.to(self._device)
        self._model.memory_capacity_hint = NUM_HANDS  # memory_idx = hand_idx < NUM_HANDS (tracker.py:353)
        self._input_size = np.array(self._model.getInputImageSizes())
        self._num_crop_points = opts.num_crop_points
        self._enable_memory = opts.enable_memory
        self._hand_ratio_in_crop: float = opts.hand_ratio_in_crop
        self._min_required_vis_landmarks: int = opts.min_required_vis_landmarks
        self._map_mode = getattr(opts, "map_mode", "fp64")
        self._valid_tracking_history = np.zeros(2, dtype=bool)
        self._crop_cameras_mode = getattr(opts, "crop_cameras", "device")
        if self._crop_cameras_mode not in ("device", "host"):
            raise ValueError("HandTrackerOpts.crop_cameras must be 'device' or 'host'")
        self._rig_cache = {}

    def _gen_crop_cameras_device(self, cameras, camera_angles, hand_model, gt_tracking, min_num_crops):
        """tracker.py:222-260 in one at_gen_crop_cameras call for all hands of the frame."""
        from .crop_cameras import CropCameraOpts, DeviceRig, crop_cameras_to_dict, gen_crop_inputs_cuda

        key = (id(hand_model), tuple(float(a) for a in camera_angles),
               tuple((c.width, c.height, tuple(c.f), tuple(c.c), tuple(c.distort)) for c in cameras))
        rig = self._rig_cache.get(key)
        if rig is None:
            self._rig_cache.clear()
            rig = self._rig_cache[key] = DeviceRig(cameras, camera_angles, hand_model, self._device)
            rig._hand_model_ref = hand_model  # keeps id(hand_model) unique while cached
        hands = list(gt_tracking.keys())
        dev = self._device
        ja = torch.from_numpy(np.stack([np.asarray(gt_tracking[h].joint_angles, np.float64) for h in hands])).float().to(dev)
        wx = torch.from_numpy(np.stack([np.asarray(gt_tracking[h].wrist_xform, np.float64) for h in hands])).float().to(dev)
        valid = torch.tensor([gt_tracking[h].hand_confidence >= CONFIDENCE_THRESHOLD for h in hands], dtype=torch.bool).to(dev)
        c2w = torch.from_numpy(np.stack([np.asarray(c.camera_to_world_xf, np.float64) for c in cameras])[None]).contiguous().to(dev)
        opts = CropCameraOpts(
            num_crop_points=self._num_crop_points, min_required_vis_landmarks=self._min_required_vis_landmarks,
            max_view_num=MAX_VIEW_NUM, min_num_crops=min_num_crops, sort_camera_index=True, mirror_right_hand=True,
            crop_size=tuple(int(s) for s in self._input_size), focal_multiplier=self._hand_ratio_in_crop,
        )
        res = gen_crop_inputs_cuda(
            rig, ja, wx, torch.zeros(len(hands), dtype=torch.int32, device=dev),
            torch.tensor(hands, dtype=torch.int32).to(dev), c2w, opts, pose_valid=valid,
        )
        return crop_cameras_to_dict(res.resolve(), opts.crop_size)

    def reset_history(self) -> None:
        self._valid_tracking_history[:] = False

    def gen_crop_cameras(
        self,
        cameras: List[camera.CameraModel],
        camera_angles: List[float],
        hand_model: HandModel,
        gt_tracking: Dict[int, SingleHandPose],
        min_num_crops: int,
    ) -> Dict[int, Dict[int, camera.PinholePlaneCameraModel]]:
        crop_cameras: Dict[int, Dict[int, camera.PinholePlaneCameraModel]] = {}
        if not gt_tracking:
            return crop_cameras
        if self._crop_cameras_mode == "device":
            return self._gen_crop_cameras_device(cameras, camera_angles, hand_model, gt_tracking, min_num_crops)
        for hand_idx, pose in gt_tracking.items():
            if pose.hand_confidence < CONFIDENCE_THRESHOLD:
                continue
            per_hand = gen_crop_cameras_from_pose(
                cameras,
                camera_angles,
                hand_model,
                pose,
                hand_idx,
                self._num_crop_points,
                self._input_size,
                max_view_num=MAX_VIEW_NUM,
                sort_camera_index=True,
                focal_multiplier=self._hand_ratio_in_crop,
                mirror_right_hand=True,
                min_required_vis_landmarks=self._min_required_vis_landmarks,
            )
            if per_hand and len(per_hand) >= min_num_crops:
                crop_cameras[hand_idx] = per_hand
        return crop_cameras

    def track_frame(
        self,
        sample: InputFrame,
        hand_model: HandModel,
        crop_cameras: Dict[int, Dict[int, camera.PinholePlaneCameraModel]],
    ) -> TrackingResult:
        if not crop_cameras:
            self.reset_history()
            return TrackingResult()
        frame_data, frame_desc, skeleton_data = self._make_inputs(sample, hand_model, crop_cameras)
        out = self._model.regress_pose_use_skeleton(frame_data, frame_desc, skeleton_data)
        return self._gen_tracking_result(out, frame_desc.hand_idx.cpu().numpy(), crop_cameras)

    def track_frame_and_calibrate_scale(
        self,
        sample: InputFrame,
        crop_cameras: Dict[int, Dict[int, camera.PinholePlaneCameraModel]],
    ) -> TrackingResult:
        if not crop_cameras:
            self.reset_history()
            return TrackingResult()
        frame_data, frame_desc, _ = self._make_inputs(sample, None, crop_cameras)
        out = self._model.regress_pose_pred_skel_scale(frame_data, frame_desc)
        return self._gen_tracking_result(out, frame_desc.hand_idx.cpu().numpy(), crop_cameras)

    def _make_inputs(self, sample: InputFrame, hand_model_mm: Optional[HandModel], crop_cameras):
        """tracker.py:315-368, with the warp + /255 done by the CUDA resampler."""
        records, intrinsics, extrinsics, ranges, hands = [], [], [], [], []
        used_views: Dict[int, int] = {}
        for hand_idx, per_hand in crop_cameras.items():
            start = len(records)
            for cam_idx, crop_cam in per_hand.items():
                view = sample.views[cam_idx]
                slot = used_views.setdefault(cam_idx, len(used_views))
                records.append((0, slot, camera.crop_kernel_params(view.camera, crop_cam)))
                intrinsics.append(crop_cam.uv_to_window_matrix())
                w2e = np.linalg.inv(crop_cam.camera_to_world_xf)
                w2e[:3, 3] *= MM_TO_M
                extrinsics.append(w2e)
            if len(records) > start:
                hands.append(hand_idx)
                ranges.append((start, len(records)))
        dev = self._device
        first = sample.views[next(iter(used_views))].image
        H, W = first.shape
        stage = torch.empty((len(used_views), H, W), dtype=torch.uint8, pin_memory=True)
        for cam_idx, slot in used_views.items():
            stage[slot].copy_(torch.from_numpy(np.ascontiguousarray(sample.views[cam_idx].image)))
        src = stage.to(dev, non_blocking=True)
        params = torch.from_numpy(pack_crop_params(records)).to(dev, non_blocking=True)
        crops = crop_resample_cuda(
            src, 0, H * W, W, params, tuple(int(s) for s in self._input_size), torch.float32, self._map_mode
        )
        hands_np = np.array(hands)
        frame_data = InputFrameData(
            left_images=crops,
            intrinsics=torch.from_numpy(np.stack(intrinsics)).float().to(dev),
            extrinsics_xf=torch.from_numpy(np.stack(extrinsics)).float().to(dev),
        )
        frame_desc = InputFrameDesc(
            sample_range=torch.from_numpy(np.array(ranges)).long().to(dev),
            memory_idx=torch.from_numpy(hands_np).long().to(dev),
            use_memory=torch.from_numpy(self._valid_tracking_history[hands_np]).bool().to(dev),
            hand_idx=torch.from_numpy(hands_np).long().to(dev),
        )
        skeleton_data = None
        if hand_model_mm is not None:
            hm = scaled_hand_model(hand_model_mm, MM_TO_M)
            skeleton_data = InputSkeletonData(
                joint_rotation_axes=hm.joint_rotation_axes.float().to(dev),
                joint_rest_positions=hm.joint_rest_positions.float().to(dev),
            )
        return frame_data, frame_desc, skeleton_data

    def _gen_tracking_result(self, regressor_output: RegressorOutput, hand_indices: np.ndarray, crop_cameras):
        """tracker.py:370-412."""
        angles = regressor_output.joint_angles.to("cpu").numpy()
        wrists = regressor_output.wrist_xfs.to("cpu").numpy()
        wrists[..., :3, 3] *= M_TO_MM
        scales = None
        if regressor_output.skel_scales is not None:
            scales = regressor_output.skel_scales.to("cpu").numpy()
        hand_poses, num_views, predicted_scales = {}, {}, {}
        for i, hand_idx in enumerate(hand_indices):
            hand_poses[hand_idx] = SingleHandPose(joint_angles=angles[i], wrist_xform=wrists[i], hand_confidence=1.0)
            num_views[hand_idx] = len(crop_cameras[hand_idx])
            if scales is not None:
                predicted_scales[hand_idx] = scales[i]
        for hand_idx in range(NUM_HANDS):
            self._valid_tracking_history[hand_idx] = hand_idx in hand_poses
        return TrackingResult(hand_poses=hand_poses, num_views=num_views, predicted_scales=predicted_scales)
