"""Crop cameras and model inputs built on the GPU (B200 extension of lib/tracker).

The reference builds crop cameras on the host, one hand at a time: HandTracker.gen_crop_cameras
(lib/tracker/tracker.py:222-260) -> perspective_crop.gen_crop_cameras_from_pose (:136-180) ->
crop.gen_crop_parameters_from_points (crop.py:31-82), then HandTracker._make_inputs
(tracker.py:315-368) turns them into tensors.  `gen_crop_inputs_cuda` does both for a whole batch
of poses in three small launches (csrc/crop_cameras.cu, C-ABI at_gen_crop_cameras) and leaves
everything on the device: AtCropParams records for the resampler, intrinsics / extrinsics /
sample_range / hand_idx / memory_idx / use_memory for the model.  There is no CPU path.
"""
import ctypes
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _cabi
from ..common import camera
from ..common.hand import HandModel


@dataclass
class CropCameraOpts:
    """The options gen_crop_cameras passes down (tracker.py:49-58,238-253)."""

    num_crop_points: int = 63
    min_required_vis_landmarks: int = 19
    max_view_num: int = 2
    min_num_crops: int = 1
    sort_camera_index: bool = True
    mirror_right_hand: bool = True
    crop_size: tuple = (96, 96)
    focal_multiplier: float = 0.8
    min_focal: float = 5.0

    def to_c(self) -> _cabi.AtCropCameraOpts:
        return _cabi.AtCropCameraOpts(
            int(self.num_crop_points), int(self.min_required_vis_landmarks), int(self.max_view_num),
            int(self.min_num_crops), int(bool(self.sort_camera_index)), int(bool(self.mirror_right_hand)),
            int(self.crop_size[0]), int(self.crop_size[1]), float(self.focal_multiplier), float(self.min_focal),
        )


def pack_rig_cameras(cameras: Sequence[camera.CameraModel], camera_angles: Sequence[float]) -> np.ndarray:
    """uint8 [n_cams,112]: AtRigCamera records of the headset's Fisheye62 cameras."""
    arr = (_cabi.AtRigCamera * len(cameras))()
    for a, cam, ang in zip(arr, cameras, camera_angles):
        if not isinstance(cam.distort, camera.Fisheye62Distortion):
            raise TypeError("the crop-camera kernel implements Fisheye62 source cameras (reference eval configs)")
        a.f[:] = [float(v) for v in cam.f]
        a.c[:] = [float(v) for v in cam.c]
        a.k[:] = [float(v) for v in tuple(cam.distort)]
        a.angle_deg = float(ang)
        a.width, a.height = int(cam.width), int(cam.height)
    return np.frombuffer(arr, dtype=np.uint8).reshape(len(cameras), _cabi.RIG_CAMERA_BYTES).copy()


class DeviceRig:
    """Per-recording constants on the device: the rig's cameras, the user's hand model (mm), the
    neutral pose (perspective_crop.neutral_joint_angles :17-24)."""

    def __init__(self, cameras, camera_angles, hand_model: HandModel, device=None):
        if not torch.cuda.is_available():
            raise _cabi.AbsoluteTrackError("DeviceRig needs a CUDA device (no CPU fallback)")
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.device = dev
        self.n_cams = len(cameras)
        self.cameras = torch.from_numpy(pack_rig_cameras(cameras, camera_angles)).to(dev)
        f32 = dict(device=dev, dtype=torch.float32)
        if hand_model.joint_rest_positions.dim() != 2:
            raise ValueError("one shared hand model per rig")
        K = int(hand_model.landmark_rest_bone_indices.shape[-1])
        self._hm_tensors = (
            hand_model.joint_rotation_axes.to(**f32).contiguous(),
            hand_model.joint_rest_positions.to(**f32).contiguous(),
            hand_model.landmark_rest_positions.to(**f32).contiguous(),
            hand_model.landmark_rest_bone_weights.to(**f32).contiguous(),
            hand_model.landmark_rest_bone_indices.to(device=dev, dtype=torch.int32).contiguous(),
        )
        self.hm = _cabi.AtHandModel(*[t.data_ptr() for t in self._hm_tensors], K, 0)
        lim = hand_model.joint_limits
        if lim is None:
            self.neutral = None
        else:
            self.neutral = (lim[..., 0] * 0.5 + lim[..., 1] * (1 - 0.5)).to(**f32).contiguous()


@dataclass
class CropInputs:
    """Device outputs of at_gen_crop_cameras at CAPACITY size; `counts` = [n, c, errors, 0].
    `resolve()` reads the counts (one small D2H + sync) and returns the sized views."""

    crop_params: torch.Tensor      # uint8 [cap_c,256]
    intrinsics: torch.Tensor       # f32 [cap_c,3,3]
    extrinsics: torch.Tensor       # f32 [cap_c,4,4]
    crop_camera_to_world: torch.Tensor  # f64 [cap_c,4,4]
    crop_camera_index: torch.Tensor     # i32 [cap_c]
    sample_range: torch.Tensor     # i64 [cap_n,2]
    hand_idx: torch.Tensor         # i64 [cap_n]
    memory_idx: torch.Tensor       # i64 [cap_n]
    use_memory: torch.Tensor       # bool [cap_n]
    pose_index: torch.Tensor       # i32 [cap_n]
    counts: torch.Tensor           # i32 [4]
    n: Optional[int] = None
    c: Optional[int] = None

    def resolve(self, counts_host: Optional[torch.Tensor] = None) -> "CropInputs":
        cnt = self.counts.cpu() if counts_host is None else counts_host
        n, c, err = int(cnt[0]), int(cnt[1]), int(cnt[2])
        if err:
            raise ValueError("Unable to create crop camera", err)  # crop.py:25-26
        return CropInputs(
            self.crop_params[:c], self.intrinsics[:c], self.extrinsics[:c], self.crop_camera_to_world[:c],
            self.crop_camera_index[:c], self.sample_range[:n], self.hand_idx[:n], self.memory_idx[:n],
            self.use_memory[:n], self.pose_index[:n], self.counts, n, c,
        )


def alloc_crop_inputs(n_poses: int, max_view_num: int, device) -> CropInputs:
    cap_c = n_poses * max_view_num
    e = lambda shape, dt: torch.empty(shape, dtype=dt, device=device)  # noqa: E731
    return CropInputs(
        e((cap_c, _cabi.CROP_PARAMS_BYTES), torch.uint8), e((cap_c, 3, 3), torch.float32), e((cap_c, 4, 4), torch.float32),
        e((cap_c, 4, 4), torch.float64), e((cap_c,), torch.int32), e((n_poses, 2), torch.int64), e((n_poses,), torch.int64),
        e((n_poses,), torch.int64), e((n_poses,), torch.bool), e((n_poses,), torch.int32),
        torch.zeros((4,), dtype=torch.int32, device=device),
    )


def gen_crop_inputs_cuda(
    rig: DeviceRig,
    joint_angles: torch.Tensor,      # f32 [P,22] device
    wrist_xforms: torch.Tensor,      # f32 [P,4,4] device (mm, right hands not mirrored)
    pose_frame: torch.Tensor,        # i32 [P] device
    pose_hand: torch.Tensor,         # i32 [P] device
    camera_to_world: torch.Tensor,   # f64 [F,n_cams,4,4] device
    opts: CropCameraOpts,
    pose_valid: Optional[torch.Tensor] = None,        # bool/u8 [P] device
    tracking_history: Optional[torch.Tensor] = None,  # bool/u8 [F*2] device, updated in place
    out: Optional[CropInputs] = None,
    workspace: Optional[torch.Tensor] = None,
) -> CropInputs:
    """Launch at_gen_crop_cameras on the current stream.  Nothing synchronises; call
    `.resolve()` on the result to learn (n, c)."""
    lib = _cabi.lib()
    dev = rig.device
    P = int(joint_angles.shape[0])
    for t, dt in ((joint_angles, torch.float32), (wrist_xforms, torch.float32), (pose_frame, torch.int32),
                  (pose_hand, torch.int32), (camera_to_world, torch.float64)):
        if not (t.is_cuda and t.dtype == dt and t.is_contiguous()):
            raise _cabi.AbsoluteTrackError(f"gen_crop_inputs_cuda: expected a contiguous CUDA {dt} tensor")
    if opts.num_crop_points > 21 and rig.neutral is None:
        raise AssertionError("hand model without joint_limits")  # perspective_crop.py:19
    F = int(camera_to_world.shape[0])
    if out is None:
        out = alloc_crop_inputs(P, opts.max_view_num, dev)
    need = lib.at_crop_cameras_workspace_bytes(P, opts.max_view_num)
    if need < 0:
        raise _cabi.AbsoluteTrackError("at_crop_cameras_workspace_bytes: invalid arguments")
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty((max(int(need), 16),), dtype=torch.uint8, device=dev)
    cin = _cabi.AtCropCameraInputs(
        joint_angles.data_ptr(), wrist_xforms.data_ptr(), pose_frame.data_ptr(), pose_hand.data_ptr(),
        _cabi.ptr(pose_valid), _cabi.ptr(rig.neutral), camera_to_world.data_ptr(), rig.cameras.data_ptr(),
        P, F, rig.n_cams, 0,
    )
    cout = _cabi.AtCropCameraOutputs(
        out.crop_params.data_ptr(), out.intrinsics.data_ptr(), out.extrinsics.data_ptr(),
        out.crop_camera_to_world.data_ptr(), out.crop_camera_index.data_ptr(), out.sample_range.data_ptr(),
        out.hand_idx.data_ptr(), out.memory_idx.data_ptr(), out.use_memory.data_ptr(), out.pose_index.data_ptr(),
        _cabi.ptr(tracking_history), out.counts.data_ptr(),
    )
    copts = opts.to_c()
    with torch.cuda.device(dev):
        rc = lib.at_gen_crop_cameras(ctypes.byref(cin), ctypes.byref(rig.hm), ctypes.byref(copts), ctypes.byref(cout),
                                     workspace.data_ptr(), workspace.numel(), _cabi.stream_ptr(dev))
    _cabi.check(rc, "at_gen_crop_cameras")
    out._workspace = workspace  # keep alive until the stream has consumed it
    return out


def crop_cameras_to_dict(res: CropInputs, crop_size=(96, 96)) -> Dict[int, Dict[int, camera.PinholePlaneCameraModel]]:
    """Resolved single-frame result -> the Dict[hand][camera] of PinholePlaneCameraModel that
    HandTracker.gen_crop_cameras returns (tracker.py:229)."""
    r = res if res.n is not None else res.resolve()
    rng = r.sample_range.cpu().numpy()
    hands = r.hand_idx.cpu().numpy()
    cams = r.crop_camera_index.cpu().numpy()
    c2w = r.crop_camera_to_world.cpu().numpy()
    par = r.crop_params.cpu().numpy()
    out: Dict[int, Dict[int, camera.PinholePlaneCameraModel]] = {}
    for i, h in enumerate(hands):
        per = {}
        for c in range(int(rng[i, 0]), int(rng[i, 1])):
            rec = _cabi.AtCropParams.from_buffer_copy(par[c].tobytes())
            per[int(cams[c])] = camera.PinholePlaneCameraModel(
                width=int(crop_size[0]), height=int(crop_size[1]), f=np.array(rec.f_dst[:]), c=np.array(rec.c_dst[:]), distort_coeffs=[],
                camera_to_world_xf=c2w[c].copy(),
            )
        out[int(h)] = per
    return out
