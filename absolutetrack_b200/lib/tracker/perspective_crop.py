"""Crop-camera selection and landmarks from poses (mirror of the reference's
lib/tracker/perspective_crop.py API).  FK/skinning runs on the GPU (csrc/fk_skin.cu)."""
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..common import camera
from ..common.crop import gen_crop_parameters_from_points
from ..common.hand import HandModel, NUM_JOINTS_PER_HAND, RIGHT_HAND_INDEX
from ..common.hand_skinning import skin_landmarks_cuda
from .tracking_result import SingleHandPose


def neutral_joint_angles(up: HandModel, lower_factor: float = 0.5) -> torch.Tensor:
    limits = up.joint_limits
    assert limits is not None
    return limits[..., 0] * lower_factor + limits[..., 1] * (1 - lower_factor)


def skin_landmarks_np(hand_model: HandModel, joint_angles: np.ndarray, wrist_transforms: np.ndarray) -> np.ndarray:
    """perspective_crop.py:27-37: numpy in, numpy out, float32 FK."""
    out = skin_landmarks_cuda(
        hand_model, torch.from_numpy(np.asarray(joint_angles)).float(), torch.from_numpy(np.asarray(wrist_transforms)).float()
    )
    return out.cpu().numpy()


def landmarks_from_hand_pose(hand_model: HandModel, hand_pose: SingleHandPose, hand_idx: int) -> np.ndarray:
    """World-space landmarks [21,3] of a (left-model) hand pose; right hands are mirrored
    (perspective_crop.py:40-51)."""
    xf = np.array(hand_pose.wrist_xform, copy=True)
    if hand_idx == RIGHT_HAND_INDEX:
        xf[:, 0] *= -1
    return skin_landmarks_np(hand_model, hand_pose.joint_angles, xf)


def _landmarks_multi(hand_model: HandModel, angle_sets: List[np.ndarray], wrist_xform: np.ndarray, hand_idx: int):
    """Landmarks of several joint-angle sets sharing one wrist transform, in ONE kernel launch."""
    xf = np.array(wrist_xform, dtype=np.float64, copy=True)
    if hand_idx == RIGHT_HAND_INDEX:
        xf[:, 0] *= -1
    ja = torch.from_numpy(np.stack([np.asarray(a, dtype=np.float64) for a in angle_sets])).float()
    wx = torch.from_numpy(np.broadcast_to(xf, (len(angle_sets), 4, 4)).copy()).float()
    per = hand_model
    if len(hand_model.joint_rest_positions.shape) == 2:
        out = skin_landmarks_cuda(per, ja, wx)
    else:
        raise ValueError("batched hand models are not supported here")
    return out.cpu().numpy()


def rank_hand_visibility_in_cameras(
    cameras: List[camera.CameraModel],
    hand_model: HandModel,
    hand_pose: SingleHandPose,
    hand_idx: int,
    min_required_vis_landmarks: int,
    landmarks_world: Optional[np.ndarray] = None,
) -> List[int]:
    """Cameras that see at least `min_required_vis_landmarks` landmarks, most first
    (perspective_crop.py:54-86)."""
    if landmarks_world is None:
        landmarks_world = landmarks_from_hand_pose(hand_model, hand_pose, hand_idx)
    counts, keep = [], []
    for cam_idx, cam in enumerate(cameras):
        eye = cam.world_to_eye(landmarks_world)
        win = cam.eye_to_window(eye)
        n_vis = int(
            (
                (win[..., 0] >= 0)
                & (win[..., 0] <= cam.width - 1)
                & (win[..., 1] >= 0)
                & (win[..., 1] <= cam.height - 1)
                & (eye[..., 2] > 0)
            ).sum()
        )
        counts.append(n_vis)
        if n_vis >= min_required_vis_landmarks:
            keep.append(cam_idx)
    keep.sort(reverse=True, key=lambda i: counts[i])
    return keep


def _get_crop_points_from_hand_pose(hand_model, gt_hand_pose, hand_idx, num_crop_points) -> np.ndarray:
    """Landmarks of the given pose (+ neutral pose, + open pose) under the same wrist transform
    (perspective_crop.py:89-133); 21, 42 or 63 points."""
    assert num_crop_points in [21, 42, 63]
    sets = [np.asarray(gt_hand_pose.joint_angles)]
    if num_crop_points > 21:
        sets.append(neutral_joint_angles(hand_model).numpy())
    if num_crop_points > 42:
        sets.append(np.zeros(NUM_JOINTS_PER_HAND, dtype=np.float32))
    lm = _landmarks_multi(hand_model, sets, gt_hand_pose.wrist_xform, hand_idx)
    return lm.reshape(-1, 3)


def gen_crop_cameras_from_pose(
    cameras: List[camera.CameraModel],
    camera_angles: List[float],
    hand_model: HandModel,
    hand_pose: SingleHandPose,
    hand_idx: int,
    num_crop_points: int,
    new_image_size: Tuple[int, int],
    max_view_num: Optional[int] = None,
    sort_camera_index: bool = False,
    focal_multiplier: float = 0.95,
    mirror_right_hand: bool = True,
    min_required_vis_landmarks: int = 19,
) -> Dict[int, camera.PinholePlaneCameraModel]:
    """perspective_crop.py:136-180."""
    crop_points = _get_crop_points_from_hand_pose(hand_model, hand_pose, hand_idx, num_crop_points)
    cam_indices = rank_hand_visibility_in_cameras(
        cameras, hand_model, hand_pose, hand_idx, min_required_vis_landmarks, landmarks_world=crop_points[:21]
    )
    if sort_camera_index:
        cam_indices = sorted(cam_indices)
    crop_cameras: Dict[int, camera.PinholePlaneCameraModel] = {}
    for cam_idx in cam_indices:
        crop_cameras[cam_idx] = gen_crop_parameters_from_points(
            cameras[cam_idx],
            crop_points,
            new_image_size,
            mirror_img_x=(mirror_right_hand and hand_idx == 1),
            camera_angle=camera_angles[cam_idx],
            focal_multiplier=focal_multiplier,
        )
        if len(crop_cameras) == max_view_num:
            break
    return crop_cameras
