"""Label + video stream (mirror of the reference's lib/tracker/video_pose_data.py API).

The JSON label loader and the per-frame (InputFrame, gt_tracking) iterator keep the reference's
semantics (:62-150).  Video decode still goes through PyAV, imported lazily: it is outside the
GPU hot path and `av` is not installed in this image.
"""
import json
from dataclasses import dataclass
from typing import Iterator, List

import numpy as np
import torch

from ..common.camera import CameraModel, read_camera_from_json
from ..common.hand import HandModel
from .tracker import InputFrame, ViewData
from .tracking_result import SingleHandPose


@dataclass
class HandPoseLabels:
    cameras: List[CameraModel]
    camera_angles: List[float]
    camera_to_world_transforms: np.ndarray
    hand_model: HandModel
    joint_angles: np.ndarray
    wrist_transforms: np.ndarray
    hand_confidences: np.ndarray

    def __len__(self):
        return len(self.joint_angles)


class VideoStream:
    def __init__(self, data_path: str):
        self._data_path = data_path

    def _open(self):
        import av  # lazy: only needed when a real mp4 is decoded

        container = av.open(self._data_path)
        return container, container.streams.video[0]

    def __len__(self) -> int:
        return self._open()[1].frames

    def __iter__(self) -> Iterator[np.ndarray]:
        container, stream = self._open()
        for frame in container.decode(stream):
            yield np.array(frame.to_image())[..., 0]


def load_hand_model_from_dict(hand_model_dict) -> HandModel:
    return HandModel(**{k: (torch.Tensor(v) if isinstance(v, list) else v) for k, v in hand_model_dict.items()})


def _load_json(p: str):
    with open(p, "rb") as f:
        return json.load(f)


def _load_hand_pose_labels(p: str) -> HandPoseLabels:
    labels = _load_json(p)
    return HandPoseLabels(
        cameras=[read_camera_from_json(c) for c in labels["cameras"]],
        camera_angles=labels["camera_angles"],
        camera_to_world_transforms=np.array(labels["camera_to_world_transforms"]),
        hand_model=load_hand_model_from_dict(labels["hand_model"]),
        joint_angles=np.array(labels["joint_angles"]),
        wrist_transforms=np.array(labels["wrist_transforms"]),
        hand_confidences=np.array(labels["hand_confidences"]),
    )


def frame_from_labels(labels: HandPoseLabels, frame_idx: int, raw_mono: np.ndarray):
    """One (InputFrame, gt_tracking) pair from a packed [H, n_cams*W] mono frame (:112-150)."""
    gt_tracking = {}
    for hand_idx in range(2):
        conf = labels.hand_confidences[frame_idx, hand_idx]
        if conf > 0:
            gt_tracking[hand_idx] = SingleHandPose(
                joint_angles=labels.joint_angles[frame_idx, hand_idx],
                wrist_xform=labels.wrist_transforms[frame_idx, hand_idx],
                hand_confidence=conf,
            )
    n_cams = len(labels.cameras)
    images = raw_mono.reshape(raw_mono.shape[0], n_cams, -1)
    if labels.camera_to_world_transforms[frame_idx].sum() == 0:
        assert not gt_tracking, "Cameras are not tracked, expecting no ground truth tracking!"
    views = [
        ViewData(
            image=images[:, i, :],
            camera=labels.cameras[i].copy(camera_to_world_xf=labels.camera_to_world_transforms[frame_idx, i]),
            camera_angle=labels.camera_angles[i],
        )
        for i in range(n_cams)
    ]
    return InputFrame(views=views), gt_tracking


class SyncedImagePoseStream:
    def __init__(self, data_path: str):
        self._hand_pose_labels = _load_hand_pose_labels(data_path[:-4] + ".json")
        self._image_stream = VideoStream(data_path)
        assert len(self._hand_pose_labels) == len(self._image_stream)

    def __len__(self) -> int:
        return len(self._image_stream)

    def __iter__(self):
        for frame_idx, raw_mono in enumerate(self._image_stream):
            yield frame_from_labels(self._hand_pose_labels, frame_idx, raw_mono)
