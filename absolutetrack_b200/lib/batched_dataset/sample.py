"""One raw torch_data sequence (mirror of the reference's lib/batched_dataset/sample.py)."""
from dataclasses import dataclass
from typing import Any, Dict

import numpy as np
import torch

from ..common.hand import HandModel, scaled_hand_model


@dataclass
class RawSample:
    """Field names and order of sample.py:17-32: images [T,V,H,W], world->eye extrinsics [T,V,4,4],
    intrinsics [T,V,3,3], enclosing_points [T,P,3], hand [T], the user's and the generic skeleton
    with their poses, pinch labels.  Lengths in mm until `scaled(0.001)`."""

    images: np.ndarray
    extrinsics: np.ndarray
    intrinsics: np.ndarray
    enclosing_points: np.ndarray
    hand: np.ndarray
    hand_model: HandModel
    wrist: np.ndarray
    joint_angles: np.ndarray
    solved_wrist_xfs: np.ndarray
    solved_joint_angles: np.ndarray
    generic_hand_model: HandModel
    pinch: np.ndarray

    def scaled(self, factor: float) -> None:
        """In-place unit change of every length (sample.py:34-40)."""
        self.extrinsics[..., :3, 3] *= factor
        self.enclosing_points *= factor
        self.wrist[..., :3, 3] *= factor
        self.solved_wrist_xfs[..., :3, 3] *= factor
        self.hand_model = scaled_hand_model(self.hand_model, factor)
        self.generic_hand_model = scaled_hand_model(self.generic_hand_model, factor)


def parse_raw_buffers(mono: np.ndarray, labels: Dict[str, Any]) -> RawSample:
    """The decoded frames of one sequence -> RawSample (sample.py:43-53): hand models become
    HandModel tuples of tensors, every other label a float32 array."""
    typed: Dict[str, Any] = {}
    for name, value in labels.items():
        if "hand_model" in name:
            typed[name] = HandModel(**{k: torch.tensor(v) for k, v in value.items()})
        else:
            typed[name] = np.array(value, dtype=np.float32)
    return RawSample(images=mono, **typed)
