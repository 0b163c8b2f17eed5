"""One raw torch_data sequence: the decoded frames of a `.torch.bin` pair as typed arrays
(counterpart of the reference's lib/batched_dataset/sample.py, same names so that
`RawSample(**...)` / `parse_raw_buffers(**data)` call sites keep working)."""
from dataclasses import dataclass
from typing import Any, Dict

import numpy as np
import torch

from ..common.hand import HandModel, scaled_hand_model

# label arrays whose translation column / coordinates carry a length (mm in the files)
_XF_FIELDS = ("extrinsics", "wrist", "solved_wrist_xfs")
_MODEL_FIELDS = ("hand_model", "generic_hand_model")


@dataclass
class RawSample:
    """T frames, V views, P crop points (sample.py:17-32)."""

    images: np.ndarray               # uint8 [T,V,H,W] mono frames
    extrinsics: np.ndarray           # f32 [T,V,4,4] world -> eye of every view
    intrinsics: np.ndarray           # f32 [T,V,3,3]
    enclosing_points: np.ndarray     # f32 [T,P,3] world points the crop must contain
    hand: np.ndarray                 # f32 [T] 0 = left, 1 = right
    hand_model: HandModel            # the user's skeleton (ground truth branch)
    wrist: np.ndarray                # f32 [T,4,4]
    joint_angles: np.ndarray         # f32 [T,22]
    solved_wrist_xfs: np.ndarray     # f32 [T,4,4] pose solved on the generic skeleton
    solved_joint_angles: np.ndarray  # f32 [T,22]
    generic_hand_model: HandModel
    pinch: np.ndarray                # f32 [T,...] pinch labels

    def scaled(self, factor: float) -> None:
        """Unit change of every length, in place (sample.py:34-40; the loader calls it with 0.001)."""
        for name in _XF_FIELDS:
            getattr(self, name)[..., :3, 3] *= factor
        self.enclosing_points *= factor
        for name in _MODEL_FIELDS:
            setattr(self, name, scaled_hand_model(getattr(self, name), factor))


def parse_raw_buffers(mono: np.ndarray, labels: Dict[str, Any]) -> RawSample:
    """Frames + msgpack labels of one sequence -> RawSample (sample.py:43-53): skeleton dicts become
    HandModel tuples of tensors, every other label a float32 array."""
    def typed(name: str, value: Any):
        if "hand_model" in name:
            return HandModel(**{k: torch.tensor(v) for k, v in value.items()})
        return np.array(value, dtype=np.float32)

    return RawSample(images=mono, **{name: typed(name, value) for name, value in labels.items()})
