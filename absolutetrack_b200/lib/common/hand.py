"""Hand model container (mirror of the reference's lib/common/hand.py API)."""
from enum import Enum
from typing import Any, Dict, NamedTuple, Optional

import torch

NUM_HANDS = 2
NUM_LANDMARKS_PER_HAND = 21
NUM_FINGERTIPS_PER_HAND = 5
NUM_JOINTS_PER_HAND = 22
LEFT_HAND_INDEX = 0
RIGHT_HAND_INDEX = 1
NUM_DIGITS: int = 5
NUM_JOINT_FRAMES: int = 1 + 1 + 3 * 5  # root + wrist + 3 per finger
DOF_PER_FINGER: int = 4

LANDMARK = Enum(
    "LANDMARK",
    {
        "THUMB_FINGERTIP": "Thumb fingertip",
        "INDEX_FINGER_FINGERTIP": "Index finger fingertip",
        "MIDDLE_FINGER_FINGERTIP": "Middle finger fingertip",
        "RING_FINGER_FINGERTIP": "Ring finger fingertip",
        "PINKY_FINGER_FINGERTIP": "Pinky finger fingertip",
        "WRIST_JOINT": "Wrist joint",
        "THUMB_INTERMEDIATE_FRAME": "Thumb intermediate frame",
        "THUMB_DISTAL_FRAME": "Thumb distal frame",
        "INDEX_PROXIMAL_FRAME": "Index proximal frame",
        "INDEX_INTERMEDIATE_FRAME": "Index intermediate frame",
        "INDEX_DISTAL_FRAME": "Index distal frame",
        "MIDDLE_PROXIMAL_FRAME": "Middle proximal frame",
        "MIDDLE_INTERMEDIATE_FRAME": "Middle intermediate frame",
        "MIDDLE_DISTAL_FRAME": "Middle distal frame",
        "RING_PROXIMAL_FRAME": "Ring proximal frame",
        "RING_INTERMEDIATE_FRAME": "Ring intermediate frame",
        "RING_DISTAL_FRAME": "Ring distal frame",
        "PINKY_PROXIMAL_FRAME": "Pinky proximal frame",
        "PINKY_INTERMEDIATE_FRAME": "Pinky intermediate frame",
        "PINKY_DISTAL_FRAME": "Pinky distal frame",
        "PALM_CENTER": "Palm center",
    },
)


class HandModel(NamedTuple):
    """Field names and order are the reference's (hand.py:48-62): the label JSONs and
    generic_hand_model.json are loaded with HandModel(**dict)."""

    joint_rotation_axes: torch.Tensor
    joint_rest_positions: torch.Tensor
    joint_frame_index: torch.Tensor
    joint_parent: torch.Tensor
    joint_first_child: torch.Tensor
    joint_next_sibling: torch.Tensor
    landmark_rest_positions: torch.Tensor
    landmark_rest_bone_weights: torch.Tensor
    landmark_rest_bone_indices: torch.Tensor
    hand_scale: Optional[torch.Tensor]
    mesh_vertices: Optional[torch.Tensor] = None
    mesh_triangles: Optional[torch.Tensor] = None
    dense_bone_weights: Optional[torch.Tensor] = None
    joint_limits: Optional[torch.Tensor] = None

    @classmethod
    def from_json(cls, json_data: Dict[str, Any]) -> "HandModel":
        return cls(**{k: (torch.tensor(v) if v is not None else None) for k, v in json_data.items()})

    def to_json(self) -> Dict[str, Any]:
        return {k: (v.tolist() if isinstance(v, torch.Tensor) else v) for k, v in self._asdict().items()}


def scaled_hand_model(hand: HandModel, multiplier) -> HandModel:
    """Scale every length of the model (hand.py:78-111); `multiplier` is a float or a tensor
    shaped like the model's leading dims."""
    lead = hand.joint_rest_positions.shape[:-2]
    mult = torch.ones(lead, dtype=hand.joint_rest_positions.dtype, device=hand.joint_rest_positions.device) * multiplier
    m = mult[..., None, None]
    verts = hand.mesh_vertices
    return hand._replace(
        joint_rest_positions=hand.joint_rest_positions * m,
        landmark_rest_positions=hand.landmark_rest_positions * m,
        mesh_vertices=None if verts is None else verts * m,
    )


def mirrored_hand_model(hand: HandModel, to_mirror: torch.Tensor) -> HandModel:
    """Left <-> right model for the samples selected by `to_mirror` (hand.py:114-147):
    rotation-axis y,z and rest-position x change sign."""
    flat = to_mirror.reshape(-1)
    end = len(to_mirror.shape) - 1
    axes = hand.joint_rotation_axes.clone()
    rest = hand.joint_rest_positions.clone()
    lms = hand.landmark_rest_positions.clone()
    torch.flatten(axes, 0, end)[flat, ..., 1:] *= -1
    torch.flatten(rest, 0, end)[flat, ..., 0] *= -1
    torch.flatten(lms, 0, end)[flat, ..., 0] *= -1
    verts = hand.mesh_vertices
    if verts is not None:
        verts = verts.clone()  # the reference multiplies the mesh by +1 (hand.py:128), i.e. leaves it
    return hand._replace(
        joint_rotation_axes=axes, joint_rest_positions=rest, landmark_rest_positions=lms, mesh_vertices=verts
    )
