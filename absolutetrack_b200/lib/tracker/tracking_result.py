"""Result tuples (mirror of the reference's lib/tracker/tracking_result.py:14-30)."""
from typing import Dict, NamedTuple

import numpy as np

from ..common.hand import NUM_JOINTS_PER_HAND


class SingleHandPose(NamedTuple):
    """joint angles (one per DoF) + wrist-to-world rigid transform."""

    joint_angles: np.ndarray = np.zeros(NUM_JOINTS_PER_HAND, dtype=np.float32)
    wrist_xform: np.ndarray = np.eye(4, dtype=np.float32)
    hand_confidence: float = 1.0


class TrackingResult(NamedTuple):
    """hand index -> pose / number of views used / predicted skeleton scale."""

    hand_poses: Dict[int, SingleHandPose] = {}
    num_views: Dict[int, int] = {}
    predicted_scales: Dict[int, float] = {}
