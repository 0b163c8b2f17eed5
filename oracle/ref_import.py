"""TEST INFRASTRUCTURE ONLY - never imported by the product path.

Imports the UNMODIFIED reference (read-only, /root/reference) in THIS container so
that (a) the oracle restatements can be pinned against it and (b) golden fixtures
under tests/golden/ can be generated (oracle/make_golden.py).  /root/reference does
not exist on the GPU box; everything that runs there uses the committed fixtures.

Two import shims, no source edits:
  * pytorch3d.transforms.so3_exp_map  (hand_skinning.py:12,46)  - pytorch3d is absent;
    restated from its published algorithm (Rodrigues with eps=1e-4 clamp on |w|^2).
  * av (video_pose_data.py:12) - only used by VideoStream; no mp4 exists.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("ABSTRACK_REFERENCE", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "lib", "tracker"))


def _so3_exp_map(log_rot, eps: float = 1e-4):
    import torch

    nrms = (log_rot * log_rot).sum(1)
    rot_angles = torch.clamp(nrms, eps).sqrt()
    rot_angles_inv = 1.0 / rot_angles
    fac1 = rot_angles_inv * rot_angles.sin()
    fac2 = rot_angles_inv * rot_angles_inv * (1.0 - rot_angles.cos())
    z = torch.zeros_like(log_rot[:, 0])
    x, y, w = log_rot[:, 0], log_rot[:, 1], log_rot[:, 2]
    skews = torch.stack([z, -w, y, w, z, -x, -y, x, z], dim=1).reshape(-1, 3, 3)
    skews_square = torch.bmm(skews, skews)
    return (
        fac1[:, None, None] * skews
        + fac2[:, None, None] * skews_square
        + torch.eye(3, dtype=log_rot.dtype, device=log_rot.device)[None]
    )


def import_reference():
    """Put the reference's `lib` package on sys.path (as `lib`) and return it."""
    if not reference_available():
        raise RuntimeError("reference tree not present at " + REFERENCE_ROOT)
    if "pytorch3d" not in sys.modules:
        p3d = types.ModuleType("pytorch3d")
        tr = types.ModuleType("pytorch3d.transforms")
        tr.so3_exp_map = _so3_exp_map
        p3d.transforms = tr
        sys.modules["pytorch3d"] = p3d
        sys.modules["pytorch3d.transforms"] = tr
    if "av" not in sys.modules:
        sys.modules["av"] = types.ModuleType("av")
    # the product ships its own `lib` mirror; make sure the reference one wins here
    for k in [k for k in sys.modules if k == "lib" or k.startswith("lib.")]:
        mod = sys.modules[k]
        f = getattr(mod, "__file__", "") or ""
        if not f.startswith(REFERENCE_ROOT):
            del sys.modules[k]
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import lib  # noqa: F401

    return lib
