"""Deterministic test scenarios built from tests/golden/geometry.npz - TEST INFRASTRUCTURE.

Shared by oracle/make_golden_model.py (runs the reference on them), the parity tests and
bench.py's CPU baseline.  Uses only numpy + the committed geometry fixture, so it runs on the
GPU box where /root/reference does not exist.
"""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
_G = None


def geometry():
    global _G
    if _G is None:
        _G = dict(np.load(os.path.join(GOLD, "geometry.npz")))
    return _G


def synth_frame(kind, seed, n_views=4, H=480, W=636):
    """uint8 [H, n_views, W] multi-view frame (layout of video_pose_data.py:123-125)."""
    if kind == "noise":
        return np.random.default_rng(seed).integers(0, 256, (H, n_views, W), dtype=np.uint8)
    y, v, x = np.meshgrid(np.arange(H), np.arange(n_views), np.arange(W), indexing="ij")
    val = 127 + 100 * np.sin((x + 7 * seed + 31 * v) / 15.0) * np.cos((y + 3 * seed) / 11.0)
    return np.clip(np.rint(val), 0, 255).astype(np.uint8)


def synth_image(kind, seed, H=480, W=636):
    """Single-view image used by tests/golden/warp.npz (must match oracle/make_golden.py)."""
    if kind == "noise":
        return np.random.default_rng(seed).integers(0, 256, (H, W), dtype=np.uint8)
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    v = 127 + 100 * np.sin((x + 7 * seed) / 15.0) * np.cos((y + 3 * seed) / 11.0)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


def src_camera(k, cam):
    G = geometry()
    return dict(f=G["cam_f"][cam], c=G["cam_c"][cam], coeffs=tuple(G["cam_coeffs"][cam]),
                c2w=G["camera_to_world"][k, cam], width=int(G["cam_wh"][cam, 0]), height=int(G["cam_wh"][cam, 1]))


def dst_camera(k, h, s):
    G = geometry()
    return dict(f=G["crop_f"][k, h, s], c=G["crop_c"][k, h, s], c2w=G["crop_c2w"][k, h, s], width=96, height=96)


def hand_model(prefix="hm_"):
    G = geometry()
    keys = ["joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions",
            "landmark_rest_bone_weights", "landmark_rest_bone_indices", "joint_limits"]
    return {k: G[prefix + k] for k in keys if prefix + k in G}


def frame_inputs(k, drop_view_of_hand=None):
    """Model inputs of geometry frame k as tracker._make_inputs builds them (tracker.py:315-368),
    minus the images: list of (hand, slot, cam) crops, intrinsics [c,3,3], extrinsics [c,4,4] (m),
    sample_range [n,2], hand_idx [n].  `drop_view_of_hand=h` keeps only the first view of hand h
    (exercises the single-view branch, umetrack_model.py:153-157)."""
    G = geometry()
    crops, intr, extr, ranges, hands = [], [], [], [], []
    for h in range(2):
        start = len(crops)
        for s in range(2):
            cam = int(G["crop_cam"][k, h, s])
            if cam < 0 or (drop_view_of_hand == h and s == 1):
                continue
            crops.append((h, s, cam))
            f, c = G["crop_f"][k, h, s], G["crop_c"][k, h, s]
            intr.append(np.array([[f[0], 0, c[0]], [0, f[1], c[1]], [0, 0, 1]]))
            w2e = np.linalg.inv(G["crop_c2w"][k, h, s])
            w2e[:3, 3] *= 0.001
            extr.append(w2e)
        if len(crops) > start:
            hands.append(h)
            ranges.append((start, len(crops)))
    return dict(crops=crops, intrinsics=np.array(intr, np.float32).reshape(-1, 3, 3),
                extrinsics=np.array(extr, np.float32).reshape(-1, 4, 4),
                sample_range=np.array(ranges, np.int64).reshape(-1, 2), hand_idx=np.array(hands, np.int64))


def model_sequence(n_frames=3, start=0, kind="smooth", single_view_frame=None):
    """A short tracked sequence: per frame the oracle-warped crops [c,96,96] float32 plus the
    model inputs; use_memory follows HandTracker's valid-history rule (tracker.py:355-357)."""
    import np_oracle as O

    seq = []
    valid = np.zeros(2, bool)
    for t in range(n_frames):
        k = start + t
        fi = frame_inputs(k, drop_view_of_hand=0 if single_view_frame == t else None)
        frame = synth_frame(kind, 1000 + k)
        imgs = [O.warp_image(src_camera(k, cam), dst_camera(k, h, s), np.ascontiguousarray(frame[:, cam, :]))
                for (h, s, cam) in fi["crops"]]
        fi["frame"] = frame
        fi["images"] = (np.stack(imgs).astype(np.float32) / 255.0) if imgs else np.zeros((0, 96, 96), np.float32)
        fi["memory_idx"] = fi["hand_idx"].copy()
        fi["use_memory"] = valid[fi["hand_idx"]].copy()
        valid[:] = False
        valid[fi["hand_idx"]] = True
        seq.append(fi)
    return seq
