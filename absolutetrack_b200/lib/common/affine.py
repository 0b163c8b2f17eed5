"""Rigid-transform helpers (mirror of the reference's lib/common/affine.py API).

Host-side fp64 numpy; used only to BUILD crop cameras and kernel parameters, never per pixel.
"""
import numpy as np


def transform_vec3(m, v):
    """Rotate vectors by the 3x3 block of `m` (reference affine.py:15-19)."""
    m = np.asarray(m)
    v = np.asarray(v)
    if m.ndim == 2:
        return (v.reshape(-1, 3) @ m[:3, :3].T).reshape(v.shape)
    return np.einsum("...ij,...j->...i", m[..., :3, :3], v)


def transform3(m, v):
    """Apply the rigid transform `m` to points (reference affine.py:11-12)."""
    return transform_vec3(m, v) + np.asarray(m)[..., :3, 3]


def normalized(v, axis=-1, eps=5.43e-20):
    """Unit vectors with the reference's norm floor (affine.py:22-24)."""
    v = np.asarray(v)
    return v / np.maximum(eps, np.sqrt((v * v).sum(axis=axis, keepdims=True)))


def skew_matrix(v):
    x, y, z = v
    return np.array([[0, -z, y], [z, 0, -x], [-y, x, 0]], dtype=np.asarray(v).dtype)


def from_two_vectors(a_orig, b_orig):
    """Smallest rotation taking direction a to direction b (affine.py:34-44)."""
    a = normalized(a_orig)
    b = normalized(b_orig)
    axis = np.cross(a, b)
    s2 = float(axis @ axis)
    c = float(a @ b)
    K = skew_matrix(axis)
    return np.eye(3) + K + (K @ K) * ((1 - c) / max(s2, 1e-15))


def rotation_z_deg(angle_deg):
    """scipy Rotation.from_euler('z', angle, degrees=True).as_matrix() (affine.py:73)."""
    t = np.deg2rad(angle_deg)
    c, s = np.cos(t), np.sin(t)
    return np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])


def make_look_at_matrix(orig_world_to_eye, center, camera_angle=0.0):
    """World->eye transform of a camera at the same position looking at `center`, rolled by
    `camera_angle` degrees (affine.py:47-76)."""
    orig_world_to_eye = np.asarray(orig_world_to_eye, dtype=np.float64)
    local = transform3(orig_world_to_eye, center)
    z_dir = local / np.linalg.norm(local)
    dR = from_two_vectors(np.array([0.0, 0.0, 1.0], dtype=np.asarray(center).dtype), z_dir)
    eye_to_world = np.linalg.inv(orig_world_to_eye)
    out = eye_to_world.copy()
    out[:3, :3] = eye_to_world[:3, :3] @ dR
    out[:3, :3] = out[:3, :3] @ rotation_z_deg(camera_angle)
    return np.linalg.inv(out)
