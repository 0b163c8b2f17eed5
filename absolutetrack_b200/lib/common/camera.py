"""Camera models (mirror of the reference's lib/common/camera.py API for the hot path).

Same class names, constructor arguments, attributes and methods as the reference
(`PinholePlaneCameraModel`, `Fisheye62CameraModel`, `read_camera_from_json`, window_to_eye /
eye_to_world / world_to_eye / eye_to_window / copy / crop / uv_to_window_matrix) so that
`HandTracker.gen_crop_cameras` and the entry scripts work unchanged.  These numpy methods are
host-side helpers for building crop cameras; the per-pixel warp runs in the CUDA resampler
(csrc/crop_resample.cu), which gets its parameters from `crop_kernel_params`.
"""
import json
from typing import NamedTuple

import numpy as np

from . import affine


class NoDistortion(NamedTuple):
    """Identity distortion (reference camera.py:98-107)."""

    def evaluate(self, p):
        return p

    def undistort(self, q):
        return q


class Fisheye62Distortion(NamedTuple):
    """6 radial + 2 tangential coefficients, field order of reference camera.py:111-118."""

    k1: float
    k2: float
    k3: float
    k4: float
    p1: float
    p2: float
    k5: float
    k6: float

    def evaluate(self, p):
        """Forward distortion (reference camera.py:122-143)."""
        k1, k2, k3, k4, p1, p2, k5, k6 = self
        p = np.asarray(p, dtype=np.float64)
        r2 = np.clip((p * p).sum(axis=-1, keepdims=True), -np.pi ** 2, np.pi ** 2)
        r4 = r2 * r2
        r6 = r2 * r4
        radial = 1 + k1 * r2 + k2 * r4 + k3 * r6 + k4 * (r4 * r4) + k5 * (r4 * r6) + k6 * (r6 * r6)
        uv = p * radial
        x, y = uv[..., 0], uv[..., 1]
        x2, y2, xy = x * x, y * y, x * y
        rr = x2 + y2
        return np.stack(
            (x + (2 * p2 * xy + p1 * (rr + 2 * x2)), y + (2 * p1 * xy + p2 * (rr + 2 * y2))), axis=-1
        )

    def undistort(self, q, solver_iters=50):
        """5 fixed-point iterations on the radial factor only (reference camera.py:146-181)."""
        k1, k2, k3, k4, _p1, _p2, k5, k6 = self
        q = np.asarray(q, dtype=np.float64)
        u = q.copy()
        for _ in range(5):
            r2 = (u * u).sum(axis=-1, keepdims=True)
            radial = 1 + k1 * r2 + k2 * r2 ** 2 + k3 * r2 ** 3 + k4 * r2 ** 4 + k5 * r2 ** 5 + k6 * r2 ** 6
            u = q / radial
        return u


class CameraModel:
    """Intrinsics + rigid camera_to_world_xf (reference camera.py:203-282)."""

    distortion_model = NoDistortion

    def __init__(self, width, height, f, c, distort_coeffs, camera_to_world_xf=None):
        self.width = width
        self.height = height
        self.f = tuple(np.broadcast_to(f, 2))
        self.c = tuple(c)
        self.camera_to_world_xf = np.eye(4) if camera_to_world_xf is None else camera_to_world_xf
        if hasattr(distort_coeffs, "evaluate"):
            self.distort = distort_coeffs
        else:
            self.distort = self.distortion_model(*distort_coeffs)

    def __repr__(self):
        return f"{type(self).__name__}({self.width}x{self.height}, f={self.f} c={self.c}"

    # --- projection (overridden) ---
    @staticmethod
    def project(v):
        raise NotImplementedError

    @staticmethod
    def unproject(p):
        raise NotImplementedError

    # --- frames ---
    def world_to_eye(self, v):
        xf = np.asarray(self.camera_to_world_xf)
        return affine.transform_vec3(xf.T, np.asarray(v) - xf[:3, 3])

    def eye_to_world(self, v):
        return affine.transform3(np.asarray(self.camera_to_world_xf), v)

    def eye_to_window(self, v):
        return self.distort.evaluate(self.project(v)) * self.f + self.c

    def eye_to_window_undistorted(self, v):
        return self.project(v) * self.f + self.c

    def window_to_eye(self, w):
        q = (np.asarray(w) - self.c) / self.f
        return self.unproject(self.distort.undistort(q))

    def copy(self, camera_to_world_xf=None):
        return self.crop(0, 0, self.width, self.height, camera_to_world_xf=camera_to_world_xf)

    def crop(self, src_x, src_y, target_width, target_height, scale=1, camera_to_world_xf=None):
        return type(self)(
            target_width,
            target_height,
            np.asarray(self.f) * scale,
            (np.array(self.c) - (src_x, src_y) + 0.5) * scale - 0.5,
            self.distort,
            self.camera_to_world_xf if camera_to_world_xf is None else camera_to_world_xf,
        )


class PinholePlaneCameraModel(CameraModel):
    distortion_model = NoDistortion

    @staticmethod
    def project(v):
        v = np.asarray(v)
        return v[..., :2] / v[..., 2, None]

    @staticmethod
    def unproject(p):
        p = np.asarray(p)
        v = np.concatenate([p, np.ones_like(p[..., :1])], axis=-1)
        return affine.normalized(v, axis=-1)

    def uv_to_window_matrix(self):
        return np.array([[self.f[0], 0, self.c[0]], [0, self.f[1], self.c[1]], [0, 0, 1]])


class Fisheye62CameraModel(CameraModel):
    distortion_model = Fisheye62Distortion

    @staticmethod
    def project(p, eps=2.0 ** -128):
        p = np.asarray(p)
        x, y, z = p[..., 0], p[..., 1], p[..., 2]
        r = np.sqrt(x * x + y * y)
        s = np.arctan2(r, z) / np.maximum(r, eps)
        return np.stack((x * s, y * s), axis=-1)

    @staticmethod
    def unproject(uv):
        uv = np.asarray(uv)
        u, v = uv[..., 0], uv[..., 1]
        r = np.sqrt(u * u + v * v)
        s = np.sinc(r / np.pi)
        return np.stack([u * s, v * s, np.cos(r)], axis=-1)


def read_camera_from_json(js):
    """Reference camera.py:421-444."""
    if isinstance(js, str):
        js = json.loads(js)
    js = js.get("Camera", js)
    model = js["DistortionModel"]
    if model == "PinholePlane":
        cls = PinholePlaneCameraModel
    elif model == "FishEye62":
        cls = Fisheye62CameraModel
    else:
        raise ValueError(f"unsupported DistortionModel {model!r}")
    coeffs = [js[name] for name in cls.distortion_model._fields]
    return cls(js["ImageSizeX"], js["ImageSizeY"], (js["fx"], js["fy"]), (js["cx"], js["cy"]), coeffs)


def crop_kernel_params(src_camera, dst_camera):
    """Parameters of one AtCropParams record (include/absolutetrack_b200.h) for warping
    `src_camera` (Fisheye62) into `dst_camera` (pinhole): M = R_src^T R_dst and
    d = R_src^T (t_dst - t_src) precomposed in fp64 (the per-pixel chain of
    tracker.py:68-76 collapses to src_eye = M dst_eye + d)."""
    if not isinstance(src_camera.distort, Fisheye62Distortion):
        raise TypeError("the CUDA resampler implements Fisheye62 sources only (reference eval configs)")
    cs = np.asarray(src_camera.camera_to_world_xf, dtype=np.float64)
    cd = np.asarray(dst_camera.camera_to_world_xf, dtype=np.float64)
    M = cs[:3, :3].T @ cd[:3, :3]
    d = cs[:3, :3].T @ (cd[:3, 3] - cs[:3, 3])
    return dict(
        M=M.reshape(9),
        d=d,
        f_dst=np.asarray(dst_camera.f, np.float64),
        c_dst=np.asarray(dst_camera.c, np.float64),
        f_src=np.asarray(src_camera.f, np.float64),
        c_src=np.asarray(src_camera.c, np.float64),
        k=np.asarray(tuple(src_camera.distort), np.float64),
        src_w=int(src_camera.width),
        src_h=int(src_camera.height),
    )
