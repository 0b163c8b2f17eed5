"""CPU ORACLE (numpy) - TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product path (absolutetrack_b200/) never does: it fails
loudly when the CUDA library is missing.

Restates, in plain numpy, the reference's per-frame geometry arithmetic.  Every function
cites the reference file:line (relative to /root/reference) whose algorithm it follows.
Pinned (tests/test_oracle_cpu.py, oracle/make_golden.py) against
  * the reference itself imported in the build container (oracle/ref_import.py),
  * the reference's own golden pickles sample_data/user05/recording_*.npy (FK),
  * the installed cv2.remap (third-party arithmetic the reference calls, tracker.py:87).
"""
import numpy as np

PI2 = np.pi ** 2

# --------------------------------------------------------------------------------------
# camera math (lib/common/camera.py, lib/common/affine.py)
# --------------------------------------------------------------------------------------


def pinhole_window_to_eye(w, f, c):
    """camera.py:320-329 (window_to_eye, NoDistortion) + :69-75 (unproject) +
    affine.py:22-24 (normalized, eps 5.43e-20)."""
    q = (np.asarray(w, dtype=np.float64) - np.asarray(c)) / np.asarray(f)
    v = np.concatenate([q, np.ones_like(q[..., :1])], axis=-1)
    d = np.maximum(5.43e-20, np.sqrt((v * v).sum(-1, keepdims=True)))
    return v / d


def eye_to_world(c2w, v):
    """camera.py:302-306 / affine.py:11-19: v @ R^T + t."""
    return v @ c2w[:3, :3].T + c2w[:3, 3]


def world_to_eye(c2w, p):
    """camera.py:296-300: R^T (p - t)   (transform_vec3 with camera_to_world.T)."""
    return (p - c2w[:3, 3]) @ c2w[:3, :3]


def arctan_project(v, eps=2.0 ** -128):
    """camera.py:80-85 ArctanProjection.project."""
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    r = np.sqrt(x * x + y * y)
    s = np.arctan2(r, z) / np.maximum(r, eps)
    return np.stack((x * s, y * s), axis=-1)


def fisheye62_distort(p, coeffs):
    """camera.py:122-143 Fisheye62CameraModel.evaluate; coeffs = (k1,k2,k3,k4,p1,p2,k5,k6)."""
    k1, k2, k3, k4, p1, p2, k5, k6 = coeffs
    r2 = (p * p).sum(-1, keepdims=True)
    r2 = np.clip(r2, -PI2, PI2)
    r4 = r2 * r2
    r6 = r2 * r4
    r8 = r4 * r4
    r10 = r4 * r6
    r12 = r6 * r6
    radial = 1 + k1 * r2 + k2 * r4 + k3 * r6 + k4 * r8 + k5 * r10 + k6 * r12
    uv = p * radial
    x, y = uv[..., 0].copy(), uv[..., 1].copy()
    x2, y2, xy = x * x, y * y, x * y
    rr = x2 + y2
    xo = x + (2 * p2 * xy + p1 * (rr + 2 * x2))
    yo = y + (2 * p1 * xy + p2 * (rr + 2 * y2))
    return np.stack((xo, yo), axis=-1)


def fisheye62_eye_to_window(v, f, c, coeffs):
    """camera.py:308-312 eye_to_window for the Fisheye62 camera."""
    return fisheye62_distort(arctan_project(v), coeffs) * np.asarray(f) + np.asarray(c)


def warp_map(src, dst, depth_check=True):
    """tracker.py:61-85: dst pixel grid -> float32 (map_x, map_y) in the source image.

    `src`/`dst` are dicts: f(2), c(2), c2w(4,4), width, height; src also has coeffs(8).
    """
    W, H = int(dst["width"]), int(dst["height"])
    px, py = np.meshgrid(np.arange(W), np.arange(H))
    win = np.column_stack((px.flatten(), py.flatten()))
    eye_d = pinhole_window_to_eye(win, dst["f"], dst["c"])
    world = eye_to_world(np.asarray(dst["c2w"], np.float64), eye_d)
    eye_s = world_to_eye(np.asarray(src["c2w"], np.float64), world)
    w = fisheye62_eye_to_window(eye_s, src["f"], src["c"], src["coeffs"])
    if depth_check:
        w[eye_s[:, 2] < 0] = -1
    w = w.astype(np.float32)
    return w[:, 0].reshape(H, W), w[:, 1].reshape(H, W)


# --------------------------------------------------------------------------------------
# cv2.remap(INTER_LINEAR, BORDER_CONSTANT 0) on uint8  (third-party: OpenCV, called at
# tracker.py:87; opencv-python unpinned in README.md:10, 4.13.0 in this image).
# Published algorithm (imgproc/imgwarp.cpp remap -> remapBilinear<FixedPtCast<int,uchar,15>>):
#   sx = cvRound(x*32), sy = cvRound(y*32)  (round-half-even, float32 product is exact)
#   ix = saturate_short(sx>>5), iy likewise; fx = sx&31, fy = sy&31
#   weights = (32-fy|fy)*(32-fx|fx)*32   (== rint(32768*wy*wx) exactly, sum 32768)
#   out = (sum_k w_k * tap_k + 16384) >> 15, taps outside the image read 0.
# --------------------------------------------------------------------------------------


def remap_bilinear_u8(img, map_x, map_y):
    img = np.ascontiguousarray(img)
    H, W = img.shape
    mx = np.asarray(map_x, np.float32)
    my = np.asarray(map_y, np.float32)
    with np.errstate(invalid="ignore", over="ignore"):
        sx = np.rint(mx.astype(np.float64) * 32.0)
        sy = np.rint(my.astype(np.float64) * 32.0)
    sx = np.clip(np.nan_to_num(sx, nan=-2.0 ** 31), -2.0 ** 31, 2.0 ** 31 - 1).astype(np.int64)
    sy = np.clip(np.nan_to_num(sy, nan=-2.0 ** 31), -2.0 ** 31, 2.0 ** 31 - 1).astype(np.int64)
    ix = np.clip(sx >> 5, -32768, 32767)
    iy = np.clip(sy >> 5, -32768, 32767)
    fx = sx & 31
    fy = sy & 31

    def tap(yy, xx):
        ok = (yy >= 0) & (yy < H) & (xx >= 0) & (xx < W)
        v = img[np.clip(yy, 0, H - 1), np.clip(xx, 0, W - 1)].astype(np.int64)
        return np.where(ok, v, 0)

    acc = (
        (32 - fy) * (32 - fx) * 32 * tap(iy, ix)
        + (32 - fy) * fx * 32 * tap(iy, ix + 1)
        + fy * (32 - fx) * 32 * tap(iy + 1, ix)
        + fy * fx * 32 * tap(iy + 1, ix + 1)
    )
    return ((acc + 16384) >> 15).astype(np.uint8)


def warp_image(src, dst, image):
    """tracker.py:61-89 _warp_image (INTER_LINEAR, depth_check=True)."""
    mx, my = warp_map(src, dst)
    return remap_bilinear_u8(image, mx, my)


# --------------------------------------------------------------------------------------
# FK + linear-blend skinning (lib/common/hand_skinning.py, lib/tracker/perspective_crop.py)
# float32 throughout, as the reference casts (perspective_crop.py:32-35).
# --------------------------------------------------------------------------------------


def so3_exp_map_f32(w, eps=1e-4):
    """pytorch3d.transforms.so3_exp_map (third-party, un-pinned `@stable`, README.md:10):
    theta = sqrt(clamp(|w|^2, eps)); R = I + sin(theta)/theta K + (1-cos(theta))/theta^2 K^2."""
    w = np.asarray(w, np.float32)
    n2 = (w * w).sum(-1)
    th = np.sqrt(np.maximum(n2, np.float32(eps))).astype(np.float32)
    inv = (np.float32(1.0) / th).astype(np.float32)
    f1 = inv * np.sin(th)
    f2 = inv * inv * (np.float32(1.0) - np.cos(th))
    z = np.zeros_like(w[..., 0])
    x, y, zz = w[..., 0], w[..., 1], w[..., 2]
    K = np.stack([z, -zz, y, zz, z, -x, -y, x, z], -1).reshape(w.shape[:-1] + (3, 3))
    K2 = K @ K
    return (f1[..., None, None] * K + f2[..., None, None] * K2 + np.eye(3, dtype=np.float32)).astype(np.float32)


def skin_landmarks(model, joint_angles, wrist_xf):
    """hand_skinning.py:189-209 skin_landmarks, for leading dims [N].

    model: dict with joint_rotation_axes [22,3] or [N,22,3], joint_rest_positions,
    landmark_rest_positions [21,3], landmark_rest_bone_indices [21,K], ..._weights [21,K]
    (a leading [N] is allowed on every field).  Returns float32 [N,21,3].
    """
    ja = np.asarray(joint_angles, np.float32).reshape(-1, 22)
    N = ja.shape[0]
    W = np.asarray(wrist_xf, np.float32).reshape(N, 4, 4)

    def b(name, tail):
        a = np.asarray(model[name])
        return np.broadcast_to(a, (N,) + tail) if a.ndim == len(tail) else a.reshape((N,) + tail)

    axes = b("joint_rotation_axes", (22, 3)).astype(np.float32)
    rest = b("joint_rest_positions", (22, 3)).astype(np.float32)
    lm = b("landmark_rest_positions", (21, 3)).astype(np.float32)
    K = np.asarray(model["landmark_rest_bone_indices"]).shape[-1]
    bi = b("landmark_rest_bone_indices", (21, K)).astype(np.int64)
    bw = b("landmark_rest_bone_weights", (21, K)).astype(np.float32)

    # _joint_local_transform (hand_skinning.py:35-53): T_j = [R_j, p_j - R_j p_j]
    R = so3_exp_map_f32(axes[:, :20] * ja[:, :20, None])
    T = np.zeros((N, 20, 4, 4), np.float32)
    T[..., 3, 3] = 1
    T[..., :3, :3] = R
    T[..., :3, 3] = rest[:, :20] - (R @ rest[:, :20, :, None])[..., 0]

    # _hand_skinning_transform (:100-127) + _finger_fk (:17-32): 17 frames
    frames = [W, W]
    for fng in range(5):
        cur = W
        for i in range(4):
            cur = (cur @ T[:, 4 * fng + i]).astype(np.float32)
            if i >= 1:
                frames.append(cur)
    F = np.stack(frames, 1)  # [N,17,4,4]

    # _get_skinning_weights (:70-97): scatter non-zero weights into [21,17]
    skin = np.zeros((N, 21, 17), np.float32)
    for k in range(K):
        nz = bw[..., k] != 0
        n_i, l_i = np.nonzero(nz)
        skin[n_i, l_i, bi[n_i, l_i, k]] = bw[n_i, l_i, k]

    # _get_skinned_vertices (:130-154) + _lbs (:56-67): sum_b F_b (w_lb [p_l,1])
    ph = np.concatenate([lm, np.ones((N, 21, 1), np.float32)], -1)
    wp = ph[:, :, None, :] * skin[..., None]  # [N,21,17,4]
    out = np.einsum("nbij,nlbj->nli", F, wp, dtype=np.float32)
    return out[..., :3].astype(np.float32)


def landmarks_from_hand_pose(model, joint_angles, wrist_xform, hand_idx):
    """perspective_crop.py:40-51: mirror column 0 for the right hand, then skin."""
    xf = np.array(wrist_xform, dtype=np.float64, copy=True).reshape(-1, 4, 4)
    hi = np.broadcast_to(np.asarray(hand_idx), (xf.shape[0],))
    xf[hi == 1, :, 0] *= -1
    return skin_landmarks(model, joint_angles, xf)


# --------------------------------------------------------------------------------------
# torch_data pinhole->pinhole resampler (lib/batched_dataset/data_transform.py:78-144)
# --------------------------------------------------------------------------------------


def resample_homography(images, xfs, out_hw):
    """data_transform.py:78-144 _resample_images_batched, restated with explicit loops over
    images (vectorised over pixels).  images [N,H,W] float32, xfs [N,4,4] float32."""
    images = np.asarray(images, np.float32)
    xfs = np.asarray(xfs, np.float32)
    N, H, W = images.shape
    oh, ow = out_hw
    v, u = np.meshgrid(np.arange(oh), np.arange(ow), indexing="ij")
    # NOTE numpy promotes float32 (matrix) x int32 (grid) to float64 in the reference's
    # tensordot (:99), so everything downstream of it is float64 until the final store.
    uv1 = np.stack([u, v, np.ones_like(u)], -1).reshape(-1, 3).astype(np.float64)
    out = np.zeros((N, oh, ow), np.float32)
    for n in range(N):
        p = uv1 @ xfs[n, :3, :3].astype(np.float64).T + xfs[n, :3, 3].astype(np.float32)
        x = p[:, 0] / p[:, 2]
        y = p[:, 1] / p[:, 2]
        m = (x >= 0) & (x < W - 1) & (y >= 0) & (y < H - 1)
        xs, ys = x[m], y[m]
        x0 = xs.astype(np.int32)
        y0 = ys.astype(np.int32)
        x1, y1 = x0 + 1, y0 + 1
        Ia, Ib = images[n, y0, x0], images[n, y1, x0]
        Ic, Id = images[n, y0, x1], images[n, y1, x1]
        val = (Ia * (x1 - xs) * (y1 - ys) + Ic * (xs - x0) * (y1 - ys)
               + Ib * (x1 - xs) * (ys - y0) + Id * (xs - x0) * (ys - y0)) / ((x1 - x0) * (y1 - y0))
        o = np.zeros(oh * ow, np.float32)
        o[m] = val
        out[n] = o.reshape(oh, ow)
    return out


# --------------------------------------------------------------------------------------
# metrics (run_eval_known_skeleton.py:92-93, load_eval.py:26-44)
# --------------------------------------------------------------------------------------


def mpjpe(tracked, gt, valid):
    """Mean keypoint error (mm) over valid hand-frames: run_eval_known_skeleton.py:92-93."""
    d = np.linalg.norm(tracked - gt, axis=-1).mean(-1)
    return float(d[valid].mean()) if valid.any() else 0.0


def mean_keypoint_accel(kp, valid):
    """load_eval.py:29-44: |p[t-1] + p[t+1] - 2 p[t]| averaged over keypoints and valid triples.
    kp [T,21,3], valid [T] bool."""
    ok = valid[:-2] & valid[1:-1] & valid[2:]
    a = np.linalg.norm(kp[:-2] + kp[2:] - 2 * kp[1:-1], axis=-1).mean(-1)
    return float(a[ok].mean()) if ok.any() else 0.0
