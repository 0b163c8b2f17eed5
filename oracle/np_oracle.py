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
# crop-camera generation (lib/tracker/perspective_crop.py:54-180, lib/common/crop.py:15-82,
# lib/common/affine.py:34-76, lib/tracker/tracker.py:222-260)
# --------------------------------------------------------------------------------------


def neutral_joint_angles(joint_limits, lower_factor=0.5):
    """perspective_crop.py:17-24: midpoint of the joint limits (float32, like the torch tensor)."""
    lim = np.asarray(joint_limits, np.float32)
    return lim[..., 0] * np.float32(lower_factor) + lim[..., 1] * np.float32(1 - lower_factor)


def crop_points_from_hand_pose(model, joint_angles, wrist_xform, hand_idx, num_crop_points):
    """perspective_crop.py:89-133: landmarks of the pose, the neutral pose and the open (zero)
    pose under the same wrist transform -> float32 [21 | 42 | 63, 3]."""
    assert num_crop_points in (21, 42, 63)
    sets = [np.asarray(joint_angles, np.float32)]
    if num_crop_points > 21:
        sets.append(neutral_joint_angles(model["joint_limits"]))
    if num_crop_points > 42:
        sets.append(np.zeros(22, np.float32))
    ja = np.stack(sets)
    xf = np.broadcast_to(np.asarray(wrist_xform, np.float64), (len(sets), 4, 4))
    return landmarks_from_hand_pose(model, ja, xf, hand_idx).reshape(-1, 3)


def rank_hand_visibility(cams, landmarks_world, min_required_vis_landmarks):
    """perspective_crop.py:54-86: cameras (dicts with c2w,f,c,coeffs,width,height) that see at
    least `min_required_vis_landmarks` landmarks, most first (stable sort)."""
    counts, keep = [], []
    for i, cam in enumerate(cams):
        eye = world_to_eye(cam["c2w"], np.asarray(landmarks_world))
        win = fisheye62_eye_to_window(eye, cam["f"], cam["c"], cam["coeffs"])
        n_vis = int(((win[..., 0] >= 0) & (win[..., 0] <= cam["width"] - 1) & (win[..., 1] >= 0)
                     & (win[..., 1] <= cam["height"] - 1) & (eye[..., 2] > 0)).sum())
        counts.append(n_vis)
        if n_vis >= min_required_vis_landmarks:
            keep.append(i)
    keep.sort(reverse=True, key=lambda i: counts[i])
    return keep, counts


def from_two_vectors(a_orig, b_orig):
    """affine.py:34-44."""
    a = a_orig / np.maximum(5.43e-20, np.sqrt((a_orig * a_orig).sum()))
    b = b_orig / np.maximum(5.43e-20, np.sqrt((b_orig * b_orig).sum()))
    v = np.cross(a, b)
    s = np.linalg.norm(v)
    c = np.dot(a, b)
    K = np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=v.dtype)
    return np.eye(3) + K + (K @ K) * (1 - c) / max(s * s, 1e-15)


def make_look_at_matrix(orig_world_to_eye, center, camera_angle=0.0):
    """affine.py:47-76 (scipy Rotation.from_euler('z', angle, degrees=True) written out)."""
    local = center @ orig_world_to_eye[:3, :3].T + orig_world_to_eye[:3, 3]
    z_dir = local / np.linalg.norm(local)
    dR = from_two_vectors(np.array([0, 0, 1], dtype=np.asarray(center).dtype), z_dir)
    e2w = np.linalg.inv(orig_world_to_eye)
    out = e2w.copy()
    out[:3, :3] = e2w[:3, :3] @ dR
    t = np.deg2rad(camera_angle)
    rz = np.array([[np.cos(t), -np.sin(t), 0.0], [np.sin(t), np.cos(t), 0.0], [0.0, 0.0, 1.0]])
    out[:3, :3] = out[:3, :3] @ rz
    return np.linalg.inv(out)


def gen_crop_parameters_from_points(c2w_orig, pts_world, new_image_size, mirror_img_x, camera_angle, focal_multiplier,
                                    min_focal=5):
    """crop.py:31-82 + :15-28 -> (f [2], c [2], camera_to_world [4,4]) of the pinhole crop camera."""
    pts_world = np.asarray(pts_world)
    w2e = np.linalg.inv(np.asarray(c2w_orig, np.float64))
    center = (pts_world.min(axis=0) + pts_world.max(axis=0)) / 2.0
    new_w2e = make_look_at_matrix(w2e, center, camera_angle)
    if mirror_img_x:
        flip = np.eye(4, dtype=np.float32)
        flip[0, 0] = -1
        new_w2e = flip @ new_w2e
    pts_eye = pts_world @ new_w2e[:3, :3].T + new_w2e[:3, 3]
    ndc = pts_eye[..., :2] / pts_eye[..., 2:]
    cx_cy = (np.array(new_image_size, dtype=pts_eye.dtype) - 1) / 2
    fx_fy = cx_cy / np.abs(ndc).max()
    if np.any(pts_eye[..., 2:] < 0.0001) or np.any(fx_fy < min_focal):
        raise ValueError("Unable to create crop camera", fx_fy)
    return focal_multiplier * fx_fy, cx_cy, np.linalg.inv(new_w2e)


def gen_crop_cameras_from_pose(cams, camera_angles, model, joint_angles, wrist_xform, hand_idx, num_crop_points=63,
                               new_image_size=(96, 96), max_view_num=2, sort_camera_index=True, focal_multiplier=0.8,
                               mirror_right_hand=True, min_required_vis_landmarks=19):
    """perspective_crop.py:136-180 -> ordered list of (cam_idx, f, c, camera_to_world)."""
    pts = crop_points_from_hand_pose(model, joint_angles, wrist_xform, hand_idx, num_crop_points)
    keep, _ = rank_hand_visibility(cams, pts[:21], min_required_vis_landmarks)
    if sort_camera_index:
        keep = sorted(keep)
    out = []
    for ci in keep:
        f, c, c2w = gen_crop_parameters_from_points(cams[ci]["c2w"], pts, new_image_size,
                                                    mirror_right_hand and hand_idx == 1, camera_angles[ci],
                                                    focal_multiplier)
        out.append((ci, f, c, c2w))
        if len(out) == max_view_num:
            break
    return out


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


# --------------------------------------------------------------------------------------
# evaluation table (load_eval.py:26-92, metric_utils.py:16-44,95-121)
# --------------------------------------------------------------------------------------
PCK_THRESHOLDS = np.linspace(0, 50, 101)  # metric_utils.py:16-17 (MAX_LANDMARK_ERROR_MM = 50)


def recording_metric_sums(gt, tracked, valid, thresholds=PCK_THRESHOLDS):
    """load_eval._compute_metrics (:26-44) reduced to what aggregate_metrics (:47-92) needs, as sums
    that add up across recordings: gt / tracked [H,T,21,3], valid [H,T] bool ->
    dict(sum_error, sum_accel, sum_gt_accel, n_valid, n_total, n_accel, pck_counts[len(thresholds)])."""
    gt = np.asarray(gt, np.float64)
    tracked = np.asarray(tracked, np.float64)
    valid = np.asarray(valid, bool)

    def accel(p):
        return np.linalg.norm(p[:, 0:-2] + p[:, 2:] - 2 * p[:, 1:-1], axis=-1).mean(axis=-1)

    err = np.linalg.norm(gt - tracked, axis=-1).mean(axis=-1)[valid]
    if valid.shape[1] >= 3:
        ok = valid[:, 0:-2] & valid[:, 1:-1] & valid[:, 2:]
        a, ga = accel(tracked)[ok], accel(gt)[ok]
    else:
        a = ga = np.zeros(0)
    return dict(
        sum_error=float(err.sum()), sum_accel=float(a.sum()), sum_gt_accel=float(ga.sum()),
        n_valid=int(valid.sum()), n_total=int(valid.size), n_accel=int(a.size),
        pck_counts=np.array([(err <= t).sum() for t in thresholds], np.int64),  # metric_utils._PCK_curve :34-44
    )


def normalized_auc(x, y, y_max=1.0):
    """metric_utils.normalized_AUC (:95-121): trapezoid area over (x[-1]-x[0]) * y_max."""
    x, y = np.asarray(x, np.float64), np.asarray(y, np.float64)
    return float(((x[1:] - x[:-1]) * (y[1:] + y[:-1]) * 0.5).sum() / ((x[-1] - x[0]) * y_max))


def eval_table(sums, thresholds=PCK_THRESHOLDS):
    """The numbers aggregate_metrics prints (:71-92) from summed `recording_metric_sums`; PCK is in
    percent and the AUC is normalised with y_max = 1 exactly as the reference calls it (:68-70)."""
    n = max(sums["n_valid"], 1)
    pck = np.where(sums["n_valid"] >= 1e-6, sums["pck_counts"] / n, 0.0) * 100.0  # _safe_div :20-31
    return dict(
        success_rate=sums["n_valid"] / sums["n_total"] * 100 if sums["n_total"] else 0.0,
        mean_error=sums["sum_error"] / n,
        pck_curve=pck,
        auc=normalized_auc(thresholds, pck),
        mean_accel=sums["sum_accel"] / max(sums["n_accel"], 1),
        mean_gt_accel=sums["sum_gt_accel"] / max(sums["n_accel"], 1),
    )


# --------------------------------------------------------------------------------------
# config-3 preprocessing (lib/batched_dataset/data_transform.py:57-277)
# --------------------------------------------------------------------------------------
def gen_crop_matrices(orig_extrinsics, orig_intrinsics, crop_points, mirror_image, crop_size, focal_multiplier=0.95):
    """data_transform._gen_crop_matrices (:145-209) + _compute_resample_matrix (:57-75): for every
    pinhole view (world->eye [V,4,4], K [V,3,3]) the crop camera that looks at `crop_points`
    -> (extrinsics_xf [V,4,4], new_intrinsics [V,3,3], resample_xf [V,4,4]) float32."""
    V = orig_extrinsics.shape[0]
    ext, K, xf = np.empty((V, 4, 4), np.float32), np.empty((V, 3, 3), np.float32), np.empty((V, 4, 4), np.float32)
    for v in range(V):
        c2w_orig = np.linalg.inv(orig_extrinsics[v])                      # :191 (float32 in, float32 out)
        f, c, c2w_new = gen_crop_parameters_from_points(c2w_orig, crop_points, crop_size, mirror_image, 0, focal_multiplier)
        ext[v] = np.linalg.inv(c2w_new)                                    # :203-204
        Kn = np.array([[f[0], 0, c[0]], [0, f[1], c[1]], [0, 0, 1.0]])
        K[v] = Kn
        Ko44, Kni44 = np.eye(4), np.eye(4)
        Ko44[:3, :3] = [[orig_intrinsics[v][0, 0], 0, orig_intrinsics[v][0, 2]],
                        [0, orig_intrinsics[v][1, 1], orig_intrinsics[v][1, 2]], [0, 0, 1]]
        Kni44[:3, :3] = np.linalg.inv(Kn)
        xf[v] = (Ko44 @ np.linalg.inv(c2w_orig) @ c2w_new @ Kni44).astype(np.float32)  # :64-73
    return ext, K, xf


def perspective_crop_images(orig_images, orig_extrinsics, orig_intrinsics, crop_points, hand_idx, crop_size):
    """data_transform._perspective_crop_images (:212-277): images [F,V,H,W] float32 grey levels ->
    (crops [F,V,h,w] in 0..1, extrinsics_xf, new_intrinsics)."""
    F, V = orig_images.shape[:2]
    mats = [gen_crop_matrices(orig_extrinsics[f], orig_intrinsics[f], crop_points[f], hand_idx == 1, crop_size) for f in range(F)]
    ext, K, xf = (np.stack([m[i] for m in mats]) for i in range(3))
    crops = resample_homography(orig_images.reshape((F * V,) + orig_images.shape[2:]), xf.reshape(F * V, 4, 4), crop_size[::-1])
    return (crops / 255).reshape(F, V, crop_size[1], crop_size[0]).astype(np.float32), ext, K
