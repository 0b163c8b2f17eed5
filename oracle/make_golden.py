"""Generate tests/golden/*.npz by running the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE.  Run as `python oracle/make_golden.py [fk] [geometry] [warp] [cropcams] [cropmats] [torchdata] [eval] [model]`
(no argument = all).  /root/reference is read-only and absent on the GPU box, so the
outputs are committed; this script is the provenance record for every fixture.

Fixtures
  fk_user05.npz   poses + hand model of sample_data/user05/recording_{00,02,11}.json and the
                  reference's own golden `gt_keypoints` from the sibling .npy pickles
                  (written by run_eval_known_skeleton.py:96-104) -> FK known-answer test.
  geometry.npz    real calibration / per-frame camera poses / GT hand poses from
                  sample_data/recording_00.json (every 3rd frame) + the crop cameras the
                  reference HandTracker.gen_crop_cameras (tracker.py:222-260) produced.
  warp.npz        reference tracker._warp_image (tracker.py:61-89, numpy fp64 + cv2.remap)
                  outputs for crops of that geometry on seeded synthetic images.
  crop_cameras.npz  reference perspective_crop.gen_crop_cameras_from_pose (:136-180) on frames of
                  all four sample recordings, default HandTracker options plus variants (rank
                  by visibility instead of camera index, 21/42 crop points, thresholds, up to
                  4 views, poses pushed towards the image borders) -> selected cameras + crop
                  camera intrinsics / extrinsics.
  eval_user05.npz the reference's own tracking results (tracked_keypoints of the same three user05
                  pickles) and what the reference's load_eval._compute_metrics (:26-44) +
                  metric_utils.PCK_curve / normalized_AUC (:34-44,95-121) make of them per recording
                  and combined (aggregate_metrics :47-92): the README-style table.
  crop_matrices.npz  reference data_transform._gen_crop_matrices (:145-209) on 40 seeded two-view
                  frames (pinhole views, left and mirrored right hands) and
                  _perspective_crop_images (:212-277) on 3 of them per hand (seeded images).
  torchdata/      two synthetic config-3 sequences written as `.torch.idx/.torch.bin` pairs (frames:
                  uint8 tensors, labels: msgpack) with absolutetrack_b200's fixture writer, and
  torchdata.npz   what the REFERENCE makes of those files: TorchIdx (idxbinfile.py:48-347) decoding
                  and data_transform.preprocess (:388-397) -> ModelInput / ModelTarget.
  model.npz       reference UmeTrackModel outputs (umetrack_model.py:188-242) for the seeded
                  state dict of absolutetrack_b200 `make_random_state_dict(0)`.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, HERE)

import ref_import  # noqa: E402

ref_import.import_reference()
REF = ref_import.REFERENCE_ROOT

from lib.tracker import tracker as RT  # noqa: E402
from lib.tracker.perspective_crop import landmarks_from_hand_pose  # noqa: E402
from lib.tracker.tracking_result import SingleHandPose  # noqa: E402
from lib.tracker.video_pose_data import _load_hand_pose_labels  # noqa: E402

HM_KEYS = [
    "joint_rotation_axes",
    "joint_rest_positions",
    "landmark_rest_positions",
    "landmark_rest_bone_weights",
    "landmark_rest_bone_indices",
    "joint_limits",
]


def hand_model_arrays(hm, prefix=""):
    out = {}
    for k in HM_KEYS:
        v = getattr(hm, k)
        if v is not None:
            out[prefix + k] = v.numpy()
    return out


def synth_image(kind, seed, H=480, W=636):
    """Synthetic source view: 'noise' = iid uniform u8, 'smooth' = band-limited."""
    if kind == "noise":
        return np.random.default_rng(seed).integers(0, 256, (H, W), dtype=np.uint8)
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    v = 127 + 100 * np.sin((x + 7 * seed) / 15.0) * np.cos((y + 3 * seed) / 11.0)
    return np.clip(np.rint(v), 0, 255).astype(np.uint8)


class _SizeOnlyModel:
    def getInputImageSizes(self):
        return (96, 96)

    def to(self, d):
        return self


def make_fk():
    out = {}
    for rec in ["00", "02", "11"]:
        L = _load_hand_pose_labels(f"{REF}/sample_data/user05/recording_{rec}.json")
        g = np.load(f"{REF}/sample_data/user05/recording_{rec}.npy", allow_pickle=True)
        g = g if isinstance(g, dict) else g.item()
        valid = g["valid_tracking"]  # [2,T]
        # sanity: the reference FK run now reproduces its own stored golden
        for h in range(2):
            for t in np.nonzero(valid[h])[0][::50]:
                lm = landmarks_from_hand_pose(
                    L.hand_model, SingleHandPose(L.joint_angles[t, h], L.wrist_transforms[t, h], 1.0), h
                )
                assert np.abs(lm - g["gt_keypoints"][h, t]).max() < 1e-3
        p = f"r{rec}_"
        out[p + "joint_angles"] = L.joint_angles.astype(np.float32)  # [T,2,22]
        out[p + "wrist_transforms"] = L.wrist_transforms.astype(np.float64)  # [T,2,4,4]
        out[p + "valid"] = valid
        out[p + "gt_keypoints"] = g["gt_keypoints"].astype(np.float32)  # [2,T,21,3]
        out.update(hand_model_arrays(L.hand_model, p))
    np.savez_compressed(os.path.join(GOLD, "fk_user05.npz"), **out)
    print("fk_user05.npz", sum(int(out[k].sum()) for k in out if k.endswith("valid")), "poses")


def make_eval():
    """Known-answer fixture of the evaluation table: the reference's own aggregation run on its own
    stored tracking results."""
    import load_eval as RL
    from lib.common import metric_utils as RM

    out, errs, accs, gaccs, valids = {}, [], [], [], []
    for rec in ["00", "02", "11"]:
        g = np.load(f"{REF}/sample_data/user05/recording_{rec}.npy", allow_pickle=True)
        g = g if isinstance(g, dict) else g.item()
        m = RL._compute_metrics(g["gt_keypoints"], g["tracked_keypoints"], g["valid_tracking"])
        p = f"r{rec}_"
        tr32 = g["tracked_keypoints"].astype(np.float32)
        assert np.array_equal(tr32.astype(np.float64), g["tracked_keypoints"])  # float32 values in a float64 array
        assert np.array_equal(g["gt_keypoints"].astype(np.float32).astype(np.float64), g["gt_keypoints"])
        out[p + "tracked_keypoints"] = tr32  # gt_keypoints / valid: fk_user05.npz
        out[p + "sum_error"] = np.float64(m.keypoint_errors.sum())
        out[p + "sum_accel"] = np.float64(m.keypoint_accelerations.sum())
        out[p + "sum_gt_accel"] = np.float64(m.gt_keypoint_accelerations.sum())
        out[p + "n_accel"] = np.int64(m.keypoint_accelerations.size)
        out[p + "pck_counts"] = np.array([(m.keypoint_errors <= t).sum() for t in RM.PCK_THRESHOLDS], np.int64)
        errs.append(m.keypoint_errors); accs.append(m.keypoint_accelerations); gaccs.append(m.gt_keypoint_accelerations)
        valids.append(g["valid_tracking"])
    e = np.concatenate(errs)
    pck = RM.PCK_curve(e, RM.PCK_THRESHOLDS) * 100.0
    v = np.concatenate(valids, axis=1)
    out["thresholds"] = RM.PCK_THRESHOLDS
    out["pck_curve"] = pck
    out["auc"] = np.float64(RM.normalized_AUC(RM.PCK_THRESHOLDS, pck))
    out["n_valid"], out["n_total"] = np.int64(v.sum()), np.int64(v.size)
    out["mean_error"] = np.float64(e.mean())
    out["mean_accel"] = np.float64(np.concatenate(accs).mean())
    out["mean_gt_accel"] = np.float64(np.concatenate(gaccs).mean())
    np.savez_compressed(os.path.join(GOLD, "eval_user05.npz"), **out)
    print("eval_user05.npz", int(out["n_valid"]), "of", int(out["n_total"]), "mean error", float(out["mean_error"]),
          "auc", float(out["auc"]), "accel", float(out["mean_accel"]), float(out["mean_gt_accel"]))


def cropmat_inputs(seed=11, n_frames=40, n_views=2, n_pts=21):
    """Seeded config-3 style inputs in metres: world->eye of pinhole views around the origin, 3x3
    intrinsics of 160x120 images, crop points ~0.35 m in front of the cameras."""
    from scipy.spatial.transform import Rotation

    rng = np.random.default_rng(seed)
    ext = np.tile(np.eye(4, dtype=np.float32), (n_frames, n_views, 1, 1))
    ext[..., :3, :3] = Rotation.from_rotvec(rng.normal(0, 0.25, (n_frames * n_views, 3))).as_matrix().reshape(n_frames, n_views, 3, 3)
    ext[..., :3, 3] = rng.normal(0, 0.06, (n_frames, n_views, 3))
    intr = np.tile(np.eye(3, dtype=np.float32), (n_frames, n_views, 1, 1))
    intr[..., 0, 0] = rng.uniform(110, 190, (n_frames, n_views))
    intr[..., 1, 1] = intr[..., 0, 0] * rng.uniform(0.97, 1.03, (n_frames, n_views))
    intr[..., 0, 2] = 79.5 + rng.normal(0, 3, (n_frames, n_views))
    intr[..., 1, 2] = 59.5 + rng.normal(0, 3, (n_frames, n_views))
    ctr = np.array([0.0, 0.0, 0.35]) + rng.normal(0, 0.04, (n_frames, 1, 3))
    pts = (ctr + rng.normal(0, 0.035, (n_frames, n_pts, 3))).astype(np.float32)
    return ext, intr, pts


def cropmat_images(seed, n, H=120, W=160):
    out = np.empty((n, H, W), np.float32)
    for i in range(n):
        out[i] = synth_image("smooth" if i % 2 == 0 else "noise", seed + i, H, W)
    return out


def make_cropmats():
    from lib.batched_dataset import data_transform as DT

    ext, intr, pts = cropmat_inputs()
    F, V = ext.shape[:2]
    out = {"orig_extrinsics": ext, "orig_intrinsics": intr, "crop_points": pts}
    for hand in (0, 1):
        res = [DT._gen_crop_matrices(ext[f], intr[f], pts[f], hand == 1, (96, 96)) for f in range(F)]
        out[f"h{hand}_extrinsics_xf"] = np.stack([r[0] for r in res])
        out[f"h{hand}_new_intrinsics"] = np.stack([r[1] for r in res])
        out[f"h{hand}_resample_xf"] = np.stack([r[2] for r in res])
        imgs = cropmat_images(100 * hand, 3 * V).reshape(3, V, 120, 160)
        crops, e2, i2 = DT._perspective_crop_images(imgs, ext[:3], intr[:3], pts[:3], hand, (96, 96))
        assert np.array_equal(e2, out[f"h{hand}_extrinsics_xf"][:3]) and np.array_equal(i2, out[f"h{hand}_new_intrinsics"][:3])
        out[f"h{hand}_crops"] = crops.astype(np.float32)
        assert crops.dtype == np.float32 and crops.max() <= 1.0 and crops.max() > 0.5
    # a point behind the new camera: the reference raises (crop.py:25-26)
    bad = pts[0].copy()
    bad[3] = -bad[3]
    try:
        DT._gen_crop_matrices(ext[0], intr[0], bad, False, (96, 96))
        raise AssertionError("expected ValueError")
    except ValueError:
        pass
    np.savez_compressed(os.path.join(GOLD, "crop_matrices.npz"), **out)
    print("crop_matrices.npz", F, "frames x", V, "views x 2 hands; focal range",
          float(out["h0_new_intrinsics"][..., 0, 0].min()), float(out["h0_new_intrinsics"][..., 0, 0].max()))


def make_torchdata():
    """Fixture files written here, decoded and preprocessed by the unmodified reference."""
    import json

    from lib.batched_dataset import data_transform as DT
    from lib.data_utils.idxbinfile import TorchIdx as RefIdx

    sys.path.insert(0, ROOT)
    from absolutetrack_b200.lib.data_utils.idxbinfile import write_torch_idx_bin

    L = _load_hand_pose_labels(f"{REF}/sample_data/recording_00.json")
    big = ("mesh_vertices", "mesh_triangles", "dense_bone_weights")  # left out: keeps the fixture small
    user = {k: v.tolist() for k, v in L.hand_model._asdict().items() if isinstance(v, torch.Tensor) and k not in big}
    generic = {k: v for k, v in json.load(open(f"{REF}/dataset/generic_hand_model.json")).items()
               if v is not None and k not in big}
    for d in (user, generic):
        d.setdefault("hand_scale", 1.0)
    ext, intr, pts = cropmat_inputs(seed=23, n_frames=6)
    rng = np.random.default_rng(23)
    T, V = 3, 2
    monos, labels = [], []
    for si, hand in enumerate((0, 1)):
        fr = slice(3 * si, 3 * si + 3)
        e_mm = ext[fr].copy()
        e_mm[..., :3, 3] *= 1000.0
        wrist = np.tile(np.eye(4, dtype=np.float32), (T, 1, 1))
        wrist[:, :3, 3] = rng.normal(0, 200, (T, 3))
        monos.append(cropmat_images(300 + 10 * si, T * V).reshape(T, V, 120, 160).astype(np.uint8))
        labels.append({
            "extrinsics": e_mm.tolist(), "intrinsics": intr[fr].tolist(), "enclosing_points": (pts[fr] * 1000.0).tolist(),
            "hand": [float(hand)] * T, "hand_model": user, "wrist": wrist.tolist(),
            "joint_angles": rng.uniform(-0.5, 0.5, (T, 22)).astype(np.float32).tolist(),
            "solved_wrist_xfs": (wrist + np.float32(0.5)).tolist(),
            "solved_joint_angles": rng.uniform(-0.5, 0.5, (T, 22)).astype(np.float32).tolist(),
            "generic_hand_model": generic, "pinch": rng.uniform(0, 1, (T, 4)).astype(np.float32).tolist(),
        })
    d = os.path.join(GOLD, "torchdata")
    write_torch_idx_bin(os.path.join(d, "seq.mono.torch.idx"), monos)
    write_torch_idx_bin(os.path.join(d, "seq.labels.torch.idx"), labels)
    # ---- the reference reads them back
    out = {}
    im, il = RefIdx(os.path.join(d, "seq.mono.torch.idx")), RefIdx(os.path.join(d, "seq.labels.torch.idx"))
    out["mono_shape"] = np.array(im.shape, np.int64)
    out["mono_is_uniform"], out["labels_is_uniform"] = np.array(im.is_uniform), np.array(il.is_uniform)
    out["labels_byte_offsets"] = np.asarray(il.byte_offsets(0, len(il) + 1), np.int64)
    out["mono_frame_bytes"] = np.int64(im.frame_size_bytes())
    bm, bl = open(im.bin_path, "rb").read(), open(il.bin_path, "rb").read()
    for i in range(len(im)):
        mono = np.array(im.view_buffer(bm)[i])  # (view_buffer_at on a uniform file trips over numpy-2 promotion in the reference)
        lab = il.view_buffer_at(i, bl)
        assert np.array_equal(mono, monos[i]) and lab["hand"] == labels[i]["hand"]
        out[f"s{i}_mono_sum"] = np.int64(mono.astype(np.int64).sum())
        mi, mt = DT.preprocess({"mono": mono, "labels": lab}, (96, 96))
        p = f"s{i}_"
        out[p + "left_images"] = mi.left_images.numpy()
        out[p + "intrinsics"], out[p + "extrinsics_xf"] = mi.intrinsics.numpy(), mi.extrinsics_xf.numpy()
        out[p + "hand_idx"] = mi.hand_idx.numpy()
        for name, pd in (("orig", mi.orig_pose_data), ("solved", mi.s_solved_pose_data)):
            out[p + name + "_joint_angles"], out[p + name + "_wrist_xfs"] = pd.joint_angles.numpy(), pd.wrist_xfs.numpy()
            out[p + name + "_axes"] = pd.left_hand_model.joint_rotation_axes.numpy()
            out[p + name + "_rest"] = pd.left_hand_model.joint_rest_positions.numpy()
            out[p + name + "_landmarks"] = pd.left_hand_model.landmark_rest_positions.numpy()
        out[p + "gt_skel_scales"] = np.asarray(mt.gt_skel_targets.skel_scales)
        out[p + "pinch"] = mt.preds_targets.pinch_prediction.numpy()
        assert np.array_equal(mt.intrinsics.numpy(), out[p + "intrinsics"])
    np.savez_compressed(os.path.join(GOLD, "torchdata.npz"), **out)
    print("torchdata", {k: os.path.getsize(os.path.join(d, k)) for k in sorted(os.listdir(d))}, out["s0_left_images"].shape)


def _frames():
    L = _load_hand_pose_labels(f"{REF}/sample_data/recording_00.json")
    idx = np.arange(0, len(L), 3)
    return L, idx


def make_geometry():
    L, idx = _frames()
    ht = RT.HandTracker(_SizeOnlyModel(), RT.HandTrackerOpts())
    out = {
        "frame_index": idx,
        "cam_f": np.array([c.f for c in L.cameras], np.float64),
        "cam_c": np.array([c.c for c in L.cameras], np.float64),
        "cam_coeffs": np.array([tuple(c.distort) for c in L.cameras], np.float64),
        "cam_wh": np.array([(c.width, c.height) for c in L.cameras], np.int64),
        "camera_angles": np.array(L.camera_angles, np.float64),
        "camera_to_world": L.camera_to_world_transforms[idx].astype(np.float64),
        "joint_angles": L.joint_angles[idx].astype(np.float64),
        "wrist_transforms": L.wrist_transforms[idx].astype(np.float64),
        "hand_confidences": L.hand_confidences[idx].astype(np.float64),
    }
    out.update(hand_model_arrays(L.hand_model, "hm_"))
    gen = _load_generic()
    out.update(hand_model_arrays(gen, "generic_"))
    # crop cameras from the reference: [F,2 hands,2 slots]: cam index (-1 = none), f, c, c2w
    F = len(idx)
    cc_cam = -np.ones((F, 2, 2), np.int64)
    cc_f = np.zeros((F, 2, 2, 2))
    cc_c = np.zeros((F, 2, 2, 2))
    cc_c2w = np.zeros((F, 2, 2, 4, 4))
    for k, fi in enumerate(idx):
        cams = [c.copy(camera_to_world_xf=L.camera_to_world_transforms[fi, i]) for i, c in enumerate(L.cameras)]
        gt = {
            h: SingleHandPose(L.joint_angles[fi, h], L.wrist_transforms[fi, h], L.hand_confidences[fi, h])
            for h in range(2)
            if L.hand_confidences[fi, h] > 0
        }
        cc = ht.gen_crop_cameras(cams, L.camera_angles, L.hand_model, gt, 1)
        for h, d in cc.items():
            for s, (ci, cam) in enumerate(d.items()):
                cc_cam[k, h, s] = ci
                cc_f[k, h, s] = cam.f
                cc_c[k, h, s] = cam.c
                cc_c2w[k, h, s] = cam.camera_to_world_xf
    out.update(crop_cam=cc_cam, crop_f=cc_f, crop_c=cc_c, crop_c2w=cc_c2w)
    np.savez_compressed(os.path.join(GOLD, "geometry.npz"), **out)
    print("geometry.npz", F, "frames", int((cc_cam >= 0).sum()), "crops")


CROPCAM_VARIANTS = [
    # (num_crop_points, min_vis, max_views, sort_camera_index, focal_multiplier, lateral shift mm)
    (63, 19, 2, True, 0.8, 0.0),       # HandTrackerOpts defaults (tracker.py:49-58,222-260)
    (63, 19, 4, False, 0.95, 0.0),     # gen_crop_cameras_from_pose defaults: ranked by visibility
    (21, 12, 3, False, 0.8, 120.0),    # hands pushed sideways: fewer cameras see them
    (42, 19, 2, True, 0.9, -150.0),
    (63, 5, 4, False, 0.8, 250.0),
]


def make_cropcams():
    from lib.tracker.perspective_crop import gen_crop_cameras_from_pose

    recs = ["sample_data/recording_00.json"] + [f"sample_data/user05/recording_{r}.json" for r in ("00", "02", "11")]
    rows = {k: [] for k in ("rec", "frame", "hand", "variant", "joint_angles", "wrist", "c2w", "n_sel", "sel_cam",
                            "sel_f", "sel_c", "sel_c2w")}
    out = {}
    for ri, rec in enumerate(recs):
        L = _load_hand_pose_labels(f"{REF}/{rec}")
        out.update({f"r{ri}_cam_f": np.array([c.f for c in L.cameras]), f"r{ri}_cam_c": np.array([c.c for c in L.cameras]),
                    f"r{ri}_cam_coeffs": np.array([tuple(c.distort) for c in L.cameras]),
                    f"r{ri}_cam_wh": np.array([(c.width, c.height) for c in L.cameras], np.int64),
                    f"r{ri}_camera_angles": np.array(L.camera_angles, np.float64)})
        out.update(hand_model_arrays(L.hand_model, f"r{ri}_"))
        for fi in range(ri, len(L), 23):
            for vi, (ncp, minvis, maxv, sort_idx, fmul, shift) in enumerate(CROPCAM_VARIANTS):
                for h in range(2):
                    if L.hand_confidences[fi, h] <= 0:
                        continue
                    wrist = L.wrist_transforms[fi, h].astype(np.float64).copy()
                    # shift along the x axis of camera 1 (world units: mm)
                    wrist[:3, 3] += shift * L.camera_to_world_transforms[fi, 1][:3, 0]
                    cams = [c.copy(camera_to_world_xf=L.camera_to_world_transforms[fi, i]) for i, c in enumerate(L.cameras)]
                    pose = SingleHandPose(L.joint_angles[fi, h], wrist, 1.0)
                    try:
                        cc = gen_crop_cameras_from_pose(cams, L.camera_angles, L.hand_model, pose, h, ncp, (96, 96),
                                                        max_view_num=maxv, sort_camera_index=sort_idx, focal_multiplier=fmul,
                                                        mirror_right_hand=True, min_required_vis_landmarks=minvis)
                        n_sel = len(cc)
                    except ValueError:
                        cc, n_sel = {}, -1  # "Unable to create crop camera" (crop.py:25-26)
                    sel_cam = -np.ones(4, np.int64); sel_f = np.zeros((4, 2)); sel_c = np.zeros((4, 2)); sel_c2w = np.zeros((4, 4, 4))
                    for s, (ci, cam) in enumerate(cc.items()):
                        sel_cam[s], sel_f[s], sel_c[s], sel_c2w[s] = ci, cam.f, cam.c, cam.camera_to_world_xf
                    for k, v in (("rec", ri), ("frame", fi), ("hand", h), ("variant", vi),
                                 ("joint_angles", L.joint_angles[fi, h].astype(np.float64)), ("wrist", wrist),
                                 ("c2w", L.camera_to_world_transforms[fi].astype(np.float64)), ("n_sel", n_sel),
                                 ("sel_cam", sel_cam), ("sel_f", sel_f), ("sel_c", sel_c), ("sel_c2w", sel_c2w)):
                        rows[k].append(v)
    out.update({k: np.array(v) for k, v in rows.items()})
    out["variants"] = np.array([[a, b, c, int(d), e, f] for a, b, c, d, e, f in CROPCAM_VARIANTS], np.float64)
    np.savez_compressed(os.path.join(GOLD, "crop_cameras.npz"), **out)
    import collections
    print("crop_cameras.npz", len(out["rec"]), "poses; views per pose:", sorted(collections.Counter(out["n_sel"].tolist()).items()),
          "; camera sets:", collections.Counter(tuple(int(c) for c in r if c >= 0) for r in out["sel_cam"]).most_common(12))


def _load_generic():
    import json

    from lib.common.hand import HandModel

    with open(f"{REF}/dataset/generic_hand_model.json") as f:
        js = json.load(f)
    keep = {k: (torch.tensor(v) if isinstance(v, list) else v) for k, v in js.items()}
    for k in ("mesh_vertices", "mesh_triangles", "dense_bone_weights"):
        keep[k] = None
    return HandModel(**keep)


def make_warp():
    from lib.common import camera as RC

    G = np.load(os.path.join(GOLD, "geometry.npz"))
    crops, meta = [], []
    n = 0
    for k in range(0, G["crop_cam"].shape[0], 10):
        for h in range(2):
            for s in range(2):
                ci = int(G["crop_cam"][k, h, s])
                if ci < 0:
                    continue
                src = RC.Fisheye62CameraModel(
                    636, 480, G["cam_f"][ci], G["cam_c"][ci], G["cam_coeffs"][ci], G["camera_to_world"][k, ci]
                )
                dst = RC.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [], G["crop_c2w"][k, h, s])
                kind = "noise" if n % 2 == 0 else "smooth"
                img = synth_image(kind, n)
                crops.append(RT._warp_image(src, dst, img))
                meta.append((k, h, s, n % 2, n))
                n += 1
    np.savez_compressed(os.path.join(GOLD, "warp.npz"), crops=np.stack(crops), meta=np.array(meta, np.int64))
    print("warp.npz", len(crops), "crops")


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    what = sys.argv[1:] or ["fk", "geometry", "warp", "cropcams", "cropmats", "torchdata", "eval", "model"]
    if "fk" in what:
        make_fk()
    if "eval" in what:
        make_eval()
    if "cropmats" in what:
        make_cropmats()
    if "torchdata" in what:
        make_torchdata()
    if "geometry" in what:
        make_geometry()
    if "cropcams" in what:
        make_cropcams()
    if "warp" in what:
        make_warp()
    if "model" in what:
        from make_golden_model import make_model

        make_model()
