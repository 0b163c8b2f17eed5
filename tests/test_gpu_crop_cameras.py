"""GPU parity tests of the device crop-camera generator (csrc/crop_cameras.cu, C-ABI
at_gen_crop_cameras) against the outputs of the REFERENCE gen_crop_cameras_from_pose
(tests/golden/crop_cameras.npz, geometry.npz) and of the host path it replaces."""
import numpy as np
import pytest
import torch

import scenario
from test_cpu import cropcam_case

pytestmark = pytest.mark.gpu


def _hand_model(C, pre):
    from absolutetrack_b200.lib.common.hand import HandModel

    g = lambda k: torch.from_numpy(np.asarray(C[pre + k]))  # noqa: E731
    z = torch.zeros(1)
    return HandModel(
        joint_rotation_axes=g("joint_rotation_axes").float(), joint_rest_positions=g("joint_rest_positions").float(),
        joint_frame_index=z, joint_parent=z, joint_first_child=z, joint_next_sibling=z,
        landmark_rest_positions=g("landmark_rest_positions").float(),
        landmark_rest_bone_weights=g("landmark_rest_bone_weights").float(),
        landmark_rest_bone_indices=g("landmark_rest_bone_indices"), hand_scale=None,
        joint_limits=g("joint_limits").float(),
    )


def test_crop_camera_kernel_matches_reference_golden(cuda_lib, gold):
    """770 poses x 5 option variants over the four sample recordings (1-4 views, ranked by
    visibility or by camera index, 21/42/63 crop points, hands pushed out of view): the selected
    cameras and their order are exact; intrinsics / extrinsics agree to the round-off of the float32
    FK landmarks (same bar as the numpy oracle, tests/test_cpu.py)."""
    from absolutetrack_b200.lib.common import camera
    from absolutetrack_b200.lib.tracker.crop_cameras import CropCameraOpts, DeviceRig, gen_crop_inputs_cuda

    C = gold["crop_cameras"]
    groups = {}
    for i in range(len(C["rec"])):
        groups.setdefault((int(C["rec"][i]), int(C["variant"][i])), []).append(i)
    n_checked = 0
    for (ri, vi), rows in sorted(groups.items()):
        pre = f"r{ri}_"
        cams_np, angles, _, o = cropcam_case(C, rows[0])
        cams = [camera.Fisheye62CameraModel(c["width"], c["height"], c["f"], c["c"], list(c["coeffs"])) for c in cams_np]
        rig = DeviceRig(cams, list(angles), _hand_model(C, pre))
        opts = CropCameraOpts(min_num_crops=1, **o)
        rows = np.array(rows)
        P = len(rows)
        res = gen_crop_inputs_cuda(
            rig,
            torch.from_numpy(C["joint_angles"][rows]).float().cuda(),
            torch.from_numpy(C["wrist"][rows]).float().cuda(),
            torch.arange(P, dtype=torch.int32).cuda(),              # every pose has its own frame (camera poses)
            torch.from_numpy(C["hand"][rows].astype(np.int32)).cuda(),
            torch.from_numpy(C["c2w"][rows]).contiguous().cuda(),
            opts,
        ).resolve()
        pose_of = res.pose_index.cpu().numpy()
        rng = res.sample_range.cpu().numpy()
        cam_idx = res.crop_camera_index.cpu().numpy()
        c2w = res.crop_camera_to_world.cpu().numpy()
        par = res.crop_params.cpu().numpy().view(np.float64).reshape(-1, 32)
        intr = res.intrinsics.cpu().numpy()
        extr = res.extrinsics.cpu().numpy()
        expect = [j for j in range(P) if C["n_sel"][rows[j]] > 0]
        assert list(pose_of) == expect                                         # dropped hands, input order kept
        assert np.array_equal(res.hand_idx.cpu().numpy(), C["hand"][rows[expect]])
        assert np.array_equal(res.memory_idx.cpu().numpy(), np.array(expect) * 2 + C["hand"][rows[expect]])
        assert not res.use_memory.any()
        off = 0
        for s_i, j in enumerate(pose_of):
            i = rows[j]
            want = [int(c) for c in C["sel_cam"][i] if c >= 0]
            assert tuple(rng[s_i]) == (off, off + len(want))
            assert list(cam_idx[off:off + len(want)]) == want, (ri, vi, i)
            for s in range(len(want)):
                c = off + s
                f_dst, c_dst = par[c, 14:16], par[c, 16:18]
                assert np.allclose(f_dst, C["sel_f"][i, s], rtol=3e-6)
                assert np.allclose(c_dst, C["sel_c"][i, s])
                assert np.abs(c2w[c][:3, :3] - C["sel_c2w"][i, s][:3, :3]).max() < 3e-6
                assert np.abs(c2w[c][:3, 3] - C["sel_c2w"][i, s][:3, 3]).max() < 1e-9
                # tensors of tracker._make_inputs (:338-345)
                assert np.allclose(intr[c], [[f_dst[0], 0, c_dst[0]], [0, f_dst[1], c_dst[1]], [0, 0, 1]], rtol=1e-6)
                w2e = np.linalg.inv(C["sel_c2w"][i, s])
                w2e[:3, 3] *= 0.001
                assert np.abs(extr[c] - w2e).max() < 5e-6
                # resampler record: M = R_src^T R_dst, d = 0 (same optical centre)
                src = C["c2w"][i, want[s]]
                assert np.abs(par[c, 2:11].reshape(3, 3) - src[:3, :3].T @ C["sel_c2w"][i, s][:3, :3]).max() < 3e-6
                assert np.abs(par[c, 11:14]).max() < 1e-9
                n_checked += 1
            off += len(want)
        assert res.c == off
    assert n_checked > 1400


def test_crop_camera_kernel_history_validity_and_empty(cuda_lib, gold):
    """use_memory / tracking-history semantics of tracker.py:105,353-358,409-410; low-confidence hands
    are skipped (:236-237); min_num_crops drops single-view hands (:254-260); empty input."""
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.lib.tracker.crop_cameras import CropCameraOpts, DeviceRig, gen_crop_inputs_cuda

    cams, angles = synthetic.rig_cameras()
    rig = DeviceRig(cams, angles, synthetic.hand_model())
    S = 5
    pb = synthetic.make_pose_batch(list(range(S)), 0)
    dev = "cuda"
    hist = torch.zeros(2 * S, dtype=torch.bool, device=dev)
    args = [getattr(pb, f).to(dev) for f in ("joint_angles", "wrist_xforms", "pose_frame", "pose_hand", "camera_to_world")]
    r1 = gen_crop_inputs_cuda(rig, *args, CropCameraOpts(), tracking_history=hist).resolve()
    assert r1.n == 2 * S and r1.c == 4 * S and not r1.use_memory.any() and hist.all()
    valid = torch.ones(2 * S, dtype=torch.uint8, device=dev)
    valid[3] = 0
    r2 = gen_crop_inputs_cuda(rig, *args, CropCameraOpts(), pose_valid=valid, tracking_history=hist).resolve()
    assert r2.n == 2 * S - 1 and r2.use_memory.all()
    assert 3 not in r2.pose_index.tolist() and not hist[3] and hist.sum() == 2 * S - 1
    r3 = gen_crop_inputs_cuda(rig, *args, CropCameraOpts(), tracking_history=hist).resolve()
    um = r3.use_memory.cpu().numpy()
    assert r3.n == 2 * S and not um[3] and um.sum() == 2 * S - 1
    # every hand of this recording is seen by exactly cameras (1, 2): asking for 3 views drops all
    r4 = gen_crop_inputs_cuda(rig, *args, CropCameraOpts(max_view_num=3, min_num_crops=3)).resolve()
    assert r4.n == 0 and r4.c == 0
    e = lambda *s, dt=torch.float32: torch.empty(s, dtype=dt, device=dev)  # noqa: E731
    r5 = gen_crop_inputs_cuda(rig, e(0, 22), e(0, 4, 4), e(0, dt=torch.int32), e(0, dt=torch.int32),
                              e(0, 4, 4, 4, dt=torch.float64), CropCameraOpts()).resolve()
    assert r5.n == 0 and r5.c == 0


def test_crops_from_device_crop_cameras_match_host_crop_cameras(cuda_lib, gold):
    """The AtCropParams records generated on the device drive the resampler to the same crops as
    the records packed on the host from the reference's crop cameras (geometry.npz): fp64 map,
    smooth images -> at most 1 grey level on the few pixels whose 1/32-px bin flips."""
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.lib.tracker.crop_cameras import CropCameraOpts, DeviceRig, gen_crop_inputs_cuda
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

    S = 24
    seq = list(range(S))
    yy, xx = np.mgrid[0:synthetic.H, 0:synthetic.W]
    img = np.clip(np.rint(127 + 100 * np.sin(xx / 15.0) * np.cos(yy / 11.0)), 0, 255).astype(np.uint8)
    frames = torch.from_numpy(np.broadcast_to(img[None, :, None, :], (S, synthetic.H, synthetic.V, synthetic.W)).copy())
    hb = synthetic.make_batch(seq, 0, frames=frames)
    pb = synthetic.make_pose_batch(seq, 0, frames=frames)
    cams, angles = synthetic.rig_cameras()
    rig = DeviceRig(cams, angles, synthetic.hand_model())
    dev = "cuda"
    res = gen_crop_inputs_cuda(rig, pb.joint_angles.to(dev), pb.wrist_xforms.to(dev), pb.pose_frame.to(dev),
                               pb.pose_hand.to(dev), pb.camera_to_world.to(dev), CropCameraOpts(),
                               pose_valid=pb.pose_valid.to(dev)).resolve()
    assert res.c == hb.crop_params.shape[0] and res.n == hb.hand_idx.shape[0]
    assert torch.equal(res.sample_range.cpu(), hb.sample_range) and torch.equal(res.hand_idx.cpu(), hb.hand_idx)
    assert torch.equal(res.memory_idx.cpu(), hb.memory_idx)
    assert np.abs(res.intrinsics.cpu().numpy() - hb.intrinsics.numpy()).max() < 2e-4
    assert np.abs(res.extrinsics.cpu().numpy() - hb.extrinsics.numpy()).max() < 5e-6
    fd = frames.to(dev)
    strides = (synthetic.H * synthetic.V * synthetic.W, synthetic.W, synthetic.V * synthetic.W)
    a = crop_resample_cuda(fd, *strides, res.crop_params, (96, 96), torch.uint8, "fp64").cpu().numpy().astype(int)
    b = crop_resample_cuda(fd, *strides, hb.crop_params.to(dev), (96, 96), torch.uint8, "fp64").cpu().numpy().astype(int)
    assert np.abs(a - b).max() <= 1
    assert (a != b).mean() < 0.01


def test_batched_tracker_step_poses_matches_step_with_host_crop_cameras(cuda_lib, gold):
    """BatchedHandTracker.step_poses (crop cameras + inputs generated on the device each step) vs
    step() on batches built from the reference's crop cameras, 3 time steps with carried memory."""
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict
    from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker

    sd = make_random_state_dict(0)
    hm = synthetic.hand_model()
    S, T = 6, 3
    seq = list(range(S))
    yy, xx = np.mgrid[0:synthetic.H, 0:synthetic.W]
    frames = []
    for t in range(T):
        img = np.clip(np.rint(127 + 100 * np.sin((xx + 7 * t) / 15.0) * np.cos((yy + 3 * t) / 11.0)), 0, 255).astype(np.uint8)
        frames.append(torch.from_numpy(np.broadcast_to(img[None, :, None, :], (S, synthetic.H, synthetic.V, synthetic.W)).copy()).pin_memory())
    ta = BatchedHandTracker(model_from_state_dict(sd).cuda(), hm, max_slots=2 * S, map_mode="fp64")
    tb = BatchedHandTracker(model_from_state_dict(sd).cuda(), hm, max_slots=2 * S, map_mode="fp64")
    tb.attach_rig(*synthetic.rig_cameras())
    for t in range(T):
        oa = ta.step(synthetic.make_batch(seq, t, frames=frames[t], first_step=(t == 0)).pin())
        ob = tb.step_poses(synthetic.make_pose_batch(seq, t, frames=frames[t]).pin())
        torch.cuda.synchronize()
        n = ob["n"]
        assert n == oa["joint_angles"].shape[0] == 2 * S and ob["c"] == 4 * S
        assert np.array_equal(ob["memory_idx"][:n].numpy(), np.arange(2 * S))
        assert np.abs(ob["joint_angles"][:n].numpy() - oa["joint_angles"].numpy()).max() < 2e-3   # rad
        assert np.abs(ob["landmarks"][:n].numpy() - oa["landmarks"].numpy()).max() < 0.5          # mm
    ta.check(); tb.check()
