"""GPU parity tests of the network + tracker through the reference-facing API, against the
committed outputs of the REFERENCE model (tests/golden/model.npz) and the CPU oracle."""
import numpy as np
import pytest
import torch

import model_oracle as MO
import np_oracle as O
import scenario

pytestmark = pytest.mark.gpu

# float32 path: differences are accumulation-order only
TOL_ANGLE_RAD = 2e-4
TOL_WRIST_M = 2e-5
TOL_FEAT = 2e-4


@pytest.fixture(scope="module")
def state_dict():
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    return make_random_state_dict(0)


@pytest.fixture()
def model(cuda_lib, state_dict):
    from absolutetrack_b200.lib.models.model_loader import model_from_state_dict

    return model_from_state_dict(state_dict).cuda()


def _skel():
    hm = scenario.hand_model()
    return (torch.from_numpy(hm["joint_rotation_axes"]).float(), torch.from_numpy(hm["joint_rest_positions"] * 0.001).float())


def _inputs(fi, dev="cuda"):
    from absolutetrack_b200.lib.models.umetrack_model import InputFrameData, InputFrameDesc

    fd = InputFrameData(torch.from_numpy(fi["images"]).to(dev), torch.from_numpy(fi["intrinsics"]).to(dev),
                        torch.from_numpy(fi["extrinsics"]).to(dev))
    de = InputFrameDesc(torch.from_numpy(fi["sample_range"]).to(dev), torch.from_numpy(fi["memory_idx"]).to(dev),
                        torch.from_numpy(fi["use_memory"]).to(dev), torch.from_numpy(fi["hand_idx"]).to(dev))
    return fd, de


def test_encoder_features_match_reference(model, gold, state_dict):
    M = gold["model"]
    seq = scenario.model_sequence(n_frames=1, start=10)
    feats = model.encode(torch.from_numpy(seq[0]["images"]).cuda())          # [c,6,6,72]
    nchw = feats.permute(0, 3, 1, 2).cpu().numpy()
    assert np.abs(nchw[:2] - M["k0_features"]).max() < TOL_FEAT                # reference golden
    ref = MO.backbone(state_dict, torch.from_numpy(seq[0]["images"])).numpy()  # oracle, all crops
    assert np.abs(nchw - ref).max() < TOL_FEAT
    # uint8 crops in == float crops in (the /255 is fused in the stem)
    u8 = torch.from_numpy(np.rint(seq[0]["images"] * 255).astype(np.uint8)).cuda()
    assert torch.equal(model.encode(u8), feats)


@pytest.mark.parametrize("head", ["k", "u"])
@pytest.mark.parametrize("tail_mode", ["f16x3", "tf32x3"])
def test_model_sequence_matches_reference_golden(model, gold, head, tail_mode):
    """4-frame known-skeleton sequence (frame 2 has a single-view sample, temporal memory carried
    across frames) and 3-frame scale-head sequence vs the reference's outputs - with both
    fp32-grade tensor-core kernels of the tail (error-compensated fp16 x 3, the default, and TF32 x 3)."""
    from absolutetrack_b200.lib.models.umetrack_model import InputSkeletonData

    model.tail_mode = tail_mode
    M = gold["model"]
    n_frames, start, sv = (int(v) for v in M[f"{head}_case"])
    seq = scenario.model_sequence(n_frames=n_frames, start=start, single_view_frame=None if sv < 0 else sv)
    axes, rest = _skel()
    for t, fi in enumerate(seq):
        fd, de = _inputs(fi)
        if head == "k":
            out = model.regress_pose_use_skeleton(fd, de, InputSkeletonData(axes.cuda(), rest.cuda()))
            assert out.skel_scales is None
        else:
            out = model.regress_pose_pred_skel_scale(fd, de)
            assert np.abs(out.skel_scales.cpu().numpy() - M[f"u{t}_skel_scales"]).max() < 1e-4
        ja, wx = out.joint_angles.cpu().numpy(), out.wrist_xfs.cpu().numpy()
        assert ja.shape == M[f"{head}{t}_joint_angles"].shape
        assert np.abs(ja - M[f"{head}{t}_joint_angles"]).max() < TOL_ANGLE_RAD, (t, np.abs(ja - M[f"{head}{t}_joint_angles"]).max())
        assert np.abs(wx - M[f"{head}{t}_wrist_xfs"]).max() < TOL_WRIST_M, (t, np.abs(wx - M[f"{head}{t}_wrist_xfs"]).max())
        assert np.abs(out.landmark_uncertainty_sigmas.cpu().numpy() - M[f"{head}{t}_sigmas"]).max() < 2e-4
        assert np.all(ja[:, 20:] == 0)


def test_model_batched_samples_and_per_sample_skeletons_match_oracle(model, state_dict):
    """Config-3 style call: many samples per call, memory_idx = arange(bs), [bs,22,3] skeletons,
    2 time steps (run_inference_torch_data.py:39-85)."""
    from absolutetrack_b200.lib.models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData

    frames = [scenario.model_sequence(n_frames=2, start=s) for s in (0, 20, 50)]
    axes, rest = _skel()
    st = MO.TemporalState()
    for t in range(2):
        imgs, intr, extr, rng_, hidx = [], [], [], [], []
        for seq in frames:
            fi = seq[t]
            off = sum(len(x) for x in imgs)
            imgs.append(fi["images"]); intr.append(fi["intrinsics"]); extr.append(fi["extrinsics"])
            rng_.append(fi["sample_range"] + off); hidx.append(fi["hand_idx"])
        imgs, intr, extr = (torch.from_numpy(np.concatenate(a)) for a in (imgs, intr, extr))
        srange, hidx = torch.from_numpy(np.concatenate(rng_)), torch.from_numpy(np.concatenate(hidx))
        n = srange.shape[0]
        midx = torch.arange(n)
        usem = torch.full((n,), t != 0)
        scale = torch.linspace(0.9, 1.1, n).reshape(n, 1, 1)
        sk = (axes.expand(n, 22, 3).contiguous(), rest * scale)
        ref = MO.forward(state_dict, st, imgs, intr, extr, srange, midx, usem, hidx, sk)
        out = model.regress_pose_use_skeleton(
            InputFrameData(imgs.cuda(), intr.cuda(), extr.cuda()), InputFrameDesc(srange.cuda(), midx.cuda(), usem.cuda(), hidx.cuda()),
            InputSkeletonData(sk[0].cuda(), sk[1].cuda()))
        assert np.abs(out.joint_angles.cpu().numpy() - ref["joint_angles"].numpy()).max() < TOL_ANGLE_RAD
        assert np.abs(out.wrist_xfs.cpu().numpy() - ref["wrist_xfs"].numpy()).max() < TOL_WRIST_M
    # temporal memory state equals the oracle's (reference layout [cap,18,6,6])
    assert np.abs(model.mem_features_nchw.cpu().numpy() - st.mem.numpy()).max() < 5e-4


def test_model_error_conventions(model):
    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.lib.models.umetrack_model import InputSkeletonData

    seq = scenario.model_sequence(n_frames=1, start=10, single_view_frame=0)
    fd, de = _inputs(seq[0])
    with pytest.raises(AssertionError):      # umetrack_model.py:224-229
        model.regress_pose_pred_skel_scale(fd, de)
    de.sample_range = de.sample_range.clone()
    de.sample_range[0, 1] = 99               # outside the crop batch
    axes, rest = _skel()
    with pytest.raises(_cabi.AbsoluteTrackError):
        model.regress_pose_use_skeleton(fd, de, InputSkeletonData(axes.cuda(), rest.cuda()))
    with pytest.raises(_cabi.AbsoluteTrackError):   # no CPU path
        model.cpu().encode(torch.zeros(1, 96, 96))


def _tracker_objects(k):
    """InputFrame + crop_cameras + hand model of geometry frame k, built from fixtures."""
    from absolutetrack_b200.lib.common import camera as C
    from absolutetrack_b200.lib.tracker.tracker import InputFrame, ViewData
    from test_gpu_kernels import _hand_model_tuple

    G = scenario.geometry()
    frame = scenario.synth_frame("smooth", 1000 + k)
    views = [ViewData(frame[:, i, :], C.Fisheye62CameraModel(636, 480, G["cam_f"][i], G["cam_c"][i], G["cam_coeffs"][i],
                                                            G["camera_to_world"][k, i]), float(G["camera_angles"][i]))
             for i in range(4)]
    cc = {}
    for h in range(2):
        for s in range(2):
            cam = int(G["crop_cam"][k, h, s])
            if cam >= 0:
                cc.setdefault(h, {})[cam] = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [],
                                                                    G["crop_c2w"][k, h, s])
    return InputFrame(views), cc, _hand_model_tuple(G, "hm_")


def test_hand_tracker_track_frame_matches_reference_golden(model, gold):
    """Full reference-facing path: HandTracker.track_frame (host uint8 views -> crops -> model ->
    TrackingResult in mm) + landmarks_from_hand_pose vs the reference's outputs; dMPJPE < 0.1 mm."""
    from absolutetrack_b200.lib.tracker.perspective_crop import landmarks_from_hand_pose
    from absolutetrack_b200.lib.tracker.tracker import HandTracker, HandTrackerOpts
    from absolutetrack_b200.lib.tracker.tracking_result import SingleHandPose

    M = gold["model"]
    n_frames, start, sv = (int(v) for v in M["k_case"])
    tracker = HandTracker(model, HandTrackerOpts())
    hm_np = scenario.hand_model()
    for t in range(2):  # frames before the single-view one
        frame, cc, hm = _tracker_objects(start + t)
        res = tracker.track_frame(frame, hm, cc)
        assert sorted(res.hand_poses) == [0, 1] and res.num_views == {0: 2, 1: 2}
        for i, h in enumerate(sorted(res.hand_poses)):
            pose = res.hand_poses[h]
            ref_w = M[f"k{t}_wrist_xfs"][i].copy()
            ref_w[:3, 3] *= 1000.0
            assert np.abs(pose.joint_angles - M[f"k{t}_joint_angles"][i]).max() < TOL_ANGLE_RAD
            lm = landmarks_from_hand_pose(hm, pose, h)
            lm_ref = O.landmarks_from_hand_pose(hm_np, M[f"k{t}_joint_angles"][i], ref_w, h)[0]
            assert np.linalg.norm(lm - lm_ref, axis=-1).mean() < 0.1       # mm
            assert np.abs(np.degrees(pose.joint_angles - M[f"k{t}_joint_angles"][i])).mean() < 0.05
    # no hands -> empty result and history reset (tracker.py:268-271)
    res = tracker.track_frame(frame, hm, {})
    assert res.hand_poses == {} and not tracker._valid_tracking_history.any()


@pytest.mark.parametrize("mode", ["device", "host"])
def test_gen_crop_cameras_matches_reference_golden(model, gold, mode):
    """HandTracker.gen_crop_cameras (tracker.py:222-260) through the device kernel (default) and through the
    numpy mirror of perspective_crop.py / crop.py, against the reference's own crop cameras."""
    from absolutetrack_b200.lib.tracker.tracker import HandTracker, HandTrackerOpts
    from absolutetrack_b200.lib.tracker.tracking_result import SingleHandPose

    G = gold["geometry"]
    tracker = HandTracker(model, HandTrackerOpts(crop_cameras=mode))
    for k in range(0, G["crop_cam"].shape[0], 7):
        frame, _, hm = _tracker_objects(k)
        gt = {h: SingleHandPose(G["joint_angles"][k, h], G["wrist_transforms"][k, h], G["hand_confidences"][k, h])
              for h in range(2) if G["hand_confidences"][k, h] > 0}
        cc = tracker.gen_crop_cameras([v.camera for v in frame.views], list(G["camera_angles"]), hm, gt, 1)
        for h in range(2):
            cams = [int(c) for c in G["crop_cam"][k, h] if c >= 0]
            assert list(cc.get(h, {}).keys()) == cams
            for s, cam in enumerate(cams):
                got = cc[h][cam]
                assert np.allclose(got.f, G["crop_f"][k, h, s], rtol=1e-5)
                assert np.allclose(got.c, G["crop_c"][k, h, s])
                assert np.allclose(got.camera_to_world_xf, G["crop_c2w"][k, h, s], atol=1e-4)


def test_batched_tracker_ingest_modes_agree_and_match_oracle(cuda_lib, state_dict):
    """BatchedHandTracker (the throughput entry bench.py times): S sequences in lock-step over 3
    time steps.  (1) Host ingest by source-box pull, by zero-copy staging and by whole-frame copy
    all give bit-identical results to device-resident frames; (2) with the bit-exact fp64 crop map
    the poses/landmarks equal the CPU oracle run on the same crops with the same memory slots."""
    from absolutetrack_b200 import synthetic
    from absolutetrack_b200.lib.models.model_loader import model_from_state_dict
    from absolutetrack_b200.lib.tracker.batch_tracker import BatchedHandTracker
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda

    S, T = 3, 3
    seq = np.array([0, 7, 31])
    hm = synthetic.hand_model()
    g = torch.Generator().manual_seed(5)
    batches = []
    for t in range(T):
        fr = torch.randint(0, 256, (S, synthetic.H, synthetic.V, synthetic.W), dtype=torch.uint8, generator=g)
        batches.append(synthetic.make_batch(seq, t, frames=fr, first_step=(t == 0)).pin())

    def run(ingest, map_mode):
        tr = BatchedHandTracker(model_from_state_dict(state_dict).cuda(), hm, max_slots=2 * S, map_mode=map_mode,
                                ingest=ingest if ingest != "device" else "pull")
        outs = []
        for hb in batches:
            if ingest == "device":
                r = tr.step_device(hb.to("cuda"))
                o = {k: v.cpu() for k, v in vars(r).items() if v is not None}
            else:
                o = tr.step(hb)
            torch.cuda.synchronize()
            outs.append({k: v.clone() for k, v in o.items()})
        tr.check()
        return outs, tr

    base, _ = run("device", "fp32")
    for ingest in ("pull", "zero_copy", "copy"):
        outs, tr = run(ingest, "fp32")
        for a, b in zip(outs, base):
            for k in a:
                assert torch.equal(a[k], b[k]), (ingest, k)
        if ingest != "copy":
            assert 0 < tr.pulled_bytes.item() < T * batches[0].frames.numel() / 4   # boxes, not frames

    # oracle: same crops (fp64 map is bit-exact vs the reference warp), batched samples, slot = seq*2+hand
    outs, _ = run("pull", "fp64")
    hmn = scenario.hand_model()
    st = MO.TemporalState()
    for t, hb in enumerate(batches):
        crops = crop_resample_cuda(hb.frames.cuda(), synthetic.H * synthetic.V * synthetic.W, synthetic.W,
                                   synthetic.V * synthetic.W, hb.crop_params.cuda(), (96, 96), torch.float32, "fp64").cpu()
        ref = MO.forward(state_dict, st, crops, hb.intrinsics, hb.extrinsics, hb.sample_range, hb.memory_idx,
                         hb.use_memory, hb.hand_idx, _skel())
        o = outs[t]
        assert np.abs(o["joint_angles"].numpy() - ref["joint_angles"].numpy()).max() < TOL_ANGLE_RAD
        for i in range(o["landmarks"].shape[0]):
            w = ref["wrist_xfs"][i].numpy().copy()
            w[:3, 3] *= 1000.0
            lm_ref = O.landmarks_from_hand_pose(hmn, ref["joint_angles"][i].numpy(), w, int(hb.hand_idx[i]))[0]
            assert np.abs(o["landmarks"][i].numpy() - lm_ref).max() < 0.05   # mm


def test_config3_files_to_poses_match_oracle_on_reference_inputs(model, state_dict, gold):
    """Config 3 end to end: .torch.idx/.torch.bin fixture pairs -> TorchIdx -> preprocess (GPU crops)
    -> the per-frame calls of run_inference_torch_data._unpack_batched_data (:39-85) on the CUDA
    model, against the CPU oracle network fed with the REFERENCE's preprocess outputs of the same
    files.  Tolerances are looser than the network's own (2e-4 rad): the inputs differ by the
    crop-matrix round-off (crops within 2e-3)."""
    import os

    from absolutetrack_b200.lib.batched_dataset import data_transform as DT
    from absolutetrack_b200.lib.data_utils.idxbinfile import TorchIdx
    from absolutetrack_b200.lib.models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData

    G = gold["torchdata"]
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "torchdata")
    mono = TorchIdx(os.path.join(d, "seq.mono.torch.idx")).read_bin()
    labels = TorchIdx(os.path.join(d, "seq.labels.torch.idx")).read_bin()
    mis = [DT.preprocess({"mono": np.array(mono[i]), "labels": labels[i]}, (96, 96))[0] for i in range(2)]
    bs, seq_len, nv = 2, 3, 2
    st = MO.TemporalState()
    srange = torch.tensor([(i * nv, (i + 1) * nv) for i in range(bs)])
    midx = torch.arange(bs)
    for t in range(seq_len):
        usem = torch.full((bs,), t != 0)
        hidx = torch.stack([m.hand_idx[t] for m in mis]).long()
        axes = torch.stack([m.orig_pose_data.left_hand_model.joint_rotation_axes[t] for m in mis])
        rest = torch.stack([m.orig_pose_data.left_hand_model.joint_rest_positions[t] for m in mis])
        out = model.regress_pose_use_skeleton(
            InputFrameData(torch.cat([m.left_images[t] for m in mis]), torch.cat([m.intrinsics[t] for m in mis]),
                           torch.cat([m.extrinsics_xf[t] for m in mis])),
            InputFrameDesc(srange.cuda(), midx.cuda(), usem.cuda(), hidx.cuda()), InputSkeletonData(axes.cuda(), rest.cuda()))
        g = lambda k: torch.from_numpy(np.concatenate([G[f"s{i}_{k}"][t] for i in range(bs)]))  # noqa: E731
        ref = MO.forward(state_dict, st, g("left_images"), g("intrinsics"), g("extrinsics_xf"), srange, midx, usem, hidx,
                         (torch.from_numpy(np.stack([G[f"s{i}_orig_axes"][t] for i in range(bs)])),
                          torch.from_numpy(np.stack([G[f"s{i}_orig_rest"][t] for i in range(bs)]))))
        assert np.abs(out.joint_angles.cpu().numpy() - ref["joint_angles"].numpy()).max() < 2e-3
        assert np.abs(out.wrist_xfs.cpu().numpy() - ref["wrist_xfs"].numpy()).max() < 2e-4
