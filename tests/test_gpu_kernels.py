"""GPU parity tests of the individual kernels, through the C-ABI, against the CPU oracle and the
committed reference goldens.  Run with `pytest -m gpu` on a B200."""
import numpy as np
import pytest
import torch

import np_oracle as O
import scenario

pytestmark = pytest.mark.gpu


def _hand_model_tuple(d, prefix=""):
    from absolutetrack_b200.lib.common.hand import HandModel

    g = lambda k: torch.from_numpy(np.asarray(d[prefix + k]))  # noqa: E731
    z = torch.zeros(1)
    return HandModel(
        joint_rotation_axes=g("joint_rotation_axes").float(), joint_rest_positions=g("joint_rest_positions").float(),
        joint_frame_index=z, joint_parent=z, joint_first_child=z, joint_next_sibling=z,
        landmark_rest_positions=g("landmark_rest_positions").float(),
        landmark_rest_bone_weights=g("landmark_rest_bone_weights").float(),
        landmark_rest_bone_indices=g("landmark_rest_bone_indices"), hand_scale=None,
        joint_limits=g("joint_limits").float() if prefix + "joint_limits" in d else None,
    )


# ------------------------------------------------------------------ FK + skinning (a11)
@pytest.mark.parametrize("rec", ["00", "02", "11"])
def test_fk_matches_reference_golden_keypoints(cuda_lib, gold, rec):
    """Known-answer test: the reference's own gt_keypoints pickles, tolerance 1e-3 mm."""
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    G = gold["fk_user05"]
    p = f"r{rec}_"
    hm = _hand_model_tuple(G, p)
    ja = torch.from_numpy(G[p + "joint_angles"])            # [T,2,22]
    wx = torch.from_numpy(G[p + "wrist_transforms"])        # [T,2,4,4] f64
    T = ja.shape[0]
    hidx = torch.tensor([0, 1]).expand(T, 2)
    out = skin_landmarks_cuda(hm, ja, wx, hand_idx=hidx).cpu().numpy()  # [T,2,21,3]
    valid = G[p + "valid"]                                   # [2,T]
    gt = G[p + "gt_keypoints"]                               # [2,T,21,3]
    n = 0
    for h in range(2):
        d = np.abs(out[:, h][valid[h]] - gt[h][valid[h]])
        n += int(valid[h].sum())
        assert d.max() < 1e-3, f"hand {h}: max |dlandmark| = {d.max()} mm"
    assert n > 800


def test_fk_matches_oracle_random_poses_and_per_sample_models(cuda_lib, gold):
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks, skin_landmarks_cuda

    G = gold["geometry"]
    hm = _hand_model_tuple(G, "hm_")
    rng = np.random.default_rng(0)
    for n in (1, 63, 64, 65, 1000):
        ja = rng.uniform(-1.5, 1.5, (n, 22)).astype(np.float32)
        wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
        wx[:, :3, :3] = O.so3_exp_map_f32(rng.normal(0, 1, (n, 3)).astype(np.float32))
        wx[:, :3, 3] = rng.normal(0, 300, (n, 3))
        hidx = rng.integers(0, 2, n)
        ref = O.landmarks_from_hand_pose(scenario.hand_model(), ja, wx, hidx)
        out = skin_landmarks_cuda(hm, torch.from_numpy(ja), torch.from_numpy(wx), hand_idx=torch.from_numpy(hidx))
        assert np.abs(out.cpu().numpy() - ref).max() < 1e-3
    # per-sample (batched) hand models with leading dims [bs, seq] (run_inference_torch_data.py:123-130)
    bs, seq = 3, 5
    scale = torch.linspace(0.8, 1.2, bs * seq).reshape(bs, seq, 1, 1)
    from absolutetrack_b200.lib.common.hand import HandModel

    hmb = hm._replace(
        joint_rotation_axes=hm.joint_rotation_axes.expand(bs, seq, 22, 3).contiguous(),
        joint_rest_positions=hm.joint_rest_positions * scale,
        landmark_rest_positions=hm.landmark_rest_positions * scale,
        landmark_rest_bone_weights=hm.landmark_rest_bone_weights.expand(bs, seq, 21, 3).contiguous(),
        landmark_rest_bone_indices=hm.landmark_rest_bone_indices.expand(bs, seq, 21, 3).contiguous(),
    )
    ja = torch.from_numpy(rng.uniform(-1, 1, (bs, seq, 22)).astype(np.float32))
    wx = torch.eye(4).expand(bs, seq, 4, 4).contiguous()
    out = skin_landmarks(hmb, ja, wx)
    assert out.shape == (bs, seq, 21, 3) and out.device.type == "cpu"
    md = {k: getattr(hmb, k).reshape((bs * seq,) + tuple(getattr(hmb, k).shape[2:])).numpy() for k in
          ("joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions", "landmark_rest_bone_weights",
           "landmark_rest_bone_indices")}
    ref = O.skin_landmarks(md, ja.reshape(-1, 22).numpy(), wx.reshape(-1, 4, 4).numpy())
    assert np.abs(out.reshape(-1, 21, 3).numpy() - ref).max() < 1e-3
    # scalar leading dims () as in perspective_crop.py:27-37
    one = skin_landmarks(hm, ja[0, 0], wx[0, 0])
    assert one.shape == (21, 3)


def test_fk_ragged_tiles_and_unaligned_pointers(cuda_lib, gold):
    """Several 256-hand tiles with a ragged tail, and plain-`float*` callers whose pointers are only
    4-byte aligned (the kernel's 16-byte copies must fall back to scalar ones): identical results."""
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    hm = _hand_model_tuple(gold["geometry"], "hm_")
    rng = np.random.default_rng(5)
    for n in (256 * 3 + 7, 2049):
        ja = rng.uniform(-1.5, 1.5, (n, 22)).astype(np.float32)
        wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
        wx[:, :3, :3] = O.so3_exp_map_f32(rng.normal(0, 1, (n, 3)).astype(np.float32))
        wx[:, :3, 3] = rng.normal(0, 300, (n, 3))
        hidx = torch.from_numpy(rng.integers(0, 2, n)).cuda()
        ref = O.landmarks_from_hand_pose(scenario.hand_model(), ja, wx, hidx.cpu().numpy())
        a = skin_landmarks_cuda(hm, torch.from_numpy(ja).cuda(), torch.from_numpy(wx).cuda(), hand_idx=hidx)
        assert np.abs(a.cpu().numpy() - ref).max() < 1e-3
        for off in (1, 2, 3):
            bj = torch.zeros(n * 22 + 4, device="cuda")
            bw = torch.zeros(n * 16 + 4, device="cuda")
            bo = torch.full((n * 63 + 8,), -7.0, device="cuda")
            vj = bj[off:off + n * 22].view(n, 22).copy_(torch.from_numpy(ja))
            vw = bw[off:off + n * 16].view(n, 4, 4).copy_(torch.from_numpy(wx))
            vo = bo[off:off + n * 63].view(n, 21, 3)
            assert vj.data_ptr() % 16 != 0 and vo.data_ptr() % 16 != 0
            b = skin_landmarks_cuda(hm, vj, vw, hand_idx=hidx, out=vo)
            assert b.data_ptr() == vo.data_ptr()
            assert torch.equal(a, vo)
            assert bool((bo[:off] == -7.0).all()) and bool((bo[off + n * 63:] == -7.0).all())  # nothing outside


def test_fk_scaled_entry_folds_the_metre_to_mm_step(cuda_lib, gold):
    """at_fk_skin_scaled_f32 = scale the wrist translations (tracker.py:379), then at_fk_skin_f32;
    the scaled transforms come back in the same launch.  Bit-identical to doing it in two steps."""
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    hm = _hand_model_tuple(gold["geometry"], "hm_")
    rng = np.random.default_rng(9)
    n = 700
    ja = torch.from_numpy(rng.uniform(-1.2, 1.2, (n, 22)).astype(np.float32)).cuda()
    wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    wx[:, :3, :3] = O.so3_exp_map_f32(rng.normal(0, 1, (n, 3)).astype(np.float32))
    wx[:, :3, 3] = rng.normal(0, 0.3, (n, 3))
    wx = torch.from_numpy(wx).cuda()
    hidx = torch.from_numpy(rng.integers(0, 2, n)).cuda()
    scaled = wx.clone()
    scaled[:, :3, 3] *= 1000.0
    per_sample = hm._replace(**{k: getattr(hm, k).expand((n,) + tuple(getattr(hm, k).shape)).contiguous() for k in
                                ("joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions",
                                 "landmark_rest_bone_weights", "landmark_rest_bone_indices")})
    for model in (hm, per_sample):
        ref = skin_landmarks_cuda(model, ja, scaled, hand_idx=hidx)
        for off in (0, 1):  # 16-byte aligned and 4-byte aligned wrist_out
            buf = torch.full((n * 16 + 4,), -3.0, device="cuda")
            wo = buf[off:off + n * 16].view(n, 4, 4)
            lm = skin_landmarks_cuda(model, ja, wx, hand_idx=hidx, translation_scale=1000.0, wrist_out=wo)
            assert torch.equal(lm, ref) and torch.equal(wo, scaled)
            assert bool((buf[:off] == -3.0).all()) and bool((buf[off + n * 16:] == -3.0).all())
    with pytest.raises(Exception):
        skin_landmarks_cuda(hm, ja, wx, wrist_out=wx)  # in place is refused


def test_fk_landmark_without_weights_is_zero(cuda_lib, gold):
    """A landmark whose bone weights are all zero contributes nothing (hand_skinning.py:92-94 drops zero
    weights): it comes out as the origin, through the fast path's zero-fill branch."""
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    hm = _hand_model_tuple(gold["geometry"], "hm_")
    w = hm.landmark_rest_bone_weights.clone()
    w[5] = 0.0
    w[20, 1:] = 0.0
    hm0 = hm._replace(landmark_rest_bone_weights=w)
    rng = np.random.default_rng(4)
    n = 300
    ja = rng.uniform(-1, 1, (n, 22)).astype(np.float32)
    wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    wx[:, :3, 3] = rng.normal(0, 100, (n, 3))
    out = skin_landmarks_cuda(hm0, torch.from_numpy(ja), torch.from_numpy(wx)).cpu().numpy()
    md = {k: np.broadcast_to(getattr(hm0, k).numpy(), (n,) + tuple(getattr(hm0, k).shape)) for k in
          ("joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions", "landmark_rest_bone_weights",
           "landmark_rest_bone_indices")}
    ref = O.skin_landmarks(md, ja, wx)
    assert np.all(out[:, 5] == 0.0) and np.abs(out - ref).max() < 1e-3 and np.abs(out[:, 4]).max() > 1.0


def test_fk_empty_batch(cuda_lib, gold):
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    hm = _hand_model_tuple(gold["geometry"], "hm_")
    out = skin_landmarks_cuda(hm, torch.zeros(0, 22), torch.zeros(0, 4, 4))
    assert out.shape == (0, 21, 3)


# ------------------------------------------------------------------ crop resampler (a1)
def _crop_records(meta):
    from absolutetrack_b200.lib.common import camera as C

    G = scenario.geometry()
    recs = []
    for (k, h, s, kind, seed) in meta:
        cam = int(G["crop_cam"][k, h, s])
        src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
        dst = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [], G["crop_c2w"][k, h, s])
        recs.append(C.crop_kernel_params(src, dst))
    return recs


def _run_crops(meta, map_mode, out_dtype=torch.uint8):
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda, pack_crop_params

    imgs = np.stack([scenario.synth_image("noise" if kind == 0 else "smooth", seed) for (_, _, _, kind, seed) in meta])
    recs = [(i, 0, r) for i, r in enumerate(_crop_records(meta))]
    params = torch.from_numpy(pack_crop_params(recs)).cuda()
    src = torch.from_numpy(imgs).cuda()
    out = crop_resample_cuda(src, 480 * 636, 0, 636, params, (96, 96), out_dtype, map_mode)
    return (out.float() if out_dtype == torch.bfloat16 else out).cpu().numpy()


def test_crop_fp64_map_is_bit_exact_vs_reference_golden(cuda_lib, gold):
    """tests/golden/warp.npz holds the reference tracker._warp_image (numpy fp64 + cv2.remap)
    outputs; the fp64-map mode must reproduce every pixel."""
    W = gold["warp"]
    out = _run_crops(W["meta"], "fp64")
    mism = int((out != W["crops"]).sum())
    assert mism == 0, f"{mism} of {out.size} pixels differ"


def test_crop_fp32_lattice_map_is_bit_exact_vs_reference_golden(cuda_lib, gold):
    """The fp32 lattice-map mode (the one bench.py times) re-evaluates every pixel whose interpolated
    coordinate lies next to a 1/32-px rounding boundary through the float64 map: all 52 golden crops,
    noise images included, must be identical to tracker._warp_image (north_star: crops within 1/255 per
    pixel - here 0), in every output dtype."""
    W = gold["warp"]
    out = _run_crops(W["meta"], "fp32")
    mism = int((out != W["crops"]).sum())
    assert mism == 0, f"{mism} of {out.size} pixels differ"
    f = _run_crops(W["meta"][:6], "fp32", torch.float32)
    assert np.array_equal(f, W["crops"][:6].astype(np.float32) / 255.0)


def test_crop_fp32_lattice_map_is_bit_exact_on_every_golden_frame(cuda_lib, gold):
    """Wider sweep than warp.npz: every crop of the 123 geometry frames (up to 4 per frame) on a noise and a
    smooth multi-view frame, fp32 lattice mode == fp64 mode (itself pinned bit-exact to the reference above)."""
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda, pack_crop_params

    G = gold["geometry"]
    K = G["crop_cam"].shape[0]
    meta = [(k, h, s, 0, 0) for k in range(K) for h in range(2) for s in range(2) if G["crop_cam"][k, h, s] >= 0]
    recs = [(i % 2, int(G["crop_cam"][m[0], m[1], m[2]]), r) for i, (m, r) in enumerate(zip(meta, _crop_records(np.array(meta))))]
    frames = torch.from_numpy(np.stack([scenario.synth_frame("noise", 31), scenario.synth_frame("smooth", 32)])).cuda()
    params = torch.from_numpy(pack_crop_params(recs)).cuda()
    args = (frames, 480 * 2544, 636, 2544, params, (96, 96), torch.uint8)
    a = crop_resample_cuda(*args, "fp32")
    b = crop_resample_cuda(*args, "fp64")
    mism = int((a != b).sum())
    assert mism == 0, f"{mism} of {a.numel()} pixels differ over {len(meta)} crops"


def test_crop_float_output_is_level_over_255(cuda_lib, gold):
    W = gold["warp"]
    meta = W["meta"][:6]
    f = _run_crops(meta, "fp64", torch.float32)
    assert np.array_equal(f, W["crops"][:6].astype(np.float32) / 255.0)
    b = _run_crops(meta, "fp64", torch.bfloat16)
    assert np.array_equal(b, torch.from_numpy(f).bfloat16().float().numpy())


def test_crop_edge_cases(cuda_lib):
    """Source behind the camera (z<0 -> (-1,-1) -> 0), fully outside, strided multi-view source."""
    from absolutetrack_b200.lib.common import camera as C
    from absolutetrack_b200.lib.tracker.tracker import _warp_image, crop_resample_cuda, pack_crop_params

    G = scenario.geometry()
    k, h, s = 5, 0, 0
    cam = int(G["crop_cam"][k, h, s])
    src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
    c2w = G["crop_c2w"][k, h, s].copy()
    back = c2w.copy()
    back[:3, :3] = c2w[:3, :3] @ np.diag([1.0, -1.0, -1.0])  # look the other way
    dst_back = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [], back)
    img = scenario.synth_image("noise", 3)
    sd = lambda cam_: dict(f=cam_.f, c=cam_.c, c2w=cam_.camera_to_world_xf, coeffs=tuple(src.distort), width=cam_.width, height=cam_.height)  # noqa: E731
    out = _warp_image(src, dst_back, img)
    assert np.array_equal(out, O.warp_image(sd(src), sd(dst_back), img))
    # wide crop (small focal) reaching beyond the source image borders
    dst_wide = C.PinholePlaneCameraModel(96, 96, (20.0, 20.0), G["crop_c"][k, h, s], [], c2w)
    out = _warp_image(src, dst_wide, img)
    assert np.array_equal(out, O.warp_image(sd(src), sd(dst_wide), img))
    # the recording layout [480, 4, 636] addressed with strides, 2 frames x 4 views
    frames = np.stack([scenario.synth_frame("noise", 11), scenario.synth_frame("smooth", 12)])
    dst = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [], c2w)
    recs = [(f, v, C.crop_kernel_params(src, dst)) for f in range(2) for v in range(4)]
    out = crop_resample_cuda(torch.from_numpy(frames).cuda(), 480 * 2544, 636, 2544,
                             torch.from_numpy(pack_crop_params(recs)).cuda(), (96, 96), torch.uint8, "fp64").cpu().numpy()
    for i, (f, v, _) in enumerate(recs):
        assert np.array_equal(out[i], O.warp_image(sd(src), sd(dst), np.ascontiguousarray(frames[f, :, v, :])))
    # empty batch
    e = crop_resample_cuda(torch.from_numpy(frames).cuda(), 0, 0, 636, torch.zeros((0, 256), dtype=torch.uint8).cuda())
    assert e.shape == (0, 96, 96)


def test_crop_staged_path_equals_direct_gather_and_zero_copy(cuda_lib, gold):
    """AT_MAP_FP32: the shared-memory-staged kernel (bounding box from the boundary pixels) must
    reproduce the direct global-memory gather bit for bit - on real geometry, on a crop too wide
    to stage, on a crop that looks away from the camera - and give the same crops when the source
    frames stay in PINNED HOST memory (zero-copy ingest)."""
    from absolutetrack_b200.lib.common import camera as C
    from absolutetrack_b200.lib.tracker.tracker import crop_resample_cuda, pack_crop_params

    G = scenario.geometry()
    frames = np.stack([scenario.synth_frame("noise", 21), scenario.synth_frame("smooth", 22)])
    recs = []
    for k in (0, 5, 40, 77, 120):
        for h in range(2):
            for s_ in range(2):
                cam = int(G["crop_cam"][k, h, s_])
                if cam < 0:
                    continue
                src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
                dst = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s_], G["crop_c"][k, h, s_], [], G["crop_c2w"][k, h, s_])
                recs.append((k % 2, cam, C.crop_kernel_params(src, dst)))
    k, h, s_ = 5, 0, 0
    cam = int(G["crop_cam"][k, h, s_])
    src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
    c2w = G["crop_c2w"][k, h, s_]
    back = c2w.copy()
    back[:3, :3] = c2w[:3, :3] @ np.diag([1.0, -1.0, -1.0])
    for dst in (C.PinholePlaneCameraModel(96, 96, (20.0, 20.0), G["crop_c"][k, h, s_], [], c2w),        # wider than the image
                C.PinholePlaneCameraModel(96, 96, (45.0, 45.0), G["crop_c"][k, h, s_], [], c2w),        # box too big to stage
                C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s_], G["crop_c"][k, h, s_], [], back),   # behind the camera
                C.PinholePlaneCameraModel(96, 96, (300.0, 300.0), (4000.0, 48.0), [], c2w)):            # far outside the image
        recs.append((0, cam, C.crop_kernel_params(src, dst)))
    params = torch.from_numpy(pack_crop_params(recs)).cuda()
    host = torch.from_numpy(frames).pin_memory()
    dev = host.cuda()
    args = (480 * 2544, 636, 2544, params, (96, 96), torch.uint8, "fp32")
    counter = torch.zeros(1, dtype=torch.int64, device="cuda")
    staged = crop_resample_cuda(dev, *args, staged_bytes=counter)   # a byte counter selects the staged kernel
    zero_copy = crop_resample_cuda(host, *args)                      # so does a pinned host source
    direct = crop_resample_cuda(dev, *args)                          # device source: direct gather
    torch.cuda.synchronize()
    assert torch.equal(staged, direct)
    assert torch.equal(zero_copy, direct)
    assert direct.float().std() > 1
    per_crop = counter.item() / len(recs)
    assert 8_000 < per_crop < 80_000, per_crop                     # ~30 KB boxes, not 305 KB views
    with pytest.raises(Exception):
        crop_resample_cuda(torch.from_numpy(frames), *args)          # pageable host memory is refused

    # source-box pull: only the boxes are copied into a poisoned device buffer of the same layout;
    # the ordinary resampler on that buffer must still give the same crops, in both map modes
    from absolutetrack_b200.lib.tracker.tracker import crop_pull_source_boxes_cuda
    for ctas in (0, 3):
        pulled = torch.full_like(dev, 0xA5)
        counter.zero_()
        crop_pull_source_boxes_cuda(host, pulled, 480 * 2544, 636, 2544, params, (96, 96), ctas, counter)
        assert torch.equal(crop_resample_cuda(pulled, *args), direct)
        args64 = args[:-1] + ("fp64",)
        assert torch.equal(crop_resample_cuda(pulled, *args64), crop_resample_cuda(dev, *args64))
        per_crop = counter.item() / len(recs)
        assert 8_000 < per_crop < 120_000, per_crop                 # two of the 24 crops pull a whole 305 KB view
        assert (pulled != dev).float().mean() > 0.5                 # most of the frame was never copied
    with pytest.raises(Exception):
        crop_pull_source_boxes_cuda(torch.from_numpy(frames), pulled, 480 * 2544, 636, 2544, params)


# ------------------------------------------------------------------ homography resampler (a12)
def test_homography_resampler_matches_oracle(cuda_lib):
    from absolutetrack_b200 import _cabi

    rng = np.random.default_rng(1)
    n, H, W = 5, 120, 160
    imgs = (rng.random((n, H, W), dtype=np.float32) * 255).astype(np.float32)
    xf = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    xf[:, 0, 0] = rng.uniform(0.8, 1.6, n); xf[:, 1, 1] = rng.uniform(0.8, 1.6, n)
    xf[:, 0, 1] = rng.normal(0, 0.05, n); xf[:, 1, 0] = rng.normal(0, 0.05, n)
    xf[:, 0, 2] = rng.normal(0, 20, n); xf[:, 1, 2] = rng.normal(0, 20, n)
    xf[:, 2, 0] = rng.normal(0, 1e-4, n); xf[:, 0, 3] = rng.normal(0, 2, n)
    xf[0] = np.eye(4)  # identity: integer grid, exercises the x == W-1 exclusion
    ref = O.resample_homography(imgs, xf, (96, 96)) / np.float32(255)
    d_img, d_xf = torch.from_numpy(imgs).cuda(), torch.from_numpy(xf).cuda()
    out = torch.empty(n, 96, 96, device="cuda")
    rc = cuda_lib.at_resample_homography_f32(d_img.data_ptr(), n, H, W, d_xf.data_ptr(), 96, 96, out.data_ptr(), 255.0,
                                             _cabi.stream_ptr())
    _cabi.check(rc)
    assert np.abs(out.cpu().numpy() - ref).max() <= 1e-6
    assert (ref != 0).mean() > 0.5


# ------------------------------------------------------------------ evaluation sums
def test_eval_sums_match_oracle_mpjpe(cuda_lib):
    from absolutetrack_b200.distributed import accumulate_metric_sums_cuda

    g = torch.Generator().manual_seed(5)
    n = 3000
    lm = torch.randn(n, 21, 3, generator=g) * 50
    gt = lm + torch.randn(n, 21, 3, generator=g)
    ja = torch.randn(n, 22, generator=g)
    gja = ja + 0.01 * torch.randn(n, 22, generator=g)
    valid = torch.rand(n, generator=g) > 0.2
    sums = torch.zeros(3, dtype=torch.float64, device="cuda")
    for lo, hi in ((0, 1000), (1000, n)):   # accumulates across calls
        accumulate_metric_sums_cuda(lm[lo:hi].cuda(), gt[lo:hi].cuda(), sums, valid[lo:hi].cuda(), ja[lo:hi].cuda(), gja[lo:hi].cuda())
    s = sums.cpu().numpy()
    assert s[2] == int(valid.sum())
    assert abs(s[0] / s[2] - O.mpjpe(lm.numpy(), gt.numpy(), valid.numpy())) < 1e-4
    dth = np.degrees(np.abs(ja.numpy()[:, :20] - gja.numpy()[:, :20])).mean(-1)[valid.numpy()].mean()
    assert abs(s[1] / s[2] - dth) < 1e-4
    accumulate_metric_sums_cuda(lm[:0].cuda(), gt[:0].cuda(), sums)   # empty batch
    assert sums.cpu().numpy()[2] == s[2]


# ------------------------------------------------------------------ evaluation table (f4)
def test_eval_recording_table_matches_reference_aggregation(cuda_lib, gold, tmp_path):
    """at_eval_recording_f32 against the reference's load_eval / metric_utils run on its own user05
    result pickles: integer counts exact (the per-frame error is bit-identical: thresholds placed ON
    error values still count exactly), sums to 1e-12, the printed table to 1e-10."""
    import pickle

    from absolutetrack_b200.load_eval import EvalSums, aggregate_metrics, compute_metric_sums_cuda

    F, E = gold["fk_user05"], gold["eval_user05"]
    total = EvalSums()
    for rec in ("00", "02", "11"):
        p = f"r{rec}_"
        gt, tr, valid = F[p + "gt_keypoints"], E[p + "tracked_keypoints"], F[p + "valid"]
        one = compute_metric_sums_cuda(gt, tr, valid, EvalSums())
        s, c = one.sums.cpu().numpy(), one.counts.cpu().numpy()
        assert np.array_equal(c[3:], E[p + "pck_counts"])
        assert c[0] == valid.sum() and c[1] == valid.size and c[2] == int(E[p + "n_accel"])
        for got, k in zip(s, ("sum_error", "sum_accel", "sum_gt_accel")):
            assert abs(got - float(E[p + k])) <= 1e-12 * abs(float(E[p + k])), k
        # thresholds equal to actual per-frame errors: exact counts need bit-identical errors
        err = np.linalg.norm(gt.astype(np.float64) - tr.astype(np.float64), axis=-1).mean(axis=-1)[valid]
        thr = np.sort(err)[:: max(1, err.size // 128)][:128]
        on = compute_metric_sums_cuda(gt, tr, valid, EvalSums(thresholds=thr))
        assert np.array_equal(on.counts.cpu().numpy()[3:], [(err <= t).sum() for t in thr])
        compute_metric_sums_cuda(torch.from_numpy(gt).double(), tr, torch.from_numpy(valid), total)  # float64 in: converted
        d = tmp_path / "real" / f"user{rec}"
        d.mkdir(parents=True)
        with open(d / f"recording_{rec}.npy", "wb") as fp:  # run_eval_known_skeleton.py:96-104
            pickle.dump({"tracked_keypoints": tr.astype(np.float64), "gt_keypoints": gt.astype(np.float64), "valid_tracking": valid}, fp)
    for t in (total.table(), aggregate_metrics(str(tmp_path), verbose=False)):
        assert (t["n_valid"], t["n_total"]) == (int(E["n_valid"]), int(E["n_total"]))
        assert abs(t["mean_keypoint_error"] - float(E["mean_error"])) < 1e-10
        assert np.abs(t["pck_curve"] - E["pck_curve"].reshape(-1)).max() < 1e-12
        assert abs(t["auc_score"] - float(E["auc"])) < 1e-10
        assert abs(t["mean_keypoint_acceleration"] - float(E["mean_accel"])) < 1e-10
        assert abs(t["gt_mean_keypoint_acceleration"] - float(E["mean_gt_accel"])) < 1e-10
    assert aggregate_metrics(str(tmp_path / "nothing_here"), verbose=False) is None


def test_eval_recording_edge_cases(cuda_lib):
    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.load_eval import EvalSums, compute_metric_sums_cuda

    rng = np.random.default_rng(3)
    # sequences shorter than a triple, all-invalid hands, triples that straddle an invalid frame or the hand boundary
    for H, T in ((2, 2), (2, 1), (1, 3), (3, 40), (2, 1000)):
        gt = rng.normal(0, 80, (H, T, 21, 3)).astype(np.float32)
        tr = (gt + rng.normal(0, 12, gt.shape)).astype(np.float32)
        valid = rng.random((H, T)) > 0.15
        valid[0, :] = T > 3 and H > 2          # one hand all valid (3 x 40 case) or all invalid
        if T >= 3:
            valid[-1, -3:] = True               # the last triple of the last hand ends exactly at the buffer end
        acc = compute_metric_sums_cuda(gt, tr, valid, EvalSums())
        ref = O.recording_metric_sums(gt, tr, valid)
        c, s = acc.counts.cpu().numpy(), acc.sums.cpu().numpy()
        assert [int(c[0]), int(c[1]), int(c[2])] == [ref["n_valid"], ref["n_total"], ref["n_accel"]], (H, T)
        assert np.array_equal(c[3:], ref["pck_counts"])
        assert np.allclose(s, [ref["sum_error"], ref["sum_accel"], ref["sum_gt_accel"]], rtol=1e-12, atol=1e-9)
    # empty recording: nothing launched, accumulators untouched
    acc = compute_metric_sums_cuda(np.zeros((2, 0, 21, 3)), np.zeros((2, 0, 21, 3)), np.zeros((2, 0), bool), EvalSums())
    assert int(acc.counts.sum()) == 0
    with pytest.raises(ValueError):
        EvalSums(thresholds=np.linspace(0, 1, 129))
    with pytest.raises(ValueError):
        compute_metric_sums_cuda(np.zeros((2, 4, 21, 3)), np.zeros((2, 5, 21, 3)), np.zeros((2, 4), bool), EvalSums())
    z = torch.zeros(8, device="cuda")
    rc = _cabi.lib().at_eval_recording_f32(z.data_ptr(), z.data_ptr(), z.data_ptr(), 1, 1, None, 200, z.data_ptr(), z.data_ptr(), None)
    assert rc != 0 and "128" in _cabi.lib().at_last_error_string().decode()


# ------------------------------------------------------------------ config-3 preprocessing (f3, GPU part)
def test_config3_crop_matrices_and_crops_match_reference_golden(cuda_lib, gold):
    """at_gen_crop_matrices_f32 + at_resample_homography_f32 through the data_transform mirror against
    the reference's _gen_crop_matrices / _perspective_crop_images outputs (float32 round-off: the
    reference inverts float32 matrices with numpy, the kernel works in float64)."""
    from absolutetrack_b200.lib.batched_dataset import data_transform as DT

    G = gold["crop_matrices"]
    ext, intr, pts = G["orig_extrinsics"], G["orig_intrinsics"], G["crop_points"]
    F = ext.shape[0]
    for hand in (0, 1):
        e, K, x = (t.cpu().numpy() for t in DT.gen_crop_matrices_cuda(ext, intr, pts, hand == 1, (96, 96)))
        assert np.abs(e - G[f"h{hand}_extrinsics_xf"]).max() < 5e-7
        assert np.allclose(K, G[f"h{hand}_new_intrinsics"], rtol=2e-6, atol=0)
        assert np.abs(x - G[f"h{hand}_resample_xf"]).max() < 2e-6 * np.abs(x).max()
        o = [O.gen_crop_matrices(ext[f], intr[f], pts[f], hand == 1, (96, 96)) for f in range(F)]
        assert np.abs(x - np.stack([a[2] for a in o])).max() < 2e-6 * np.abs(x).max()
        imgs = np.stack([scenario.synth_image("smooth" if i % 2 == 0 else "noise", 100 * hand + i, 120, 160) for i in range(6)])
        crops, e3, K3 = DT._perspective_crop_images(imgs.astype(np.float32).reshape(3, 2, 120, 160), ext[:3], intr[:3], pts[:3], hand, (96, 96))
        assert crops.is_cuda and tuple(crops.shape) == (3, 2, 96, 96)
        assert np.abs(crops.cpu().numpy() - G[f"h{hand}_crops"]).max() < 2e-3
        assert np.array_equal(e3.cpu().numpy(), e[:3]) and np.array_equal(K3.cpu().numpy(), K[:3])
    # per-frame mirror flags = the two uniform calls interleaved
    flags = np.arange(F) % 2 == 1
    e, K, x = (t.cpu().numpy() for t in DT.gen_crop_matrices_cuda(ext, intr, pts, flags, (96, 96)))
    for hand in (0, 1):
        ref = DT.gen_crop_matrices_cuda(ext, intr, pts, hand == 1, (96, 96))[2].cpu().numpy()
        assert np.array_equal(x[flags == (hand == 1)], ref[flags == (hand == 1)])
    # crop.py:25-26: a point behind the crop camera -> ValueError, as in the reference
    bad = pts.copy()
    bad[7, 3] = -bad[7, 3]
    with pytest.raises(ValueError, match="Unable to create crop camera"):
        DT.gen_crop_matrices_cuda(ext, intr, bad, False, (96, 96))
    # empty sequence / shape errors
    e, K, x = DT.gen_crop_matrices_cuda(ext[:0], intr[:0], pts[:0], False, (96, 96))
    assert tuple(x.shape) == (0, 2, 4, 4)
    with pytest.raises(ValueError):
        DT.gen_crop_matrices_cuda(ext, intr[:, :1], pts, False, (96, 96))


def test_config3_preprocess_from_torch_idx_files_matches_reference(cuda_lib, gold):
    """The committed .torch.idx/.torch.bin fixture pairs -> TorchIdx -> data_transform.preprocess
    (crops on the GPU) against the reference's own preprocess of the same files."""
    import os

    from absolutetrack_b200.lib.batched_dataset import data_transform as DT
    from absolutetrack_b200.lib.data_utils.idxbinfile import TorchIdx

    G = gold["torchdata"]
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "torchdata")
    mono = TorchIdx(os.path.join(d, "seq.mono.torch.idx")).read_bin()
    labels = TorchIdx(os.path.join(d, "seq.labels.torch.idx")).read_bin()
    for i in range(2):
        mi, mt = DT.preprocess({"mono": np.array(mono[i]), "labels": labels[i]}, (96, 96))
        p = f"s{i}_"
        assert mi.left_images.is_cuda and tuple(mi.left_images.shape) == (3, 2, 96, 96)
        assert np.abs(mi.left_images.cpu().numpy() - G[p + "left_images"]).max() < 2e-3
        assert np.allclose(mi.intrinsics.cpu().numpy(), G[p + "intrinsics"], rtol=2e-6, atol=0)
        assert np.abs(mi.extrinsics_xf.cpu().numpy() - G[p + "extrinsics_xf"]).max() < 5e-7
        assert np.array_equal(mi.hand_idx.numpy(), G[p + "hand_idx"])
        for name, pd in (("orig", mi.orig_pose_data), ("solved", mi.s_solved_pose_data)):
            assert np.array_equal(pd.joint_angles.numpy(), G[p + name + "_joint_angles"])
            assert np.array_equal(pd.wrist_xfs.numpy(), G[p + name + "_wrist_xfs"])       # mm -> m on the translations
            assert np.array_equal(pd.left_hand_model.joint_rotation_axes.numpy(), G[p + name + "_axes"])  # mirrored for the right hand
            assert np.array_equal(pd.left_hand_model.joint_rest_positions.numpy(), G[p + name + "_rest"])
            assert np.array_equal(pd.left_hand_model.landmark_rest_positions.numpy(), G[p + name + "_landmarks"])
        assert np.array_equal(np.asarray(mt.gt_skel_targets.skel_scales), G[p + "gt_skel_scales"])
        assert np.array_equal(mt.preds_targets.pinch_prediction.numpy(), G[p + "pinch"])
        assert mt.intrinsics is mi.intrinsics and mt.extrinsics_xf is mi.extrinsics_xf


@pytest.mark.gpu
def test_fk_per_sample_models_warp_kernel_five_weights(cuda_lib, gold):
    """Per-sample hand models run one warp per hand (fk_skin_warp_kernel).  Five skinning weights per landmark
    (the KMAX = 8 instantiation; the recordings use three) with per-sample skeleton scales, right and left hands,
    against the numpy oracle (hand_skinning.py:56-127), and against the shared-model path for the unscaled model."""
    from absolutetrack_b200.lib.common.hand_skinning import skin_landmarks_cuda

    hm = _hand_model_tuple(gold["geometry"], "hm_")
    rng = np.random.default_rng(21)
    n = 777
    w3 = hm.landmark_rest_bone_weights.numpy()
    i3 = hm.landmark_rest_bone_indices.numpy()
    w5 = np.zeros((21, 5), np.float32)
    i5 = np.zeros((21, 5), i3.dtype)
    w5[:, :3], i5[:, :3] = w3, i3
    w5[:, 3] = 0.25 * w5[:, 0]; w5[:, 0] *= 0.75; i5[:, 3] = rng.integers(0, 17, 21)     # a fourth bone takes a share
    w5[::2, 4] = 0.5 * w5[::2, 1]; w5[::2, 1] *= 0.5; i5[:, 4] = rng.integers(0, 17, 21)  # a fifth on every other landmark
    scale = np.linspace(0.8, 1.25, n).astype(np.float32).reshape(n, 1, 1)
    md = {
        "joint_rotation_axes": np.broadcast_to(hm.joint_rotation_axes.numpy(), (n, 22, 3)).copy(),
        "joint_rest_positions": hm.joint_rest_positions.numpy()[None] * scale,
        "landmark_rest_positions": hm.landmark_rest_positions.numpy()[None] * scale,
        "landmark_rest_bone_weights": np.broadcast_to(w5, (n, 21, 5)).copy(),
        "landmark_rest_bone_indices": np.broadcast_to(i5, (n, 21, 5)).copy(),
    }
    ja = rng.uniform(-1.3, 1.3, (n, 22)).astype(np.float32)
    wx = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    wx[:, :3, :3] = O.so3_exp_map_f32(rng.normal(0, 1, (n, 3)).astype(np.float32))
    wx[:, :3, 3] = rng.normal(0, 300, (n, 3))
    ref = O.skin_landmarks(md, ja, wx)
    per = hm._replace(**{k: torch.from_numpy(v) for k, v in md.items()})
    out = skin_landmarks_cuda(per, torch.from_numpy(ja), torch.from_numpy(wx)).cpu().numpy()
    assert np.abs(out - ref).max() < 1e-3
    # three weights, unscaled, mirrored hands: the warp kernel against the shared-model kernel
    hidx = torch.from_numpy(rng.integers(0, 2, n))
    same = hm._replace(**{k: getattr(hm, k).expand((n,) + tuple(getattr(hm, k).shape)).contiguous() for k in
                          ("joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions",
                           "landmark_rest_bone_weights", "landmark_rest_bone_indices")})
    a = skin_landmarks_cuda(same, torch.from_numpy(ja), torch.from_numpy(wx), hand_idx=hidx)
    b = skin_landmarks_cuda(hm, torch.from_numpy(ja), torch.from_numpy(wx), hand_idx=hidx)
    assert float((a - b).abs().max()) < 5e-4
