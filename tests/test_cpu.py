"""CPU suite (`pytest -m "not gpu"`): the oracle against the reference's golden vectors, the host
logic, the C-ABI surface and the world_size-2 sharding path over gloo.  No CUDA calls."""
import ctypes
import os
import re
import sys

import numpy as np
import pytest
import torch

import model_oracle as MO
import np_oracle as O
import scenario

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ------------------------------------------------------------------ C-ABI surface
def _declared_functions():
    src = open(os.path.join(ROOT, "include", "absolutetrack_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(at_[a-z0-9_]+)\s*\(", src)))


def test_cabi_library_exports_every_declared_symbol():
    from absolutetrack_b200 import _cabi

    names = _declared_functions()
    assert len(names) >= 20
    assert os.path.exists(_cabi.LIB_PATH), "run __graft_entry__.build() first"
    raw = ctypes.CDLL(_cabi.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/absolutetrack_b200.h but not exported"
    # the ctypes binding covers the whole header and nothing else
    assert sorted(_cabi.SIGNATURES) == names


def test_every_exported_symbol_is_documented_here():
    """INTEGRATION.md names every entry point that include/absolutetrack_b200.h declares."""
    import re

    hdr = open(os.path.join(ROOT, "include", "absolutetrack_b200.h")).read()
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    syms = sorted(set(re.findall(r"\b(at_[a-z0-9_]+)\s*\(", hdr)))
    assert len(syms) > 25
    missing = [s for s in syms if f"`{s}`" not in doc and f"{s}(" not in doc]
    assert not missing, missing


def test_cabi_host_side_queries_need_no_gpu():
    from absolutetrack_b200 import _cabi

    lib = _cabi.lib()
    assert lib.at_version() >= 100
    table = _cabi.layer_table()
    assert len(table) == 46 and table[0]["name"] == "backbone.stem" and table[28]["name"] == "backbone.proj"
    assert sum(1 for L in table if L["name"].startswith("backbone.")) == 29          # 29 convs, SURVEY 2.1
    flops = sum(2 * L["cin"] * L["cout"] * L["ksize"] ** 2 * hw for L, hw in zip(table[:29], _backbone_hw()))
    assert flops == 969_228_288
    assert lib.at_encoder_workspace_bytes(4, _cabi.AT_PREC_BF16) > 0
    assert lib.at_encoder_chunk_crops() >= 256
    assert C_sizeof_crop_params() == 256


def _backbone_hw():
    hw = [96 * 96]                       # stem conv runs at 96x96
    hw += [48 * 48] * 4                  # layer1
    hw += [24 * 24] * 7                  # layer2 (conv1 s2, conv2, down, 2 blocks)
    hw += [12 * 12] * 11
    hw += [6 * 6] * 5
    hw += [6 * 6]                        # proj
    return hw


def C_sizeof_crop_params():
    from absolutetrack_b200 import _cabi

    return ctypes.sizeof(_cabi.AtCropParams)


def test_product_path_fails_loudly_without_cuda():
    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, model_from_state_dict

    model = model_from_state_dict(make_random_state_dict(0))
    with pytest.raises(_cabi.AbsoluteTrackError):
        model.encode(torch.zeros(1, 96, 96))


# ------------------------------------------------------------------ oracle vs reference goldens
@pytest.mark.parametrize("rec", ["00", "02", "11"])
def test_oracle_fk_matches_reference_golden_keypoints(gold, rec):
    """sample_data/user05/recording_*.npy['gt_keypoints'] (written by the reference's eval script)
    vs the numpy FK restatement, every 4th frame; 1e-3 mm."""
    G = gold["fk_user05"]
    p = f"r{rec}_"
    hm = {k[len(p):]: v for k, v in G.items() if k.startswith(p)}
    n = 0
    for h in range(2):
        idx = np.nonzero(G[p + "valid"][h])[0][::4]
        lm = O.landmarks_from_hand_pose(hm, G[p + "joint_angles"][idx, h], G[p + "wrist_transforms"][idx, h], h)
        d = np.abs(lm - G[p + "gt_keypoints"][h][idx])
        assert d.max() < 1e-3, f"hand {h}: {d.max()} mm"
        n += len(idx)
    assert n > 150


def test_oracle_warp_is_bit_exact_vs_reference_golden(gold):
    """tests/golden/warp.npz = reference tracker._warp_image (numpy fp64 map + cv2.remap)."""
    W = gold["warp"]
    for i in range(0, len(W["meta"]), 5):
        k, h, s, kind, seed = (int(v) for v in W["meta"][i])
        cam = int(scenario.geometry()["crop_cam"][k, h, s])
        img = scenario.synth_image("noise" if kind == 0 else "smooth", seed)
        out = O.warp_image(scenario.src_camera(k, cam), scenario.dst_camera(k, h, s), img)
        assert np.array_equal(out, W["crops"][i]), f"crop {i}"


def test_oracle_remap_matches_installed_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (120, 160), dtype=np.uint8)
    mx = rng.uniform(-3, 163, (96, 96)).astype(np.float32)
    my = rng.uniform(-3, 123, (96, 96)).astype(np.float32)
    mx[:4] = -1.0
    my[:4] = -1.0
    assert np.array_equal(O.remap_bilinear_u8(img, mx, my), cv2.remap(img, mx, my, cv2.INTER_LINEAR))


@pytest.mark.parametrize("head", ["k", "u"])
def test_model_oracle_matches_reference_golden(gold, head):
    """tests/golden/model.npz = outputs of the REFERENCE UmeTrackModel (oracle/make_golden_model.py)."""
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    M = gold["model"]
    sd = make_random_state_dict(0)
    n_frames, start, sv = (int(v) for v in M[f"{head}_case"])
    seq = scenario.model_sequence(n_frames=n_frames, start=start, single_view_frame=None if sv < 0 else sv)
    hm = scenario.hand_model()
    skel = (torch.from_numpy(hm["joint_rotation_axes"]).float(), torch.from_numpy(hm["joint_rest_positions"] * 0.001).float())
    st = MO.TemporalState()
    for t, fi in enumerate(seq):
        args = [torch.from_numpy(fi[n]) for n in ("images", "intrinsics", "extrinsics", "sample_range", "memory_idx", "use_memory", "hand_idx")]
        out = MO.forward(sd, st, *args, skel if head == "k" else None)
        assert np.abs(out["joint_angles"].numpy() - M[f"{head}{t}_joint_angles"]).max() < 2e-5
        assert np.abs(out["wrist_xfs"].numpy() - M[f"{head}{t}_wrist_xfs"]).max() < 2e-6
        assert np.abs(out["sigmas"].numpy() - M[f"{head}{t}_sigmas"]).max() < 2e-5


def cropcam_case(C, i):
    """Inputs of golden row i of crop_cameras.npz for the oracle / the kernel."""
    ri = int(C["rec"][i])
    pre = f"r{ri}_"
    cams = [dict(c2w=C["c2w"][i, v], f=C[pre + "cam_f"][v], c=C[pre + "cam_c"][v], coeffs=C[pre + "cam_coeffs"][v],
                 width=int(C[pre + "cam_wh"][v, 0]), height=int(C[pre + "cam_wh"][v, 1])) for v in range(4)]
    model = {k: C[pre + k] for k in ("joint_rotation_axes", "joint_rest_positions", "landmark_rest_positions",
                                     "landmark_rest_bone_weights", "landmark_rest_bone_indices", "joint_limits")}
    ncp, minvis, maxv, sort_idx, fmul, _ = C["variants"][int(C["variant"][i])]
    opts = dict(num_crop_points=int(ncp), min_required_vis_landmarks=int(minvis), max_view_num=int(maxv),
                sort_camera_index=bool(sort_idx), focal_multiplier=float(fmul))
    return cams, C[pre + "camera_angles"], model, opts


def test_oracle_crop_cameras_match_reference_golden(gold):
    """gen_crop_cameras_from_pose restatement vs the reference's own outputs on 770 poses of the four
    sample recordings x 5 option variants: selection exact; cameras to the round-off of the float32
    FK landmarks they are fitted to (1e-4 mm on ~300 mm: 1e-6 relative)."""
    C = gold["crop_cameras"]
    for i in range(len(C["rec"])):
        cams, angles, model, opts = cropcam_case(C, i)
        got = O.gen_crop_cameras_from_pose(cams, angles, model, C["joint_angles"][i], C["wrist"][i], int(C["hand"][i]), **opts)
        assert [g[0] for g in got] == [int(c) for c in C["sel_cam"][i] if c >= 0], i
        for s, (_, f, c, c2w) in enumerate(got):
            assert np.allclose(f, C["sel_f"][i, s], rtol=3e-6)
            assert np.allclose(c, C["sel_c"][i, s])
            assert np.abs(c2w[:3, :3] - C["sel_c2w"][i, s][:3, :3]).max() < 3e-6
            assert np.abs(c2w[:3, 3] - C["sel_c2w"][i, s][:3, 3]).max() < 1e-9


def test_oracle_metrics():
    rng = np.random.default_rng(0)
    gt = rng.normal(size=(2, 10, 21, 3))
    tr = gt + np.array([3.0, 0.0, 4.0])
    valid = np.ones((2, 10), bool)
    valid[1, 3] = False
    assert abs(O.mpjpe(tr, gt, valid) - 5.0) < 1e-12
    lin = np.arange(10)[None, :, None, None] * np.ones((2, 1, 21, 3))
    assert O.mean_keypoint_accel(lin, np.ones((2, 10), bool)) < 1e-12      # constant velocity


def _oracle_eval_arrays(F, E, recs):
    """Oracle sums of the given user05 recordings in the EvalSums layout (f64 [3], i64 [3 + 101])."""
    sums, counts = np.zeros(3), np.zeros(3 + 101, np.int64)
    for rec in recs:
        p = f"r{rec}_"
        s = O.recording_metric_sums(F[p + "gt_keypoints"], E[p + "tracked_keypoints"], F[p + "valid"])
        sums += [s["sum_error"], s["sum_accel"], s["sum_gt_accel"]]
        counts[:3] += [s["n_valid"], s["n_total"], s["n_accel"]]
        counts[3:] += s["pck_counts"]
    return sums, counts


def _check_eval_table(t, E):
    assert (t["n_valid"], t["n_total"]) == (int(E["n_valid"]), int(E["n_total"]))
    assert abs(t["mean_keypoint_error"] - float(E["mean_error"])) < 1e-11
    assert np.abs(t["pck_curve"] - E["pck_curve"].reshape(-1)).max() < 1e-12
    assert abs(t["auc_score"] - float(E["auc"])) < 1e-11
    assert abs(t["mean_keypoint_acceleration"] - float(E["mean_accel"])) < 1e-11
    assert abs(t["gt_mean_keypoint_acceleration"] - float(E["mean_gt_accel"])) < 1e-11


def test_oracle_eval_table_matches_reference_aggregation(gold):
    """The reference's load_eval / metric_utils run on its own user05 result pickles: per-recording
    sums and PCK counts exact, the combined table (success rate, error, PCK, AUC, accelerations)."""
    from absolutetrack_b200.lib.common import metric_utils as M
    from absolutetrack_b200.load_eval import table_from_sums

    F, E = gold["fk_user05"], gold["eval_user05"]
    assert np.array_equal(O.PCK_THRESHOLDS, E["thresholds"]) and np.array_equal(M.PCK_THRESHOLDS, E["thresholds"])
    tot = None
    for rec in ("00", "02", "11"):
        p = f"r{rec}_"
        s = O.recording_metric_sums(F[p + "gt_keypoints"], E[p + "tracked_keypoints"], F[p + "valid"])
        assert np.array_equal(s["pck_counts"], E[p + "pck_counts"]) and s["n_accel"] == int(E[p + "n_accel"])
        for k in ("sum_error", "sum_accel", "sum_gt_accel"):
            assert abs(s[k] - float(E[p + k])) <= 1e-12 * abs(float(E[p + k]))
        tot = s if tot is None else {k: tot[k] + s[k] for k in s}
    t = O.eval_table(tot)
    assert abs(t["mean_error"] - float(E["mean_error"])) < 1e-11 and abs(t["auc"] - float(E["auc"])) < 1e-11
    assert np.abs(t["pck_curve"] - E["pck_curve"].reshape(-1)).max() < 1e-12
    # the host arithmetic of the product (counts -> PCK -> AUC) on the same sums
    _check_eval_table(table_from_sums(*_oracle_eval_arrays(F, E, ("00", "02", "11")), M.PCK_THRESHOLDS), E)
    # short sequences: no triples, no valid frames
    z = O.recording_metric_sums(np.zeros((2, 2, 21, 3)), np.ones((2, 2, 21, 3)), np.zeros((2, 2), bool))
    assert z["n_accel"] == 0 and z["n_valid"] == 0 and z["n_total"] == 4
    e = table_from_sums(np.zeros(3), np.zeros(104, np.int64), M.PCK_THRESHOLDS)
    assert e["success_rate"] == 0.0 and np.isnan(e["mean_keypoint_error"]) and e["auc_score"] == 0.0


def test_oracle_config3_crop_matrices_and_crops_match_reference_golden(gold):
    """data_transform._gen_crop_matrices / _perspective_crop_images of the reference run here on 40
    seeded two-view frames.  The reference mixes float32 and float64 (float32 numpy inverses of
    float32 matrices); the restatement is float64 -> agreement to float32 round-off."""
    G = gold["crop_matrices"]
    ext, intr, pts = G["orig_extrinsics"], G["orig_intrinsics"], G["crop_points"]
    for hand in (0, 1):
        m = [O.gen_crop_matrices(ext[f], intr[f], pts[f], hand == 1, (96, 96)) for f in range(ext.shape[0])]
        e, K, x = (np.stack([a[i] for a in m]) for i in range(3))
        assert np.abs(e - G[f"h{hand}_extrinsics_xf"]).max() < 5e-7
        assert np.allclose(K, G[f"h{hand}_new_intrinsics"], rtol=2e-6, atol=0)
        assert np.abs(x - G[f"h{hand}_resample_xf"]).max() < 2e-6 * np.abs(x).max()
        imgs = np.stack([scenario.synth_image("smooth" if i % 2 == 0 else "noise", 100 * hand + i, 120, 160) for i in range(6)])
        crops, e3, K3 = O.perspective_crop_images(imgs.astype(np.float32).reshape(3, 2, 120, 160), ext[:3], intr[:3], pts[:3], hand, (96, 96))
        d = np.abs(crops - G[f"h{hand}_crops"])
        assert d.max() < 2e-3 and (G[f"h{hand}_crops"] > 0).mean() > 0.3   # 1e-4 px of map round-off on noise images
    bad = pts[0].copy()
    bad[3] = -bad[3]
    with pytest.raises(ValueError):
        O.gen_crop_matrices(ext[0], intr[0], bad, False, (96, 96))


def test_torch_idx_reader_matches_reference_decoding(gold, tmp_path):
    """lib/data_utils/idxbinfile.TorchIdx on the committed fixture pairs against what the reference's
    TorchIdx made of the same files (tests/golden/torchdata.npz), a write/read round trip over dtypes
    and ragged frames, and the reference's error cases."""
    from absolutetrack_b200.lib.batched_dataset.sample import parse_raw_buffers
    from absolutetrack_b200.lib.data_utils.idxbinfile import MAGIC, TorchIdx, write_torch_idx_bin

    G = gold["torchdata"]
    d = os.path.join(ROOT, "tests", "golden", "torchdata")
    im, il = TorchIdx(os.path.join(d, "seq.mono.torch.idx")), TorchIdx(os.path.join(d, "seq.labels.torch.idx"))
    assert im.bin_path.endswith("seq.mono.torch.bin")
    assert im.is_uniform == bool(G["mono_is_uniform"]) and il.is_uniform == bool(G["labels_is_uniform"])
    assert tuple(im.shape) == tuple(G["mono_shape"]) and im.dtype == np.uint8 and il.shape is None
    assert im.frame_size_bytes() == int(G["mono_frame_bytes"]) and im.item_shape() == tuple(G["mono_shape"][1:])
    assert np.array_equal(np.asarray(il.byte_offsets(0, len(il) + 1), np.int64), G["labels_byte_offsets"])
    mono, labels = im.read_bin(), il.read_bin()
    assert len(im) == len(il) == len(labels) == 2
    bm = open(im.bin_path, "rb").read()
    for i in range(2):
        assert int(mono[i].astype(np.int64).sum()) == int(G[f"s{i}_mono_sum"])
        assert np.array_equal(im.view_buffer_at(i, bm), mono[i])
        s = parse_raw_buffers(np.array(mono[i]), labels[i])
        assert s.images.shape == (3, 2, 120, 160) and s.extrinsics.dtype == np.float32 and s.hand.tolist() == [float(i)] * 3
        assert np.array_equal(s.hand, G[f"s{i}_hand_idx"])
        assert tuple(s.hand_model.joint_rest_positions.shape) == (22, 3) and s.generic_hand_model.joint_limits is not None
    with pytest.raises(ValueError):
        il.frame_size_bytes()
    with pytest.raises(ValueError):
        im.view_buffer(bm[:-1])
    # round trip: dtypes, ragged frames, msgpack objects
    rng = np.random.default_rng(2)
    for name in ("uint8", "int8", "int16", "int32", "int64", "float32", "float64"):
        fr = [(rng.normal(0, 50, (4, 5))).astype(name) for _ in range(3)]
        p = str(tmp_path / f"u_{name}.torch.idx")
        write_torch_idx_bin(p, fr)
        t = TorchIdx(p)
        assert t.is_uniform and t.shape == (3, 4, 5) and t.dtype == np.dtype(name) and np.array_equal(t.read_bin(), np.stack(fr))
        assert t.byte_offsets_array is None and t.byte_offset(-1) == t.data_size_bytes() == 3 * 20 * np.dtype(name).itemsize
    ragged = [np.zeros((2, 3), np.float32), np.ones((4,), np.float32), np.full((1, 2, 2), 7, np.float32)]
    p = str(tmp_path / "r.torch.idx")
    write_torch_idx_bin(p, ragged)
    t = TorchIdx(p)
    assert not t.is_uniform and t.shape is None and [a.shape for a in t.read_bin()] == [(2, 3), (4,), (1, 2, 2)]
    assert t.item_shape(2) == (1, 2, 2) and np.array_equal(t.read_bin()[-1], ragged[2])
    with pytest.raises(ValueError):
        t.item_shape()
    # header errors (idxbinfile.py:137-157,329-347)
    good = np.frombuffer(open(p, "rb").read(), np.int64).copy()
    for pos, val, exc in ((0, 1, ValueError), (1, 2, ValueError), (2, 99, KeyError), (3, 8, ValueError)):
        bad = good.copy()
        bad[pos] = val
        with pytest.raises(exc):
            TorchIdx(p, buffer=bad.tobytes())
    with pytest.raises(ValueError):
        TorchIdx(p, buffer=good.tobytes()[:-3])
    v0 = good.copy()
    v0[0], v0[1] = 0, 0                      # version 0 files carry no magic
    assert len(TorchIdx(p, buffer=v0.tobytes())) == 3 and good[0] == MAGIC


def test_config3_preprocess_host_side_matches_reference(gold, monkeypatch):
    """Everything of data_transform.preprocess except the GPU crop stage (stubbed here; the GPU test
    covers it): unit change, left-hand skeletons, pose tensors, targets - exact against the
    reference's preprocess of the same fixture files."""
    from absolutetrack_b200.lib.batched_dataset import data_transform as DT
    from absolutetrack_b200.lib.data_utils.idxbinfile import TorchIdx

    G = gold["torchdata"]
    d = os.path.join(ROOT, "tests", "golden", "torchdata")
    mono = TorchIdx(os.path.join(d, "seq.mono.torch.idx")).read_bin()
    labels = TorchIdx(os.path.join(d, "seq.labels.torch.idx")).read_bin()
    seen = {}

    def fake_crops(images, ext, intr, pts, hand_idx, crop_size):
        seen.update(ext=np.array(ext), pts=np.array(pts), hand=hand_idx, images=images)
        return [torch.zeros(3, 2, 96, 96), torch.zeros(3, 2, 4, 4), torch.zeros(3, 2, 3, 3)]

    monkeypatch.setattr(DT, "_perspective_crop_images", fake_crops)
    for i in range(2):
        mi, mt = DT.preprocess({"mono": np.array(mono[i]), "labels": labels[i]}, (96, 96))
        p = f"s{i}_"
        assert seen["hand"] == i and seen["images"].shape == (3, 2, 120, 160)
        assert np.abs(seen["ext"][..., :3, 3]).max() < 1.0 and np.abs(seen["pts"]).max() < 1.0   # metres now
        assert np.array_equal(mi.hand_idx.numpy(), G[p + "hand_idx"])
        for name, pd in (("orig", mi.orig_pose_data), ("solved", mi.s_solved_pose_data)):
            assert np.array_equal(pd.joint_angles.numpy(), G[p + name + "_joint_angles"])
            assert np.array_equal(pd.wrist_xfs.numpy(), G[p + name + "_wrist_xfs"])
            assert np.array_equal(pd.left_hand_model.joint_rotation_axes.numpy(), G[p + name + "_axes"])
            assert np.array_equal(pd.left_hand_model.joint_rest_positions.numpy(), G[p + name + "_rest"])
            assert np.array_equal(pd.left_hand_model.landmark_rest_positions.numpy(), G[p + name + "_landmarks"])
        assert np.array_equal(np.asarray(mt.gt_skel_targets.skel_scales), G[p + "gt_skel_scales"])
        assert np.array_equal(mt.preds_targets.pinch_prediction.numpy(), G[p + "pinch"])
        assert mt.intrinsics is mi.intrinsics and mt.extrinsics_xf is mi.extrinsics_xf
    # the per-rank iterator over a file pair: rank striding of the reference Sampler, spans read one by one
    paths = (os.path.join(d, "seq.mono.torch.idx"), os.path.join(d, "seq.labels.torch.idx"))
    got = {r: [(i, float(mi.hand_idx[0])) for i, mi, _ in DT.iter_preprocessed_sequences(*paths, rank=r, world=2)] for r in (0, 1)}
    assert got == {0: [(0, 0.0)], 1: [(1, 1.0)]}
    assert [i for i, _, _ in DT.iter_preprocessed_sequences(*paths)] == [0, 1]


# ------------------------------------------------------------------ host logic
def test_state_dict_spec_is_the_reference_key_set():
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict, pack_state_dict, state_dict_spec

    spec = state_dict_spec()
    assert len(spec) == 252
    sd = make_random_state_dict(1)
    assert sum(v.numel() for k, v in sd.items() if not k.endswith("num_batches_tracked")) == 4_259_371
    bad = dict(sd)
    bad.pop("_temporal._temporal_module.0.bias")
    with pytest.raises(RuntimeError, match="state_dict"):
        pack_state_dict(bad)
    bad = dict(sd)
    bad["extra.weight"] = torch.zeros(1)
    with pytest.raises(RuntimeError, match="unexpected"):
        pack_state_dict(bad)


def test_bn_folding_matches_conv_plus_batchnorm():
    """pack_state_dict folds eval-mode BN (eps 1e-5) into [kh][kw][Cin][CoutP] + bias."""
    import torch.nn.functional as F

    from absolutetrack_b200 import _cabi
    from absolutetrack_b200.lib.models.model_loader import _layer_keys, make_random_state_dict, pack_state_dict

    sd = make_random_state_dict(3)
    blob = pack_state_dict(sd)
    g = torch.Generator().manual_seed(0)
    for lid in (0, 2, 7, 13, 28, 30, 37):
        L = _cabi.layer_table()[lid]
        conv, bn = _layer_keys(L["name"])
        k, cin, cout = L["ksize"], L["cin"], L["cout"]
        x = torch.randn(2, cin, 8, 8, generator=g)
        ref = F.conv2d(x, sd[conv + ".weight"], sd.get(conv + ".bias"), 1, k // 2)
        if bn:
            ref = F.batch_norm(ref, sd[bn + ".running_mean"], sd[bn + ".running_var"], sd[bn + ".weight"], sd[bn + ".bias"], False, 0.0, 1e-5)
        coutp = (cout + 3) // 4 * 4
        w = blob[L["weight_offset"]: L["weight_offset"] + k * k * cin * coutp].reshape(k, k, cin, coutp)[..., :cout]
        b = blob[L["bias_offset"]: L["bias_offset"] + cout]
        got = F.conv2d(x, w.permute(3, 2, 0, 1).contiguous(), b, 1, k // 2)
        assert (got - ref).abs().max() < 2e-5 * max(1.0, ref.abs().max().item()), L["name"]


def test_camera_mirror_matches_oracle_projection():
    from absolutetrack_b200.lib.common import camera as C

    G = scenario.geometry()
    k, cam = 7, 2
    src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
    rng = np.random.default_rng(2)
    pts_eye = np.concatenate([rng.uniform(-200, 200, (50, 2)), rng.uniform(150, 600, (50, 1))], axis=1)
    got = src.eye_to_window(pts_eye)
    ref = O.fisheye62_eye_to_window(pts_eye, G["cam_f"][cam], G["cam_c"][cam], tuple(G["cam_coeffs"][cam]))
    assert np.abs(got - ref).max() < 1e-9
    world = src.eye_to_world(pts_eye)
    assert np.abs(src.world_to_eye(world) - pts_eye).max() < 1e-3         # camera.py:36-38 contract (float32 calibration: R is orthonormal to ~1e-7)


def test_crop_kernel_params_precompose_the_rotation_in_fp64():
    from absolutetrack_b200.lib.common import camera as C
    from absolutetrack_b200.lib.tracker.tracker import pack_crop_params

    G = scenario.geometry()
    k, h, s = 3, 1, 0
    cam = int(G["crop_cam"][k, h, s])
    src = C.Fisheye62CameraModel(636, 480, G["cam_f"][cam], G["cam_c"][cam], G["cam_coeffs"][cam], G["camera_to_world"][k, cam])
    dst = C.PinholePlaneCameraModel(96, 96, G["crop_f"][k, h, s], G["crop_c"][k, h, s], [], G["crop_c2w"][k, h, s])
    rec = C.crop_kernel_params(src, dst)
    buf = pack_crop_params([(5, 2, rec)])
    assert buf.shape == (1, 256) and buf.dtype == np.uint8
    from absolutetrack_b200 import _cabi

    P = _cabi.AtCropParams.from_buffer_copy(buf.tobytes())
    assert (P.frame, P.view, P.src_w, P.src_h) == (5, 2, 636, 480)
    M = np.array(P.M).reshape(3, 3)
    Rs, Rd = G["camera_to_world"][k, cam][:3, :3], G["crop_c2w"][k, h, s][:3, :3]
    assert np.abs(M - Rs.T @ Rd).max() < 1e-12
    assert abs(abs(np.linalg.det(M)) - 1) < 1e-5                              # right hand: det -1 (mirrorx); float32 calibration
    assert np.abs(np.array(P.d)).max() < 1e-3                                 # same optical centre, mm


def test_bundles_collate_and_to_device_keep_dataclasses():
    from absolutetrack_b200.lib.data_utils import bundles
    from absolutetrack_b200.lib.models.umetrack_model import InputFrameData

    a = InputFrameData(torch.zeros(2, 96, 96), torch.zeros(2, 3, 3), torch.zeros(2, 4, 4))
    b = bundles.to_device(a, torch.device("cpu"))
    assert isinstance(b, InputFrameData) and b.left_images.shape == (2, 96, 96)
    c = bundles.collate([a, a])
    assert isinstance(c, InputFrameData) and c.left_images.shape == (2, 2, 96, 96)


def test_hand_model_mirror_and_scale():
    from absolutetrack_b200.lib.common import hand as H
    from test_gpu_kernels import _hand_model_tuple

    hm = _hand_model_tuple(scenario.geometry(), "hm_")
    sc = H.scaled_hand_model(hm, 0.001)
    assert torch.allclose(sc.joint_rest_positions, hm.joint_rest_positions * 0.001)
    assert torch.equal(sc.joint_rotation_axes, hm.joint_rotation_axes)
    assert H.NUM_HANDS == 2 and H.NUM_LANDMARKS_PER_HAND == 21


# ------------------------------------------------------------------ multi-process sharding (gloo, world_size 2)
def _gloo_worker(rank, world, port, out_dir):
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    from absolutetrack_b200 import distributed as D

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n_seq = 11
    mine = list(D.shard_indices(n_seq, rank, world))
    rng = np.random.default_rng(0)
    gt = torch.from_numpy(rng.normal(size=(n_seq, 21, 3)))
    tracked = gt + torch.from_numpy(rng.normal(size=(n_seq, 21, 3)))
    valid = torch.from_numpy(rng.random(n_seq) > 0.2)
    sums = torch.zeros(2, dtype=torch.float64)
    D.mpjpe_accumulate(sums, tracked[mine], gt[mine], valid[mine])
    D.all_reduce_metric_sums(sums)
    full = torch.zeros(2, dtype=torch.float64)
    D.mpjpe_accumulate(full, tracked, gt, valid)
    # evaluation table: the three user05 recordings strided over the ranks, float64 sums + int64 counts
    from absolutetrack_b200.lib.common import metric_utils as M
    from absolutetrack_b200.load_eval import all_reduce_eval_sums, table_from_sums

    F = dict(np.load(os.path.join(ROOT, "tests", "golden", "fk_user05.npz")))
    E = dict(np.load(os.path.join(ROOT, "tests", "golden", "eval_user05.npz")))
    es, ec = (torch.from_numpy(a) for a in _oracle_eval_arrays(F, E, D.shard(["00", "02", "11"], rank, world)))
    all_reduce_eval_sums(es, ec)
    table = table_from_sums(es.numpy(), ec.numpy(), M.PCK_THRESHOLDS)
    torch.save({"mine": mine, "sums": sums, "full": full, "table": table}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_sequences_shard_across_two_ranks_and_metric_sums_all_reduce(tmp_path):
    import socket

    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(tmp_path / f"r{r}.pt", weights_only=False) for r in range(2))
    assert sorted(r0["mine"] + r1["mine"]) == list(range(11)) and not set(r0["mine"]) & set(r1["mine"])
    assert r0["mine"] == list(range(0, 11, 2))                            # reference Sampler striding
    for r in (r0, r1):
        assert torch.allclose(r["sums"], r["full"], rtol=1e-12)           # every rank holds the global sums
        _check_eval_table(r["table"], dict(np.load(os.path.join(ROOT, "tests", "golden", "eval_user05.npz"))))


def test_shard_indices_validate_rank():
    from absolutetrack_b200 import distributed as D

    with pytest.raises(ValueError):
        D.shard_indices(4, 2, 2)
    assert list(D.shard_indices(5, 1, 2)) == [1, 3]
    assert D.shard(["a", "b", "c"], 0, 2) == ["a", "c"]
    with pytest.raises(TypeError):
        D.all_reduce_metric_sums(torch.zeros(2))


def test_ctypes_structs_match_the_c_header(tmp_path):
    """include/absolutetrack_b200.h compiled as plain C (gcc): every struct the Python binding mirrors has the
    same size there as in ctypes, so by-value / by-pointer arguments cannot silently drift apart."""
    import shutil
    import subprocess

    from absolutetrack_b200 import _cabi

    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    names = ["AtCropParams", "AtHandModel", "AtLayerInfo", "AtTemporalState", "AtTailInputs", "AtTailOutputs",
             "AtRigCamera", "AtCropCameraOpts", "AtCropCameraInputs", "AtCropCameraOutputs"]
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "absolutetrack_b200.h"\nint main(void){\n'
                   + "".join(f'printf("{n} %zu\\n", sizeof({n}));\n' for n in names) + "return 0;}\n")
    exe = tmp_path / "sizes"
    subprocess.check_call([gcc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = dict(line.split() for line in subprocess.check_output([str(exe)], text=True).splitlines())
    for n in names:
        assert int(out[n]) == ctypes.sizeof(getattr(_cabi, n)), n


def test_layer1_space_to_depth_weight_mapping_is_the_same_convolution():
    """The encoder runs layer 1 (3x3, 32 -> 32, 48 x 48) as a 3x3 conv with 64 -> 64 channels over the x-space-to-depth
    view (rows of 24 pixel pairs, channels [even column | odd column]); csrc/conv_tc.cu::prepack_s2d_kernel builds the
    weights as W'[dy][di][(pi, ci)][(po, co)] = W[dy][dx = 2 di + pi - po][ci][co] for |dx| <= 1, else 0.  Restated in
    torch on the CPU: both forms are the same convolution, the side taps (di = +-1) only connect one input parity to
    one output parity (the K steps / column halves the kernel does not issue), and the bias wraps modulo 32."""
    import torch
    import torch.nn.functional as F

    g = torch.Generator().manual_seed(3)
    w = torch.randn(32, 32, 3, 3, generator=g, dtype=torch.float64)   # [co, ci, ky, kx]
    b = torch.randn(32, generator=g, dtype=torch.float64)
    x = torch.randn(2, 32, 48, 48, generator=g, dtype=torch.float64)
    ref = F.conv2d(x, w, b, padding=1)

    w2 = torch.zeros(64, 64, 3, 3, dtype=torch.float64)               # [(po, co), (pi, ci), ky, kdi]
    for di in (-1, 0, 1):
        for pi in (0, 1):
            for po in (0, 1):
                dx = 2 * di + pi - po
                if -1 <= dx <= 1:
                    w2[po * 32:(po + 1) * 32, pi * 32:(pi + 1) * 32, :, di + 1] = w[:, :, :, dx + 1]
    # structural zeros the kernel skips: di = -1 uses only odd inputs -> even outputs, di = +1 even inputs -> odd outputs
    assert not w2[:, :32, :, 0].any() and not w2[32:, :, :, 0].any()
    assert not w2[:, 32:, :, 2].any() and not w2[:32, :, :, 2].any()
    xs = torch.cat([x[:, :, :, 0::2], x[:, :, :, 1::2]], dim=1)        # [n, 64, 48, 24]
    ys = F.conv2d(xs, w2, b.repeat(2), padding=1)
    out = torch.empty_like(ref)
    out[:, :, :, 0::2], out[:, :, :, 1::2] = ys[:, :32], ys[:, 32:]
    assert torch.allclose(out, ref, rtol=0, atol=1e-12)
