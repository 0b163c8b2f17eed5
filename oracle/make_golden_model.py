"""tests/golden/model.npz: outputs of the UNMODIFIED reference UmeTrackModel on seeded inputs.

Run in the build container only (needs /root/reference).  Weights = absolutetrack_b200
make_random_state_dict(0) loaded STRICTLY into the reference model (the real checkpoint is not
distributed).  Also asserts that oracle/model_oracle.py reproduces the reference.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)


def build_reference_model(sd):
    import ref_import

    ref_import.import_reference()
    from lib.models import feature_extractor as fe, model_loader as rml, skeleton_encoder as se, temporal as tem
    from lib.models.model_opts import ModelOpts
    from lib.models.umetrack_model import UmeTrackModel

    mo = ModelOpts()
    f = fe.FeatureExtractor((96, 96), mo)
    m = UmeTrackModel(f, tem.create_temporal_model(mo, f.output_feature_sizes),
                      se.SkeletonEncoder([4, *f.output_feature_sizes]),
                      rml._create_regressor(mo, f.output_feature_sizes, True, False),
                      rml._create_regressor(mo, f.output_feature_sizes, False, True))
    m.load_state_dict(sd, strict=True)
    return m.eval()


def run_reference(model, seq, skel, head):
    from lib.models.umetrack_model import InputFrameData, InputFrameDesc, InputSkeletonData

    outs = []
    for fi in seq:
        fd = InputFrameData(torch.from_numpy(fi["images"]), torch.from_numpy(fi["intrinsics"]), torch.from_numpy(fi["extrinsics"]))
        de = InputFrameDesc(torch.from_numpy(fi["sample_range"]), torch.from_numpy(fi["memory_idx"]),
                            torch.from_numpy(fi["use_memory"]), torch.from_numpy(fi["hand_idx"]))
        with torch.no_grad():
            if head == "k":
                o = model.regress_pose_use_skeleton(fd, de, InputSkeletonData(skel[0], skel[1]))
            else:
                o = model.regress_pose_pred_skel_scale(fd, de)
            feats = model._feature_extractor._image_backbone(fd.left_images.unsqueeze(1))
        outs.append(dict(joint_angles=o.joint_angles.numpy(), wrist_xfs=o.wrist_xfs.numpy(),
                         sigmas=o.landmark_uncertainty_sigmas.numpy(),
                         skel_scales=None if o.skel_scales is None else o.skel_scales.numpy(),
                         features=feats.numpy()))
    return outs


def make_model():
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    sd = make_random_state_dict(0)
    import scenario
    import model_oracle as MO

    hm = scenario.hand_model()
    skel = (torch.from_numpy(hm["joint_rotation_axes"]).float(), torch.from_numpy(hm["joint_rest_positions"] * 0.001).float())
    out = {}
    cases = {"k": dict(n_frames=4, start=10, single_view_frame=2), "u": dict(n_frames=3, start=40)}
    for head, kw in cases.items():
        seq = scenario.model_sequence(kind="smooth", **kw)
        ref = run_reference(build_reference_model(sd), seq, skel, head)
        st = MO.TemporalState()
        for t, (fi, r) in enumerate(zip(seq, ref)):
            o = MO.forward(sd, st, torch.from_numpy(fi["images"]), torch.from_numpy(fi["intrinsics"]),
                           torch.from_numpy(fi["extrinsics"]), torch.from_numpy(fi["sample_range"]),
                           torch.from_numpy(fi["memory_idx"]), torch.from_numpy(fi["use_memory"]),
                           torch.from_numpy(fi["hand_idx"]), skel if head == "k" else None)
            for key in ("joint_angles", "wrist_xfs", "sigmas"):
                d = np.abs(o[key].numpy() - r[key]).max()
                assert d < 2e-5, (head, t, key, d)
            assert np.abs(o["features"].numpy() - r["features"]).max() < 1e-4
            for key, v in r.items():
                if v is not None and key != "features":
                    out[f"{head}{t}_{key}"] = v
            out[f"{head}{t}_features"] = r["features"][:2].astype(np.float32)  # first two crops only (size)
            out[f"{head}{t}_n_crops"] = np.array(len(fi["crops"]))
        out[f"{head}_case"] = np.array([kw["n_frames"], kw["start"], kw.get("single_view_frame", -1) if kw.get("single_view_frame") is not None else -1])
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "model.npz"), **out)
    print("model.npz", len(out), "arrays; oracle == reference within 2e-5")


if __name__ == "__main__":
    make_model()
