"""tests/golden/scripts/: script-level fixtures and goldens (BASELINE.json configs[0..2]).

TEST INFRASTRUCTURE - run in the build container only (needs /root/reference).  It runs the
UNMODIFIED entry scripts of the reference, through the reference's own `lib`, on CPU:

  * run_eval_known_skeleton._track_sequence            (run_eval_known_skeleton.py:48-106)
  * run_eval_unknown_skeleton._track_sequence (which calls _track_sequence_and_calibrate, :49-128)
  * run_inference_torch_data._eval_batch               (run_inference_torch_data.py:88-135), batch 256

on inputs that can be rebuilt bit for bit on the GPU box:

  rec_{a,b}.json   label files cut from sample_data/recording_00.json and sample_data/user05/recording_02.json
                   (N consecutive frames, mesh fields dropped: video_pose_data.py:62-93 does not need them)
  rec_{a,b}.mp4    fake-av descriptors (tests/fake_av/av.py synthesises the frames deterministically)
  generic_hand_model.json   dataset/generic_hand_model.json without the mesh fields
  the checkpoint   torch.save(make_random_state_dict(0)) - rebuilt by the test, not stored
  config 3 batch   256 samples derived from tests/golden/torchdata/seq.* by `batch256_sample`

and stores what the reference produced (scripts_golden.npz).  It also packs the entry scripts and the
reference's host-plumbing modules they import but the product does not mirror (lib/data_utils/{fs,split,
dataset_util,async_dataset,async_utils,nested_async}.py) into ref_sources.tar.gz: the script-level GPU
tests unpack it into a temporary "reference checkout", put absolutetrack_b200/ in front of it on
PYTHONPATH and execute the scripts' functions unchanged.  The archive is a test fixture; the product
never imports from it.
"""
import hashlib
import io
import json
import os
import pickle
import sys
import tarfile
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLD = os.path.join(ROOT, "tests", "golden")
OUT = os.path.join(GOLD, "scripts")
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(ROOT, "tests", "fake_av"))   # `import av` -> the deterministic stand-in
sys.path.insert(0, os.path.join(ROOT, "tests"))

import ref_import  # noqa: E402

BIG = ("mesh_vertices", "mesh_triangles", "dense_bone_weights")
RECORDINGS = {  # name -> (label file under the reference root, first frame, frames, video seed)
    "rec_a": ("sample_data/recording_00.json", 30, 20, 7),
    "rec_b": ("sample_data/user05/recording_02.json", 200, 20, 8),
}
REF_SOURCES = [
    "run_eval_known_skeleton.py", "run_eval_unknown_skeleton.py", "run_inference_torch_data.py", "load_eval.py",
    "lib/data_utils/fs.py", "lib/data_utils/split.py", "lib/data_utils/dataset_util.py",
    "lib/data_utils/async_dataset.py", "lib/data_utils/async_utils.py", "lib/data_utils/nested_async.py",
]
# The sample recordings label both hands with confidence 1 in every frame.  To reach the branches the scripts
# have for missing hands (video_pose_data.py:116-128 drops hands with confidence 0, tracker.py:236 skips those
# below 0.5, :268-271 returns an empty result and resets the history), a few confidences of rec_b are edited:
CONFIDENCE_EDITS = {"rec_b": {(5, 1): 0.0, (6, 1): 0.0, (9, 0): 0.0, (9, 1): 0.0, (13, 0): 0.3}}
N_CALIBRATION_SAMPLES = 12   # the script's __main__ uses 30; 12 makes the 20-frame fixtures take the early `break`


def write_label_fixtures():
    os.makedirs(OUT, exist_ok=True)
    for name, (src, start, n, seed) in RECORDINGS.items():
        with open(os.path.join(ref_import.REFERENCE_ROOT, src)) as f:
            L = json.load(f)
        cut = {k: L[k] for k in ("cameras", "camera_angles")}
        cut["hand_model"] = {k: v for k, v in L["hand_model"].items() if k not in BIG}
        for k in ("joint_angles", "hand_confidences", "wrist_transforms", "camera_to_world_transforms"):
            cut[k] = L[k][start:start + n]
        for (frame, hand), conf in CONFIDENCE_EDITS.get(name, {}).items():
            cut["hand_confidences"][frame][hand] = conf
        with open(os.path.join(OUT, name + ".json"), "w") as f:
            json.dump(cut, f, separators=(",", ":"))
        cam = L["cameras"][0]
        with open(os.path.join(OUT, name + ".mp4"), "w") as f:
            json.dump({"absolutetrack_fake_mp4": 1, "seed": seed, "n_frames": n, "height": int(cam["ImageSizeY"]),
                       "width": int(cam["ImageSizeX"]) * len(L["cameras"])}, f)
    with open(os.path.join(ref_import.REFERENCE_ROOT, "dataset", "generic_hand_model.json")) as f:
        g = json.load(f)
    with open(os.path.join(OUT, "generic_hand_model.json"), "w") as f:
        json.dump({k: v for k, v in g.items() if k not in BIG}, f, separators=(",", ":"))


def pack_reference_sources():
    buf = io.BytesIO()
    digests = {}
    with tarfile.open(fileobj=buf, mode="w:gz") as tar:
        for rel in REF_SOURCES:
            data = open(os.path.join(ref_import.REFERENCE_ROOT, rel), "rb").read()
            digests[rel] = hashlib.sha256(data).hexdigest()
            info = tarfile.TarInfo(rel)
            info.size, info.mtime, info.mode = len(data), 0, 0o644
            tar.addfile(info, io.BytesIO(data))
    with open(os.path.join(OUT, "ref_sources.tar.gz"), "wb") as f:
        f.write(buf.getvalue())
    with open(os.path.join(OUT, "ref_sources.sha256.json"), "w") as f:
        json.dump(digests, f, indent=1)


def batch256_sample(mono, labels, i):
    """Sample i of the 256-sample config-3 batch: fixture sequence i % 2 with its grey levels shifted by
    3 * (i // 2) (mod 256) - integer arithmetic, identical here and on the GPU box."""
    s = i % 2
    m = ((np.asarray(mono[s]).astype(np.int32) + 3 * (i // 2)) % 256).astype(np.uint8)
    return {"mono": m, "labels": labels[s]}


def main():
    write_label_fixtures()
    import av  # noqa: F401  the stand-in of tests/fake_av, before ref_import installs its empty stub

    assert hasattr(av, "synth_video_frame")
    ref_import.import_reference()
    pack_reference_sources()
    sys.path.insert(0, ROOT)
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict  # pure torch: no GPU needed

    # (the product package is only needed for the seeded checkpoint; top-level `lib` is the reference's)
    import run_eval_known_skeleton as RK
    import run_eval_unknown_skeleton as RU
    import run_inference_torch_data as RI
    from lib.tracker.video_pose_data import _load_json, load_hand_model_from_dict

    assert RK.HandTracker.__module__ == "lib.tracker.tracker" and RK.__file__.startswith(ref_import.REFERENCE_ROOT)
    out = {}
    torch.manual_seed(0)
    with tempfile.TemporaryDirectory() as tmp:
        ckpt = os.path.join(tmp, "pretrained_weights.torch")
        torch.save(make_random_state_dict(0), ckpt)
        generic = load_hand_model_from_dict(_load_json(os.path.join(OUT, "generic_hand_model.json")))
        for name in RECORDINGS:
            data_path = os.path.join(OUT, name + ".mp4")
            for mode, fn in (("known", lambda io_: RK._track_sequence(io_, model_path=ckpt)),
                             ("unknown", lambda io_: RU._track_sequence(io_, model_path=ckpt, generic_hand_model=generic,
                                                                        n_calibration_samples=N_CALIBRATION_SAMPLES))):
                res_path = os.path.join(tmp, mode, name + ".npy")
                err = fn((data_path, res_path))
                with open(res_path, "rb") as f:
                    res = pickle.load(f)
                p = f"{mode}_{name}_"
                out[p + "tracked_keypoints"] = res["tracked_keypoints"].astype(np.float32)
                out[p + "gt_keypoints"] = res["gt_keypoints"].astype(np.float32)
                out[p + "valid_tracking"] = res["valid_tracking"]
                out[p + "per_frame_mean_error"] = err
                print(mode, name, "valid hand-frames", int(res["valid_tracking"].sum()), "mean error", float(err.mean()))
        # ---- config 3, batch 256
        from lib.batched_dataset.data_transform import preprocess
        from lib.data_utils import bundles
        from lib.data_utils.idxbinfile import TorchIdx

        d = os.path.join(GOLD, "torchdata")
        im, il = TorchIdx(os.path.join(d, "seq.mono.torch.idx")), TorchIdx(os.path.join(d, "seq.labels.torch.idx"))
        bm, bl = open(im.bin_path, "rb").read(), open(il.bin_path, "rb").read()
        mono = [np.array(im.view_buffer(bm)[i]) for i in range(len(im))]
        labels = [il.view_buffer_at(i, bl) for i in range(len(il))]
        batch = [preprocess(batch256_sample(mono, labels, i), (96, 96)) for i in range(256)]
        model_input, model_target = bundles.collate(batch)
        model = RI.load_pretrained_model(ckpt)
        model.eval()
        with torch.inference_mode():
            for mode, use_skel in (("k", True), ("u", False)):
                err = RI._eval_batch(model, model_input, model_target, "multiv", use_skel=use_skel, device="cpu")
                out[f"evalbatch256_{mode}_keypoint_errors_mm"] = err.numpy()
                print("config 3 batch 256, use_skel", use_skel, "mean", float(err.mean()))
    np.savez_compressed(os.path.join(OUT, "scripts_golden.npz"), **out)
    print({k: os.path.getsize(os.path.join(OUT, k)) for k in sorted(os.listdir(OUT))})


if __name__ == "__main__":
    main()
