"""Script-level parity (BASELINE.json configs[0], [1], [2]): the UNMODIFIED
functions of the reference's entry scripts -

  run_eval_known_skeleton._track_sequence             (run_eval_known_skeleton.py:48-106)
  run_eval_unknown_skeleton._track_sequence           (run_eval_unknown_skeleton.py:80-142, with
                                                        _track_sequence_and_calibrate :49-78)
  run_inference_torch_data._eval_batch                (run_inference_torch_data.py:88-135), batch 256

- executed over absolutetrack_b200 by package shadowing and compared with what
the REFERENCE itself produced from the same scripts, label files, fake-av "videos" and seeded checkpoint in
the build container (oracle/make_golden_scripts.py -> tests/golden/scripts/scripts_golden.npz)."""
import os
import pickle
import subprocess
import sys
import tarfile

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
FIX = os.path.join(HERE, "golden", "scripts")
N_CALIBRATION_SAMPLES = 12  # = oracle/make_golden_scripts.py

# The encoder the scripts get by default (AT_PRECISION unset) is the fp16 tcgen05 one.  north_star's end-to-end
# tolerance is on the METRIC: |MPJPE(new vs GT) - MPJPE(reference vs GT)| <= 0.1 mm, per recording.  The stricter
# pose-to-pose distance mean ||new - reference|| is bounded too: 0.1 mm is met as a mean over many hand-frames
# (tests/test_gpu_parity.py: 0.088 mm over 4,096); a 40-hand-frame recording scatters around it (0.12 measured).
TOL_MPJPE_MM = 0.1
TOL_POSE_DISTANCE_MM = 0.2
TOL_MAX_MM = 0.5


@pytest.fixture(scope="module")
def ref_checkout(tmp_path_factory):
    """The temporary "reference checkout": entry scripts + the reference's un-mirrored plumbing."""
    d = tmp_path_factory.mktemp("reference_checkout")
    with tarfile.open(os.path.join(FIX, "ref_sources.tar.gz")) as tar:
        tar.extractall(d, filter="data")
    return str(d)


@pytest.fixture(scope="module")
def checkpoint(tmp_path_factory):
    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    p = str(tmp_path_factory.mktemp("ckpt") / "pretrained_weights.torch")
    torch.save(make_random_state_dict(0), p)
    return p


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(os.path.join(FIX, "scripts_golden.npz")))


def run_driver(ref_checkout, *args, env_extra=None):
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "absolutetrack_b200"), ref_checkout, os.path.join(HERE, "fake_av")])
    env.pop("AT_PRECISION", None)
    env.update(env_extra or {})
    r = subprocess.run([sys.executable, os.path.join(HERE, "script_driver.py"), *args], cwd=ref_checkout, env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    launches = [int(l.split()[1]) for l in r.stdout.splitlines() if l.startswith("launches ")]
    assert launches and launches[-1] > 0, "the script ran without launching a kernel of libabsolutetrack_b200.so"
    return r


def compare_recording(result_path, G, prefix):
    with open(result_path, "rb") as f:
        res = pickle.load(f)
    assert np.array_equal(res["valid_tracking"], G[prefix + "valid_tracking"])
    v = res["valid_tracking"]
    assert np.abs(res["gt_keypoints"] - G[prefix + "gt_keypoints"]).max() < 1e-3            # FK of the labels, mm
    d = np.linalg.norm(res["tracked_keypoints"] - G[prefix + "tracked_keypoints"], axis=-1)  # [2, T, 21] mm
    per_hand_frame = d[v].mean(-1)
    assert not np.any(res["tracked_keypoints"][~v]), "hand-frames that were not tracked must stay zero"
    # the scripts' own metric (run_eval_known_skeleton.py:92-93): mean keypoint error vs the labels, per hand-frame
    mpjpe_new = np.linalg.norm((res["gt_keypoints"] - res["tracked_keypoints"])[v], axis=-1).mean(-1).mean()
    mpjpe_ref = G[prefix + "per_frame_mean_error"].mean()
    return dict(d_mpjpe_mm=float(abs(mpjpe_new - mpjpe_ref)), pose_distance_mm=float(per_hand_frame.mean()),
                max_landmark_distance_mm=float(d[v].max()), n=int(v.sum()))


def check(label, r, tol_metric=TOL_MPJPE_MM, tol_pose=TOL_POSE_DISTANCE_MM, tol_max=TOL_MAX_MM):
    print(label, r)
    assert r["d_mpjpe_mm"] <= tol_metric and r["pose_distance_mm"] <= tol_pose and r["max_landmark_distance_mm"] <= tol_max, r


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["rec_a", "rec_b"])
def test_run_eval_known_skeleton_track_sequence(ref_checkout, checkpoint, golden, tmp_path, name):
    out = str(tmp_path / "known" / (name + ".npy"))
    run_driver(ref_checkout, "known", os.path.join(FIX, name + ".mp4"), out, checkpoint)
    check(f"known-skeleton {name}:", compare_recording(out, golden, f"known_{name}_"))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["rec_a", "rec_b"])
def test_run_eval_unknown_skeleton_calibrate_and_track(ref_checkout, checkpoint, golden, tmp_path, name):
    out = str(tmp_path / "unknown" / (name + ".npy"))
    run_driver(ref_checkout, "unknown", os.path.join(FIX, name + ".mp4"), out, checkpoint,
               os.path.join(FIX, "generic_hand_model.json"), str(N_CALIBRATION_SAMPLES))
    check(f"unknown-skeleton {name}:", compare_recording(out, golden, f"unknown_{name}_"))


@pytest.mark.gpu
def test_run_eval_known_skeleton_fp32_encoder(ref_checkout, checkpoint, golden, tmp_path):
    """AT_PRECISION=fp32 selects the CUDA-core fp32 encoder without editing a script: accumulation-order noise only."""
    out = str(tmp_path / "known32" / "rec_a.npy")
    run_driver(ref_checkout, "known", os.path.join(FIX, "rec_a.mp4"), out, checkpoint, env_extra={"AT_PRECISION": "fp32"})
    check("known-skeleton rec_a, fp32 encoder:", compare_recording(out, golden, "known_rec_a_"), 0.01, 0.02, 0.1)


@pytest.mark.gpu
def test_run_inference_torch_data_eval_batch_256(ref_checkout, checkpoint, golden, tmp_path):
    """Config 3 at batch 256 (BASELINE.json configs[2]): preprocess (GPU crops) -> bundles.collate -> the script's
    _eval_batch (known-skeleton and scale head).  The reference preprocessed the same 256 samples on the CPU; its
    crop matrices are float32 inverses, so the inputs already differ by 2e-3 of the grey range."""
    out = str(tmp_path / "evalbatch.npz")
    run_driver(ref_checkout, "evalbatch", os.path.join(HERE, "golden", "torchdata"), out, checkpoint, "256")
    res = np.load(out)
    for k in ("k", "u"):
        g = golden[f"evalbatch256_{k}_keypoint_errors_mm"]
        assert res[k].shape == g.shape == (256,)
        d = np.abs(res[k] - g)
        print(f"config 3 batch 256 head {k}: keypoint error {g.mean():.2f} mm, |delta| mean {d.mean():.4f} max {d.max():.4f} mm")
        assert d.mean() <= 0.1 and d.max() <= 0.5


def test_entry_scripts_import_under_package_shadowing(ref_checkout):
    """CPU: all three entry scripts (and load_eval.py) import with absolutetrack_b200/ shadowing `lib`; the hot-path
    modules resolve to the mirror, the un-mirrored plumbing to the reference's files."""
    code = (
        "import lib, lib.data_utils.fs as fs\n"
        "import run_eval_known_skeleton as a, run_eval_unknown_skeleton as b, run_inference_torch_data as c, load_eval\n"
        "import sys\n"
        "mine = lambda m: 'absolutetrack_b200' in sys.modules[m].__file__\n"
        "assert mine('lib') and mine('lib.tracker.tracker') and mine('lib.models.model_loader') and mine('lib.batched_dataset.data_transform')\n"
        "assert mine('lib.common.hand_skinning') and mine('lib.data_utils.bundles') and mine('lib.data_utils.idxbinfile')\n"
        "for m in ('lib.data_utils.fs', 'lib.data_utils.split', 'lib.data_utils.dataset_util', 'lib.data_utils.async_dataset'):\n"
        "    assert not mine(m), m\n"
        "assert a.HandTracker is b.HandTracker and c.load_pretrained_model is a.load_pretrained_model\n"
        "print('imports ok')\n"
    )
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "absolutetrack_b200"), ref_checkout, os.path.join(HERE, "fake_av")])
    r = subprocess.run([sys.executable, "-c", code], cwd=ref_checkout, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "imports ok" in r.stdout, r.stdout[-1000:] + r.stderr[-3000:]
