"""TEST INFRASTRUCTURE - runs ONE function of an UNMODIFIED reference entry script in a fresh process.

tests/test_gpu_scripts.py starts it with PYTHONPATH = <repo>/absolutetrack_b200 : <temporary reference
checkout unpacked from tests/golden/scripts/ref_sources.tar.gz> : tests/fake_av, i.e. exactly the
"package shadowing" route of INTEGRATION.md: the scripts import `lib.*`, the hot path resolves to
absolutetrack_b200/lib, the un-mirrored plumbing (`lib.data_utils.fs`, ...) to the reference's files.

usage: script_driver.py known   <rec.mp4> <out.npy> <ckpt>
       script_driver.py unknown <rec.mp4> <out.npy> <ckpt> <generic_hand_model.json> <n_calibration_samples>
       script_driver.py evalbatch <torchdata dir> <out.npz> <ckpt> <batch> [time]
"""
import os
import sys

import numpy as np
import torch


def batch_sample(mono, labels, i):
    """= oracle/make_golden_scripts.py::batch256_sample (integer arithmetic only)."""
    s = i % 2
    m = ((np.asarray(mono[s]).astype(np.int32) + 3 * (i // 2)) % 256).astype(np.uint8)
    return {"mono": m, "labels": labels[s]}


def main():
    mode = sys.argv[1]
    import lib

    assert os.path.basename(os.path.dirname(lib.__file__)) == "lib" and "absolutetrack_b200" in lib.__file__, lib.__file__
    if mode == "known":
        import run_eval_known_skeleton as R

        assert R.HandTracker.__module__ == "lib.tracker.tracker" and "absolutetrack_b200" in sys.modules[R.HandTracker.__module__].__file__
        assert "absolutetrack_b200" not in R.fs.__file__, "lib.data_utils.fs must be the reference's own file"
        err = R._track_sequence((sys.argv[2], sys.argv[3]), model_path=sys.argv[4])
        print("mean error", float(np.mean(err)))
    elif mode == "unknown":
        import run_eval_unknown_skeleton as R
        from lib.tracker.video_pose_data import _load_json, load_hand_model_from_dict

        generic = load_hand_model_from_dict(_load_json(sys.argv[5]))
        err = R._track_sequence((sys.argv[2], sys.argv[3]), model_path=sys.argv[4], generic_hand_model=generic,
                                n_calibration_samples=int(sys.argv[6]))
        print("mean error", float(np.mean(err)))
    elif mode == "evalbatch":
        import run_inference_torch_data as R
        from lib.batched_dataset.data_transform import preprocess
        from lib.data_utils import bundles
        from lib.data_utils.idxbinfile import TorchIdx

        assert R.find_dataset.__module__ == "lib.data_utils.async_dataset" and "absolutetrack_b200" not in sys.modules[R.find_dataset.__module__].__file__
        d, bs = sys.argv[2], int(sys.argv[5])
        mono = TorchIdx(os.path.join(d, "seq.mono.torch.idx")).read_bin()
        labels = TorchIdx(os.path.join(d, "seq.labels.torch.idx")).read_bin()
        mono = [np.array(mono[i]) for i in range(2)]
        labels = [labels[i] for i in range(2)]
        batch = [preprocess(batch_sample(mono, labels, i), (96, 96)) for i in range(bs)]
        model_input, model_target = bundles.collate(batch)
        model = R.load_pretrained_model(sys.argv[4])
        model.eval()
        model.to("cuda")
        out = {}
        with torch.inference_mode():
            for name, use_skel in (("k", True), ("u", False)):
                err = R._eval_batch(model, model_input, model_target, "multiv", use_skel=use_skel, device="cuda")
                out[name] = err.cpu().numpy()
        np.savez(sys.argv[3], **out)
        if len(sys.argv) > 6 and sys.argv[6] == "time":   # scripts/config3_throughput.py: the same calls, timed
            import json
            import time

            T = int(model_input.left_images.shape[1])   # [batch, frames, views, 96, 96]
            reps = 5
            t0 = time.perf_counter()
            for _ in range(reps):
                batch = [preprocess(batch_sample(mono, labels, i), (96, 96)) for i in range(bs)]
                model_input, model_target = bundles.collate(batch)
            torch.cuda.synchronize()
            t_pre = (time.perf_counter() - t0) / reps
            times = {}
            with torch.inference_mode():
                for name, use_skel in (("known_skeleton", True), ("predict_scale", False)):
                    R._eval_batch(model, model_input, model_target, "multiv", use_skel=use_skel, device="cuda")
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    for _ in range(reps):
                        R._eval_batch(model, model_input, model_target, "multiv", use_skel=use_skel, device="cuda")
                    torch.cuda.synchronize()
                    times[name] = (time.perf_counter() - t0) / reps
            shapes = {k: list(v.shape) for k, v in vars(model_input).items() if hasattr(v, "shape")} if hasattr(model_input, "__dict__") else {}
            print("timing " + json.dumps({"batch": bs, "frames_per_sequence": T, "preprocess_collate_s": t_pre,
                                          "eval_batch_s": times, "input_shapes": shapes}))
    else:
        raise SystemExit(f"unknown mode {mode}")
    from absolutetrack_b200 import _cabi

    print("launches", int(_cabi.lib().at_launch_count()))


if __name__ == "__main__":
    main()
