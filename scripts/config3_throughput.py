"""BASELINE.json configs[2] - run_inference_torch_data.py at batch 256 - as a throughput line: the reference's UNMODIFIED
_eval_batch (run_inference_torch_data.py:88-135) over absolutetrack_b200 by package shadowing, on the 256-sample batch of
tests/test_gpu_scripts.py (the fixture sequences of tests/golden/torchdata, grey levels shifted per sample), preprocess
(crop matrices + homography resampling on the GPU) + bundles.collate timed separately.  Wall-clock around synchronised
calls (the script's own structure: per-frame model calls from Python).  Prints one JSON line.
"""
import json
import os
import subprocess
import sys
import tarfile
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TESTS = os.path.join(ROOT, "tests")
sys.path.insert(0, ROOT)


def main():
    import torch

    from absolutetrack_b200.lib.models.model_loader import make_random_state_dict

    bs = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    with tempfile.TemporaryDirectory() as tmp:
        with tarfile.open(os.path.join(TESTS, "golden", "scripts", "ref_sources.tar.gz")) as tar:
            tar.extractall(tmp, filter="data")
        ckpt = os.path.join(tmp, "pretrained_weights.torch")
        torch.save(make_random_state_dict(0), ckpt)
        env = dict(os.environ)
        env["PYTHONPATH"] = os.pathsep.join([os.path.join(ROOT, "absolutetrack_b200"), tmp, os.path.join(TESTS, "fake_av")])
        r = subprocess.run([sys.executable, os.path.join(TESTS, "script_driver.py"), "evalbatch", os.path.join(TESTS, "golden", "torchdata"),
                            os.path.join(tmp, "out.npz"), ckpt, str(bs), "time"], cwd=tmp, env=env, capture_output=True, text=True, timeout=1200)
        if r.returncode != 0:
            sys.stderr.write(r.stdout[-2000:] + r.stderr[-4000:])
            raise SystemExit(r.returncode)
        t = json.loads([l for l in r.stdout.splitlines() if l.startswith("timing ")][-1][7:])
    T = t["frames_per_sequence"] or 1
    line = {"workload": f"run_inference_torch_data._eval_batch, batch {bs} sequences x {T} frames (2 views), fixture sequences of tests/golden/torchdata",
            "precision": os.environ.get("AT_PRECISION", "fp16"), **t}
    for k, s in t["eval_batch_s"].items():
        line[f"sequences_per_s_{k}"] = bs / s
        line[f"hand_frames_per_s_{k}"] = bs * T / s
    line["sequences_per_s_preprocess_collate"] = bs / t["preprocess_collate_s"]
    print(json.dumps(line))


if __name__ == "__main__":
    main()
