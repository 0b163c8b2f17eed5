"""BASELINE.json configs[3]: batch sweep B = 1, 2, 4 ... 4096 frames per step of the synthetic 4-view stream through
bench.py (device-resident value, e2e, per-stage times and roofline fractions per B).  Writes one JSON line per B.
usage: python scripts/batch_sweep.py out.jsonl [max_B]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_path = sys.argv[1]
max_b = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
with open(out_path, "w") as out:
    b = 1
    while b <= max_b:
        steps = 200 if b <= 16 else (100 if b <= 128 else (30 if b <= 1024 else 10))
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--frames", str(b), "--steps", str(steps), "--warmup", "5",
                            "--no-cpu-baseline", "--no-parity"], capture_output=True, text=True, timeout=900)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ""
        try:
            d = json.loads(line)
            keep = {"frames_per_step": b, "steps": steps, "value": d["value"], "ms_per_step": d["ms_per_step"], "e2e": d["e2e"]["value"],
                    "e2e_from_poses": d["e2e_from_poses"]["value"] if d.get("e2e_from_poses") else None,
                    "encoder_frac_of_burst_peak": d["roofline"]["frac"], "stages": d["stages"], "gpu_launches_per_step": d["gpu_launches"] / steps,
                    "unit": d["unit"], "clocks": d["clocks"]}
        except Exception as e:  # keep going: one failing size must not lose the sweep
            keep = {"frames_per_step": b, "error": f"{type(e).__name__}: {e}", "stderr": r.stderr[-600:]}
        out.write(json.dumps(keep) + "\n")
        out.flush()
        print(json.dumps({k: keep.get(k) for k in ("frames_per_step", "value", "e2e", "encoder_frac_of_burst_peak", "error")}), flush=True)
        b *= 2
