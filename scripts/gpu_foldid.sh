set -x
AT_TC_FOLD_ID=15 timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -q -x > gpurun_out/pytest_foldid.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_foldid.log
for m in 0 1 2 4 8 6 15; do
  AT_TC_FOLD_ID=$m timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_foldid_$m.json 2> gpurun_out/bench_foldid_$m.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_foldid_$m.json"))
    print("mask=$m", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", round(d["ms_per_step"],4), "enc", round(d["stages"]["encoder_ms"],4), "frac", round(d["roofline"]["frac"],4))
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_foldid_$m.err").read()[-1500:])
PY
done
