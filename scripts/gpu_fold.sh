# GPU check of the folded shortcut convs: 16-bit encoder tests, then bench A/B
set -x
timeout 300 python -m pytest tests/test_gpu_tc.py -m gpu -q -x > gpurun_out/pytest_fold.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_fold.log
for f in 1 0; do
  AT_TC_FOLD=$f timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fold_$f.json 2> gpurun_out/bench_fold_$f.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_fold_$f.json"))
    print("fold=$f", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", d["ms_per_step"], "enc", d["stages"]["encoder_ms"], "frac", d["roofline"]["frac"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_fold_$f.err").read()[-1500:])
PY
done
