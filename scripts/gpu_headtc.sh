# GPU check of the tcgen05 pose head: model parity tests with it on and off, bench A/B
set -x
timeout 600 python -m pytest tests/test_gpu_model.py tests/test_gpu_tc.py -m gpu -q -x > gpurun_out/pytest_headtc.log 2>&1; tail -15 gpurun_out/pytest_headtc.log
AT_TAIL_HEAD_TC=0 timeout 300 python -m pytest tests/test_gpu_model.py -m gpu -q -x > gpurun_out/pytest_headmma.log 2>&1; tail -3 gpurun_out/pytest_headmma.log
for cfg in "1 2" "1 1" "0 2"; do set -- $cfg
  AT_TAIL_HEAD_TC=$1 AT_TC_MT80=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_head_$1_$2.json 2> gpurun_out/bench_head_$1_$2.err
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/bench_head_$1_$2.json"))
    print("head_tc=$1 mt80=$2", "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "ms", d["ms_per_step"], d["stages"])
except Exception as e:
    print("bench failed", e); print(open("gpurun_out/bench_head_$1_$2.err").read()[-1500:])
PY
done
