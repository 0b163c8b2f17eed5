#!/bin/bash
# One parameterised gpurun driver (replaces the per-experiment gpu_*.sh files).
#   scripts/gpu_run.sh tests [pytest args]   pytest -m gpu (+ smoke)
#   scripts/gpu_run.sh bench [args]          python bench.py [args]  -> gpurun_out/bench.json
#   scripts/gpu_run.sh launches              ncu launch list of a short bench run -> gpurun_out/launches.csv
#   scripts/gpu_run.sh ncu <regex> <n>       ncu --set full of kernels matching <regex> (n launches) during a short bench
#   scripts/gpu_run.sh py <script> ...       python <script> ...    -> gpurun_out/<script>.log
#   scripts/gpu_run.sh ab <name> "<cmd>" "<ENV=..>" ...   the same command once per environment setting, back to back on
#                                            one box -> gpurun_out/<name>.log  (A/B experiments; "" = default environment)
# Everything it writes goes to gpurun_out/ (merged back by gpurun).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
what=$1; shift
case "$what" in
  tests)
    python -m pytest tests -m gpu -x -q -s "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
    python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" | tee -a gpurun_out/smoke.log
    tail -n 25 gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/smoke.log ;;
  bench)
    python bench.py "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err ;;
  launches)
    ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity "$@" > gpurun_out/ncu_bench.log 2>&1; echo "ncu exit $?" ;;
  ncu)
    regex=$1; n=$2; shift 2
    ncu --set full --clock-control none --import-source on -k "regex:$regex" -c "$n" -o gpurun_out/ncu_full -f \
      python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-parity "$@" > gpurun_out/ncu_full.log 2>&1; echo "ncu exit $?" ;;
  py)
    script=$1; shift
    python "$script" "$@" > "gpurun_out/$(basename "$script" .py).log" 2>&1; echo "exit $?"; tail -n 40 "gpurun_out/$(basename "$script" .py).log" ;;
  ab)
    name=$1; cmd=$2; shift 2
    out=gpurun_out/$name.log; : > "$out"
    for setting in "$@"; do
      echo "== ${setting:-(default)}" >> "$out"
      # shellcheck disable=SC2086
      env $setting timeout 600 $cmd >> "$out" 2> gpurun_out/$name.err || { echo "FAILED" >> "$out"; tail -n 5 gpurun_out/$name.err >> "$out"; }
    done
    cut -c1-400 "$out" ;;
  *) echo "unknown mode $what"; exit 2 ;;
esac
