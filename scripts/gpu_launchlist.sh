set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
