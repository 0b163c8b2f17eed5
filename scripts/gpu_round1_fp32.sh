set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err; echo bench rc=$?; tail -3 gpurun_out/bench_fp32.err; cat gpurun_out/bench_fp32.json
python bench.py --steps 2 --warmup 3 --frames 32 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r1_fp32.csv python bench.py --steps 2 --warmup 3 --frames 32 --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo ncu rc=$?
