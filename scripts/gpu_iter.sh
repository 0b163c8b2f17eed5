timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k crop 2>&1 | tail -3
for m in zero_copy; do timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --ingest $m > gpurun_out/bench_$m.json 2>gpurun_out/bench_$m.err; echo "rc=$?"; tail -4 gpurun_out/bench_$m.err; done
