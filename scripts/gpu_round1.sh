# full GPU check of a round: tests, smoke, bench (+ reference arm), small timing logs for profiles/
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
if [ "$1" = "full" ]; then
  timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 1500 gpurun_out/bench_reference.json
  timeout 100 python scripts/ingest_trace.py pull > gpurun_out/ingest_trace.log 2>&1
  timeout 100 python scripts/pull_time.py 512 37 74 148 296 > gpurun_out/pull_time.log 2>&1
  timeout 100 python scripts/overlap_probe.py > gpurun_out/overlap_probe.log 2>&1
  timeout 100 scripts/pcie_pull_bench > gpurun_out/pcie_pull_bench.log 2>&1
  timeout 300 python scripts/kernel_roofline.py > gpurun_out/kernel_roofline.log 2>&1
fi
