# GPU check of the device crop-camera generator: its tests, the tracker tests around it, bench A/B of the step size
set -x
timeout 600 python -m pytest tests/test_gpu_crop_cameras.py tests/test_gpu_model.py -m gpu -q > gpurun_out/pytest_cropcams.log 2>&1; tail -30 gpurun_out/pytest_cropcams.log
timeout 500 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3500 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --steps 10 --warmup 3 --frames 472 --no-cpu-baseline > gpurun_out/bench_472.json 2> gpurun_out/bench_472.err; tail -c 3500 gpurun_out/bench_472.json; tail -5 gpurun_out/bench_472.err
