# short evidence refresh after a kernel change: GPU tests, standalone rooflines, the bench line
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -2 gpurun_out/pytest_gpu.log
timeout 300 python scripts/kernel_roofline.py > gpurun_out/kernel_roofline.log 2>&1; grep -c kernel gpurun_out/kernel_roofline.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 300 gpurun_out/bench.json
