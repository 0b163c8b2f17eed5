timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k fk 2>&1 | tail -3
timeout 300 python scripts/kernel_roofline.py 2>&1 | grep fk_skin | tee gpurun_out/kernel_roofline.log
