timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k crop 2>&1 | tail -5
timeout 300 python scripts/kernel_roofline.py 2>&1 | grep crop_res | tee gpurun_out/kernel_roofline.log
