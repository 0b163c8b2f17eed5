python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "config3 or homography" 2>&1 | tail -15
