python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "eval or fk" 2>&1 | tail -15
