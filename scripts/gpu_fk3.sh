python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "fk or skin or eval" 2>&1 | tail -2
python scripts/kernel_roofline.py 2>/dev/null | grep -i fk | cut -c1-200
