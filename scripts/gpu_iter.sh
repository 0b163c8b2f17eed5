timeout 900 python -m pytest tests/test_gpu_tc.py -m gpu -q -s -k tracker 2>&1 | tail -25
