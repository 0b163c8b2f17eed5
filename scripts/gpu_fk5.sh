python -m pytest tests/test_gpu_kernels.py tests/test_gpu_crop_cameras.py -x -q -m gpu -k "fk or skin or eval or crop_camera" 2>&1 | tail -2
python scripts/kernel_roofline.py > gpurun_out/kernel_roofline.log 2>/dev/null; grep '"fk_skin"' gpurun_out/kernel_roofline.log | cut -c1-200
