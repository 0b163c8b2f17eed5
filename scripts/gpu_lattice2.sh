set -x
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_model.py tests/test_gpu_crop_cameras.py -m gpu -q > gpurun_out/pytest_lattice.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_lattice.log
AT_CROP_LATTICE=1 timeout 200 python scripts/kernel_roofline.py 2>&1 | grep crop_resample
timeout 600 ncu --set full --clock-control none --import-source on -k regex:crop_resample_lattice -s 4 -c 1 -o /tmp/lat_full -f python scripts/kernel_roofline.py > gpurun_out/ncu_lat.log 2>&1
ncu -i /tmp/lat_full.ncu-rep --page source --csv > gpurun_out/lat_source.csv 2>/dev/null
ncu -i /tmp/lat_full.ncu-rep --page raw --csv > gpurun_out/lat_raw.csv 2>/dev/null
ncu -i /tmp/lat_full.ncu-rep --page details 2>/dev/null | grep -E "Duration|Issue Slots Busy|Executed Ipc|Achieved Occupancy|way bank|Executed Instructions |wavefronts" | head
