# FK fast-path check: parity tests, standalone roofline, one ncu --set full summary
python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "fk or skin or eval" 2>&1 | tail -3
python scripts/kernel_roofline.py 2>/dev/null | grep -i fk
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fk_skin_shared -s 16 -c 1 -o /tmp/fk_full -f python scripts/kernel_roofline.py > gpurun_out/ncu_fk.log 2>&1
echo "rc=$?"
ncu -i /tmp/fk_full.ncu-rep --page source --csv > gpurun_out/fk_source.csv 2>/dev/null
ncu -i /tmp/fk_full.ncu-rep --page details 2>/dev/null | grep -E "Duration|Issue Slots Busy|Executed Ipc|Achieved Occupancy|way bank|Executed Instructions |DRAM Throughput|Registers Per|No Eligible|Block Limit" | head -20
