# ncu --set full of the FK fast kernel at 4M hands (scripts/kernel_roofline.py), summary lines + stall reasons
python scripts/kernel_roofline.py > /dev/null 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fk_skin_shared -s 16 -c 1 -o /tmp/fk_full -f python scripts/kernel_roofline.py > gpurun_out/ncu_fk.log 2>&1
echo "rc=$?"
ncu -i /tmp/fk_full.ncu-rep --page source --csv > gpurun_out/fk_source.csv 2>/dev/null
ncu -i /tmp/fk_full.ncu-rep --page details > gpurun_out/fk_details.txt 2>/dev/null
ncu -i /tmp/fk_full.ncu-rep --page raw --csv > gpurun_out/fk_raw.csv 2>/dev/null
grep -E "Duration|Issue Slots Busy|Executed Ipc Active|Achieved Occupancy|Theoretical Occupancy|Executed Instructions  |DRAM Throughput|Registers Per|No Eligible|Block Limit|Grid Size|Block Size|Dynamic Shared|Warp Cycles Per Issued" gpurun_out/fk_details.txt | head -30
