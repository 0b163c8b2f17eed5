timeout 600 ncu --set full --clock-control none --import-source on -k regex:fk_skin_shared -s 16 -c 1 -o /tmp/fk_full -f python scripts/kernel_roofline.py > gpurun_out/ncu_fk.log 2>&1
echo "rc=$?"
ncu -i /tmp/fk_full.ncu-rep --page source --csv > gpurun_out/fk_source.csv 2>/dev/null
ncu -i /tmp/fk_full.ncu-rep --page details 2>/dev/null | grep -E "Duration|Issue Slots Busy|Executed Ipc|Achieved Occupancy|Theoretical Occ|way bank|Executed Instructions |DRAM Throughput|Grid Size|Registers Per|Warp Cycles Per Issued|No Eligible|Eligible Warps|L1/TEX Hit|Mem Busy|Max Bandwidth|Block Limit" | head -30
ncu -i /tmp/fk_full.ncu-rep --page details 2>/dev/null | grep -B2 -A6 "Stall" | head -60
