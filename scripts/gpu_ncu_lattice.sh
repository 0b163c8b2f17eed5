set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:crop_resample_lattice -s 4 -c 1 -o /tmp/lat_full -f python scripts/kernel_roofline.py > gpurun_out/ncu_lat.log 2>&1
echo "rc=$?"
ncu -i /tmp/lat_full.ncu-rep --page raw --csv > gpurun_out/lat_raw.csv 2>/dev/null
ncu -i /tmp/lat_full.ncu-rep --page source --csv > gpurun_out/lat_source.csv 2>/dev/null
ncu -i /tmp/lat_full.ncu-rep --page details 2>/dev/null | grep -E "Duration|Issue Slots Busy|Executed Ipc|Theoretical Occupancy|Achieved Occupancy|Bank conflicts|shared|L1/TEX Hit|Warp Cycles Per Issued|Stall|stall|Registers|Shared Memory Config|Dynamic Shared|Block Limit|Executed Instructions |DRAM Throughput|Memory Throughput" | head -60
