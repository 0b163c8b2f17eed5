CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tf32x3|tail_|split_|decode_|post_fusion|skeleton|fk_skin|elementwise|copy|crop_res" -c 400 --csv --log-file gpurun_out/launches_tail.csv $CMD > gpurun_out/ncu_tail.log 2>&1; echo rc=$?
