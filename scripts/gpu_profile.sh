# ncu evidence of a round (every ncu pass runs only after the same command exited 0 without ncu).
# Reports are exported to raw CSV on the box; gpurun copies back at most 64 MiB.
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
timeout 200 python scripts/tc_time.py fp16 2048 > gpurun_out/tc_time.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:conv_tc -s 87 -c 29 -o /tmp/tc_full -f python scripts/tc_time.py fp16 2048 > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; cat gpurun_out/tc_time.log
ncu -i /tmp/tc_full.ncu-rep --page raw --csv > gpurun_out/tc_full_raw.csv 2>/dev/null
# the other kernels of the step (resampler, source-box pull, tail convs, FK): one full capture each from the bench
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"crop_resample_fast|crop_pull_boxes|conv_nhwc_f16x3|fk_skin_shared|tail_prepare|decode_kernel" -s 30 -c 16 -o gpurun_out/step_other_full -f $CMD > gpurun_out/ncu_other.log 2>&1
echo "other rc=$?"
ncu -i gpurun_out/step_other_full.ncu-rep --page raw --csv > gpurun_out/step_other_raw.csv 2>/dev/null
ls -la gpurun_out/ /tmp/tc_full.ncu-rep
timeout 300 python scripts/tc_seq.py 2048 > gpurun_out/tc_seq.log 2>&1; tail -3 gpurun_out/tc_seq.log
