# ncu evidence of a round (run only after the same commands exited 0 without ncu)
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
python scripts/tc_time.py fp16 2048 > gpurun_out/tc_time.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:conv_tc -s 87 -c 29 -o gpurun_out/tc_full -f python scripts/tc_time.py fp16 2048 > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; cat gpurun_out/tc_time.log; ls -la gpurun_out/*.ncu-rep
timeout 300 python scripts/tc_seq.py 2048 > gpurun_out/tc_seq.log 2>&1; tail -3 gpurun_out/tc_seq.log
timeout 600 python scripts/kernel_roofline.py > gpurun_out/kernel_roofline.log 2>&1; cat gpurun_out/kernel_roofline.log
