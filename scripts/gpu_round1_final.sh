# Round-end evidence: GPU tests, smoke, bench (+ reference arm), timing logs, ncu launch list and full captures.
# Every ncu pass runs only after the same command exited 0 without ncu.
set -x
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 400 gpurun_out/bench_reference.json
timeout 300 python scripts/kernel_roofline.py > gpurun_out/kernel_roofline.log 2>&1; tail -5 gpurun_out/kernel_roofline.log
timeout 300 python scripts/tc_seq.py 2048 > gpurun_out/tc_seq.log 2>&1; tail -3 gpurun_out/tc_seq.log
timeout 100 python scripts/ingest_trace.py pull > gpurun_out/ingest_trace.log 2>&1
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
timeout 200 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_bench.csv $CMD > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
timeout 200 python scripts/tc_time.py fp16 2048 > gpurun_out/tc_time.log 2>&1 && \
timeout 900 ncu --set full --clock-control none -k regex:conv_tc -s 78 -c 26 -o /tmp/tc_full -f python scripts/tc_time.py fp16 2048 > gpurun_out/ncu_full.log 2>&1
echo "full rc=$?"; cat gpurun_out/tc_time.log
ncu -i /tmp/tc_full.ncu-rep --page raw --csv > gpurun_out/tc_full_raw.csv 2>/dev/null
# the other kernels of one step: resampler, crop-camera generator, tail (incl. the tcgen05 head convs), FK
timeout 900 ncu --set full --clock-control none -k regex:"crop_resample_lattice|crop_cameras_|eval_accumulate|conv_nhwc_f16x3|conv_tc_kernel<9, 5|head_split|fk_skin_shared|tail_prepare|post_fusion|split_kernel|decode_kernel" -s 40 -c 24 -o /tmp/step_other_full -f $CMD > gpurun_out/ncu_other.log 2>&1
echo "other rc=$?"
ncu -i /tmp/step_other_full.ncu-rep --page raw --csv > gpurun_out/step_other_raw.csv 2>/dev/null
ls -la gpurun_out/
