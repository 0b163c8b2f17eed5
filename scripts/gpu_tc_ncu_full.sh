python scripts/tc_time.py bf16 2048 > gpurun_out/tc_time.log 2>&1 || exit 1
for idx in 1 7 14 25; do
  ncu --set full --clock-control none --import-source on -k regex:conv_tc -s $((112 + idx)) -c 1 -o gpurun_out/tc_full_$idx -f python scripts/tc_time.py bf16 2048 > gpurun_out/ncu_$idx.log 2>&1
  echo "idx $idx rc=$?"
done
ls -la gpurun_out/*.ncu-rep
