timeout 200 python scripts/tc_probe.py bf16 > gpurun_out/probe.log 2>&1; grep -E "layers ok|encoder|bad=[1-9]|rc=" gpurun_out/probe.log
AT_TC_PDL=0 timeout 100 python scripts/tc_time.py bf16 2048
timeout 100 python scripts/tc_seq.py 2048 > gpurun_out/seq.log 2>&1; cut -c1-120 gpurun_out/seq.log | tail -8
