timeout 300 python scripts/tc_probe.py bf16 > gpurun_out/probe.log 2>&1; grep -E "layers ok|encoder|bad=[1-9]|rc=" gpurun_out/probe.log
timeout 300 python scripts/tc_seq.py 2048 > gpurun_out/seq.log 2>&1; cat gpurun_out/seq.log
