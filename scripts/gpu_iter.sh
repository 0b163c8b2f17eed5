timeout 300 python scripts/tc_probe.py bf16 > gpurun_out/probe.log 2>&1; grep -E "layers ok|encoder|bad=[1-9]|rc=" gpurun_out/probe.log
timeout 300 python scripts/tc_prof.py > gpurun_out/prof.log 2>&1; tail -12 gpurun_out/prof.log
