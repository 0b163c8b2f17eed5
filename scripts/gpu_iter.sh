timeout 200 python scripts/tc_probe.py bf16 > gpurun_out/probe.log 2>&1; grep -E "layers ok|encoder|bad=[1-9]|rc=" gpurun_out/probe.log
timeout 200 python scripts/tc_probe.py bf16 2 > gpurun_out/probe2.log 2>&1; grep -E "layers ok|bad=[1-9]|rc=" gpurun_out/probe2.log
for m in 1 2; do echo "== mt=$m"; timeout 120 python scripts/tc_prof.py $m 2>&1 | grep -E "^layer"; done > gpurun_out/dbg.log 2>&1; cut -c1-150 gpurun_out/dbg.log
