timeout 200 python scripts/tc_probe.py fp16 > gpurun_out/probe.log 2>&1; grep -E "layers ok|bad=[1-9]|rc=|c=2048" gpurun_out/probe.log
AT_TC_DBG=32 timeout 100 python scripts/tc_time.py fp16 2048
timeout 100 python scripts/tc_time.py fp16 2048
AT_TC_DBG=32 timeout 100 python scripts/tc_time.py bf16 2048
timeout 100 python scripts/tc_time.py bf16 2048
