timeout 300 python scripts/tc_probe.py bf16 2>&1 | tail -40
