for cfg in "1 1" "2 1" "4 1" "2 2" "4 2" "1 2"; do set -- $cfg; echo "MT128=$1 MT256=$2"; AT_TC_MT128=$1 AT_TC_MT256=$2 python scripts/tc_time.py bf16 2048 2>&1 | tail -1; done
echo "MT64=2"; AT_TC_MT64=2 python scripts/tc_time.py bf16 2048 2>&1 | tail -1
echo "MT32=2 MT64=2"; AT_TC_MT32=2 AT_TC_MT64=2 python scripts/tc_time.py bf16 2048 2>&1 | tail -1
