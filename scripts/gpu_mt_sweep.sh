run() { echo "$@"; env "$@" python scripts/tc_time.py bf16 2048 2>&1 | tail -1; }
run AT_X=0
run AT_TC_MT32=2
run AT_TC_MT32=4
run AT_TC_MT32=4 AT_TC_MT64=2
run AT_TC_MT32=4 AT_TC_MT64=4
run AT_TC_MT32=4 AT_TC_MT64=2 AT_TC_MT128=2
run AT_TC_MT32=4 AT_TC_MT64=2 AT_TC_MT128=2 AT_TC_MT256=2
