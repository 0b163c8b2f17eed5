#!/bin/bash
# A/B matrix of the cta_group::2 conv kernel's ring configuration (encoder time per 2048 crops, scripts/tc_time.py)
cd "$(dirname "$0")/.."
run() { echo -n "$* : "; env "$@" python scripts/tc_time.py fp16 2048 2>&1 | tail -1; }
run AT_TC_CG2=0
run AT_TC_CG2=7
run AT_TC_CG2=1
run AT_TC_CG2=2
run AT_TC_CG2=4
run AT_TC_CG2=2 AT_TC_MT128=1
run AT_TC_CG2=2 AT_TC2_MINRA=3
run AT_TC_CG2=2 AT_TC2_RES=0 AT_TC2_WS=8
run AT_TC_CG2=2 AT_TC2_RES=0 AT_TC2_WS=12 AT_TC_MT128=1
run AT_TC_CG2=4 AT_TC2_WS=8
run AT_TC_CG2=4 AT_TC2_WS=9
run AT_TC_CG2=1 AT_TC_MT64=1
run AT_TC_CG2=2 AT_TC_FOLD_ID=3
run AT_TC_CG2=6 AT_TC_FOLD_ID=3 AT_TC2_WS=8
