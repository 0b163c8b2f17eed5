#!/bin/bash
# L2-resident spans of the first encoder levels (AT_TC_SUB / AT_TC_SUB_LEVELS): encoder ms per 2,048 crops per setting
mkdir -p gpurun_out; out=gpurun_out/sub_experiments.log; : > $out
run() { echo "== $*" >> $out; env "$@" timeout 300 python scripts/encoder_ab.py >> $out 2>gpurun_out/sub.err || { echo "FAILED" >> $out; tail -3 gpurun_out/sub.err >> $out; }; }
run AT_TC_SUB=0
for sub in 256 512 1024; do run AT_TC_SUB=$sub AT_TC_SUB_LEVELS=1; done
run AT_TC_SUB=0
cat $out
