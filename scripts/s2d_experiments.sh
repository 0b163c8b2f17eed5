#!/bin/bash
# layer 1 in its x-space-to-depth form (AT_TC_S2D): encoder ms per 2,048 crops and feature error against the fp32 encoder
mkdir -p gpurun_out; out=gpurun_out/s2d_experiments.log; : > $out
run() { echo "== $*" >> $out; env "$@" timeout 300 python scripts/encoder_ab.py >> $out 2>gpurun_out/s2d.err || { echo "FAILED" >> $out; tail -5 gpurun_out/s2d.err >> $out; }; }
run AT_TC_S2D=0
run AT_TC_S2D=1
run AT_TC_S2D=1 AT_TC_MT64=1
run AT_TC_S2D=0
run AT_TC_S2D=1
cat $out
