#!/bin/bash
# Coverage run of every kernel family (scripts/sanitize_small.py).  compute-sanitizer itself is closed on this pool
# (profiles/r2_sanitizer_closed_on_pool.txt); on a machine where it is available:
#   compute-sanitizer --tool memcheck  python scripts/sanitize_small.py
#   compute-sanitizer --tool racecheck python scripts/sanitize_small.py
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python scripts/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain exit $?"; tail -6 gpurun_out/sanitize_plain.log
