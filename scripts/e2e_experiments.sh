#!/bin/bash
# e2e arm of bench.py under ingest variants (same box, back to back)
cd "$(dirname "$0")/.."
run() { echo -n "$* : "; env "$@" python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-parity ${ARGS:-} 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']), 'e2e', round(d['e2e']['value']), 'poses', round(d['e2e_from_poses']['value']) if d.get('e2e_from_poses') else None, 'enc', round(d['stages']['encoder_ms'],3))"; }
run AT_INGEST_RELEASE=crops AT_INGEST_PRIORITY=-1
run AT_INGEST_RELEASE=crops AT_INGEST_PRIORITY=0
run AT_INGEST_RELEASE=encoder AT_INGEST_PRIORITY=-1
run AT_INGEST_RELEASE=encoder AT_INGEST_PRIORITY=0
run AT_INGEST_RELEASE=encoder AT_INGEST_PRIORITY=0 AT_PULL_DEPTH=2
ARGS="--pull-ctas 296" run AT_INGEST_RELEASE=encoder AT_INGEST_PRIORITY=0
AT_INGEST_RELEASE=encoder AT_INGEST_PRIORITY=0 python scripts/ingest_trace.py pull | tail -4
