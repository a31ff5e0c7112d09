#!/usr/bin/env bash
# e2e of bench.py for several chunk counts of uvmtier_service_replayable_faults_streamed
for c in 1 2 4 8 12 16 32; do
  UVMTIER_E2E_CHUNKS=$c python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-migration 2>/dev/null | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('chunks', $c, 'e2e ms', round(d['e2e']['ms_per_step'],4), 'e2e G/s', round(d['e2e']['value']/1e9,3), 'value G/s', round(d['value']/1e9,3))"
done
