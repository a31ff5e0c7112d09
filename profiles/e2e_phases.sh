#!/usr/bin/env bash
# where the e2e step's time goes: H2D alone, chunked H2D + preprocess, + service without copy-back, everything
for m in copy fetch norb ""; do
  UVMTIER_E2E_MODE=$m python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-migration 2>/dev/null | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('mode', '$m' or 'full', 'e2e ms', round(d['e2e']['ms_per_step'],4))"
done
UVMTIER_GRAPHS=0 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-migration 2>/dev/null | \
    python -c "import sys,json; d=json.loads(sys.stdin.read()); print('no graphs e2e ms', round(d['e2e']['ms_per_step'],4))"
