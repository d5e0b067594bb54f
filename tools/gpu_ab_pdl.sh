#!/bin/bash
# A/B of programmatic dependent launch on the C2 step: same box, alternating runs.
for i in 1 2 3; do
  for v in 0 1; do
    PG_PDL=$v timeout 300 python bench.py --no-cpu-baseline --legs none --steps 40 --warmup 10 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('pdl=$v', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['serial_value'])"
  done
done
