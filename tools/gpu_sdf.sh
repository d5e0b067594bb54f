#!/bin/bash
# SDF query A/B: parity tests + the C1 / map-sized legs under both kernels.
timeout 600 python -m pytest tests/test_neural_gpu.py tests/test_pinmap.py -m gpu -x -q 2>&1 | tail -3
for impl in half warp; do
  PG_SDF_IMPL=$impl timeout 300 python bench.py --no-cpu-baseline --legs sdf --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['extra']['sdf']
print('impl=$impl', round(s['value']/1e6,1), 'M q/s', s['ms_per_step'], 'large', {k:v for k,v in s.get('large_map',{}).items() if k in ('value','ms_per_step')})"
done
