#!/bin/bash
# query_feature A/B: parity tests + the training-shaped fwd+bwd sub-result of the sdf leg and the warp SDF kernel
timeout 600 python -m pytest tests/test_neural_gpu.py tests/test_pinmap.py tests/test_tracker_gpu.py -m gpu -x -q 2>&1 | tail -2
for impl in half warp; do
PG_SDF_IMPL=$impl timeout 300 python bench.py --no-cpu-baseline --legs sdf --steps 30 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d['extra']['sdf']
print('sdf impl=$impl', round(s['value']/1e6,1), 'M q/s | train fwd+bwd bs10000', s['train_fwd_bwd_bs10000'], '| large', round(s['large_map']['value']/1e6,1))"
done
