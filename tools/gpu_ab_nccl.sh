#!/bin/bash
# N-GPU C2 step under different NCCL CTA budgets (the all-reduce overlaps the next view; fewer CTAs = fewer SMs taken
# from the rasterizer kernels).  usage (under gpurun --gpus N): bash tools/gpu_ab_nccl.sh N
# B200 x2, measured: default 1593 Mpix/s (exchange 0.175 ms), 2 CTAs 1289 (0.745), 4: 1479, 8: 1575, 16: 1607 -> default kept.
N=${1:-2}
for ctas in default 2 4 8 16; do
  if [ "$ctas" = default ]; then unset NCCL_MAX_CTAS; else export NCCL_MAX_CTAS=$ctas; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N \
    bench.py --gpus $N --steps 40 --warmup 10 --no-cpu-baseline --legs none 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ctas=$ctas', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'exch_ms', d['config'].get('exchange_ms'))"
done
