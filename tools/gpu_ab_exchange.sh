#!/bin/bash
# N-GPU A/B of the gradient exchange: our multimem kernel vs ncclAllReduce, C2 step (+ C4 mapping step), same box.
# usage (under gpurun --gpus N): bash tools/gpu_ab_exchange.sh N
N=${1:-2}
for ex in ${EXS:-nccl multimem nccl multimem}; do
  PG_EXCHANGE=$ex timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2957$N \
    bench.py --gpus $N --steps 40 --warmup 10 --no-cpu-baseline --legs mapping 2>gpurun_out/ab_ex_$ex.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); m=d['extra'].get('mapping_step',{})
print('exchange=$ex', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'exch_ms', d['config'].get('exchange_ms'), '| C4', m.get('value'), m.get('ms_per_step'), m.get('exchange_backend'), m.get('replicas_in_sync'), m.get('loss_after_steps'))" || tail -5 gpurun_out/ab_ex_$ex.err
done
