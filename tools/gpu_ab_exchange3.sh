#!/bin/bash
N=${1:-2}
run() {
  env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2959$N \
    bench.py --gpus $N --steps 40 --warmup 10 --no-cpu-baseline --legs none 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$*', d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'exch_ms', d['config'].get('exchange_ms'))"
}
run PG_EXCHANGE=multimem PG_EXCHANGE_CTAS=16
run PG_EXCHANGE=multimem PG_EXCHANGE_CTAS=32
run PG_EXCHANGE=nccl
run PG_EXCHANGE=multimem PG_EXCHANGE_CTAS=8
