#!/bin/bash
# Final multi-GPU measurements of the round: bench.py at N (both arms where cheap), exchange test at N >= 2.
# usage (under gpurun --gpus N): bash tools/gpu_final.sh N
N=${1:-2}; out=gpurun_out/final; mkdir -p $out
if [ "$N" -ge 2 ]; then (timeout 300 python -m pytest tests/test_exchange_gpu.py -m gpu -q 2>&1 | tail -2) > $out/exchange_test_n$N.log; fi
run() { timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 296$1$1 \
  bench.py --gpus $1 --steps 20 --warmup 5 > $out/bench_n$1.json 2> $out/bench_n$1.err; tail -c 400 $out/bench_n$1.json; echo; }
run $N
if [ "$N" -eq 8 ]; then run 4; fi
cat $out/exchange_test_n$N.log 2>/dev/null
