#!/usr/bin/env bash
# Multi-GPU visit (gpurun --gpus N): bench.py at N ranks, then config 4 with the triple pool edge-partitioned over the
# N GPUs (frontier exchange + candidate merge over NCCL) checked against a single-shard run on rank 0.
# Usage: bash scripts/gpu_multi.sh TAG N [triples nodes queries]
set -uo pipefail
T="${1:-multi}"; N="${2:-2}"; TR="${3:-100000000}"; ND="${4:-30000000}"; Q="${5:-128}"
O=gpurun_out; mkdir -p $O
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $RUN bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > $O/${T}_bench_n$N.json 2> $O/${T}_bench_n$N.err
echo "bench exit $?"; tail -c 1500 $O/${T}_bench_n$N.json
timeout 1200 $RUN scripts/pool_dist_check.py --triples $TR --nodes $ND --queries $Q --hub-seeds --table 67108864 \
  > $O/${T}_pool_n$N.json 2> $O/${T}_pool_n$N.err
echo "pool exit $?"; cat $O/${T}_pool_n$N.json; tail -3 $O/${T}_pool_n$N.err
