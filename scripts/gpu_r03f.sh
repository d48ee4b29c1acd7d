#!/usr/bin/env bash
# 8-GPU visit: pool parity tests, config 4 at its stated size (N = 1, 2, 4, 8 in one launch), 8-rank bench of configs 2 and 3
set -uo pipefail
T="${1:-r03f}"; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_pool.py -m gpu -q -s > $O/${T}_pytest_pool.log 2>&1; echo "pytest exit $?"; tail -2 $O/${T}_pytest_pool.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 1500 $TR --master-port 29511 scripts/pool_scale.py --slab ${SLAB:-16384} --worlds 1,2,4,8 --profile-exchange \
  --out $O/${T}_pool_scale.json > $O/${T}_pool_scale.log 2>&1; echo "pool exit $?"; grep '^{' $O/${T}_pool_scale.log | cut -c1-1000; tail -5 $O/${T}_pool_scale.log | cut -c1-300
timeout 900 $TR --master-port 29512 bench.py --gpus 8 --steps 10 --warmup 3 > $O/${T}_bench_n8.json 2> $O/${T}_bench_n8.err; echo "bench exit $?"
head -c 400 $O/${T}_bench_n8.json; tail -2 $O/${T}_bench_n8.err
timeout 900 $TR --master-port 29513 bench.py --gpus 8 --config 3 --steps 6 --warmup 3 > $O/${T}_bench3_n8.json 2> $O/${T}_bench3_n8.err; echo "bench3 exit $?"
head -c 400 $O/${T}_bench3_n8.json; tail -2 $O/${T}_bench3_n8.err
