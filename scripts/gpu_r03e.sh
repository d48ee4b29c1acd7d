#!/usr/bin/env bash
# 2-GPU visit: parity tests (incl. the non-current-device test), 2-rank config-4 dry run, 2-rank bench of configs 2 and 3
set -uo pipefail
T="${1:-r03e}"; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -s > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 scripts/pool_scale.py --triples 4000000 --nodes 1200000 --queries 16384 --slab 4096 \
  --worlds 1,2 --profile-exchange --out $O/${T}_pool_small.json > $O/${T}_pool_small.log 2>&1; echo "pool exit $?"; grep '^{' $O/${T}_pool_small.log | cut -c1-900
timeout 900 $TR --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 > $O/${T}_bench_n2.json 2> $O/${T}_bench_n2.err; echo "bench exit $?"
tail -c 600 $O/${T}_bench_n2.json; tail -3 $O/${T}_bench_n2.err
timeout 900 $TR --master-port 29513 bench.py --gpus 2 --config 3 --steps 6 --warmup 3 > $O/${T}_bench3_n2.json 2> $O/${T}_bench3_n2.err; echo "bench3 exit $?"
head -c 700 $O/${T}_bench3_n2.json; tail -3 $O/${T}_bench3_n2.err
