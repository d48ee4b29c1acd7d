#!/usr/bin/env bash
# GPU visit for the configs bench.py does not time: config 3 / config 5 sweep (scripts/bench_configs.py) and config 4
# on one GPU at the full pool size of BASELINE.json (100M triples over 30M nodes; scripts/pool_dist_check.py).
# Usage (under gpurun): bash scripts/gpu_configs.sh TAG
set -uo pipefail
T="${1:-cfg}"; O=gpurun_out; mkdir -p $O
timeout 900 python scripts/bench_configs.py --steps 6 --warmup 2 --sweep > $O/${T}_configs.jsonl 2> $O/${T}_configs.err
echo "configs exit $?"; cut -c1-220 $O/${T}_configs.jsonl
# config 4, popular-entity seeds (dense 2-hop neighbourhoods: the candidate cap binds) and uniform seeds
timeout 900 python scripts/pool_dist_check.py --triples 100000000 --nodes 30000000 --queries 128 --hub-seeds \
  --table 67108864 > $O/${T}_pool_100m_hub.json 2> $O/${T}_pool_100m_hub.err
echo "pool hub exit $?"; cat $O/${T}_pool_100m_hub.json; tail -3 $O/${T}_pool_100m_hub.err
timeout 900 python scripts/pool_dist_check.py --triples 100000000 --nodes 30000000 --queries 4096 \
  > $O/${T}_pool_100m_uniform.json 2> $O/${T}_pool_100m_uniform.err
echo "pool uniform exit $?"; cat $O/${T}_pool_100m_uniform.json; tail -3 $O/${T}_pool_100m_uniform.err
nvidia-smi --query-gpu=memory.used,memory.total --format=csv
