#!/usr/bin/env bash
# 8-GPU visit: stage breakdown of the config-4 extraction at N = 1 and N = 8 (profile mode synchronises per stage)
set -uo pipefail
T="${1:-r03g}"; O=gpurun_out; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 scripts/pool_scale.py --queries 262144 --slab 16384 --worlds 1,8 --profile-exchange \
  --out $O/${T}_pool_profile.json > $O/${T}_pool_profile.log 2>&1; echo "pool exit $?"; grep '^{' $O/${T}_pool_profile.log | cut -c1-1200; tail -3 $O/${T}_pool_profile.log | cut -c1-300
