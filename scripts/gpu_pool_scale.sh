#!/usr/bin/env bash
# 8-GPU visit: pool parity tests, stage breakdown of config 4 at N = 8 (profile mode), then the full 1M-query curve
set -uo pipefail
T="${1:-r03g}"; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_pool.py -m gpu -q > $O/${T}_pytest_pool.log 2>&1; echo "pytest exit $?"; tail -2 $O/${T}_pytest_pool.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 scripts/pool_scale.py --queries 262144 --slab 16384 --worlds 8 --profile-exchange \
  --out $O/${T}_pool_profile.json > $O/${T}_pool_profile.log 2>&1; echo "profile exit $?"; grep '^{' $O/${T}_pool_profile.log | cut -c1-1300; tail -3 $O/${T}_pool_profile.log | cut -c1-300
timeout 1500 $TR --master-port 29512 scripts/pool_scale.py --slab 16384 --worlds 1,2,4,8 \
  --out $O/${T}_pool_scale.json > $O/${T}_pool_scale.log 2>&1; echo "scale exit $?"; grep '^{' $O/${T}_pool_scale.log | cut -c1-700; tail -3 $O/${T}_pool_scale.log | cut -c1-300
