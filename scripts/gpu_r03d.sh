#!/usr/bin/env bash
# GPU visit: parity tests, a small single-GPU dry run of the config-4 scaling driver, one bench run
set -uo pipefail
T="${1:-r03d}"; O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -s > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
timeout 600 python scripts/pool_scale.py --triples 4000000 --nodes 1200000 --queries 16384 --slab 4096 --worlds 1 \
  --out $O/${T}_pool_small.json > $O/${T}_pool_small.log 2>&1; echo "pool exit $?"; tail -3 $O/${T}_pool_small.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err; echo "bench exit $?"
tail -c 1500 $O/${T}_bench.json
