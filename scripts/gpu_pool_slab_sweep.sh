#!/usr/bin/env bash
# 1-GPU visit: pool parity tests with the long-row work list, then the slab sweep again
set -uo pipefail
T="${1:-r03n}"; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_pool.py -m gpu -q > $O/${T}_pytest_pool.log 2>&1; echo "pytest exit $?"; tail -2 $O/${T}_pytest_pool.log
for slab in 2048 4096 8192 16384; do
  timeout 600 python scripts/pool_scale.py --queries 131072 --slab $slab --pairs-per-query 1024 --worlds 1 --profile-exchange \
    --out $O/${T}_slab$slab.json > $O/${T}_slab$slab.log 2>&1
  python - $O/${T}_slab$slab.json $slab <<'PY'
import json, sys
r = json.load(open(sys.argv[1]))["results"][0]
print("slab", sys.argv[2], "q/s", round(r["queries_per_s"]), "extract_ms", round(r["extract_ms"]), "score_ms", round(r["score_ms"]), "slots", r["table_slots"], {k: round(v, 3) for k, v in r["stage_s"].items()})
PY
done
