#!/usr/bin/env bash
# 1-GPU visit: slab-size sweep of config 4 (the visited table is L2-resident for small slabs)
set -uo pipefail
T="${1:-r03m}"; O=gpurun_out; mkdir -p $O
for slab in 1024 2048 4096 8192 16384; do
  timeout 600 python scripts/pool_scale.py --queries 131072 --slab $slab --pairs-per-query 1024 --worlds 1 --profile-exchange \
    --out $O/${T}_slab$slab.json > $O/${T}_slab$slab.log 2>&1
  python - $O/${T}_slab$slab.json $slab <<'PY'
import json, sys
r = json.load(open(sys.argv[1]))["results"][0]
print("slab", sys.argv[2], "q/s", round(r["queries_per_s"]), "extract_ms", round(r["extract_ms"]), "score_ms", round(r["score_ms"]), "slots", r["table_slots"], {k: round(v, 3) for k, v in r["stage_s"].items()})
PY
done
