#!/usr/bin/env bash
# 1-GPU visit: chunk-count sweep of the scoring stage
set -uo pipefail
T="${1:-r03u}"; O=gpurun_out; mkdir -p $O
for rep in 1 2; do
for c in 2 3 4 6 8; do
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs --pipeline-chunks $c > $O/${T}_c${c}_$rep.json 2> $O/${T}_c${c}_$rep.err
  python - $O/${T}_c${c}_$rep.json $c <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("chunks", sys.argv[2], "value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "enqueue", round(d["host_enqueue_ms_per_step"], 2))
PY
done
done
