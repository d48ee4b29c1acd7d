#!/usr/bin/env bash
# 1-GPU visit: GPU tests, then alternating A/B of the [hi | lo] epilogue (staged TMA stores vs per-thread row stores)
set -uo pipefail
T="${1:-r03l}"; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -x > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $O/${T}_pytest.log
for i in 1 2 3; do
  for mode in tma direct; do
    if [ $mode = direct ]; then export EVI_HILO_DIRECT=1; else unset EVI_HILO_DIRECT; fi
    timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs > $O/${T}_${mode}_$i.json 2> $O/${T}_${mode}_$i.err
    python - "$O/${T}_${mode}_$i.json" $mode <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], "value", round(d["value"]), "ms", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"]))
PY
  done
done
unset EVI_HILO_DIRECT
