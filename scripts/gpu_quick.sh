#!/usr/bin/env bash
# Quick GPU visit: parity tests, two plain bench runs, ncu launch list. Usage (under gpurun): bash scripts/gpu_quick.sh TAG
set -uo pipefail
T="${1:-quick}"; O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $O/${T}_pytest.log
for i in 1 2; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > $O/${T}_bench$i.json 2> $O/${T}_bench$i.err
  python - "$O/${T}_bench$i.json" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "frac", round(d["roofline"]["frac"], 3),
      "share", round(d["roofline"]["share_of_step"], 3), d["clocks"])
PY
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/${T}_ncu_list.log 2>&1; echo "ncu exit $?"
python scripts/summarize_launches.py $O/${T}_launches.csv 2>/dev/null | head -12
