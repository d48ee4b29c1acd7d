#!/usr/bin/env bash
# A/B on one box: parity tests, then bench with each listed environment setting. Usage: bash scripts/gpu_ab.sh tag "ENV1=.." "ENV2=.."
set -uo pipefail
TAG="$1"; shift
OUT=gpurun_out; mkdir -p "$OUT"
timeout 900 python -m pytest tests -m gpu -x -q -s > "$OUT/${TAG}_pytest.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/${TAG}_pytest.log"
i=0
for envset in "$@"; do
  i=$((i+1))
  echo "== variant $i: $envset"
  env $envset timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > "$OUT/${TAG}_bench_v$i.json" 2> "$OUT/${TAG}_bench_v$i.err"; echo "bench exit $?"
  python - "$OUT/${TAG}_bench_v$i.json" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print("value %.0f q/s  ms/step %.3f  e2e %.0f  gemm %.0f TF/s  clocks %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["achieved"], d["clocks"]))
PY
done
BARGS="--steps 2 --warmup 1 --no-cpu-baseline"
timeout 600 python bench.py $BARGS > "$OUT/${TAG}_plain.log" 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file "$OUT/${TAG}_launches.csv" python bench.py $BARGS > "$OUT/${TAG}_ncu_list.log" 2>&1
echo "ncu list exit $?"
