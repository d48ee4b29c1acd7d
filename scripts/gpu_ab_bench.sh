#!/usr/bin/env bash
# Alternating bench A/B on one box (no tests): bash scripts/gpu_ab_bench.sh TAG ROUNDS "ENVA=.." "ENVB=.."
set -uo pipefail
TAG="$1"; R="$2"; shift 2
OUT=gpurun_out; mkdir -p "$OUT"
for r in $(seq 1 "$R"); do
  i=0
  for envset in "$@"; do
    i=$((i+1))
    env $envset timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > "$OUT/${TAG}_r${r}_v$i.json" 2> "$OUT/${TAG}_r${r}_v$i.err"
    python - "$OUT/${TAG}_r${r}_v$i.json" "$envset" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print("%-14s value %.0f q/s  ms/step %.3f  e2e %.0f  gemm %.0f TF/s %.3f ms  clocks %s" % (sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["achieved"], d["roofline"]["ms_per_launch"], d["clocks"]["sm_mhz"]))
PY
  done
done
BARGS="--steps 2 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file "$OUT/${TAG}_launches.csv" python bench.py $BARGS > "$OUT/${TAG}_ncu_list.log" 2>&1
python scripts/summarize_launches.py "$OUT/${TAG}_launches.csv" | head -12
