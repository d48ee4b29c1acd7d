#!/usr/bin/env bash
# Step-size A/B on one box: questions per step x chunks per step. Usage: bash scripts/gpu_batch_ab.sh TAG
set -uo pipefail
T="${1:-batch}"; O=gpurun_out; mkdir -p $O
for r in 1 2; do
  for cfg in "128 4" "256 8" "192 6" "128 3" "256 6"; do
    set -- $cfg
    timeout 600 python bench.py --steps 12 --warmup 3 --no-cpu-baseline --questions-per-step $1 --pipeline-chunks $2 > $O/${T}_q$1_c$2_r$r.json 2> $O/${T}_q$1_c$2_r$r.err
    python - "$O/${T}_q$1_c$2_r$r.json" "$cfg" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print("q/step x chunks %-8s value %.0f q/s  ms/step %.3f  e2e %.0f  gemm %.0f TF/s  clocks %s" % (sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["achieved"], d["clocks"]["sm_mhz"]))
PY
  done
done
