#!/usr/bin/env bash
# One GPU-box visit: parity tests + smoke, then an alternating bench A/B over bench.py flag sets, then the ncu launch
# list of the default run. Usage (under gpurun): bash scripts/gpu_visit_flags.sh TAG ROUNDS "FLAGS_A" "FLAGS_B" ...
# A variant may carry environment settings before a '|': "EVI_LIB_PATH=build/other.so | --struct-norm analytic".
set -uo pipefail
TAG="$1"; R="$2"; shift 2
OUT=gpurun_out; mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/${TAG}_smi.txt" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s > "$OUT/${TAG}_pytest.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/${TAG}_pytest.log"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/${TAG}_smoke.log" 2>&1; echo "smoke exit $?"
for r in $(seq 1 "$R"); do
  i=0
  for flags in "$@"; do
    i=$((i+1))
    envs=""; args="$flags"
    case "$flags" in *"|"*) envs="${flags%%|*}"; args="${flags#*|}";; esac
    env $envs timeout 600 python bench.py --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline --no-other-configs $args > "$OUT/${TAG}_r${r}_v$i.json" 2> "$OUT/${TAG}_r${r}_v$i.err"
    python - "$OUT/${TAG}_r${r}_v$i.json" "v$i [$flags]" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    nt=d.get("roofline_non_tensor",{})
    print("%-28s value %.0f q/s  ms/step %.3f  e2e %.0f  dropin %.0f  gemm %.0f TF/s %.3f ms  tail %.3f ms  clocks %s" % (sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], (d.get("e2e_dropin") or {}).get("value",0), d["roofline"]["achieved"], d["roofline"]["ms_per_launch"], nt.get("ms_per_launch",0), d["clocks"]["sm_mhz"]))
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
  done
done
BARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-other-configs"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file "$OUT/${TAG}_launches.csv" python bench.py $BARGS > "$OUT/${TAG}_ncu_list.log" 2>&1
echo "ncu list exit $?"
python scripts/summarize_launches.py "$OUT/${TAG}_launches.csv" > "$OUT/${TAG}_launches.md" 2>/dev/null; head -22 "$OUT/${TAG}_launches.md"
