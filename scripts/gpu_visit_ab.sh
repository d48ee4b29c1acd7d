#!/usr/bin/env bash
# One GPU-box visit: parity tests + smoke on the in-tree library, then an alternating bench A/B of the in-tree
# library against another build of it (EVI_LIB_PATH), then the ncu launch list of the in-tree build.
# Usage (under gpurun): bash scripts/gpu_visit_ab.sh TAG ROUNDS [path of the other build]
set -uo pipefail
TAG="$1"; R="$2"; OTHER="${3:-}"
OUT=gpurun_out; mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/${TAG}_smi.txt" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s > "$OUT/${TAG}_pytest.log" 2>&1; echo "pytest exit $?"; tail -3 "$OUT/${TAG}_pytest.log"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/${TAG}_smoke.log" 2>&1; echo "smoke exit $?"
show() {
python - "$1" "$2" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
nt=d.get("roofline_non_tensor",{})
print("%-10s value %.0f q/s  ms/step %.3f  e2e %.0f  gemm %.0f TF/s %.3f ms  tail %.3f ms  clocks %s" % (sys.argv[2], d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["achieved"], d["roofline"]["ms_per_launch"], nt.get("ms_per_launch",0), d["clocks"]["sm_mhz"]))
PY
}
for r in $(seq 1 "$R"); do
  timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs > "$OUT/${TAG}_r${r}_new.json" 2> "$OUT/${TAG}_r${r}_new.err"; show "$OUT/${TAG}_r${r}_new.json" new
  if [ -n "$OTHER" ]; then
    EVI_LIB_PATH="$OTHER" timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs > "$OUT/${TAG}_r${r}_old.json" 2> "$OUT/${TAG}_r${r}_old.err"; show "$OUT/${TAG}_r${r}_old.json" old
  fi
done
BARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-other-configs"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file "$OUT/${TAG}_launches.csv" python bench.py $BARGS > "$OUT/${TAG}_ncu_list.log" 2>&1
echo "ncu list exit $?"
python scripts/summarize_launches.py "$OUT/${TAG}_launches.csv" | head -16
