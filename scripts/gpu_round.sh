#!/usr/bin/env bash
# One GPU-box visit: parity tests, smoke, bench (plain), then ncu launch list + full capture of the top kernel.
# Usage (under gpurun): bash scripts/gpu_round.sh [tag]
set -uo pipefail
TAG="${1:-r01}"
OUT=gpurun_out
mkdir -p "$OUT"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > "$OUT/${TAG}_smi.txt" 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s > "$OUT/${TAG}_pytest.log" 2>&1; echo "pytest exit $?" | tee -a "$OUT/${TAG}_pytest.log"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > "$OUT/${TAG}_smoke.log" 2>&1; echo "smoke exit $?" | tee -a "$OUT/${TAG}_smoke.log"
timeout 1200 python bench.py --steps 10 --warmup 3 > "$OUT/${TAG}_bench.json" 2> "$OUT/${TAG}_bench.err"; echo "bench exit $?"
tail -c 3000 "$OUT/${TAG}_bench.json"
BARGS="--steps 2 --warmup 1 --no-cpu-baseline --no-other-configs"
timeout 600 python bench.py $BARGS > "$OUT/${TAG}_plain.log" 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file "$OUT/${TAG}_launches.csv" python bench.py $BARGS > "$OUT/${TAG}_ncu_list.log" 2>&1
echo "ncu list exit $?"
timeout 600 python bench.py $BARGS > "$OUT/${TAG}_plain2.log" 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"gemm2_kernel|score_tail_bulk_kernel|edge_operand_kernel" -s 10 -c 8 \
    -o "$OUT/${TAG}_gemm" -f python bench.py $BARGS > "$OUT/${TAG}_ncu_full.log" 2>&1
echo "ncu full exit $?"
ls -la "$OUT"
