#!/usr/bin/env bash
# Multi-GPU visit: the stream-pipeline tests on one GPU, then bench.py on N ranks the way the driver launches it.
# Usage (under gpurun --gpus N): bash scripts/gpu_visit_multi.sh TAG N
set -uo pipefail
TAG="$1"; N="$2"
OUT=gpurun_out; mkdir -p "$OUT"
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "run_stream or projected_entity_table or non_current_device" > "$OUT/${TAG}_pytest.log" 2>&1; echo "pytest exit $?"; tail -2 "$OUT/${TAG}_pytest.log"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus "$N" --steps 10 --warmup 3 --no-cpu-baseline > "$OUT/${TAG}_bench_n$N.json" 2> "$OUT/${TAG}_bench_n$N.err"; echo "bench exit $?"
python - "$OUT/${TAG}_bench_n$N.json" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value %.0f q/s  ms/step %.3f  e2e %.0f" % (d["value"], d["ms_per_step"], d["e2e"]["value"]), d["clocks"])
for c in d.get("other_configs") or []:
    print("  ", c["case"], "%.0f q/s" % c["questions_per_s"], c.get("imbalance_max_over_mean_edges"))
PY
tail -5 "$OUT/${TAG}_bench_n$N.err"
