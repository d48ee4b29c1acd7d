#!/usr/bin/env bash
# 1-GPU visit: all GPU tests (incl. the training path), smoke
set -uo pipefail
T="${1:-r03j}"; O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -s > $O/${T}_pytest.log 2>&1; echo "pytest exit $?"; grep -n "tc_matmul\|training" $O/${T}_pytest.log | head -20; tail -5 $O/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 $O/${T}_smoke.log
