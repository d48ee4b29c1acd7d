#!/usr/bin/env bash
# Build libevi_rag_b200.so (sm_100a) in-tree. Usage: ./build.sh [extra nvcc flags]
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
SRC="$ROOT/evi-rag_b200/csrc"
OUT="$ROOT/evi-rag_b200/libevi_rag_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
"$NVCC" -std=c++17 -O3 -lineinfo -shared -Xcompiler -fPIC \
  -gencode arch=compute_100a,code=sm_100a \
  -I "$ROOT/include" -I "$SRC" "$@" \
  "$SRC/graph_kernels.cu" "$SRC/topk_kernels.cu" "$SRC/scorer_kernels.cu" "$SRC/gagent_kernels.cu" "$SRC/pool_kernels.cu" "$SRC/loss_kernels.cu" \
  -o "$OUT"
echo "built $OUT"
