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
# host-side record builder of the writer (CPython extension, plain C; optional: writer.py falls back to its Python loop)
PYINC="$(python -c 'import sysconfig; print(sysconfig.get_paths()["include"])')"
EXT="$(python -c 'import sysconfig; print(sysconfig.get_config_var("EXT_SUFFIX"))')"
gcc -O2 -shared -fPIC -I "$PYINC" "$SRC/records.c" -o "$ROOT/evi-rag_b200/_evi_records$EXT"
echo "built $ROOT/evi-rag_b200/_evi_records$EXT"
