#!/usr/bin/env bash
# Builds libmingpt_b200.so in-tree (next to this script's parent package) for sm_100a.
# Usage: csrc/build.sh [extra nvcc flags]
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
ROOT="$(cd "$HERE/../.." && pwd)"
OUT="$HERE/../libmingpt_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17
       -Xcompiler -fPIC -I"$ROOT/include" -I"$HERE" "$@")
OBJS=()
pids=()
for f in gemm_sm100 gemm2_sm100 rowwise optim attention attention_sm100 dropout; do
  # the optimizer / norm kernels keep IEEE division and sqrt (parity with torch.optim.AdamW); the rest use fast math
  FM=(--use_fast_math); [ "$f" = optim ] && FM=()
  "$NVCC" "${FLAGS[@]}" "${FM[@]}" -c "$HERE/$f.cu" -o "$HERE/$f.o" &
  pids+=($!)
  OBJS+=("$HERE/$f.o")
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -o "$OUT" "${OBJS[@]}" -lcudart
# stamp: hash of the sources this binary was built from (_lib.py refuses to load a binary whose stamp differs)
(cd "$HERE" && LC_ALL=C ls *.cu *.cuh *.sh | LC_ALL=C sort | xargs cat; cat "$ROOT/include/mingpt_b200.h") | sha256sum | cut -d' ' -f1 > "$OUT.srchash"
echo "built $OUT"
