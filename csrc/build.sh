#!/bin/bash
# Build liberasurehead.so in-tree for sm_100a (cross-compiles without a GPU).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/../erasurehead_b200/ops/liberasurehead.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --use_fast_math -Xptxas -v"
mkdir -p "$HERE/build"
pids=()
for f in "$HERE"/*.cu; do
  o="$HERE/build/$(basename "${f%.cu}").o"
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ -n "$(find "$HERE" "$HERE/../include" -name '*.cuh' -newer "$o" -o -name '*.h' -newer "$o" | head -1)" ]; then
    ( $NVCC $FLAGS -c "$f" -o "$o" > "$o.log" 2>&1 || { cat "$o.log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o "$OUT" "$HERE"/build/*.o -lcudart
echo "built $OUT"
