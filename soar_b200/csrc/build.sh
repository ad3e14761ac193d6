#!/usr/bin/env bash
# Builds libsoar_b200.so (the C-ABI library, include/soar_b200.h) for sm_100a.
# No torch headers are involved: the host side (soar_b200/_capi.py) binds the
# library with ctypes and hands it raw device pointers of torch tensors.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libsoar_b200.so"
OBJ="$HERE/../_build"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS="-O3 -std=c++17 -Xcompiler -fPIC -gencode arch=compute_100a,code=sm_100a -lineinfo ${SOAR_NVCC_EXTRA:-}"
mkdir -p "$OBJ"
pids=()
for f in api prof preprocess binning blend_fwd blend_bwd geom_bwd knn lbs lbs_query smplx postops densify occgrid; do
  src="$HERE/$f.cu"; obj="$OBJ/$f.o"
  newest=$(ls -t "$src" "$HERE"/*.h "$HERE"/*.cuh "$HERE/../../include/soar_b200.h" | head -1)
  if [ ! -f "$obj" ] || [ "$newest" -nt "$obj" ]; then
    $NVCC $FLAGS -c "$src" -o "$obj" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT" "$OBJ"/*.o -lcudart
echo "build: ok -> $OUT"
