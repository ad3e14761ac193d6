#!/usr/bin/env bash
# On-box experiment: rebuild one kernel with other launch bounds and time the c3 frame's kernels.
# Usage (under gpurun): bash tools/gpu_bounds_sweep.sh <file.cu> '<sed pattern with N>' "<values>"
F=soar_b200/csrc/$1; PAT="$2"; VALS="$3"
cp $F /tmp/orig.cu
for n in $VALS; do
  cp /tmp/orig.cu $F
  sed -i "$(echo "$PAT" | sed "s/@N@/$n/g")" $F
  bash soar_b200/csrc/build.sh > /tmp/build.log 2>&1 || { tail -5 /tmp/build.log; continue; }
  echo "== $1 variant $n"; python tools/tile_stats.py c3 | tail -1
done
cp /tmp/orig.cu $F
