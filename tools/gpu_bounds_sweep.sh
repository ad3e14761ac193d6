#!/usr/bin/env bash
# On-box experiment: rebuild one kernel with other launch bounds / macros and time the c3 frame's kernels
# (bench.py's per-kernel table: 40 launches each, views back to back on one stream).
# Usage (under gpurun): bash tools/gpu_bounds_sweep.sh <file.cu> '<sed pattern with @N@>' "<values>"
F=soar_b200/csrc/$1; PAT="$2"; VALS="$3"
cp $F /tmp/orig.cu
for n in $VALS; do
  cp /tmp/orig.cu $F
  sed -i "$(echo "$PAT" | sed "s/@N@/$n/g")" $F
  bash soar_b200/csrc/build.sh > /tmp/build.log 2>&1 || { tail -5 /tmp/build.log; continue; }
  python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu-baseline --no-lbs --no-sequential --no-gpu-baselines > /tmp/b.json 2>/dev/null
  python - <<PY
import json
d = json.load(open("/tmp/b.json"))
print("== $1 variant $n: value %.1f" % d["value"], {k: round(v["ms"] / v["launches"] * 1000, 1) for k, v in d["kernels"].items()})
PY
done
cp /tmp/orig.cu $F
