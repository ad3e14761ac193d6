#!/usr/bin/env bash
# Sweep the number of CUDA streams the views of a step are spread over (bench.py --streams).
# Usage (under gpurun): bash tools/gpu_streams.sh <tag> [pytest files...]
TAG="${1:-rXX}"; shift
OUT=gpurun_out; mkdir -p $OUT
if [ -n "${*:-}" ]; then timeout 900 python -m pytest $* -m gpu -x -q > $OUT/${TAG}_tests.log 2>&1; tail -5 $OUT/${TAG}_tests.log; fi
FL="--no-cpu-baseline --no-lbs --no-gpu-baselines --no-sequential"
: > $OUT/${TAG}_streams.txt
for wl in ${WLS:-c3 c4}; do
  for k in ${KS:-1 2 3 4 8}; do
    timeout 300 python bench.py --workload $wl --streams $k --steps 20 --warmup 5 $FL > $OUT/${TAG}_${wl}_s$k.json 2> $OUT/${TAG}_${wl}_s$k.err || tail -3 $OUT/${TAG}_${wl}_s$k.err
    python - <<PY | tee -a $OUT/${TAG}_streams.txt
import json
try:
    d = json.load(open("$OUT/${TAG}_${wl}_s$k.json"))
    print("$wl streams=$k value %.1f frames/s (%.3f ms/step)  e2e %.1f" % (d["value"], d["ms_per_step"], d["e2e"]["value"] if d["e2e"] else -1))
except Exception as e:
    print("$wl streams=$k failed", e)
PY
  done
done
