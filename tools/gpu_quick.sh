#!/usr/bin/env bash
# Quick GPU visit: rasterizer parity tests + a short c3 bench with the per-kernel table.
# Usage (under gpurun): bash tools/gpu_quick.sh <tag> [pytest files...]
TAG="${1:-rXX}"; shift
FILES="${*:-tests/test_rasterizer_gpu.py tests/test_views_gpu.py tests/test_pipeline_gpu.py}"
timeout 900 python -m pytest $FILES -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; tail -5 gpurun_out/${TAG}_tests.log
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-gpu-baselines ${BENCH_FLAGS:-} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench.json"))
print("value", d["value"], "e2e", d["e2e"] and d["e2e"]["value"], "sequential", d["sequential"] and d["sequential"]["value"])
print({k: round(v["ms"] / v["launches"] * 1000, 1) for k, v in d["kernels"].items()})
if d.get("c5_step"): print(d["c5_step"]["ms_per_step"], d["c5_step"]["kernel_ms_per_step"])
PY
tail -3 gpurun_out/${TAG}_bench.err
