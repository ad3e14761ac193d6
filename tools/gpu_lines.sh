#!/usr/bin/env bash
# Per-source-line instruction counts of the hot kernels at c3: one `ncu --set full --import-source on`
# capture, then tools/ncu_lines.py on the box (only text comes back).
# Usage (under gpurun): bash tools/gpu_lines.sh <tag> [regex] [skip] [count]
set -uo pipefail
TAG="${1:-rXX}"; REGEX="${2:-blend_bwd2|blend_fwd2}"; SKIP="${3:-20}"; COUNT="${4:-2}"
OUT=gpurun_out; mkdir -p $OUT
SHORT="python bench.py --steps 1 --warmup 1 --views-per-rank 1 --no-e2e --no-cpu-baseline --no-lbs --no-sequential --no-gpu-baselines"
ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s $SKIP -c $COUNT \
    -o /tmp/${TAG}_lines -f $SHORT > $OUT/${TAG}_lines_ncu.log 2>&1
echo "ncu exit $?"
ncu -i /tmp/${TAG}_lines.ncu-rep --page raw --csv > $OUT/${TAG}_lines_raw.csv 2>/dev/null
NCU_KNAME=blend_fwd2_kernel_t python tools/ncu_lines.py /tmp/${TAG}_lines.ncu-rep blend_fwd2_kernel_tILb0ELb0 blend_fwd 60 > $OUT/${TAG}_lines_blend_fwd.txt 2>&1
NCU_KNAME=blend_bwd2_kernel_t python tools/ncu_lines.py /tmp/${TAG}_lines.ncu-rep blend_bwd2_kernel_tILb0ELb0ELb0 blend_bwd 70 > $OUT/${TAG}_lines_blend_bwd.txt 2>&1
head -30 $OUT/${TAG}_lines_blend_fwd.txt
