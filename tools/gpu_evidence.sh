#!/usr/bin/env bash
# One GPU-box visit for evidence that needs no timing: `ncu --set full` captures of the stage kernels
# outside the rasterizer frame (KNN, weight query, LBS, SMPL-X, post-ops, Adam/densify, occupancy grid)
# and compute-sanitizer memcheck / racecheck over small instances of every stage.
# Only text comes back (gpurun_out/ is capped at 64 MiB): the .ncu-rep is turned into its raw-page CSV
# on the box and deleted.
# Usage (under gpurun): bash tools/gpu_evidence.sh <tag>
set -uo pipefail
TAG="${1:-rXX}"
OUT=gpurun_out
mkdir -p $OUT
if [ "${SKIP_NCU:-0}" != "1" ]; then
python tools/stages.py full knn query lbs smplx post densify occgrid > $OUT/${TAG}_stages_plain.log 2>&1; echo "stages full exit $?"
timeout 1500 ncu --set full --clock-control none \
    -k regex:"knn_|lbs_|smplx_|post_|adam_|densify_|mask_|rows_|split_|occgrid_|contour_" -c 70 \
    -o /tmp/${TAG}_stages_prof -f python tools/stages.py full knn query lbs smplx post densify occgrid \
    > $OUT/${TAG}_stages_ncu.log 2>&1
echo "ncu stages exit $?"
ncu -i /tmp/${TAG}_stages_prof.ncu-rep --page raw --csv > $OUT/${TAG}_stages_raw.csv 2>/dev/null
ls -la $OUT/${TAG}_stages_raw.csv
fi
if [ "${SKIP_SAN:-0}" != "1" ]; then
python tools/stages.py small > $OUT/${TAG}_stages_small.log 2>&1; echo "stages small exit $?"
for tool in memcheck racecheck; do
  timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python tools/stages.py small \
      > $OUT/${TAG}_sanitizer_${tool}.log 2>&1
  echo "sanitizer $tool exit $?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|stage .* ok|stages: all|Error|error" $OUT/${TAG}_sanitizer_${tool}.log | head -25
  tail -5 $OUT/${TAG}_sanitizer_${tool}.log
done
fi
du -sh $OUT
