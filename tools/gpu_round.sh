#!/usr/bin/env bash
# One GPU-box visit: parity tests, both bench arms, ncu launch list + full capture of the hot kernels.
# Usage (from the repo root, under gpurun): bash tools/gpu_round.sh <tag>
set -uo pipefail
TAG="${1:-rXX}"
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.csv 2>&1
if [ "${SKIP_TESTS:-0}" != "1" ]; then
rm -f $OUT/parity_report.jsonl
python -m pytest tests -m gpu ${PYTEST_X--x} -q -rA > $OUT/${TAG}_gpu_tests.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_gpu_tests.log
tail -3 $OUT/${TAG}_gpu_tests.log
grep -E "^(FAILED|ERROR)" $OUT/${TAG}_gpu_tests.log | head -20
[ -f $OUT/parity_report.jsonl ] && mv $OUT/parity_report.jsonl $OUT/${TAG}_parity_report.jsonl
fi
if [ "${SKIP_REF:-0}" != "1" ]; then
python bench.py --impl reference --steps 10 --warmup 3 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "ref bench exit $?"
fi
python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench_soar.json 2> $OUT/${TAG}_bench_soar.err; echo "soar bench exit $?"
cat $OUT/${TAG}_bench_reference.json $OUT/${TAG}_bench_soar.json
if [ "${SKIP_NCU:-0}" != "1" ]; then
SHORT="python bench.py --steps 1 --warmup 1 --views-per-rank 1 --no-e2e --no-cpu-baseline --no-lbs --no-sequential --no-gpu-baselines"
$SHORT > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $SHORT > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
ncu --set full --clock-control none --import-source on \
    -k regex:"${NCU_REGEX:-blend_bwd|blend_fwd|onesweep|duplicate|geom_bwd|preprocess_fwd}" -s ${NCU_SKIP:-93} -c ${NCU_COUNT:-26} \
    -o $OUT/${TAG}_prof -f $SHORT > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
fi
