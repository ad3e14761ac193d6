#!/usr/bin/env bash
# Scaling runs on the GPUs of one box: for every N and workload, both arms through bench.py
# (1 process per GPU under torch.distributed.run when N > 1).
# Usage (under gpurun --gpus G): bash tools/gpu_scale.sh <tag> "<N list>" "<workloads>" [steps]
set -uo pipefail
TAG="${1:-rXX}"; NS="${2:-1 2}"; WLS="${3:-c3 c4 c5}"; STEPS="${4:-20}"
OUT=gpurun_out; mkdir -p $OUT
FL="--no-cpu-baseline --no-lbs --no-gpu-baselines"
: > $OUT/${TAG}_scale.txt
for wl in $WLS; do
  for impl in ${IMPLS:-soar reference}; do
    [ "$wl" = "c5" ] && [ "$impl" = "reference" ] && continue
    for N in $NS; do
      f=$OUT/${TAG}_${wl}_${impl}_n${N}
      if [ "$N" = "1" ]; then
        timeout 600 python bench.py --workload $wl --impl $impl --steps $STEPS --warmup 5 $FL > $f.json 2> $f.err
      else
        timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
          bench.py --gpus $N --workload $wl --impl $impl --steps $STEPS --warmup 5 $FL > $f.json 2> $f.err
      fi
      rc=$?
      python - <<PY | tee -a $OUT/${TAG}_scale.txt
import json
try:
    d = json.loads([l for l in open("$f.json") if l.startswith("{")][-1])
    e = d.get("e2e") or {}
    print("$wl $impl N=$N value %.1f %s (%.3f ms/step) e2e %.1f allreduce_ms %s share %s verified %s scaling %s" % (
        d["value"], d["unit"], d["ms_per_step"], e.get("value", -1), d.get("allreduce_ms"),
        d.get("allreduce_share_of_step"), (d.get("allreduce") or {}).get("allreduce_verified"), d.get("scaling")))
except Exception as ex:
    print("$wl $impl N=$N failed rc=$rc", ex)
PY
      [ $rc -ne 0 ] && tail -5 $f.err
    done
  done
done
exit 0
