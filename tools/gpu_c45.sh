#!/usr/bin/env bash
# BASELINE configs 4 and 5 through bench.py on the GPUs of this box (1 process per GPU when N > 1).
# Usage (under gpurun [--gpus N]): bash tools/gpu_c45.sh <tag> <N> [workloads...]
set -uo pipefail
TAG="${1:-rXX}"; N="${2:-1}"; shift 2 || true
WLS="${*:-c4 c5}"
OUT=gpurun_out; mkdir -p $OUT
for wl in $WLS; do
  for impl in soar reference; do
    [ "$wl" = "c5" ] && [ "$impl" = "reference" ] && continue
    if [ "$N" = "1" ]; then
      python bench.py --workload $wl --impl $impl --steps 10 --warmup 3 > $OUT/${TAG}_${wl}_${impl}_n${N}.json 2> $OUT/${TAG}_${wl}_${impl}_n${N}.err
    else
      python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
        bench.py --gpus $N --workload $wl --impl $impl --steps 10 --warmup 3 > $OUT/${TAG}_${wl}_${impl}_n${N}.json 2> $OUT/${TAG}_${wl}_${impl}_n${N}.err
    fi
    echo "$wl $impl n=$N exit $?"; tail -c 1500 $OUT/${TAG}_${wl}_${impl}_n${N}.json; tail -3 $OUT/${TAG}_${wl}_${impl}_n${N}.err
  done
done
