#!/usr/bin/env bash
# A/B of the flat gradient buffer's allocation under NCCL (FlatGrads, soar_b200/views.py):
# plain torch allocation (0) vs ncclMemAlloc pool registered with the communicator (1) vs symmetric window (2).
N=${1:-8}; MODES=${2:-"0 1 2"}; TAG=${3:-r03c}
for mode in $MODES; do
  SOAR_FLAT_NCCL_POOL=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port $((29510 + mode)) bench.py --gpus $N --steps 20 --warmup 5 \
    --no-sequential > gpurun_out/${TAG}_n${N}_pool${mode}.json 2> gpurun_out/${TAG}_n${N}_pool${mode}.err
  echo "mode $mode rc=$?"
  tail -2 gpurun_out/${TAG}_n${N}_pool${mode}.err
done
