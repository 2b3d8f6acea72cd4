#!/bin/bash
# exchange route at N GPUs: how scatter (next batch) and ingest (this batch) should share the SMs
cd "$(dirname "$0")/.."
N=${N:-2}; mkdir -p gpurun_out
for cfg in "1 2" "1 3" "2 2" "2 1"; do
  set -- $cfg
  FMQ_XS_BLOCKS_PER_SM=$1 FMQ_XS_INGEST_BLOCKS=$2 timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
     bench.py --gpus $N --steps 20 --warmup 3 --multi exchange --no-variants --no-e2e --no-verify > gpurun_out/x${N}_s$1_i$2.json 2> gpurun_out/x${N}_s$1_i$2.err
  python - "$1" "$2" "gpurun_out/x${N}_s$1_i$2.json" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[3]).read().strip().splitlines()[-1]); print("scatter blocks/SM", sys.argv[1], "ingest blocks/SM", sys.argv[2], "-> %.2f G ev/s, %.4f ms/step" % (l["value"]/1e9, l["ms_per_step"]))
except Exception as e: print(sys.argv[1:], "failed", e)
PY
done
