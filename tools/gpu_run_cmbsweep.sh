#!/bin/bash
# combiner route at N GPUs: ingest blocks per SM while the previous batch's fire kernel runs beside it
cd "$(dirname "$0")/.."
N=${N:-8}; mkdir -p gpurun_out
for b in 4 3 2; do
  FMQ_CMB_INGEST_BLOCKS=$b timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
     bench.py --gpus $N --steps 20 --warmup 3 --no-variants --no-e2e --no-verify > gpurun_out/c${N}_b$b.json 2> gpurun_out/c${N}_b$b.err
  python - "$b" "gpurun_out/c${N}_b$b.json" <<'PY'
import json,sys
try:
    l=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=l["roofline"]; print("ingest blocks/SM", sys.argv[1], "-> %.2f G ev/s, %.4f ms/step, ingest %.4f, fire chain %.4f" % (l["value"]/1e9, l["ms_per_step"], r["avg_launch_ms"], r["fire_chain_ms"]))
except Exception as e: print(sys.argv[1:], "failed", e)
PY
done
