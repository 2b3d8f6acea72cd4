#!/bin/bash
# exchange route at N GPUs: the default library against every library under build/variants (A/B on one box)
cd "$(dirname "$0")/.."; N=${N:-2}; mkdir -p gpurun_out
shopt -s nullglob
for lib in default build/variants/libfmq_*.so default; do
  if [ "$lib" = default ]; then unset FMQ_LIB_PATH; else export FMQ_LIB_PATH=$PWD/$lib; fi
  timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) \
     bench.py --gpus $N --steps 20 --warmup 3 --multi exchange --no-variants --no-e2e --no-verify > gpurun_out/xab.json 2> gpurun_out/xab.err
  python - "$lib" <<'PY'
import json,sys
try:
    l=json.loads(open("gpurun_out/xab.json").read().strip().splitlines()[-1]); print(sys.argv[1], "-> %.2f G ev/s, %.4f ms/step, ingest %.4f" % (l["value"]/1e9, l["ms_per_step"], l["roofline"]["avg_launch_ms"]))
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
