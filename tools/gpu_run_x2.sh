#!/bin/bash
# 2 GPUs: exchange parity tests (one process per GPU) and the bench line with the exchange variants
cd "$(dirname "$0")/.."; N=${N:-2}; mkdir -p gpurun_out
timeout -s KILL 600 python -m pytest tests/test_gpu_exchange.py tests/test_gpu_host.py -m gpu -q --timeout 300 -k "one_process_per_gpu or two_gpu or gpus" 2>&1 | tail -3
timeout -s KILL 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) bench.py --gpus $N --steps 20 --warmup 3 --no-e2e > gpurun_out/x${N}_bench.json 2> gpurun_out/x${N}_bench.err
echo "bench rc=$?"
python - <<'PY'
import json,os
N=os.environ.get("N","2")
l=json.loads(open(f"gpurun_out/x{N}_bench.json").read().strip().splitlines()[-1]); r=l["roofline"]
print("value %.2f G ev/s  step %.4f ms  ingest %.4f ms  fire %.4f" % (l["value"]/1e9, l["ms_per_step"], r["avg_launch_ms"], r["fire_chain_ms"]))
for k,v in (l.get("variants") or {}).items(): print("   ", k[:70], {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk in ("value","ms_per_step","nvlink_GBps_per_gpu_each_direction")})
print("    verify:", (l.get("verify_vs_oracle") or {}).get("equal"))
PY
