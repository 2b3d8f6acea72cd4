#!/bin/bash
# usage: tools/bench_n.sh N [extra bench.py flags]   — launches bench.py the way the driver does
N=$1; shift
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
  timeout 600 python bench.py --gpus 1 "$@" 2> gpurun_out/b$N.err | tail -1 > gpurun_out/bench_n$N.json
else
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@" 2> gpurun_out/b$N.err | tail -1 > gpurun_out/bench_n$N.json
fi
grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/b$N.err | tail -3
python - <<PY
import json
d = json.load(open('gpurun_out/bench_n$N.json'))
r = d['roofline']
print(d["n_gpus"], "GPUs", "issue", round(r.get("host_issue_ms_per_step", 0), 4), round(d['value'] / 1e9, 2), 'Gev/s', round(d['ms_per_step'], 4), 'ms/step  ingest', round(r['avg_launch_ms'], 4), 'fire', round(r['fire_chain_ms'], 4), 'frac', round(r['frac'], 4), d['clocks'])
PY
