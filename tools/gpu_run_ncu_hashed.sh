#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
S="--steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-variants --no-clocks --key-mode hashed"
timeout -s KILL 200 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu-baseline --no-variants --key-mode hashed > gpurun_out/h_bench.json 2>&1
python -c "
import json; l=json.loads(open('gpurun_out/h_bench.json').read().strip().splitlines()[-1]); r=l['roofline']; print('hashed: value %.2f G ev/s step %.4f ingest %.4f fire %.4f' % (l['value']/1e9, l['ms_per_step'], r['avg_launch_ms'], r['fire_chain_ms']))"
timeout -s KILL 200 python bench.py $S > gpurun_out/h_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"ingest_ring" -s 4 -c 1 -o gpurun_out/h_prof python bench.py $S > gpurun_out/h_ncu.log 2>&1
ls -la gpurun_out | grep " h_"
