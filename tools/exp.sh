timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-variants 2> gpurun_out/b1.err | tail -1 > gpurun_out/bench_n1.json; tail -2 gpurun_out/b1.err; python -c "
import json; d=json.load(open('gpurun_out/bench_n1.json')); print(d['value']/1e9,'Gev/s', d['ms_per_step'],'ms/step', d['e2e'])"
