timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 2> gpurun_out/b2.err | tail -1 > gpurun_out/bench_n2.json; python -c "
import json; d=json.load(open('gpurun_out/bench_n2.json')); print(d['value']/1e9,'Gev/s', d['ms_per_step'],'ms/step', d['roofline']['avg_launch_ms'], d['clocks'])"
