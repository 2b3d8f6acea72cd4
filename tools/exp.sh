N=$1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 2> gpurun_out/b$N.err | tail -1 > gpurun_out/bench_n$N.json; tail -2 gpurun_out/b$N.err; python -c "
import json; d=json.load(open('gpurun_out/bench_n$N.json')); print(d['n_gpus'], d['value']/1e9,'Gev/s', d['ms_per_step'],'ms/step', d['roofline']['avg_launch_ms'], d['roofline']['fire_chain_ms'], d['clocks'])"
