"""Developer probe: host wall-clock per phase of ShardedOperator (no added syncs)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
from monitor_systam_flink_quickstart_b200 import capi, shard
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
import bench
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
B = 1 << 24
kw = bench.workload_kwargs(capi, 2 * B); kw.update(device=lr, key_mode=1, table_capacity=500000, fired_capacity=1 << 26, walert_capacity=1 << 26, ealert_capacity=1 << 26)
eng = capi.Engine(**kw)
spec = bench.stream_spec(StreamSpec)
batches = [generate_torch(spec, (s * world + rank) * B, B, device=f"cuda:{lr}") for s in range(10)]
torch.cuda.synchronize()
op = shard.ShardedOperator(eng, rank, world, 5000)
# monkeypatch timers
import types
orig_part = op._partition
marks = []
def timed_part(*a):
    t=time.perf_counter(); r=orig_part(*a); marks.append(("split", time.perf_counter()-t)); return r
op._partition = timed_part
orig_a2a = dist.all_to_all_single
def timed_a2a(*a, **k):
    t=time.perf_counter(); r=orig_a2a(*a, **k); marks.append(("a2a", time.perf_counter()-t)); return r
dist.all_to_all_single = timed_a2a
orig_push = eng.push_batch
def timed_push(*a, **k):
    t=time.perf_counter(); r=orig_push(*a, **k); marks.append(("push", time.perf_counter()-t)); return r
eng.push_batch = timed_push
for s in range(10):
    marks.clear(); t0=time.perf_counter()
    op.process_batch(*batches[s])
    tot=time.perf_counter()-t0
    if rank==0: print(f"step {s}: total {tot*1e3:.2f} ms  " + " ".join(f"{k}={v*1e3:.2f}" for k,v in marks), flush=True)
eng.synchronize(); dist.destroy_process_group()
