"""Developer probe: wall-clock phases of ShardedOperator.process_batch on N GPUs (torchrun)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch, torch.distributed as dist
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
import bench
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
B = 1 << 24
kw = bench.workload_kwargs(capi, 2 * B); kw.update(device=lr, key_mode=1, table_capacity=500000, fired_capacity=1 << 25, walert_capacity=1 << 25, ealert_capacity=1 << 25)
eng = capi.Engine(**kw)
spec = bench.stream_spec(StreamSpec)
ext = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", lr))
batches = [generate_torch(spec, (s * world + rank) * B, B, device=f"cuda:{lr}") for s in range(6)]
torch.cuda.synchronize()
def T():
    torch.cuda.synchronize(); return time.perf_counter()
with torch.cuda.stream(ext):
    for s in range(6):
        ts, key, val = batches[s]
        t0 = T(); m = int(ts.max().item()); t1 = T()
        outs = [torch.empty_like(t) for t in (ts, key, val)]
        counts = eng.partition_by_owner(ts, key, val, world, *outs); t2 = T()
        sc = torch.as_tensor(counts.astype(np.int64), device=ts.device); rc = torch.empty_like(sc)
        dist.all_to_all_single(rc, sc); rcl = rc.cpu().tolist(); t3 = T()
        recv = []
        for col in outs:
            o = torch.empty(sum(rcl), dtype=col.dtype, device=col.device)
            dist.all_to_all_single(o, col, output_split_sizes=rcl, input_split_sizes=[int(c) for c in counts]); recv.append(o)
        t4 = T()
        mm = torch.tensor([m], dtype=torch.int64, device=ts.device); dist.all_reduce(mm, op=dist.ReduceOp.MAX); wm = int(mm.item()) - 5000; t5 = T()
        eng.push_batch(recv[0], recv[1], recv[2], wm); eng.synchronize(); t6 = T()
        if rank == 0: print(f"step {s}: max {1e3*(t1-t0):.2f} split {1e3*(t2-t1):.2f} counts {1e3*(t3-t2):.2f} a2a {1e3*(t4-t3):.2f} wm {1e3*(t5-t4):.2f} push {1e3*(t6-t5):.2f} ms", flush=True)
eng.poll_fired(); dist.destroy_process_group()
