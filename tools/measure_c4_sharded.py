"""BASELINE.json configs[3] in its named form: 60 s sliding window (10 s slide) per (host, metric) sum + Mbps alert,
10 M keys, int64 flows, hashed keys, keyBy hash-sharded across N GPUs (NCCL all-to-all or peer-pull over NVLink).
Weak scaling by event rate like bench.py.  Launch with torchrun; rank 0 prints one JSON line.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/measure_c4_sharded.py [alltoall|peerpull] [steps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.shard import PeerPullOperator, ShardedOperator
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch

route = sys.argv[1] if len(sys.argv) > 1 else "alltoall"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 6
W, B, BOUND = 3, 1 << 24, 5000
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
keys = 10_000_000
spec = StreamSpec(n_keys=keys, events_per_ms=10_000 * world, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind="host_metric")
eng = capi.Engine(device=lr, value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, window_size_ms=60_000, window_slide_ms=10_000,
                  max_out_of_orderness_ms=BOUND, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                  walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.05, table_capacity=int(keys * 1.5 / world) + 1024, pane_ring=16,
                  max_batch=2 * B, fired_capacity=1 << 26, walert_capacity=1 << 26)
op = (PeerPullOperator if route == "peerpull" else ShardedOperator)(eng, rank, world, BOUND)
batches = [generate_torch(spec, (s * world + rank) * B, B, device=f"cuda:{lr}") for s in range(W + K)]
torch.cuda.synchronize()
stream = torch.cuda.ExternalStream(eng.stream, device=torch.device("cuda", lr))
for s in range(W):
    op.process_batch(*batches[s])
eng.synchronize(); eng.poll_fired(chunk=1 << 24); eng.poll_window_alerts(chunk=1 << 24)
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for s in range(W, W + K):
    op.process_batch(*batches[s])
e1.record(stream)
eng.synchronize(); torch.cuda.synchronize(); dist.barrier()
t = torch.tensor([e0.elapsed_time(e1)], device=f"cuda:{lr}", dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
c = eng.counters()
tot = torch.tensor([c["events_in"], c["windows_fired"], c["window_alerts"], c["keys"]], device=f"cuda:{lr}", dtype=torch.int64)
dist.all_reduce(tot)
if rank == 0:
    ms = float(t.item())
    print(json.dumps({"workload": "C4 60s/10s sliding sum + Mbps alert, 10M (host,metric) keys, i64, hashed, keyBy-sharded", "route": route, "n_gpus": world,
                      "events_per_step_per_gpu": B, "steps": K, "ms_per_step": ms / K, "value": world * B * K / (ms * 1e-3), "unit": "events/s",
                      "events_in": int(tot[0]), "windows_fired": int(tot[1]), "window_alerts": int(tot[2]), "keys": int(tot[3])}))
dist.destroy_process_group()
