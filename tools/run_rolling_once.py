"""A few pushes of chapter2's rolling max (keyBy(0).max(2)), 100 K hosts, dense ids.  usage: python tools/run_rolling_once.py [pushes]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B = 1 << 24
eng = capi.Engine(max_batch=B, rolling_max_enabled=1, table_capacity=100_000, key_mode=1)
spec = StreamSpec(n_keys=100_000)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    cols = generate_torch(spec, i * B, B); torch.cuda.synchronize()
    eng.push_batch(*cols); eng.synchronize()
    print(i, eng.push_timing(0), flush=True)
