import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B = 1 << 24
def run(name, spec, **kw):
    eng = capi.Engine(max_batch=B, **kw)
    bs = [generate_torch(spec, s * B, B) for s in range(4)]
    torch.cuda.synchronize(); t = []
    for i, b in enumerate(bs):
        eng.push_batch(*b); eng.synchronize()
        if i >= 1: t.append(eng.push_timing(0)["keyed_ms"])
        eng.poll_consec_alerts()
    print(f"{name}: keyed {np.mean(t):.3f} ms", flush=True); eng.close()
run("C2 dense 100K", StreamSpec(n_keys=100_000), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=100_000, key_mode=1, calert_capacity=B // 8)
run("C5 zipf consec hashed", StreamSpec(n_keys=100_000_000, key_kind="zipf"), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=1 << 22, calert_capacity=B // 8)
run("C5 zipf rolling max hashed", StreamSpec(n_keys=100_000_000, key_kind="zipf"), rolling_max_enabled=1, table_capacity=1 << 22)
