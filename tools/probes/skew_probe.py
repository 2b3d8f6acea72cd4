import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B = 1 << 24
W10 = dict(window_enabled=1, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000, pane_ring=16)
def run(name, spec, **kw):
    eng = capi.Engine(max_batch=B, **kw)
    bs = [generate_torch(spec, s * B, B) for s in range(5)]
    torch.cuda.synchronize(); t = []
    for i, b in enumerate(bs):
        eng.push_batch(*b); eng.synchronize()
        if i >= 2: t.append(eng.push_timing(0)["ingest_ms"])
        eng.poll_fired(chunk=1 << 24)
    print(f"{name}: ingest {np.mean(t):.3f} ms", flush=True); eng.close()
run("C3 dense uniform", StreamSpec(n_keys=1_000_000, late_frac_ppm=1000), agg_mask=11, table_capacity=1_000_000, key_mode=1, fired_capacity=1 << 23, **W10)
run("C3 hashed uniform", StreamSpec(n_keys=1_000_000, late_frac_ppm=1000), agg_mask=11, table_capacity=1_500_000, fired_capacity=1 << 23, **W10)
run("C5 zipf hashed", StreamSpec(n_keys=100_000_000, key_kind="zipf", late_frac_ppm=1000), agg_mask=11, table_capacity=1 << 22, fired_capacity=1 << 24, **W10)
