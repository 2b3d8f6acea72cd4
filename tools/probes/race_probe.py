"""Developer probe: checksum conservation of the sliding int64 job under kernel-path variants."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B, nb, k = 1 << 24, 3, 6
def run(name, key_kind, key_mode, cap, misalign, size=60_000, slide=10_000):
    spec = StreamSpec(n_keys=10_000_000, events_per_ms=10_000, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind=key_kind)
    eng = capi.Engine(value_type=1, window_enabled=1, agg_mask=15, window_size_ms=size, window_slide_ms=slide, max_out_of_orderness_ms=5000,
                      key_mode=key_mode, table_capacity=cap, pane_ring=16, max_batch=B + 8, fired_capacity=120_000_000)
    tot = 0
    for b in range(nb):
        ts, key, val = generate_torch(spec, b * B, B + 1)
        o = 1 if misalign else 0
        ts, key, val = ts[o:B + o], key[o:B + o], val[o:B + o]
        torch.cuda.synchronize()
        eng.push_batch(ts, key, val); eng.synchronize()
        tot += int(val.sum().item())
    eng.flush()
    f = eng.poll_fired(chunk=1 << 24); c = eng.counters()
    kk = size // slide
    print(f"{name:40s} rows {len(f)} count_ok {int(f.count.sum()) == kk * nb * B} sum_diff {int(f.sum.sum()) - kk * tot} late {c['late_dropped']}", flush=True)
    eng.close()
run("hashed composite TMA", "host_metric", 0, 15_000_000, False)
run("hashed composite direct loads", "host_metric", 0, 15_000_000, True)
run("dense uniform TMA", "uniform", 1, 10_000_000, False)
run("dense uniform direct", "uniform", 1, 10_000_000, True)
run("dense uniform TMA tumbling", "uniform", 1, 10_000_000, False, 10_000, 10_000)
run("hashed composite TMA again", "host_metric", 0, 15_000_000, False)
