"""One GPU: does firing batch k beside the ingest of batch k+1 (the combiner's overlapped schedule with world = 1) beat the
sequential push?  usage: python tools/exp_overlap_1gpu.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B, K, W = 1 << 24, 20, 3
spec = StreamSpec(n_keys=1_000_000, events_per_ms=1000, late_frac_ppm=1000)
kw = dict(max_batch=B, window_enabled=1, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000, pane_ring=16, agg_mask=11, walert_source=2, walert_threshold=55.0,
          efilter_enabled=1, efilter_threshold=90.0, ealert_capacity=1 << 26, table_capacity=1_000_000, key_mode=1, fired_capacity=1 << 26, walert_capacity=1 << 26)
bs = [generate_torch(spec, s * B, B) for s in range(W + K)]
torch.cuda.synchronize()
for mode in ("sequential", "overlapped", "sequential", "overlapped"):
    eng = capi.Engine(**kw)
    if mode == "overlapped":
        eng.attach_peers(0, 1, [eng.table_ptr]); eng.combiner_overlap(True)
    ext = torch.cuda.ExternalStream(eng.stream)
    def step(b):
        if mode == "overlapped": eng.push_local(*b); eng.fire_global(); eng.retire()
        else: eng.push_batch(*b)
    def drain():
        eng.synchronize(); torch.cuda.synchronize()
        return len(eng.poll_fired(chunk=1 << 24)), len(eng.poll_window_alerts(chunk=1 << 24)), len(eng.poll_event_alerts(chunk=1 << 24))
    for b in bs[:W]: step(b)
    drain()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for b in bs[W:]: step(b)
    if mode == "overlapped":
        e2 = torch.cuda.ExternalStream(eng.fire_stream); ev = torch.cuda.Event(); ev.record(e2); ext.wait_event(ev)
    e1.record(ext); torch.cuda.synchronize()
    rows = drain()
    print(f"{mode:11s} {e0.elapsed_time(e1) / K:.4f} ms/step  rows {rows}  counters {eng.counters()['windows_fired']}", flush=True)
    eng.close()
