"""Developer probe: device time of the chapter1 filter path and the keyed ValueState path."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B = 1 << int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
def run(name, spec, **kw):
    eng = capi.Engine(max_batch=B, **kw)
    bs = [generate_torch(spec, s * B, B) for s in range(5)]
    torch.cuda.synchronize()
    for b in bs[:2]:
        eng.push_batch(*b); eng.synchronize()
        if kw.get("efilter_enabled"): eng.poll_event_alerts()
        if kw.get("consec_n"): eng.poll_consec_alerts()
    ts = []
    for b in bs[2:]:
        eng.push_batch(*b); eng.synchronize(); t = eng.push_timing(0); ts.append(t)
        if kw.get("efilter_enabled"): eng.poll_event_alerts()
        if kw.get("consec_n"): eng.poll_consec_alerts()
    tot = np.mean([t["total_ms"] for t in ts])
    print(f"{name}: total {tot:.3f} ms  ingest {np.mean([t['ingest_ms'] for t in ts]):.3f} keyed {np.mean([t['keyed_ms'] for t in ts]):.3f} fire {np.mean([t['fire_ms'] for t in ts]):.3f}  -> {B/tot/1e6:.2f} Gev/s", eng.counters(), flush=True)
    eng.close()
run("C1 filter", StreamSpec(n_keys=800_000), efilter_enabled=1, efilter_threshold=90.0, ealert_capacity=int(B * 0.12))
run("C2 consec3 100K keys hashed", StreamSpec(n_keys=100_000), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=150_000, calert_capacity=B // 8)
run("C2 consec3 100K keys dense", StreamSpec(n_keys=100_000), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=100_000, key_mode=1, calert_capacity=B // 8)
run("rolling max 100K keys dense", StreamSpec(n_keys=100_000), rolling_max_enabled=1, table_capacity=100_000, key_mode=1)
