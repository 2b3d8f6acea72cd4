"""Device-time measurements of every BASELINE.json config and of the SURVEY §8f rows on one B200.
Prints a markdown table (kept in profiles/).  usage: python tools/measure_configs.py [log2 batch]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
B = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 24)
rows = []

ONLY = os.environ.get("MEASURE_ONLY")                   # substring of the row name: measure just those rows

def run(name, spec, nb=5, drain=(), **kw):
    if ONLY and ONLY not in name:
        return          # (MEASURE_ONLY=A4: only the owner split at the end)
    eng = capi.Engine(max_batch=B, **kw)
    bs = [generate_torch(spec, s * B, B) for s in range(nb)]
    torch.cuda.synchronize()
    t = []
    for i, b in enumerate(bs):
        eng.push_batch(*b); eng.synchronize()
        if i >= 2: t.append(eng.push_timing(0))
        for d in drain: getattr(eng, d)(chunk=1 << 24)
    c = eng.counters()
    tot = np.mean([x["total_ms"] for x in t])
    rows.append(f"| {name} | {tot:.3f} | {np.mean([x['ingest_ms'] for x in t]):.3f} | {np.mean([x['keyed_ms'] for x in t]):.3f} | {np.mean([x['fire_ms'] for x in t]):.3f} | {B / tot / 1e6:.1f} | {24 * B / tot / 1e6 / 6550.7:.3f} |")
    print(rows[-1], c, flush=True)
    eng.close()

W10 = dict(window_enabled=1, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000, pane_ring=16)
run("C1 chapter1 filter `>90` (f64)", StreamSpec(n_keys=800_000), drain=("poll_event_alerts",), efilter_enabled=1, efilter_threshold=90.0, ealert_capacity=int(B * 0.12))
run("C2 3-consecutive, 100 K keys, dense ids", StreamSpec(n_keys=100_000), drain=("poll_consec_alerts",), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1,
    table_capacity=100_000, key_mode=1, calert_capacity=B // 8)
run("C2 3-consecutive, 100 K keys, hashed", StreamSpec(n_keys=100_000), drain=("poll_consec_alerts",), consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1,
    table_capacity=150_000, calert_capacity=B // 8)
run("chapter2 rolling max, 100 K keys, dense", StreamSpec(n_keys=100_000), rolling_max_enabled=1, table_capacity=100_000, key_mode=1)
run("C3 10 s tumbling count/avg/max + alert, 1 M keys, dense", StreamSpec(n_keys=1_000_000, events_per_ms=1000, late_frac_ppm=1000), drain=("poll_fired", "poll_window_alerts"),
    agg_mask=11, walert_source=2, walert_threshold=55.0, table_capacity=1_000_000, key_mode=1, fired_capacity=1 << 23, walert_capacity=1 << 23, **W10)
run("C3 same, hashed keys", StreamSpec(n_keys=1_000_000, events_per_ms=1000, late_frac_ppm=1000), drain=("poll_fired", "poll_window_alerts"),
    agg_mask=11, walert_source=2, walert_threshold=55.0, table_capacity=1_500_000, fired_capacity=1 << 23, walert_capacity=1 << 23, **W10)
run("C4 60 s / 10 s sliding sum + Mbps alert, 10 M (host,metric) keys, i64, hashed", StreamSpec(n_keys=10_000_000, events_per_ms=10_000, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind="host_metric"),
    drain=("poll_fired", "poll_window_alerts"), value_type=1, window_enabled=1, agg_mask=3, window_size_ms=60_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000,
    walert_source=1, walert_op=1, walert_mul=8.0, walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.05, table_capacity=15_000_000, pane_ring=16,
    fired_capacity=1 << 26, walert_capacity=1 << 26)
run("C5 Zipf(1.2) over 10^8 ids, 10 s tumbling count/sum/max, hashed", StreamSpec(n_keys=100_000_000, events_per_ms=1000, key_kind="zipf", late_frac_ppm=1000), drain=("poll_fired",),
    agg_mask=11, table_capacity=1 << 22, fired_capacity=1 << 24, **W10)
run("C5 Zipf(1.2), 3-consecutive, hashed", StreamSpec(n_keys=100_000_000, events_per_ms=1000, key_kind="zipf"), drain=("poll_consec_alerts",), consec_n=3, consec_threshold=90.0,
    consec_reset_on_fire=1, table_capacity=1 << 22, calert_capacity=B // 8)
run("N3 10 s tumbling median, 1 M keys, dense", StreamSpec(n_keys=1_000_000, events_per_ms=1000, late_frac_ppm=1000), nb=4, drain=("poll_fired", "poll_medians"),
    agg_mask=1, table_capacity=1_000_000, key_mode=1, fired_capacity=1 << 23, median_log_capacity=3 * B, median_capacity=1 << 23, **W10)
print("\n| config | step ms | ingest | keyed | plan+fire | Gev/s | 24 B x ev/s over 6550.7 GB/s |\n|---|---|---|---|---|---|---|")
print("\n".join(rows))
if ONLY and ONLY != "A4":
    sys.exit(0)
# A4: the owner split by itself (count + scan + block-staged scatter into one output), device time on the handle's stream
eng = capi.Engine(efilter_enabled=1, max_batch=B)
cols = generate_torch(StreamSpec(n_keys=1_000_000), 0, B)
outs = [torch.empty_like(c) for c in cols]
ext = torch.cuda.ExternalStream(eng.stream)
torch.cuda.synchronize()
for world in (2, 8, 64):
    ts_ = []
    for i in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(ext); eng.partition_by_owner(cols[0], cols[1], cols[2], world, *outs); e1.record(ext); torch.cuda.synchronize()
        if i >= 2: ts_.append(e0.elapsed_time(e1))
    print(f"A4 owner split, world {world}: {np.mean(ts_):.3f} ms per 2^{int(np.log2(B))} events (incl. the 1 KB read-back of the bucket starts) = {24 * B / np.mean(ts_) / 1e6:.0f} GB/s of events")
eng.close()
# N2: text ingest
lines = [f"{1563452000 + i // 1000} 10.{(i * 7919) % 250}.{(i * 104729) % 250}.{i % 250} cpu{i % 8} {((i * 2654435761) % 100000) / 1000:.3f}" for i in range(1 << 20)]
text = ("\n".join(lines) + "\n").encode() * 4
eng = capi.Engine(efilter_enabled=1, text_max_bytes=len(text) + 64, max_batch=1 << 23)
import ctypes as C
dtext = torch.frombuffer(bytearray(text), dtype=torch.uint8).cuda(); torch.cuda.synchronize()
n, bad = C.c_size_t(0), C.c_size_t(0)
lib = capi.load_library()
for _ in range(3):
    t0 = time.perf_counter()
    st = lib.fmq_parse_text(eng.handle, dtext.data_ptr(), len(text), 1, 0, 8, None, None, None, None, 1 << 23, 1, C.byref(n), C.byref(bad)); eng.synchronize()
    dt = time.perf_counter() - t0
print(f"\nN2 text ingest (device-resident text, {len(text) / 1e6:.0f} MB, {n.value} lines, {bad.value} bad): {dt * 1e3:.2f} ms wall incl. 2 host syncs -> {len(text) / dt / 1e9:.1f} GB/s, {n.value / dt / 1e6:.0f} M lines/s (status {st})")
# N4: snapshot / restore
eng = capi.Engine(agg_mask=11, table_capacity=1_000_000, key_mode=1, max_batch=B, **W10)
eng.push_batch(*generate_torch(StreamSpec(n_keys=1_000_000), 0, B)); eng.poll_fired()
t0 = time.perf_counter(); snap = eng.snapshot(); t1 = time.perf_counter(); eng.restore(snap); t2 = time.perf_counter()
print(f"N4 snapshot {snap.nbytes / 1e6:.0f} MB: {1e3 * (t1 - t0):.0f} ms (pageable host buffer), restore {1e3 * (t2 - t1):.0f} ms")
