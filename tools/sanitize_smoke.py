"""Small end-to-end run of every kernel family, for compute-sanitizer (memcheck / racecheck)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
for key_mode in (0, 1):
    e = capi.Engine(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000, key_mode=key_mode,
                    walert_source=2, walert_threshold=50.0, efilter_enabled=1, efilter_threshold=90.0, rolling_max_enabled=1, consec_n=3,
                    consec_threshold=50.0, consec_reset_on_fire=1, median_log_capacity=1 << 17, late_capacity=1 << 12, text_max_bytes=1 << 16,
                    table_capacity=1 << 12, pane_ring=64, max_batch=1 << 14)
    spec = StreamSpec(n_keys=3000, events_per_ms=1, jitter_ms=3000, late_frac_ppm=20000, late_extra_ms=20000)
    for b in range(3):
        ts, key, val = generate_numpy(spec, b * 9001, 9001)       # > 8 tiles on a few blocks, ragged tail
        e.push_batch(ts, key, val)
        e.poll_rolling_max(9001)
    e.flush()
    print(key_mode, len(e.poll_fired()), len(e.poll_medians()), len(e.poll_late()), len(e.poll_event_alerts()), len(e.poll_consec_alerts()), e.counters()["late_dropped"])
    ts, key, val, st = e.parse_text(b"1563452056 10.8.22.1 cpu0 80.5\n1563452051 10.8.22.1 cpu2 10.5\n", 0)
    snap = None
    e.close()
print("ok")
