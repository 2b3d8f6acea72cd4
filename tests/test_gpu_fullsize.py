"""Full-size runs (BASELINE.json scale per step) checked through size-independent properties, since
the oracle cannot replay 10^7..10^8 events in test time:
  * conservation: every non-late event is counted in exactly size/slide windows, so
    sum(count) == k * (events_in - late_dropped) when no event is partially late, and the int64
    sum of sums is k times the sum of the accepted values (a checksum of checksums);
  * every (key, window_end) fires exactly once; window ends are aligned to the slide;
  * min <= avg <= max, alerts == rows passing the predicate recomputed on the host;
  * a dense-id engine and a hashed-key engine fed the same stream agree row for row."""
import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
from oracle_lib import sort_fired

pytestmark = pytest.mark.gpu


def _run(eng, spec, nb, B):
    import torch
    tot_val = 0
    for b in range(nb):
        ts, key, val = generate_torch(spec, b * B, B)
        torch.cuda.synchronize()
        eng.push_batch(ts, key, val)
        tot_val += int(val.sum().item()) if val.dtype == torch.int64 else 0
        eng.synchronize()          # device columns must stay valid until the engine's stream has consumed them
    eng.flush()
    return tot_val


def test_c4_sliding_10m_composite_keys_conservation():
    """60 s / 10 s sliding sum per (host, metric), 10 M keys, int64 flows, 2^24 events per batch."""
    B, nb, k = 1 << 24, 4, 6
    spec = StreamSpec(n_keys=10_000_000, events_per_ms=10_000, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind="host_metric")
    eng = capi.Engine(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=10_000,
                      max_out_of_orderness_ms=5000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0,
                      walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.05,
                      table_capacity=15_000_000, pane_ring=16, max_batch=B, fired_capacity=120_000_000, walert_capacity=120_000_000)
    tot_val = _run(eng, spec, nb, B)
    c = eng.counters()
    assert c["events_in"] == nb * B and c["late_dropped"] == 0
    f = eng.poll_fired(chunk=1 << 24)
    assert len(f) == c["windows_fired"] > 0
    assert int(f.count.sum()) == k * nb * B                       # every event in exactly size/slide windows
    assert int(f.sum.sum()) == k * tot_val                        # checksum of checksums, exact in int64
    assert np.all(f.window_end % 10_000 == 0) and np.all(f.min <= f.max) and np.all(f.count > 0)
    pairs = np.stack([f.key, f.window_end.view(np.uint64)], 1)
    assert len(np.unique(pairs, axis=0)) == len(f)               # each (key, window) fired once
    a = eng.poll_window_alerts(chunk=1 << 24)
    mbps = f.sum.astype(np.float64) * 8.0 / 60 / 1024 / 1024
    assert len(a) == int((mbps < 0.05).sum()) == c["window_alerts"]


def test_c3_tumbling_1m_keys_dense_equals_hashed_and_conserves():
    import torch
    B, nb = 1 << 24, 3
    spec = StreamSpec(n_keys=1_000_000, events_per_ms=1000, jitter_ms=5000, late_frac_ppm=1000, late_extra_ms=25_000)
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000,
              walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=55.0, efilter_enabled=1, efilter_threshold=90.0,
              pane_ring=16, max_batch=B, fired_capacity=12_000_000, walert_capacity=12_000_000, ealert_capacity=6_000_000)
    outs = []
    for mode, cap in ((capi.FMQ_KEYS_DENSE, 1_000_000), (capi.FMQ_KEYS_HASHED, 1_500_000)):
        eng = capi.Engine(key_mode=mode, table_capacity=cap, **kw)
        n_e = 0
        for b in range(nb):
            ts, key, val = generate_torch(spec, b * B, B)
            torch.cuda.synchronize()
            eng.push_batch(ts, key, val)
            n_e += len(eng.poll_event_alerts(chunk=1 << 23))      # synchronises: the columns may be freed afterwards
        eng.flush()
        c = eng.counters()
        f = sort_fired(eng.poll_fired(chunk=1 << 24))
        assert int(f.count.sum()) == c["events_in"] - c["late_dropped"] and c["late_dropped"] > 0      # tumbling: k = 1
        assert n_e == c["event_alerts"]
        avg = f.sum / f.count
        assert np.all(f.min <= avg + 1e-9) and np.all(avg <= f.max + 1e-9) and np.all(f.max < 100.0)
        assert c["window_alerts"] == int((avg > 55.0).sum())
        outs.append((f, c))
        eng.close()
    (fd, cd), (fh, ch) = outs
    np.testing.assert_array_equal(fd.key, fh.key); np.testing.assert_array_equal(fd.window_end, fh.window_end)
    np.testing.assert_array_equal(fd.count, fh.count); np.testing.assert_array_equal(fd.max, fh.max); np.testing.assert_array_equal(fd.min, fh.min)
    np.testing.assert_allclose(fd.sum, fh.sum, rtol=1e-9)
    assert cd["late_dropped"] == ch["late_dropped"] and cd["event_alerts"] == ch["event_alerts"]


def test_c2_consecutive_100k_keys_idempotent_replay():
    """2^24 events, 100 K keys: the same batch replayed into two fresh handles gives identical alert
    sets and final state (the path is deterministic although thousands of warps share the work), and
    the alert count matches a numpy recount of non-overlapping over-threshold triples per key."""
    import torch
    B = 1 << 24
    spec = StreamSpec(n_keys=100_000, events_per_ms=1000)
    ts, key, val = generate_torch(spec, 0, B)
    torch.cuda.synchronize()
    res = []
    for _ in range(2):
        eng = capi.Engine(consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, key_mode=capi.FMQ_KEYS_DENSE, table_capacity=100_000,
                          max_batch=B, calert_capacity=1 << 22)
        eng.push_batch(ts, key, val)
        a = eng.poll_consec_alerts()
        o = np.lexsort((a.key, a.ts))
        res.append((a.ts[o], a.key[o], a.val[o], eng.key_state(np.arange(100_000, dtype=np.uint64))[1]))
        eng.close()
    for x, y in zip(res[0], res[1]):
        np.testing.assert_array_equal(x, y)
    # host recount: stable sort by key, run lengths of consecutive overs, floor(run / 3) alerts each
    k = key.cpu().numpy(); over = (val.cpu().numpy() > 90.0)
    o = np.argsort(k, kind="stable"); ks, ov = k[o], over[o]
    brk = np.flatnonzero((~ov) | np.r_[True, ks[1:] != ks[:-1]] & True)
    # run id changes at every under-event and at every key change
    change = (~ov) | np.r_[True, ks[1:] != ks[:-1]]
    rid = np.cumsum(change)
    runs = np.bincount(rid[ov])
    assert int((runs // 3).sum()) == len(res[0][0])
