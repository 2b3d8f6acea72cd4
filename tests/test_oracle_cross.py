"""Cross-checks of the test infrastructure itself: the C++ literal oracle against the independent
pure-Python model (hypothesis, small random streams), and the multi-thread CPU baseline against
the literal oracle (row for row)."""
import os
import sys

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
from oracle_lib import OracleEngine, ParallelOracle, assert_events_equal, assert_fired_equal, sort_walerts
from test_oracle_golden import NaiveAdapter

CFGS = {
    "tumbling_f64": dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=3000,
                         walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=55.0,
                         efilter_enabled=1, efilter_threshold=90.0, rolling_max_enabled=1,
                         consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1),
    "sliding_i64": dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=10_000,
                        window_offset_ms=3000, max_out_of_orderness_ms=5000,
                        walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                        walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.5,
                        consec_n=2, consec_threshold=500000.0, consec_reset_on_fire=0),
    # .allowedLateness(T): windows kept T ms after they fired, re-fired once per batch that added records to them
    "tumbling_late_f64": dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=3000,
                              allowed_lateness_ms=15_000, walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=55.0),
    "sliding_late_i64": dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=30_000, window_slide_ms=10_000,
                             window_offset_ms=3000, max_out_of_orderness_ms=5000, allowed_lateness_ms=25_000,
                             walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                             walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.5),
}


def _run(e, batches, flush=True):
    roll = []
    for ts, key, val, wm in batches:
        e.push_batch(ts, key, val, wm)
        if e.cfg.rolling_max_enabled if hasattr(e, "cfg") else e.m.rmax:
            roll.append(e.poll_rolling_max(len(ts)))
    if flush:
        e.flush()
    return roll


@st.composite
def small_stream(draw, f64):
    nb = draw(st.integers(1, 5))
    batches, t = [], 1_000_000
    for _ in range(nb):
        n = draw(st.integers(0, 40))
        ts = np.array([max(0, t + draw(st.integers(-40_000, 20_000))) for _ in range(n)], np.int64)
        key = np.array([draw(st.integers(0, 5)) for _ in range(n)], np.uint64)
        if f64:
            val = np.array([draw(st.one_of(st.floats(0, 100, allow_nan=False), st.sampled_from([0.0, -0.0, 90.0, 50.0]))) for _ in range(n)], np.float64)
        else:
            val = np.array([draw(st.integers(0, 1 << 20)) for _ in range(n)], np.int64)
        wm = draw(st.sampled_from([capi.FMQ_WM_AUTO, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP, t - 5000, t + 10_000]))
        batches.append((ts, key, val, wm))
        t += draw(st.integers(0, 30_000))
    return batches


@pytest.mark.parametrize("name", list(CFGS))
@settings(max_examples=60, deadline=None)
@given(data=st.data())
def test_cpp_oracle_matches_python_model(name, data):
    kw = CFGS[name]
    batches = data.draw(small_stream(f64=kw.get("value_type", 0) == 0))
    a, b = OracleEngine(**kw), NaiveAdapter(**kw)
    ra, rb = _run(a, batches), _run(b, batches)
    for x, y in zip(ra, rb):
        np.testing.assert_array_equal(x.view(np.uint64), y.view(np.uint64))
    assert_fired_equal(a.poll_fired(), b.poll_fired(), f64_sum_rtol=0.0)      # same order of adds: bit-exact
    wa, wb = sort_walerts(a.poll_window_alerts()), sort_walerts(b.poll_window_alerts())
    np.testing.assert_array_equal(wa.key, wb.key); np.testing.assert_array_equal(wa.window_end, wb.window_end)
    np.testing.assert_array_equal(wa.value, wb.value)
    assert_events_equal(a.poll_event_alerts(), b.poll_event_alerts())
    ca = a.poll_consec_alerts()
    assert [(t, k, v) for t, k, v in zip(ca.ts.tolist(), ca.key.tolist(), ca.val.tolist())] == b.m.calerts
    assert a.counters()["late_dropped"] == b.counters()["late_dropped"]
    assert a.counters()["windows_fired"] == b.counters()["windows_fired"]


@pytest.mark.parametrize("name,threads", [("tumbling_f64", 3), ("sliding_i64", 4)])
def test_parallel_baseline_matches_literal_oracle(name, threads):
    kw = dict(CFGS[name]); kw["rolling_max_enabled"] = 0
    spec = StreamSpec(n_keys=500, events_per_ms=2, jitter_ms=3000, late_frac_ppm=20000, late_extra_ms=20000,
                      value_type="f64" if kw.get("value_type", 0) == 0 else "i64")
    a, b = OracleEngine(**kw), ParallelOracle(threads=threads, **kw)
    for bi in range(6):
        ts, key, val = generate_numpy(spec, bi * 20000, 20000)
        wm = capi.FMQ_WM_KEEP if bi in (2, 3) else capi.FMQ_WM_AUTO       # two buffers fill up between watermarks
        a.push_batch(ts, key, val, wm); b.push_batch(ts, key, val, wm)
    a.flush(); b.flush()
    assert_fired_equal(a.poll_fired(), b.poll_fired(), f64_sum_rtol=0.0)
    wa, wb = sort_walerts(a.poll_window_alerts()), sort_walerts(b.poll_window_alerts())
    np.testing.assert_array_equal(wa.key, wb.key); np.testing.assert_array_equal(wa.value, wb.value)
    assert_events_equal(a.poll_event_alerts(), b.poll_event_alerts())
    assert_events_equal(a.poll_consec_alerts(), b.poll_consec_alerts())
    ca, cb = a.counters(), b.counters()
    assert ca["late_dropped"] == cb["late_dropped"] and ca["windows_fired"] == cb["windows_fired"] > 0
    if name == "tumbling_f64":
        assert ca["late_dropped"] > 0          # the stream's late tail really exercises the drop


def test_oracle_median_against_numpy_for_every_run_length():
    """The window median of the literal oracle (the ListState restatement of ComputeCpuMiddle.java:34-47) against
    numpy.median, for run lengths 1..35 and a few long ones, with ties, both zeros and both infinities: the same
    stream the GPU test `test_window_median_run_lengths_around_the_short_long_boundary` compares with this oracle."""
    kw = dict(window_enabled=1, agg_mask=1, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=0,
              median_log_capacity=500_000, table_capacity=1 << 12, pane_ring=128, max_batch=1 << 18)
    cpu = OracleEngine(**kw)
    lens = list(range(1, 36)) + [63, 64, 65, 255, 256, 257, 1000, 4097, 100_001, 100_000]
    rng = np.random.default_rng(5)
    key = np.repeat(np.arange(len(lens), dtype=np.uint64) + 1, lens)
    key = key[rng.permutation(len(key))]
    want = {}
    for w in range(2):
        val = np.round(rng.normal(0, 3, len(key)), 1 if w == 0 else 12)
        val[::53] = -0.0; val[1::53] = 0.0; val[7::211] = np.inf; val[11::223] = -np.inf
        ts = 1_600_000_000_000 + w * 10_000 + rng.integers(0, 10_000, len(key)).astype(np.int64)
        cpu.push_batch(ts, key, val, capi.FMQ_WM_KEEP)
        for k in np.unique(key):
            want[(int(k), 1_600_000_000_000 + (w + 1) * 10_000)] = float(np.median(val[key == k]))
    cpu.flush()
    m = cpu.poll_medians()
    got = {(int(k), int(e)): float(v) for k, e, v in zip(m.key, m.window_end, m.value)}
    assert got.keys() == want.keys()
    assert all(got[x] == want[x] for x in want), [x for x in want if got[x] != want[x]][:5]
