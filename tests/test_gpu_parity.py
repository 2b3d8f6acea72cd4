"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded
inputs.  Bit-exact for counts, integer sums, min/max, alert tuples, late-drop counts and keyed
state; double sums within 1e-9 relative (BASELINE.json north_star; atomics reorder the adds)."""
import json
import os

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
from monitor_systam_flink_quickstart_b200.text import KeyDictionary, parse_cpu_line, parse_flow_line
from oracle_lib import (OracleEngine, assert_events_equal, assert_fired_equal, sort_fired, sort_walerts)

pytestmark = pytest.mark.gpu
G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_sessions.json"), encoding="utf-8"))
SMALL = dict(table_capacity=1 << 12, pane_ring=128, max_batch=1 << 16)


def pair(**kw):
    okw = {k: v for k, v in kw.items()}
    return capi.Engine(**kw), OracleEngine(**okw)


def compare_all(gpu, cpu, agg_mask=15, rtol=1e-9):
    if gpu.cfg.late_capacity:
        assert_events_equal(gpu.poll_late(), cpu.poll_late())
    assert_fired_equal(gpu.poll_fired(), cpu.poll_fired(), f64_sum_rtol=rtol, agg_mask=agg_mask)
    if gpu.cfg.walert_source:
        a, b = sort_walerts(gpu.poll_window_alerts()), sort_walerts(cpu.poll_window_alerts())
        np.testing.assert_array_equal(a.key, b.key); np.testing.assert_array_equal(a.window_end, b.window_end)
        np.testing.assert_allclose(a.value, b.value, rtol=rtol, atol=0)
    if gpu.cfg.efilter_enabled:
        assert_events_equal(gpu.poll_event_alerts(), cpu.poll_event_alerts())
    if gpu.cfg.consec_n:
        assert_events_equal(gpu.poll_consec_alerts(), cpu.poll_consec_alerts())
    cg, cc = gpu.counters(), cpu.counters()
    for k in ("events_in", "late_dropped", "windows_fired", "window_alerts", "event_alerts", "consec_alerts", "watermark"):
        assert cg[k] == cc[k], (k, cg, cc)


# ---- the reference's own golden sessions, through the CUDA path ---------------------------------
def test_g1_filter_gpu():
    d = KeyDictionary()
    e = capi.Engine(efilter_enabled=1, efilter_op=capi.FMQ_CMP_GT, efilter_threshold=90.0, **SMALL)
    rows = [parse_cpu_line(l) for l in G["G1_filter"]["input"]]
    e.push_batch(np.array([r[0] * 1000 for r in rows], np.int64), np.array([d.encode(r[1]) for r in rows], np.uint64),
                 np.array([r[3] for r in rows]))
    out = e.poll_event_alerts()
    assert out.val.tolist() == [99.2] and d.decode(out.key[0]) == "10.8.22.1"


def test_g3_rolling_max_gpu():
    e = capi.Engine(rolling_max_enabled=1, **SMALL)
    rows = [parse_cpu_line(l) for l in G["G3_rolling_max"]["input"]]
    e.push_batch(np.array([r[0] * 1000 for r in rows], np.int64), np.zeros(3, np.uint64), np.array([r[3] for r in rows]))
    assert e.poll_rolling_max(3).tolist() == [80.5, 80.5, 99.9]


def test_g4_avg_gpu():
    d = KeyDictionary()
    e = capi.Engine(window_enabled=1, agg_mask=3, window_size_ms=60000, window_slide_ms=60000,
                    walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=-1.0, **SMALL)
    rows = [parse_cpu_line(l) for l in G["G4_avg"]["input"]]
    ts = 1563452040000 + np.arange(4, dtype=np.int64) * 1000
    e.push_batch(ts, np.array([d.encode(r[1]) for r in rows], np.uint64), np.array([r[3] for r in rows]), 0)
    e.flush()
    a = e.poll_window_alerts()
    got = {d.decode(k): v for k, v in zip(a.key, a.value)}
    # a 3-term double sum in a different order may differ in the last bit; the README value is the
    # sequential one — accept 1 ulp here, exactness of the sequential order is the oracle's job
    assert abs(got["10.8.22.1"] - 86.26666666666667) <= 2e-14 and got["10.8.22.2"] == 20.2


def test_g6_g7_bandwidth_gpu():
    kw = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, window_size_ms=300000, window_slide_ms=5000,
              max_out_of_orderness_ms=60000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0,
              walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=100.0, **SMALL)
    e = capi.Engine(**kw)
    d = KeyDictionary()
    for t, ch, flow in [parse_flow_line(l) for l in G["G7_event_time"]["input"]]:
        e.push_batch(np.array([t], np.int64), np.array([d.encode(ch)], np.uint64), np.array([flow], np.int64))
    f = sort_fired(e.poll_fired())
    assert f.sum.tolist() == [10000] * 12 + [10100] * 12 + [10200] * 36
    assert e.counters()["late_dropped"] == 1
    a = e.poll_window_alerts()
    assert {0.0012715657552083333, 0.0012969970703125} <= set(a.value.tolist()) and len(a) == 60   # bit-exact doubles
    e.flush()
    assert sort_fired(e.poll_fired()).sum.tolist() == [200] * 24 + [100] * 48
    # G6: tumbling long sum 11200 and its filter value
    kw.update(window_size_ms=60000, window_slide_ms=60000)
    e = capi.Engine(**kw)
    rows = [parse_flow_line(l) for l in G["G6_sum"]["input"]]
    e.push_batch(1566957600000 + np.arange(4, dtype=np.int64) * 1000, np.zeros(4, np.uint64), np.array([r[2] for r in rows], np.int64), 0)
    e.flush()
    assert e.poll_fired().sum.tolist() == [11200] and e.poll_window_alerts().value.tolist() == [0.0014241536458333333]


# ---- seeded streams ------------------------------------------------------------------------------
CONFIGS = {
    "c3_tumbling_f64": (dict(late_capacity=1 << 16, window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=5000,
                             walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=50.0,
                             efilter_enabled=1, efilter_threshold=90.0),
                        StreamSpec(n_keys=3000, events_per_ms=3, late_frac_ppm=20000)),
    "c4_sliding_i64": (dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=10_000,
                            window_offset_ms=2500, max_out_of_orderness_ms=5000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT,
                            walert_mul=8.0, walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=2.5),
                       StreamSpec(n_keys=4096, events_per_ms=2, value_type="i64", key_kind="host_metric", late_frac_ppm=20000, late_extra_ms=90000)),
    "count_only": (dict(window_enabled=1, agg_mask=1, window_size_ms=3000, window_slide_ms=1000, max_out_of_orderness_ms=0, emit_fired=1),
                   StreamSpec(n_keys=100, events_per_ms=1, jitter_ms=0, late_frac_ppm=0)),
}


@pytest.mark.parametrize("name", list(CONFIGS))
@pytest.mark.parametrize("batch", [1000, 37_001])
@pytest.mark.parametrize("key_mode", ["hashed", "dense"])
def test_window_parity_seeded(name, batch, key_mode):
    kw, spec = CONFIGS[name]
    if key_mode == "dense":
        if spec.key_kind != "uniform":
            pytest.skip("composite (host<<32|metric) keys are not dense ids")
        kw = dict(kw, key_mode=capi.FMQ_KEYS_DENSE)
    gpu, cpu = pair(**kw, **SMALL)
    total = 150_000
    for s in range(0, total, batch):
        n = min(batch, total - s)
        ts, key, val = generate_numpy(spec, s, n)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
        if (s // batch) % 3 == 1:                      # interleave polls with pushes
            compare_all(gpu, cpu, agg_mask=kw["agg_mask"])
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu, agg_mask=kw["agg_mask"])
    assert cpu.counters()["windows_fired"] > 0


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
def test_window_parity_when_blocks_reuse_their_staging_ring(key_mode):
    """2.2 M events per batch = ~5 tiles per resident block: every shared-memory stage of the TMA ring is
    refilled while the kernel runs (the small tests above never get that far).  Against the oracle."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000, late_capacity=1 << 16,
              efilter_enabled=1, efilter_threshold=97.0, key_mode=key_mode, table_capacity=60_000, pane_ring=64, max_batch=1 << 22,
              fired_capacity=1 << 21, ealert_capacity=1 << 20)
    spec = StreamSpec(n_keys=40_000, events_per_ms=200, jitter_ms=3000, late_frac_ppm=2000, late_extra_ms=20_000)
    gpu, cpu = pair(**kw)
    n = 2_200_000
    for b in range(3):
        ts, key, val = generate_numpy(spec, b * n, n)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)
    assert cpu.counters()["late_dropped"] > 0


@pytest.mark.parametrize("vt", ["f64", "i64"])
def test_window_parity_skewed_keys_switch_on_warp_combining(vt):
    """Zipf(1.2) keys: after the first batch the engine pre-aggregates same-cell lanes inside a warp
    (skew adaptation).  Counts, integer sums, min/max stay bit-exact; double sums within 1e-9."""
    kw = dict(value_type=capi.FMQ_VAL_F64 if vt == "f64" else capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=10_000,
              window_slide_ms=5000, max_out_of_orderness_ms=3000, table_capacity=1 << 19, pane_ring=64, max_batch=1 << 19, fired_capacity=1 << 22)
    spec = StreamSpec(n_keys=1_000_000, events_per_ms=50, key_kind="zipf", jitter_ms=3000, late_frac_ppm=5000, late_extra_ms=20_000, value_type=vt)
    gpu, cpu = pair(**kw)
    for b in range(4):
        ts, key, val = generate_numpy(spec, b * 300_000, 300_000)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)
    u, c = np.unique(generate_numpy(spec, 0, 300_000)[1], return_counts=True)
    assert c.max() > 0.1 * 300_000                      # the stream really is skewed


def test_full_buffers_between_two_upstream_watermarks_do_not_move_time():
    """FMQ_WM_KEEP: a shim behind Flink's own assigner pushes a full buffer without a watermark; records that an
    early watermark would have made late must still be counted (WindowOperator moves time in processWatermark only)."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=5000, window_slide_ms=1000, max_out_of_orderness_ms=0, late_capacity=1 << 12, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=200, events_per_ms=1, jitter_ms=6000, late_frac_ppm=0, t0=1_000_000)
    plan = [1_003_000, capi.FMQ_WM_KEEP, capi.FMQ_WM_KEEP, 1_012_500, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP]
    for i, wm in enumerate(plan):
        ts, key, val = generate_numpy(spec, i * 4000, 4000)
        for e in (gpu, cpu):
            e.push_batch(ts, key, val, wm)
        assert gpu.counters()["watermark"] == cpu.counters()["watermark"]
    compare_all(gpu, cpu)
    for e in (gpu, cpu):
        e.flush()
    compare_all(gpu, cpu)
    assert cpu.counters()["late_dropped"] > 0


def test_keep_before_any_watermark_leaves_the_stream_unanchored():
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=4000, window_slide_ms=2000, max_out_of_orderness_ms=1000, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=64, events_per_ms=1, jitter_ms=3000, late_frac_ppm=0, t0=50_000)
    for i, wm in enumerate([capi.FMQ_WM_KEEP, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP]):
        ts, key, val = generate_numpy(spec, i * 3000, 3000)
        for e in (gpu, cpu):
            e.push_batch(ts, key, val, wm)
    for e in (gpu, cpu):
        e.flush()
    compare_all(gpu, cpu)


# ---- .allowedLateness(T): option 2 of the lateness handling the tutorial lists (chapter3/README.md:211-212) ----------
def test_allowed_lateness_known_answer():
    """One 10 s window, T = 5 s: two records, the window fires; a third record arrives late but within T: the window
    fires again with all three; after the watermark passes window end - 1 + T a fourth record is dropped."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, allowed_lateness_ms=5000, late_capacity=16, **SMALL)
    e = capi.Engine(**kw)
    one = lambda t, v: (np.array([t], np.int64), np.array([7], np.uint64), np.array([v], np.float64))  # noqa: E731
    e.push_batch(np.array([1000, 2000], np.int64), np.array([7, 7], np.uint64), np.array([1.0, 2.0]), 12_000)
    f = e.poll_fired(); assert (f.window_end.tolist(), f.count.tolist(), f.sum.tolist()) == ([10_000], [2], [3.0])
    e.push_batch(*one(3000, 4.0), capi.FMQ_WM_KEEP)                        # late, but the window is still kept: it fires again
    f = e.poll_fired(); assert (f.window_end.tolist(), f.count.tolist(), f.sum.tolist(), f.max.tolist()) == ([10_000], [3], [7.0], [4.0])
    e.push_batch(*one(3500, 8.0), 14_998)                                   # window end - 1 + T = 14 999 > 14 998: still kept
    f = e.poll_fired(); assert (f.count.tolist(), f.sum.tolist()) == ([4], [15.0])
    e.push_batch(*one(4000, 16.0), 14_999)                                  # judged against 14 998: accepted, re-fired, then the state goes
    f = e.poll_fired(); assert (f.count.tolist(), f.sum.tolist()) == ([5], [31.0])
    e.push_batch(*one(5000, 32.0), capi.FMQ_WM_KEEP)                        # the window is gone: dropped
    assert len(e.poll_fired()) == 0 and e.counters()["late_dropped"] == 1 and e.poll_late().ts.tolist() == [5000]
    e.flush()
    assert len(e.poll_fired()) == 0 and e.counters()["windows_fired"] == 4


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
@pytest.mark.parametrize("name", ["tumbling_f64", "sliding_i64", "sliding_offset_short_T"])
def test_allowed_lateness_parity(name, key_mode):
    """Streams with a heavy late tail (some of it within T, some beyond), automatic / explicit / absent watermarks:
    fired rows (first firings and re-firings), alerts, the late side output and every counter against the oracle."""
    cfgs = {
        "tumbling_f64": (dict(window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=3000, allowed_lateness_ms=15_000,
                              walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=55.0), "f64"),
        "sliding_i64": (dict(value_type=capi.FMQ_VAL_I64, window_size_ms=30_000, window_slide_ms=10_000, window_offset_ms=3000, max_out_of_orderness_ms=5000,
                             allowed_lateness_ms=25_000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                             walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.5), "i64"),
        "sliding_offset_short_T": (dict(window_size_ms=20_000, window_slide_ms=5000, window_offset_ms=1234, max_out_of_orderness_ms=2000, allowed_lateness_ms=3000), "f64"),
    }
    c, vt = cfgs[name]
    kw = dict(window_enabled=1, agg_mask=15, late_capacity=1 << 17, key_mode=key_mode, table_capacity=1 << 12, pane_ring=64, max_batch=1 << 16, **c)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=700, events_per_ms=2, jitter_ms=4000, late_frac_ppm=150_000, late_extra_ms=40_000, value_type=vt)
    t0 = spec.t0
    wms = [capi.FMQ_WM_AUTO, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, t0 + 70_000, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, capi.FMQ_WM_AUTO]
    for i, wm in enumerate(wms):
        ts, key, val = generate_numpy(spec, i * 30_000, 30_000 if i != 5 else 777)
        for e in (gpu, cpu):
            e.push_batch(ts, key, val, wm)
        if i % 3 == 1:
            compare_all(gpu, cpu)
    compare_all(gpu, cpu)
    for e in (gpu, cpu):
        e.flush()
    compare_all(gpu, cpu)
    c = cpu.counters()
    assert c["windows_fired"] > 0 and (name == "sliding_i64" or c["late_dropped"] > 0)     # sliding_i64 keeps everything this stream brings
    f = cpu.poll_fired()


def test_explicit_watermarks_and_empty_batches():
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=5000, window_slide_ms=5000, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=50, events_per_ms=1, jitter_ms=8000, late_frac_ppm=0, t0=100_000)
    wms = [capi.FMQ_WM_AUTO, 95_000, 90_000, 104_999, 105_000, capi.FMQ_WM_AUTO, 130_000]
    for i, wm in enumerate(wms):
        n = 0 if i == 2 else 4001
        ts, key, val = generate_numpy(spec, i * 4001, n)
        for e in (gpu, cpu):
            e.push_batch(ts, key, val, wm)
    compare_all(gpu, cpu)
    for e in (gpu, cpu):
        e.flush(150_000); e.flush(149_000); e.flush()
    compare_all(gpu, cpu)


@pytest.mark.parametrize("t0", [0, 7, (1 << 55) + 123])
def test_timestamps_at_the_epoch_and_far_in_the_future(t0):
    """Windows that start before time 0 (negative window starts, pane -1) and timestamps near 2^55."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, window_offset_ms=1234, max_out_of_orderness_ms=2000,
              late_capacity=1 << 12, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=40, events_per_ms=1, jitter_ms=2500, late_frac_ppm=30000, late_extra_ms=9000, t0=t0)
    for b in range(5):
        ts, key, val = generate_numpy(spec, b * 7000, 7000)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
        if b == 2:
            for e in (gpu, cpu):
                e.push_batch(ts[:0], key[:0], val[:0])           # an empty batch in between
                e.flush(int(ts.max()) - 20_000)                   # a watermark that does not advance
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)
    f = cpu.poll_fired()
    assert cpu.counters()["windows_fired"] > 0


def test_device_resident_and_misaligned_inputs():
    import torch
    kw, spec = CONFIGS["c3_tumbling_f64"]
    gpu, cpu = pair(**kw, **SMALL)
    ts, key, val = generate_numpy(spec, 0, 50_003)
    dts, dkey, dval = (torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).cuda() for x in (ts, key, val))
    for lo, hi in [(0, 20_000), (20_001, 35_002), (35_002, 50_003)]:      # 8-byte-aligned-only slices too
        gpu.push_batch(dts[lo:hi], dkey[lo:hi], dval[lo:hi])
        cpu.push_batch(ts[lo:hi], key[lo:hi], val[lo:hi])
    gpu.synchronize()
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)


def test_special_doubles_in_min_max():
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=1000, window_slide_ms=1000, **SMALL)
    gpu, cpu = pair(**kw)
    val = np.array([0.0, -0.0, 5.5, -3.25, np.inf, -np.inf, -0.0, 0.0, 1e-310, -1e-310], np.float64)
    key = np.array([0, 0, 1, 1, 2, 2, 3, 3, 4, 4], np.uint64)
    ts = np.full(10, 500, np.int64)
    for e in (gpu, cpu):
        e.push_batch(ts, key, val, 0); e.flush()
    assert_fired_equal(gpu.poll_fired(), cpu.poll_fired(), agg_mask=12 | 1)      # min/max bit patterns incl. -0.0


def test_int64_sum_wraps_like_java_long():
    kw = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=1000, window_slide_ms=1000, **SMALL)
    gpu, cpu = pair(**kw)
    val = np.array([2 ** 62, 2 ** 62, 2 ** 62, -5, -(2 ** 63), 7], np.int64)
    key = np.array([1, 1, 1, 2, 2, 2], np.uint64)
    for e in (gpu, cpu):
        e.push_batch(np.full(6, 10, np.int64), key, val, 0); e.flush()
    assert_fired_equal(gpu.poll_fired(), cpu.poll_fired())


# ---- keyed ValueState operators ------------------------------------------------------------------
@pytest.mark.parametrize("vt", ["f64", "i64"])
@pytest.mark.parametrize("reset", [1, 0])
@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
@pytest.mark.parametrize("rolling", [1, 0])          # 0: the over-threshold bit rides in the sorted items, no value gathers
def test_keyed_state_parity(vt, reset, key_mode, rolling):
    thr = 50.0 if vt == "f64" else 500_000.0
    kw = dict(value_type=capi.FMQ_VAL_F64 if vt == "f64" else capi.FMQ_VAL_I64, rolling_max_enabled=rolling, key_mode=key_mode,
              consec_n=3, consec_threshold=thr, consec_reset_on_fire=reset, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=700, events_per_ms=5, value_type=vt)
    for b, n in enumerate([1, 5000, 33, 60_000, 0, 4097]):
        ts, key, val = generate_numpy(spec, b * 70_000, n)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
        if rolling:
            np.testing.assert_array_equal(gpu.poll_rolling_max(n).view(np.uint64), cpu.poll_rolling_max(n).view(np.uint64))
    assert_events_equal(gpu.poll_consec_alerts(), cpu.poll_consec_alerts())
    keys = np.arange(0, 800, dtype=np.uint64)
    fg, cg, rg = gpu.key_state(keys); fc, cc, rc = cpu.key_state(keys)
    np.testing.assert_array_equal(fg, fc); np.testing.assert_array_equal(cg, cc)
    np.testing.assert_array_equal(rg.view(np.uint64), rc.view(np.uint64))
    assert gpu.counters()["consec_alerts"] == cpu.counters()["consec_alerts"] > 0


@pytest.mark.parametrize("n_keys,thr", [(3, 50.0), (1, 20.0), (40, 90.0), (100_000, 50.0)])
@pytest.mark.parametrize("rolling", [1, 0])
def test_keyed_state_long_runs_and_sparse_keys(n_keys, thr, rolling):
    """Heavy-hitter keys (one run spans many warps' ranges and rounds) and very sparse keys."""
    kw = dict(rolling_max_enabled=rolling, consec_n=3, consec_threshold=thr, consec_reset_on_fire=1,
              table_capacity=1 << 18, max_batch=1 << 17)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=n_keys, events_per_ms=5)
    for b, n in enumerate([70_001, 100_000, 513]):
        ts, key, val = generate_numpy(spec, b * 100_000, n)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
        if rolling:
            np.testing.assert_array_equal(gpu.poll_rolling_max(n).view(np.uint64), cpu.poll_rolling_max(n).view(np.uint64))
        assert_events_equal(gpu.poll_consec_alerts(), cpu.poll_consec_alerts())
    keys = np.arange(0, min(n_keys, 5000), dtype=np.uint64)
    fg, cg, rg = gpu.key_state(keys); fc, cc, rc = cpu.key_state(keys)
    np.testing.assert_array_equal(fg, fc); np.testing.assert_array_equal(cg, cc)
    np.testing.assert_array_equal(rg.view(np.uint64), rc.view(np.uint64))


def test_consecutive_rule_known_answers():
    """6 consecutive overs -> 2 alerts with reset-on-fire; a dip resets the run."""
    e = capi.Engine(consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, **SMALL)
    v = np.array([95, 96, 97, 98, 99, 91, 10, 95, 95, 90.0, 95, 95, 95], np.float64)
    e.push_batch(np.arange(13, dtype=np.int64), np.zeros(13, np.uint64), v)
    a = e.poll_consec_alerts()
    assert sorted(a.ts.tolist()) == [2, 5, 12]
    e2 = capi.Engine(consec_n=3, consec_threshold=90.0, consec_reset_on_fire=0, **SMALL)
    e2.push_batch(np.arange(13, dtype=np.int64), np.zeros(13, np.uint64), v)
    assert sorted(e2.poll_consec_alerts().ts.tolist()) == [2, 3, 4, 5, 12]


# ---- error behaviour -----------------------------------------------------------------------------
def test_error_codes():
    e = capi.Engine(window_enabled=1, agg_mask=3, window_size_ms=1000, window_slide_ms=1000, table_capacity=64, pane_ring=4, max_batch=1 << 12)
    with pytest.raises(capi.FmqError) as ex:
        e.push_batch(np.zeros(1 << 13, np.int64), np.zeros(1 << 13, np.uint64), np.zeros(1 << 13))
    assert ex.value.status == 1                                           # larger than max_batch
    e.push_batch(np.full(200, 5, np.int64), np.arange(200, dtype=np.uint64), np.ones(200))
    with pytest.raises(capi.FmqError) as ex:
        e.counters()
    assert ex.value.status == 2 and "table_capacity" in str(ex.value)      # directory full -> sticky
    with pytest.raises(capi.FmqError) as ex:
        e.push_batch(np.zeros(1, np.int64), np.zeros(1, np.uint64), np.zeros(1))
    assert ex.value.status == 8
    e = capi.Engine(window_enabled=1, agg_mask=3, window_size_ms=1000, window_slide_ms=1000, table_capacity=64, pane_ring=4, max_batch=1 << 12)
    e.push_batch(np.array([0, 10_000], np.int64), np.zeros(2, np.uint64), np.ones(2))   # 11 panes apart, ring of 4
    with pytest.raises(capi.FmqError) as ex:
        e.counters()
    assert ex.value.status == 3
    e = capi.Engine(window_enabled=1, agg_mask=3, window_size_ms=1000, window_slide_ms=1000, table_capacity=64, pane_ring=4, max_batch=1 << 12)
    e.push_batch(np.array([-1], np.int64), np.zeros(1, np.uint64), np.ones(1))
    with pytest.raises(capi.FmqError) as ex:
        e.counters()
    assert ex.value.status == 6
    e = capi.Engine(efilter_enabled=1, efilter_threshold=0.5, ealert_capacity=10, max_batch=1 << 12)
    e.push_batch(np.zeros(100, np.int64), np.zeros(100, np.uint64), np.ones(100))
    with pytest.raises(capi.FmqError) as ex:
        e.poll_event_alerts()
    assert ex.value.status == 4


def test_dense_ids_outside_the_table_are_an_input_error():
    e = capi.Engine(window_enabled=1, agg_mask=3, window_size_ms=1000, window_slide_ms=1000, key_mode=capi.FMQ_KEYS_DENSE,
                    table_capacity=64, pane_ring=4, max_batch=1 << 12)
    e.push_batch(np.array([5, 6], np.int64), np.array([63, 64], np.uint64), np.ones(2))
    with pytest.raises(capi.FmqError) as ex:
        e.counters()
    assert ex.value.status == 6


def test_partial_polls_drain_everything():
    kw, spec = CONFIGS["c3_tumbling_f64"]
    gpu, cpu = pair(**kw, **SMALL)
    ts, key, val = generate_numpy(spec, 0, 60_000)
    gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
    gpu.flush(); cpu.flush()
    assert_fired_equal(gpu.poll_fired(chunk=777), cpu.poll_fired())
    assert_events_equal(gpu.poll_event_alerts(chunk=100), cpu.poll_event_alerts())
    assert len(gpu.poll_fired()) == 0


# ---- non-incremental window function: median --------------------------------------
def test_g5_median_gpu():
    """chapter2/README.md:236-248: medians 80.5 and 20.2 (ComputeCpuMiddle.java:41-47)."""
    d = KeyDictionary()
    e = capi.Engine(window_enabled=1, agg_mask=1, window_size_ms=60000, window_slide_ms=60000, median_log_capacity=1024, **SMALL)
    rows = [parse_cpu_line(l) for l in G["G5_median"]["input"]]
    e.push_batch(1563452040000 + np.arange(4, dtype=np.int64) * 1000, np.array([d.encode(r[1]) for r in rows], np.uint64), np.array([r[3] for r in rows]), 0)
    e.flush()
    m = e.poll_medians()
    assert {d.decode(k): v for k, v in zip(m.key, m.value)} == {"10.8.22.1": 80.5, "10.8.22.2": 20.2}


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
@pytest.mark.parametrize("slide", [10_000, 5_000])
def test_window_median_parity(key_mode, slide):
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=slide, max_out_of_orderness_ms=3000,
              median_log_capacity=400_000, key_mode=key_mode, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=37, events_per_ms=2, jitter_ms=3000, late_frac_ppm=20000, late_extra_ms=20000)   # even and odd counts, long runs
    spec2 = StreamSpec(n_keys=3000, events_per_ms=3, late_frac_ppm=20000)
    for b in range(6):
        ts, key, val = generate_numpy(spec if b % 2 == 0 else spec2, b * 30_000, 30_000)
        val = np.round(val, 1)                                  # many ties; -0.0 and 0.0 both present
        val[::97] = -0.0
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)
    gpu.flush(); cpu.flush()
    a, w = sort_walerts(gpu.poll_medians()), sort_walerts(cpu.poll_medians())
    assert len(w) == cpu.counters()["windows_fired"] > 1000
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.window_end, w.window_end)
    np.testing.assert_array_equal(a.value.view(np.uint64), w.value.view(np.uint64))          # medians are order statistics: bit-exact
    assert_fired_equal(gpu.poll_fired(), cpu.poll_fired())


def test_window_median_run_lengths_around_the_short_long_boundary():
    """Runs of 1 .. 34 values are ordered by one thread, longer ones by a block's radix select (csrc/fmq_median.cuh):
    every length around the boundary, even and odd, one run of 10^5 values, ties, both zeros and both infinities."""
    kw = dict(window_enabled=1, agg_mask=1, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=0,
              median_log_capacity=500_000, **dict(SMALL, max_batch=1 << 18))
    gpu, cpu = pair(**kw)
    lens = list(range(1, 36)) + [63, 64, 65, 255, 256, 257, 1000, 4097, 100_001, 100_000]
    rng = np.random.default_rng(5)
    key = np.repeat(np.arange(len(lens), dtype=np.uint64) + 1, lens)
    perm = rng.permutation(len(key)); key = key[perm]
    for w in range(2):
        val = np.round(rng.normal(0, 3, len(key)), 1 if w == 0 else 12)
        val[::53] = -0.0; val[1::53] = 0.0; val[7::211] = np.inf; val[11::223] = -np.inf
        ts = 1_600_000_000_000 + w * 10_000 + rng.integers(0, 10_000, len(key)).astype(np.int64)
        gpu.push_batch(ts, key, val, capi.FMQ_WM_KEEP); cpu.push_batch(ts, key, val, capi.FMQ_WM_KEEP)
    gpu.flush(); cpu.flush()
    a, w = sort_walerts(gpu.poll_medians()), sort_walerts(cpu.poll_medians())
    assert len(w) == 2 * len(lens)
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.window_end, w.window_end)
    np.testing.assert_array_equal(a.value.view(np.uint64), w.value.view(np.uint64))


def test_window_median_many_windows_per_fire_small_work_buffer_and_snapshot():
    """Watermarks that jump: a dozen sliding windows fire in one push (all of them selected and sorted in one pass, or in
    groups when the (record, window) pairs would not fit the work buffer: the log capacity here is barely above the
    log's size), and a snapshot taken in the middle carries the value log along."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=20_000, window_slide_ms=5000, max_out_of_orderness_ms=2000,
              median_log_capacity=70_000, **SMALL)
    gpu, cpu = pair(**kw)
    spec = StreamSpec(n_keys=500, events_per_ms=1, jitter_ms=2000, late_frac_ppm=10000, late_extra_ms=15000)
    med = []
    wms = [capi.FMQ_WM_KEEP, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP, capi.FMQ_WM_AUTO, capi.FMQ_WM_AUTO]
    for b, wm in enumerate(wms):
        ts, key, val = generate_numpy(spec, b * 20_000, 20_000)
        gpu.push_batch(ts, key, val, wm); cpu.push_batch(ts, key, val, wm)
        if b == 3:
            med.append(gpu.poll_medians()); fired_before = gpu.poll_fired()
            snap = gpu.snapshot()
            gpu.close()
            gpu = capi.Engine(**kw)
            gpu.restore(snap)
    gpu.flush(); cpu.flush()
    med.append(gpu.poll_medians())
    a = sort_walerts(capi.WindowAlerts(*[np.concatenate([getattr(m, f) for m in med]) for f in ("key", "window_end", "value")])); w = sort_walerts(cpu.poll_medians())
    assert len(w) > 5000
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.window_end, w.window_end)
    np.testing.assert_array_equal(a.value.view(np.uint64), w.value.view(np.uint64))
    f2 = gpu.poll_fired()
    assert_fired_equal(capi.Fired(*[np.concatenate([getattr(x, f) for x in (fired_before, f2)]) for f in ("key", "window_end", "count", "sum", "min", "max")]), cpu.poll_fired())


# ---- checkpoint / restore ---------------------------------------------------------
@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
def test_snapshot_restore_continues_exactly(key_mode):
    """Kill the operator mid-stream, restore a fresh handle from the snapshot, finish the stream: the
    concatenated output equals an uninterrupted run of the oracle."""
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=20_000, window_slide_ms=5000, max_out_of_orderness_ms=4000,
              walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=50.0, rolling_max_enabled=1,
              consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1, key_mode=key_mode, **SMALL)
    spec = StreamSpec(n_keys=2000, events_per_ms=3, late_frac_ppm=20000)
    cpu = OracleEngine(**kw)
    gpu = capi.Engine(**kw)
    fired, cal, wal = [], [], []

    def collect(e):
        fired.append(e.poll_fired()); cal.append(e.poll_consec_alerts()); wal.append(e.poll_window_alerts())
    for b in range(6):
        ts, key, val = generate_numpy(spec, b * 20_000, 20_000)
        cpu.push_batch(ts, key, val); gpu.push_batch(ts, key, val)
        np.testing.assert_array_equal(gpu.poll_rolling_max(20_000).view(np.uint64), cpu.poll_rolling_max(20_000).view(np.uint64))
        if b == 2:
            with pytest.raises(capi.FmqError):
                gpu.snapshot()                       # un-polled output rows are not state
            collect(gpu)
            snap = gpu.snapshot()
            gpu.close()
            gpu = capi.Engine(**kw)
            gpu.restore(snap)
    gpu.flush(); cpu.flush(); collect(gpu)
    cat = lambda parts, cls, fields: cls(*[np.concatenate([getattr(p, f) for p in parts]) for f in fields])  # noqa: E731
    assert_fired_equal(cat(fired, capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), cpu.poll_fired())
    assert_events_equal(cat(cal, capi.EventRows, ("ts", "key", "val")), cpu.poll_consec_alerts())
    a, w = sort_walerts(cat(wal, capi.WindowAlerts, ("key", "window_end", "value"))), sort_walerts(cpu.poll_window_alerts())
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_allclose(a.value, w.value, rtol=1e-9)
    cg, cc = gpu.counters(), cpu.counters()
    assert cg["late_dropped"] == cc["late_dropped"] and cg["windows_fired"] == cc["windows_fired"] and cg["events_in"] == cc["events_in"]
    with pytest.raises(capi.FmqError):
        capi.Engine(**dict(kw, pane_ring=64)).restore(snap)        # different geometry is refused


def test_overlapped_polls_return_the_same_rows():
    """fmq_poll_all_begin/end (copy-out on a second stream while the next batch is pushed) against the
    oracle, batch by batch."""
    import torch
    kw, spec = CONFIGS["c3_tumbling_f64"]
    kw = {k: v for k, v in kw.items() if k != "late_capacity"}
    gpu = capi.Engine(overlapped_polls=1, **kw, **SMALL)
    cpu = OracleEngine(**kw)
    cap = 1 << 16
    sets = [([torch.empty(cap, dtype=torch.int64).pin_memory() for _ in range(3)] + [torch.empty(cap, dtype=torch.float64).pin_memory() for _ in range(3)],
             [torch.empty(cap, dtype=torch.int64).pin_memory(), torch.empty(cap, dtype=torch.int64).pin_memory(), torch.empty(cap, dtype=torch.float64).pin_memory()],
             [torch.empty(cap, dtype=torch.int64).pin_memory(), torch.empty(cap, dtype=torch.int64).pin_memory(), torch.empty(cap, dtype=torch.float64).pin_memory()])
            for _ in range(2)]
    fired, eal = [], []
    pending = None

    def harvest():
        nonlocal pending
        if pending is None:
            return
        gpu.poll_all_end()
        (f, w, e), (nf, nw, ne) = pending
        fired.append(capi.Fired(f[0][:nf].numpy().view(np.uint64).copy(), f[1][:nf].numpy().copy(), f[2][:nf].numpy().copy(),
                                f[3][:nf].numpy().copy(), f[4][:nf].numpy().copy(), f[5][:nf].numpy().copy()))
        eal.append(capi.EventRows(e[0][:ne].numpy().copy(), e[1][:ne].numpy().view(np.uint64).copy(), e[2][:ne].numpy().copy()))
        pending = None
    for b in range(8):
        ts, key, val = generate_numpy(spec, b * 20_000, 20_000)
        gpu.push_batch(ts, key, val); cpu.push_batch(ts, key, val)      # the D2H of the previous batch is still in flight here
        harvest()
        bufs = sets[b % 2]
        pending = (bufs, gpu.poll_all_begin(*bufs))
    harvest()
    gpu.flush(); cpu.flush()
    fired.append(gpu.poll_fired()); eal.append(gpu.poll_event_alerts())
    cat = lambda parts, cls, fields: cls(*[np.concatenate([getattr(p, f) for p in parts]) for f in fields])  # noqa: E731
    assert_fired_equal(cat(fired, capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), cpu.poll_fired())
    assert_events_equal(cat(eal, capi.EventRows, ("ts", "key", "val")), cpu.poll_event_alerts())


# ---- keyBy exchange ------------------------------------------------------------------------------
@pytest.mark.parametrize("world", [1, 2, 3, 7, 8, 24, 32, 64, 200])      # 3-bit digits up to 8 owners, 5-bit up to 32, 8-bit beyond; powers of two and not
def test_partition_by_owner_is_a_stable_split(world):
    import torch
    e = capi.Engine(efilter_enabled=1, max_batch=1 << 17)
    spec = StreamSpec(n_keys=5000, events_per_ms=5)
    n = 100_003
    ts, key, val = generate_numpy(spec, 0, n)
    d = [torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).cuda() for x in (ts, key, val)]
    o = [torch.empty_like(x) for x in d]
    counts = e.partition_by_owner(d[0], d[1], d[2], world, o[0], o[1], o[2])
    lib = capi.load_library()
    owner = np.array([lib.fmq_owner_of(int(k), world) for k in key[:2000]])
    from monitor_systam_flink_quickstart_b200.shard import owner_of_numpy
    own = owner_of_numpy(key, world)
    np.testing.assert_array_equal(own[:2000], owner)
    order = np.argsort(own, kind="stable")
    np.testing.assert_array_equal(counts, np.bincount(own, minlength=world).astype(np.uint64))
    np.testing.assert_array_equal(o[0].cpu().numpy(), ts[order])
    np.testing.assert_array_equal(o[1].cpu().numpy().view(np.uint64), key[order])
    np.testing.assert_array_equal(o[2].cpu().numpy(), val[order])


# ---- narrow wire format (fmq_push_packed) ----------------------------------------------------------
@pytest.mark.parametrize("name", ["c3_tumbling_f64", "c4_sliding_i64"])
@pytest.mark.parametrize("where", ["host", "device"])
def test_packed_columns_give_the_results_of_wide_columns(name, where):
    """uint32 timestamp offsets / uint32 key ids / int32 values are widened on the device: same rows
    as the oracle fed the wide columns.  Seven host pushes in a row cycle both staging sets."""
    import torch
    kw, spec = CONFIGS[name]
    keyed = dict(consec_n=3, consec_threshold=60.0, consec_reset_on_fire=1) if name == "c3_tumbling_f64" else {}
    gpu, cpu = pair(**kw, **keyed, **SMALL)
    batch, total = 23_456, 7 * 23_456
    for s in range(0, total, batch):
        ts, key, val = generate_numpy(spec, s, batch)
        base = int(ts.min()) - 17
        cols = [(ts - base).astype(np.uint32), key, val]
        if key.max() < (1 << 32):
            cols[1] = key.astype(np.uint32)
        if val.dtype == np.int64 and np.abs(val).max() < (1 << 31):
            cols[2] = val.astype(np.int32)
        if s // batch == 3:
            cols[0], base = ts, 0                               # a wide timestamp column among narrow ones
        assert any(c.dtype.itemsize == 4 for c in cols)
        if where == "device":
            cols = [torch.from_numpy(c.view(np.int32) if c.dtype == np.uint32 else (c.view(np.int64) if c.dtype == np.uint64 else c)).cuda() for c in cols]
        gpu.push_packed(cols[0], cols[1], cols[2], ts_base=base)
        gpu.synchronize()
        cpu.push_batch(ts, key, val)
        if (s // batch) % 3 == 2:
            compare_all(gpu, cpu, agg_mask=kw["agg_mask"])
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu, agg_mask=kw["agg_mask"])
    assert cpu.counters()["windows_fired"] > 0


def test_packed_columns_bad_arguments():
    gpu = capi.Engine(window_enabled=1, window_size_ms=1000, window_slide_ms=1000, **SMALL)       # an f64 handle
    ts = np.arange(10, dtype=np.uint32); key = np.arange(10, dtype=np.uint32)
    with pytest.raises(capi.FmqError) as e:
        gpu.push_packed(ts, key, np.arange(10, dtype=np.int32))
    assert e.value.status == 1 and "int32" in str(e.value)                          # FMQ_ERR_BAD_ARG
    with pytest.raises(capi.FmqError):
        gpu.push_packed(ts, key, np.zeros(10), ts_base=(1 << 63) - 5)
    with pytest.raises(capi.FmqError):
        gpu.push_packed(ts, key.astype(np.uint16), np.zeros(10))
    gpu.push_packed(ts, key, np.ones(10), ts_base=5_000)                                           # and a good one still works
    gpu.flush()
    assert gpu.poll_fired().count.sum() == 10


@pytest.mark.parametrize("keyed", [0, 1])
def test_large_host_pushes_are_copied_in_pieces(keyed):
    """Host batches above 2^18 events go over in up to four pieces whose ingest overlaps the next copy;
    cutting a micro-batch must not change a single row (wide and packed columns, odd sizes)."""
    kw, spec = CONFIGS["c3_tumbling_f64"]
    extra = dict(consec_n=3, consec_threshold=60.0, consec_reset_on_fire=1) if keyed else {}
    big = dict(table_capacity=1 << 12, pane_ring=1024, max_batch=1 << 21, fired_capacity=1 << 21, walert_capacity=1 << 21, ealert_capacity=1 << 21)
    if keyed:
        big["calert_capacity"] = 1 << 21
    gpu, cpu = pair(**dict(kw, **extra), **big)
    s = 0
    for i, n in enumerate([1_000_003, 300_001, 1 << 21, 777_777]):
        ts, key, val = generate_numpy(spec, s, n); s += n
        if i % 2 == 0:
            gpu.push_batch(ts, key, val)
        else:
            base = int(ts.min())
            gpu.push_packed((ts - base).astype(np.uint32), key.astype(np.uint32), val, ts_base=base)
        cpu.push_batch(ts, key, val)
        compare_all(gpu, cpu)
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)


def test_staging_the_next_host_batch_ahead():
    """fmq_stage_host starts the copy-in of batch k+1 before batch k's results are polled; rows are
    those of plain pushes; a push of other columns while one batch is staged is refused."""
    kw, spec = CONFIGS["c3_tumbling_f64"]
    gpu, cpu = pair(**kw, **SMALL)
    n = 30_011
    batches = [generate_numpy(spec, s * n, n) for s in range(6)]
    packed = []
    for i, (ts, key, val) in enumerate(batches):
        if i % 2:
            base = int(ts.min())
            packed.append(((ts - base).astype(np.uint32), key.astype(np.uint32), val, base))
        else:
            packed.append((ts, key, val, 0))
    gpu.stage_host(*packed[0][:3], ts_base=packed[0][3])
    gpu.stage_host(*packed[1][:3], ts_base=packed[1][3])                # two batches ahead, first in first out
    for i, cols in enumerate(packed):
        if i == 2:
            with pytest.raises(capi.FmqError) as e:
                gpu.push_batch(*batches[0])                     # not the oldest staged columns
            assert e.value.status == 1 and "staged" in str(e.value)
            with pytest.raises(capi.FmqError):
                gpu.push_packed(*packed[3][:3], ts_base=packed[3][3])   # staged, but not the oldest
            with pytest.raises(capi.FmqError):
                gpu.stage_host(*packed[4][:3], ts_base=packed[4][3])    # only two batches ahead
        gpu.push_packed(*cols[:3], ts_base=cols[3])
        if i + 2 < len(packed):
            gpu.stage_host(*packed[i + 2][:3], ts_base=packed[i + 2][3])
        cpu.push_batch(*batches[i])
        compare_all(gpu, cpu)
    gpu.flush(); cpu.flush()
    compare_all(gpu, cpu)
