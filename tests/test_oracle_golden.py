"""Pins the CPU oracle (and the independent Python model) against every golden vector the
reference holds for this path: the README sessions G1-G8 of SURVEY.md §4, extracted verbatim by
tests/golden/make_readme_sessions.py into tests/golden/readme_sessions.json."""
import json
import os
import sys

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.text import KeyDictionary, parse_cpu_line, parse_flow_line
from oracle_lib import OracleEngine, lib, sort_fired

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import naive_model  # noqa: E402

G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_sessions.json"), encoding="utf-8"))


class NaiveAdapter:
    """Gives NaiveFlinkModel the same surface as EngineBase for these tests."""

    def __init__(self, **kw):
        self.m = naive_model.NaiveFlinkModel(**kw)
        self.f64 = kw.get("value_type", 0) == 0

    def push_batch(self, ts, key, val, wm=capi.FMQ_WM_AUTO):
        self.m.push_batch(ts, key, val, wm)

    def flush(self, wm=capi.FMQ_WM_FINAL):
        self.m.flush(wm)

    def poll_fired(self):
        r, self.m.fired = self.m.fired, []
        vd = np.float64 if self.f64 else np.int64
        cols = list(zip(*r)) if r else [[]] * 6
        return capi.Fired(np.array(cols[0], np.uint64), np.array(cols[1], np.int64), np.array(cols[2], np.int64),
                          np.array(cols[3], vd), np.array(cols[4], vd), np.array(cols[5], vd))

    def poll_window_alerts(self):
        r, self.m.walerts = self.m.walerts, []
        cols = list(zip(*r)) if r else [[]] * 3
        return capi.WindowAlerts(np.array(cols[0], np.uint64), np.array(cols[1], np.int64), np.array(cols[2], np.float64))

    def poll_event_alerts(self):
        r, self.m.ealerts = self.m.ealerts, []
        cols = list(zip(*r)) if r else [[]] * 3
        return capi.EventRows(np.array(cols[0], np.int64), np.array(cols[1], np.uint64), np.array(cols[2], np.float64 if self.f64 else np.int64))

    def poll_rolling_max(self, n):
        return np.array(self.m.rolling, np.float64 if self.f64 else np.int64)

    def counters(self):
        return {"late_dropped": self.m.late_dropped, "events_in": self.m.events_in, "windows_fired": self.m.windows_fired}


def engines(**kw):
    return [OracleEngine(**kw), NaiveAdapter(**kw)]


def _cpu_cols(lines, d):
    rows = [parse_cpu_line(l) for l in lines]
    ts = np.array([r[0] * 1000 for r in rows], np.int64)
    key = np.array([d.encode(r[1]) for r in rows], np.uint64)
    val = np.array([r[3] for r in rows], np.float64)
    return ts, key, val


@pytest.mark.parametrize("which", [0, 1])
def test_g1_threshold_filter(which):
    """chapter1/README.md:114-123: only the 99.2 reading passes `value.f2 > 90` (Main.java:31)."""
    d = KeyDictionary()
    e = engines(efilter_enabled=1, efilter_op=capi.FMQ_CMP_GT, efilter_threshold=90.0)[which]
    ts, key, val = _cpu_cols(G["G1_filter"]["input"], d)
    e.push_batch(ts, key, val)
    out = e.poll_event_alerts()
    assert [f"({d.decode(k)},cpu2,{v})" for k, v in zip(out.key, out.val)] == [G["G1_filter"]["output"][0].split("> ")[1]]
    # strictness: exactly 90 does not alert
    e.push_batch(np.array([1], np.int64), np.array([0], np.uint64), np.array([90.0]))
    assert len(e.poll_event_alerts()) == 0


@pytest.mark.parametrize("which", [0, 1])
def test_g3_rolling_max(which):
    """chapter2/README.md:54-66: keyBy(0).max(2) emits 80.5, 80.5, 99.9."""
    d = KeyDictionary()
    e = engines(rolling_max_enabled=1)[which]
    ts, key, val = _cpu_cols(G["G3_rolling_max"]["input"], d)
    e.push_batch(ts, key, val)
    want = [float(o.split(",")[-1].rstrip(")")) for o in G["G3_rolling_max"]["output"]]
    assert list(e.poll_rolling_max(len(ts))) == want


@pytest.mark.parametrize("which", [0, 1])
def test_g4_window_avg(which):
    """chapter2/README.md:152-164: 1-minute tumbling count/sum -> avg, sequential double sum
    (ComputeCpuAvg.java:41-42,49): 86.26666666666667 and 20.2 exactly."""
    d = KeyDictionary()
    e = engines(window_enabled=1, agg_mask=3, window_size_ms=60000, window_slide_ms=60000,
                walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=-1.0)[which]
    _, key, val = _cpu_cols(G["G4_avg"]["input"], d)
    # the job runs on processing time: the four lines are typed within one wall-clock minute
    ts = 1563452040000 + np.arange(4, dtype=np.int64) * 1000
    e.push_batch(ts, key, val, 0)
    e.flush()
    a = e.poll_window_alerts()
    got = {d.decode(k): v for k, v in zip(a.key, a.value)}
    want = [float(o.split("> ")[1]) for o in G["G4_avg"]["output"]]
    assert got == {"10.8.22.1": want[0], "10.8.22.2": want[1]}
    f = e.poll_fired()
    assert sorted(f.count.tolist()) == [1, 3]


@pytest.mark.parametrize("which", [0, 1])
def test_g6_tumbling_sum_and_filter(which):
    """chapter3/README.md:71-81: long sum 11200 (BandwidthMonitor.java:37) and the alert filter
    11200*8.0/60/1024/1024 < 100 (BandwidthMonitor.java:39) lets it through."""
    d = KeyDictionary()
    e = engines(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, window_size_ms=60000, window_slide_ms=60000,
                walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                walert_div1=1024.0, walert_div2=1024.0, walert_threshold=100.0)[which]
    rows = [parse_flow_line(l) for l in G["G6_sum"]["input"]]
    key = np.array([d.encode(r[1]) for r in rows], np.uint64)
    val = np.array([r[2] for r in rows], np.int64)
    ts = 1566957600000 + np.arange(4, dtype=np.int64) * 1000        # processing time: same minute
    e.push_batch(ts, key, val, 0)
    e.flush()
    f = e.poll_fired()
    assert [f"({d.decode(k)},{s})" for k, s in zip(f.key, f.sum)] == G["G6_sum"]["output"]
    a = e.poll_window_alerts()
    assert a.value.tolist() == [11200 * 8.0 / 60 / 1024 / 1024] == [0.0014241536458333333]


@pytest.mark.parametrize("which", [0, 1])
def test_g7_event_time_sliding_late_drop(which):
    """chapter3/README.md:284-297: sliding 5 min / 5 s, bound 1 min.  With a watermark after every
    record: 60 fired windows (12 x 10000, 12 x 10100, 36 x 10200), the 09:01 record is dropped as
    late, and the two printed Mbps doubles reproduce bit-exactly."""
    d = KeyDictionary()
    e = engines(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, window_size_ms=300000, window_slide_ms=5000,
                max_out_of_orderness_ms=60000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT,
                walert_mul=8.0, walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=100.0)[which]
    rows = [parse_flow_line(l) for l in G["G7_event_time"]["input"]]
    assert rows[0][0] == 1566957600000                               # epoch check of SURVEY.md §4
    for t, ch, flow in rows:
        e.push_batch(np.array([t], np.int64), np.array([d.encode(ch)], np.uint64), np.array([flow], np.int64))
    f = sort_fired(e.poll_fired())
    assert len(f) == 60
    assert f.sum.tolist() == [10000] * 12 + [10100] * 12 + [10200] * 36
    assert f.window_end[0] == 1566957605000 and f.window_end[-1] == 1566957900000
    assert e.counters()["late_dropped"] == 1
    a = e.poll_window_alerts()
    want = [float(o.split(",")[1].rstrip(")")) for o in G["G7_event_time"]["output"]]
    assert want == [0.0012715657552083333, 0.0012969970703125]
    assert set(a.value.tolist()) >= set(want) and len(a) == 60
    # end of input: Watermark(Long.MAX_VALUE) fires what is still open: window ends 10:05:05..10:06:00
    # hold 10:01+10:02 (200), 10:06:05..10:07:00 hold 10:02+10:06 (200), 10:07:05..10:11:00 hold 10:06 (100)
    e.flush()
    f2 = sort_fired(e.poll_fired())
    assert f2.sum.tolist() == [200] * 24 + [100] * 48


def test_g7_same_watermark_period_is_not_late():
    """If all five lines land inside one watermark period the 09:01 record is NOT late."""
    for e in engines(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, window_size_ms=300000,
                     window_slide_ms=5000, max_out_of_orderness_ms=60000):
        rows = [parse_flow_line(l) for l in G["G7_event_time"]["input"]]
        e.push_batch(np.array([r[0] for r in rows], np.int64), np.zeros(5, np.uint64), np.array([r[2] for r in rows], np.int64))
        assert e.counters()["late_dropped"] == 0
        e.flush()
        f = e.poll_fired()
        # 09:01 opens 60 windows of its own (ends 09:01:05..09:06:00), none overlapping the 10:xx ones
        assert int((f.window_end <= 1566954360000).sum()) == 60 and set(f.sum[f.window_end <= 1566954360000].tolist()) == {100}


def test_g8_key_group_routing():
    """Print prefixes `3>` `3>` `2>` (chapter2/README.md:63-65,162-163; chapter3/README.md:295-296):
    subtask = murmurHash(String.hashCode()) % 128 * 4 / 128 -> key groups 78, 90, 58."""
    l = lib()
    want = {"10.8.22.1": (78, 2), "10.8.22.2": (90, 2), "www.163.com": (58, 1)}
    for name, (kg, sub) in want.items():
        h = l.orc_java_string_hash(name.encode(), len(name))
        assert h == naive_model.java_string_hash(name)
        assert l.orc_key_group(h, 128) == kg == naive_model.key_group(h, 128)
        assert l.orc_subtask(kg, 4, 128) == sub == naive_model.subtask(kg, 4, 128)
    prefixes = [int(o.split(">")[0]) for o in G["G3_rolling_max"]["output"] + G["G4_avg"]["output"] + G["G7_event_time"]["output"]]
    assert prefixes == [3, 3, 3, 3, 3, 2, 2]


def test_g2_passthrough_uses_four_subtasks():
    """chapter1/README.md:72-84: un-keyed print prefixes 3> 4> 1> — consistent with p = 4 round-robin."""
    p = [int(o.split(">")[0]) for o in G["G2_passthrough"]["output"]]
    assert p == [3, 4, 1] and all(((p[i] - 1 + 1) % 4) + 1 == p[i + 1] for i in range(2))


def test_g5_median_oracle():
    """chapter2/README.md:236-248 through the oracle's ListState restatement (ComputeCpuMiddle.java:36-47)."""
    d = KeyDictionary()
    e = OracleEngine(window_enabled=1, agg_mask=1, window_size_ms=60000, window_slide_ms=60000, median_log_capacity=1)
    _, key, val = _cpu_cols(G["G5_median"]["input"], d)
    e.push_batch(1563452040000 + np.arange(4, dtype=np.int64) * 1000, key, val, 0)
    e.flush()
    m = e.poll_medians()
    assert {d.decode(k): v for k, v in zip(m.key, m.value)} == {"10.8.22.1": 80.5, "10.8.22.2": 20.2}


def test_g5_median_is_out_of_scope_value_check():
    """chapter2/README.md:236-248 (ComputeCpuMiddle.java:41-47): median 80.5 / 20.2 — SURVEY §8f N3,
    restated here only so the fixture is exercised."""
    rows = [parse_cpu_line(l) for l in G["G5_median"]["input"]]
    by = {}
    for _, host, _, v in rows:
        by.setdefault(host, []).append(v)
    med = []
    for host in ("10.8.22.1", "10.8.22.2"):
        v = sorted(by[host]); n = len(v)
        med.append(v[n // 2] if n % 2 else (v[n // 2] + v[n // 2 - 1]) / 2)
    assert med == [float(o.split("> ")[1]) for o in G["G5_median"]["output"]]
