"""Device text ingest against the host restatement of the reference's map functions
(text.py: Main.java:21-25, BandwidthMonitorWithEventTime.java:33-44), bit for bit."""
import json
import os

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.text import parse_cpu_line, parse_flow_line, text_key_hash
from oracle_lib import OracleEngine, assert_fired_equal

pytestmark = pytest.mark.gpu
G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "readme_sessions.json"), encoding="utf-8"))


def test_key_hash_host_and_library_agree():
    lib = capi.load_library()
    for name in ("10.8.22.1", "www.163.com", "", "h" * 300):
        assert lib.fmq_text_key_hash(name.encode(), len(name)) == text_key_hash(name)


def test_readme_lines_parse_like_the_reference():
    cpu_lines = G["G2_passthrough"]["input"] + G["G4_avg"]["input"] + ["1563452051 10.8.22.1 cpu2 99.2"]
    e = capi.Engine(efilter_enabled=1, text_max_bytes=1 << 16, max_batch=1 << 12)
    ts, key, val, st = e.parse_text(("\n".join(cpu_lines) + "\n").encode(), capi.FMQ_TEXT_CPU)
    want = [parse_cpu_line(l) for l in cpu_lines]
    assert st.tolist() == [0] * len(want)
    assert ts.tolist() == [w[0] * 1000 for w in want] and val.tolist() == [w[3] for w in want]
    assert key.tolist() == [text_key_hash(w[1]) for w in want]
    flow_lines = G["G7_event_time"]["input"]
    e = capi.Engine(value_type=capi.FMQ_VAL_I64, efilter_enabled=1, text_max_bytes=1 << 16, max_batch=1 << 12)
    ts, key, val, st = e.parse_text("\n".join(flow_lines).encode(), capi.FMQ_TEXT_FLOW)       # no trailing newline
    want = [parse_flow_line(l) for l in flow_lines]
    assert st.tolist() == [0] * 5 and ts.tolist() == [w[0] for w in want] and val.tolist() == [w[2] for w in want]
    assert ts[0] == 1566957600000 and key.tolist() == [text_key_hash("www.163.com")] * 5


def test_random_lines_bit_exact_and_edge_cases():
    rng = np.random.default_rng(7)
    n = 200_000
    secs = rng.integers(1_500_000_000, 1_700_000_000, n)
    hosts = rng.integers(0, 5000, n)
    # values printed with 1..15 significant digits and assorted exponents: all inside the exact path
    vals = []
    for i in range(n):
        k = i % 6
        x = rng.random() * 100
        vals.append([f"{x:.1f}", f"{x:.6f}", f"{x:.12f}", f"{x * 1e-3:.3e}", f"{int(x)}", f"-{x:.2f}"][k])
    lines = [f"{s} 10.{h >> 8}.{h & 255}.7 cpu{h % 8} {v}" for s, h, v in zip(secs, hosts, vals)]
    text = ("\n".join(lines) + "\n").encode()
    e = capi.Engine(efilter_enabled=1, text_max_bytes=len(text) + 100, max_batch=1 << 18)
    ts, key, val, st = e.parse_text(text, capi.FMQ_TEXT_CPU)
    assert len(ts) == n and st.sum() == 0
    np.testing.assert_array_equal(ts, secs.astype(np.int64) * 1000)
    np.testing.assert_array_equal(val.view(np.uint64), np.array([float(v) for v in vals]).view(np.uint64))   # == Double.parseDouble
    np.testing.assert_array_equal(key[:1000], np.array([text_key_hash(l.split(" ")[1]) for l in lines[:1000]], np.uint64))
    # edge cases: host-side values, malformed lines, CRLF, empty line
    odd = ["1 h c 0.1234567890123456789", "2 h c 1e400", "3 h c NaN", "oops", "4 h c", "5 h c 7.25\r", "", "6 h c -0.0"]
    ts, key, val, st = e.parse_text("\n".join(odd).encode(), capi.FMQ_TEXT_CPU)
    assert st.tolist() == [2, 2, 2, 1, 1, 0, 1, 0]
    assert val[5] == 7.25 and np.signbit(val[7]) and ts[5] == 5000


def test_push_text_equals_push_of_host_parsed_columns():
    """Text straight into the operator chain == host parse (the reference's map) + fmq_push_batch; and
    against the oracle."""
    lines = []
    rng = np.random.default_rng(3)
    t0 = 1566957600
    for i in range(30_000):
        sec = t0 + i // 20 - int(rng.integers(0, 4))
        hh, mm, ss = (sec - t0) // 3600 + 10, (sec - t0) % 3600 // 60, sec % 60
        lines.append(f"2019-08-28T{hh:02d}:{mm:02d}:{ss:02d} site{int(rng.integers(0, 300))}.example.com {int(rng.integers(0, 1 << 20))}")
    kw = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=15_000, max_out_of_orderness_ms=2000,
              table_capacity=1024, pane_ring=64, max_batch=1 << 15)
    gpu = capi.Engine(text_max_bytes=1 << 22, **kw)
    cpu = OracleEngine(**kw)
    for lo in range(0, len(lines), 10_000):
        chunk = lines[lo:lo + 10_000]
        n = gpu.push_text(("\n".join(chunk) + "\n").encode(), capi.FMQ_TEXT_FLOW)
        assert n == len(chunk)
        rows = [parse_flow_line(l) for l in chunk]
        cpu.push_batch(np.array([r[0] for r in rows], np.int64), np.array([text_key_hash(r[1]) for r in rows], np.uint64), np.array([r[2] for r in rows], np.int64))
    gpu.flush(); cpu.flush()
    assert_fired_equal(gpu.poll_fired(), cpu.poll_fired())
    with pytest.raises(capi.FmqError):
        gpu.push_text(b"not a line\n", capi.FMQ_TEXT_FLOW)
