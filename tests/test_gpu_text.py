"""Device text ingest against the host restatement of the reference's map functions
(text.py: Main.java:21-25, BandwidthMonitorWithEventTime.java:33-44), bit for bit."""
import json
import os

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.text import KeyDictionary, parse_cpu_line, parse_flow_line, text_key_hash
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


def test_device_dictionary_gives_dense_ids_and_names_back():
    """text_dict_capacity > 0: the key column of the text parser carries dense ids (0, 1, 2, ... one per distinct
    string, no gaps) that index a FMQ_KEYS_DENSE table directly, and fmq_dict_lookup returns the strings, so that
    results can be printed with the name the reference prints (BandwidthMonitorWithEventTime.java:55).  The fired
    rows, re-keyed by NAME, equal the oracle fed the host dictionary's ids re-keyed by name."""
    rng = np.random.default_rng(11)
    t0 = 1566957600
    names = [f"site{i}.example.com" for i in range(700)] + ["a", "bb", "www.163.com", "x" * 40]
    lines = []
    for i in range(40_000):
        sec = t0 + i // 25 - int(rng.integers(0, 3))
        hh, mm, ss = (sec - t0) // 3600 + 10, (sec - t0) % 3600 // 60, sec % 60
        lines.append(f"2019-08-28T{hh:02d}:{mm:02d}:{ss:02d} {names[int(rng.integers(0, len(names)))]} {int(rng.integers(0, 1 << 20))}")
    kw = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=20_000, max_out_of_orderness_ms=2000,
              table_capacity=1024, pane_ring=64, max_batch=1 << 15)
    gpu = capi.Engine(text_max_bytes=1 << 22, text_dict_capacity=1024, key_mode=capi.FMQ_KEYS_DENSE, **kw)
    cpu = OracleEngine(**kw)
    host_dict = KeyDictionary()
    for lo in range(0, len(lines), 8_000):
        chunk = lines[lo:lo + 8_000]
        assert gpu.push_text(("\n".join(chunk) + "\n").encode(), capi.FMQ_TEXT_FLOW) == len(chunk)
        rows = [parse_flow_line(l) for l in chunk]
        cpu.push_batch(np.array([r[0] for r in rows], np.int64), np.array([host_dict.encode(r[1]) for r in rows], np.uint64), np.array([r[2] for r in rows], np.int64))
    gpu.flush(); cpu.flush()
    n_ids = gpu.dict_size()
    assert n_ids == len(host_dict) == len(set(l.split(" ")[1] for l in lines))
    dev_names = gpu.dict_lookup(np.arange(n_ids, dtype=np.uint64))
    assert sorted(dev_names) == sorted(set(l.split(" ")[1] for l in lines)) and gpu.dict_lookup([n_ids + 5]) == [None]
    g, c = gpu.poll_fired(), cpu.poll_fired()
    assert int(g.key.max()) < n_ids                                            # dense: every id below the number of names
    by_name = lambda f, nm: sorted(zip([nm(int(k)) for k in f.key], f.window_end.tolist(), f.count.tolist(), f.sum.tolist(), f.min.tolist(), f.max.tolist()))  # noqa: E731
    assert by_name(g, lambda k: dev_names[k]) == by_name(c, host_dict.decode)
    # the parser alone: ids are stable across calls (a string keeps the id of its first sight)
    ts, key, val, st = gpu.parse_text(b"2019-08-28T10:00:00 www.163.com 5\n2019-08-28T10:00:01 brand.new.host 6\n", capi.FMQ_TEXT_FLOW)
    assert st.tolist() == [0, 0] and dev_names[int(key[0])] == "www.163.com" and int(key[1]) == n_ids
    assert gpu.dict_lookup([n_ids]) == ["brand.new.host"]


def test_device_dictionary_full_hands_strings_to_the_host():
    e = capi.Engine(text_max_bytes=1 << 16, text_dict_capacity=3, efilter_enabled=1, max_batch=64)
    text = "".join(f"{100 + i} host{i % 5} cpu0 {10.5 + i}\n" for i in range(10)).encode()
    ts, key, val, st = e.parse_text(text, capi.FMQ_TEXT_CPU)
    assert e.dict_size() == 3
    assert sorted(set(key[st == 0].tolist())) == [0, 1, 2] and (st == capi.FMQ_TEXT_KEY_ON_HOST).sum() == 4     # two of the five hosts did not fit
