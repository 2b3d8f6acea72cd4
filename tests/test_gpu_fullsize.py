"""Full-size runs (BASELINE.json scale per step), compared ROW FOR ROW with the CPU oracle.

The checker is `ParallelOracle` (oracle/cpu_parallel.cpp): the key-partitioned multi-thread restatement
of the Flink semantics, itself pinned row-for-row to the literal single-thread oracle by
tests/test_oracle_cross.py.  It replays 10^7..10^8 events in seconds on the GPU box's host cores, so the
CUDA path is compared with it on exactly the workloads bench.py and BASELINE.json name:

  C3  bench.py's own configuration: 1 M hosts, 10 s tumbling, bound 5 s + late tail, avg alert and the
      per-event filter, 3 x 2^24 events, dense ids AND hashed keys;
  C4  10 M (host << 32 | metric) keys, 60 s / 10 s sliding, int64 flows, 2 x 2^24 events;
  C5  Zipf(1.2) over 10^8 ids, tumbling windows plus the consecutive-readings rule, 2^24 events;
  C2  100 K hosts, 3 consecutive readings > 90, 2^24 records: alerts and every key's final counter.

Bit-exact: keys, window ends, counts, integer sums, min / max, alert tuples, late-drop counts, final keyed
state; double sums within 1e-9 relative (BASELINE.json's tolerance; the atomics reorder the adds).  Rows are
brought into one order and compared on the GPU (tens of millions of rows), with torch as plumbing only.
The size-independent properties (conservation of counts and of the int64 checksum, one row per (key, window))
are still asserted: they localise a failure faster than a row diff does."""
import os

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
from oracle_lib import ParallelOracle

pytestmark = pytest.mark.gpu
THREADS = max(1, min(64, os.cpu_count() or 1))
BIG = 1 << 26


def _canon(cols, order):
    """cols: numpy columns of one row set.  Returns them as CUDA int64 tensors (bit patterns), rows sorted
    by the columns named in `order` (least significant first; stable sorts on the GPU)."""
    import torch
    t = [torch.from_numpy(np.ascontiguousarray(c).view(np.int64)).cuda() for c in cols]
    idx = torch.arange(len(t[0]), device="cuda")
    for ci in order:
        idx = idx[torch.sort(t[ci][idx], stable=True)[1]]
    return [c[idx] for c in t]


def _assert_rows(got_cols, want_cols, order, names, f64_cols=()):
    import torch
    assert len(got_cols[0]) == len(want_cols[0]), f"{names}: got {len(got_cols[0])} rows, want {len(want_cols[0])}"
    g, w = _canon(got_cols, order), _canon(want_cols, order)
    for i, name in enumerate(names):
        if i in f64_cols:
            a, b = g[i].view(torch.float64), w[i].view(torch.float64)
            bad = (a - b).abs() > 1e-9 * b.abs()
            assert not bool(bad.any()), f"{name}: {int(bad.sum())} rows beyond 1e-9 relative"
        else:
            assert torch.equal(g[i], w[i]), f"{name}: {int((g[i] != w[i]).sum())} rows differ"


def _compare(gpu, cpu, f64):
    f, o = gpu.poll_fired(chunk=BIG), cpu.poll_fired(chunk=BIG)
    m = gpu.cfg.agg_mask                                        # columns the job did not ask for are not defined: left out
    names = ["key", "window_end", "count"] + [n for n, bit in (("sum", capi.FMQ_AGG_SUM), ("min", capi.FMQ_AGG_MIN), ("max", capi.FMQ_AGG_MAX)) if m & bit]
    fcols = lambda r: [getattr(r, n) for n in names]  # noqa: E731
    _assert_rows(fcols(f), fcols(o), [0, 1], names, f64_cols=(names.index("sum"),) if (f64 and "sum" in names) else ())
    if gpu.cfg.walert_source:
        a, b = gpu.poll_window_alerts(chunk=BIG), cpu.poll_window_alerts(chunk=BIG)
        _assert_rows([a.key, a.window_end, a.value], [b.key, b.window_end, b.value], [0, 1], ["alert key", "alert window_end", "alert value"], f64_cols=(2,))
    if gpu.cfg.efilter_enabled:
        a, b = gpu.poll_event_alerts(chunk=BIG), cpu.poll_event_alerts(chunk=BIG)
        _assert_rows([a.ts, a.key, a.val], [b.ts, b.key, b.val], [2, 1, 0], ["event ts", "event key", "event val"])
    if gpu.cfg.consec_n:
        a, b = gpu.poll_consec_alerts(chunk=BIG), cpu.poll_consec_alerts(chunk=BIG)
        _assert_rows([a.ts, a.key, a.val], [b.ts, b.key, b.val], [2, 1, 0], ["consec ts", "consec key", "consec val"])
    cg, cc = gpu.counters(), cpu.counters()
    for k in ("events_in", "late_dropped", "windows_fired", "watermark"):
        assert cg[k] == cc[k], (k, cg, cc)
    return f, cg


def _feed(gpu, cpu, spec, nb, B):
    """The same nb x B events into both engines; the device columns are generated on the GPU, the oracle
    reads their host copy (bit-identical by construction of the generator)."""
    import torch
    tot = 0
    for b in range(nb):
        ts, key, val = generate_torch(spec, b * B, B)
        torch.cuda.synchronize()
        gpu.push_batch(ts, key, val)
        cpu.push_batch(ts.cpu().numpy(), key.cpu().numpy().view(np.uint64), val.cpu().numpy())
        if val.dtype == torch.int64:
            tot += int(val.sum().item())
        gpu.synchronize()            # the device columns are freed when this iteration ends
    gpu.flush(); cpu.flush()
    return tot


@pytest.mark.parametrize("key_mode", ["dense", "hashed"])
def test_c3_bench_configuration_rows_equal_the_oracle(key_mode):
    """Exactly what bench.py times (same kwargs, same stream), 3 x 2^24 events, both key modes."""
    import bench
    B, nb = 1 << 24, 3
    spec = bench.stream_spec(StreamSpec)
    kw = bench.workload_kwargs(capi, B)
    kw.update(fired_capacity=12_000_000, walert_capacity=12_000_000, ealert_capacity=6_000_000)
    gkw = dict(kw)
    if key_mode == "dense":
        gkw.update(key_mode=capi.FMQ_KEYS_DENSE, table_capacity=bench.N_KEYS)
    gpu, cpu = capi.Engine(**gkw), ParallelOracle(threads=THREADS, **kw)
    _feed(gpu, cpu, spec, nb, B)
    f, c = _compare(gpu, cpu, f64=True)
    assert c["late_dropped"] > 0 and int(f.count.sum()) == c["events_in"] - c["late_dropped"]      # tumbling: every accepted event in one window
    gpu.close(); cpu.close()


def test_c4_sliding_10m_composite_keys_rows_equal_the_oracle():
    """60 s / 10 s sliding sum per (host, metric), 10 M keys, int64 flows, 2 x 2^24 events, hashed keys."""
    B, nb, k = 1 << 24, 2, 6
    spec = StreamSpec(n_keys=10_000_000, events_per_ms=10_000, jitter_ms=4000, late_frac_ppm=0, value_type="i64", key_kind="host_metric")
    kw = dict(value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=15, window_size_ms=60_000, window_slide_ms=10_000,
              max_out_of_orderness_ms=5000, walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0,
              walert_div0=60.0, walert_div1=1024.0, walert_div2=1024.0, walert_threshold=0.05,
              table_capacity=15_000_000, pane_ring=16, max_batch=B, fired_capacity=80_000_000, walert_capacity=80_000_000)
    gpu, cpu = capi.Engine(**kw), ParallelOracle(threads=THREADS, **kw)
    tot_val = _feed(gpu, cpu, spec, nb, B)
    f, c = _compare(gpu, cpu, f64=False)
    assert c["late_dropped"] == 0
    assert int(f.count.sum()) == k * nb * B                       # every event in exactly size/slide windows
    assert int(f.sum.sum()) == k * tot_val                        # checksum of checksums, exact in int64
    assert np.all(f.window_end % 10_000 == 0)
    gpu.close(); cpu.close()


def test_c5_zipf_100m_ids_tumbling_and_consecutive_rule_rows_equal_the_oracle():
    """BASELINE.json configs[4] on one GPU: Zipf(1.2) over 10^8 ids (one id gets 18 % of the events), tumbling
    count/sum/min/max plus 3 consecutive readings > 50 per id, hashed keys, 2^24 events in two pushes."""
    B, nb = 1 << 23, 2
    spec = StreamSpec(n_keys=100_000_000, events_per_ms=1000, jitter_ms=3000, late_frac_ppm=500, late_extra_ms=20_000, key_kind="zipf")
    kw = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=10_000, max_out_of_orderness_ms=3000,
              walert_source=capi.FMQ_ALERT_MAX, walert_op=capi.FMQ_CMP_GT, walert_threshold=99.0,
              consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1,
              table_capacity=6_000_000, pane_ring=16, max_batch=B, fired_capacity=16_000_000, walert_capacity=16_000_000, calert_capacity=4_000_000)
    gpu, cpu = capi.Engine(**kw), ParallelOracle(threads=THREADS, **kw)
    _feed(gpu, cpu, spec, nb, B)
    f, c = _compare(gpu, cpu, f64=True)
    assert c["late_dropped"] > 0 and len(f) > 1_000_000
    sample = np.unique(f.key)[:: 97][:200_000]
    _assert_key_state(gpu, cpu, sample)
    gpu.close(); cpu.close()


def _assert_key_state(gpu, cpu, keys):
    import ctypes as C
    fg, cg, rg = gpu.key_state(keys)
    n = len(keys)
    fo, co, ro = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.uint64)
    k = np.ascontiguousarray(keys, dtype=np.uint64)
    assert cpu._lib.orcp_get_key_state(cpu._h, k.ctypes.data, n, fo.ctypes.data, co.ctypes.data, ro.ctypes.data) == 0
    np.testing.assert_array_equal(fg, fo)
    np.testing.assert_array_equal(cg, co)


def test_c2_consecutive_100k_keys_alerts_and_final_state_equal_the_oracle():
    """BASELINE.json configs[1]: 2^24 records, 100 K hosts, 3 consecutive readings > 90.  The alert rows and
    every host's final counter against the oracle; a replay into a fresh handle gives the same rows again
    (thousands of warps share one key's run, the result must not depend on how)."""
    import torch
    B = 1 << 24
    spec = StreamSpec(n_keys=100_000, events_per_ms=1000)
    kw = dict(consec_n=3, consec_threshold=90.0, consec_reset_on_fire=1, table_capacity=100_000, max_batch=B, calert_capacity=1 << 22)
    cpu = ParallelOracle(threads=THREADS, **kw)
    ts, key, val = generate_torch(spec, 0, B)
    torch.cuda.synchronize()
    cpu.push_batch(ts.cpu().numpy(), key.cpu().numpy().view(np.uint64), val.cpu().numpy())
    want = cpu.poll_consec_alerts(chunk=BIG)
    assert len(want) > 1000
    all_keys = np.arange(100_000, dtype=np.uint64)
    for mode in (capi.FMQ_KEYS_DENSE, capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE):
        gpu = capi.Engine(key_mode=mode, **dict(kw, table_capacity=100_000 if mode == capi.FMQ_KEYS_DENSE else 150_000))
        gpu.push_batch(ts, key, val)
        got = gpu.poll_consec_alerts(chunk=BIG)
        _assert_rows([got.ts, got.key, got.val], [want.ts, want.key, want.val], [2, 1, 0], ["consec ts", "consec key", "consec val"])
        _assert_key_state(gpu, cpu, all_keys)
        assert gpu.counters()["consec_alerts"] == len(want)
        gpu.close()
    cpu.close()
