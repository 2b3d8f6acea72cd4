"""GPU tests of the host-side mirrors of the reference's operator interface: the C++ operator
(compiled against the C ABI only) and the Python operator, both replaying the reference's
event-time session, and the NCCL-sharded operator on 2 GPUs when the box has them."""
import json
import os
import pickle
import subprocess
import tempfile

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.operator import GpuWindowAggOperator
from monitor_systam_flink_quickstart_b200.text import parse_flow_line
from oracle_lib import OracleEngine, assert_events_equal, assert_fired_equal

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = json.load(open(os.path.join(ROOT, "tests", "golden", "readme_sessions.json"), encoding="utf-8"))


def test_cpp_operator_mirror_replays_g7():
    exe = os.path.join(ROOT, "monitor_systam_flink_quickstart_b200", "host", "operator_demo")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", os.path.dirname(exe)])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "(www.163.com,0.0012715657552083333)" in r.stdout and "(www.163.com,0.0012969970703125)" in r.stdout
    assert "alerts=132 late_dropped=1 OK" in r.stdout


def test_python_operator_mirror_replays_g7_on_gpu():
    out = []
    op = GpuWindowAggOperator(lambda kind, row: out.append((kind, row)), batch_capacity=64,
                              value_type=capi.FMQ_VAL_I64, window_enabled=1, agg_mask=3, emit_fired=1,
                              window_size_ms=300000, window_slide_ms=5000, max_out_of_orderness_ms=60000,
                              walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0,
                              walert_div1=1024.0, walert_div2=1024.0, walert_threshold=100.0, table_capacity=256, pane_ring=128).open()
    for line in G["G7_event_time"]["input"]:
        t, ch, flow = parse_flow_line(line)
        op.process_element(t, ch, flow)
        op.process_watermark()
    assert op.metrics()["late_dropped"] == 1
    printed = [f"({row[0]},{row[2]!r})" for kind, row in out if kind == "window_alert"]
    assert all(w.split("> ")[1] in printed for w in G["G7_event_time"]["output"]) and len(printed) == 60
    sums = [row[3] for kind, row in out if kind == "window"]
    assert sums == [10000] * 12 + [10100] * 12 + [10200] * 36
    op.close()


# ---- 2 GPUs over NCCL ------------------------------------------------------------------------------
KW = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000,
          efilter_enabled=1, efilter_threshold=90.0, consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1,
          table_capacity=1 << 12, pane_ring=64, max_batch=1 << 16)
NB, B = 5, 20_000
NBC = 12           # batches of the 2-GPU combiner test (enough steps for fire / retire to run beside a later ingest)


def _nccl_worker(rank, world, port, outdir, key_mode=0):
    import torch
    import torch.distributed as dist
    from monitor_systam_flink_quickstart_b200.shard import ShardedOperator
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    eng = capi.Engine(device=rank, key_mode=key_mode, **KW)
    op = ShardedOperator(eng, rank, world, KW["max_out_of_orderness_ms"])
    for b in range(NB):
        ts, key, val = generate_torch(spec, (b * world + rank) * B, B, device=f"cuda:{rank}")
        torch.cuda.synchronize()
        op.process_batch(ts, key, val)
    op.flush()
    res = dict(fired=eng.poll_fired(), ealerts=eng.poll_event_alerts(), calerts=eng.poll_consec_alerts(), counters=eng.counters())
    for name in ("fired", "ealerts", "calerts"):
        res[name].key = op.global_keys(res[name].key)
    pickle.dump(res, open(os.path.join(outdir, f"r{rank}.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
def test_two_gpu_sharded_equals_oracle(key_mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
    world, port = 2, 29600 + os.getpid() % 2000
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_nccl_worker, args=(world, port + key_mode, d, key_mode), nprocs=world, join=True)
        res = [pickle.load(open(os.path.join(d, f"r{r}.pkl"), "rb")) for r in range(world)]
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    okw = {k: v for k, v in KW.items()}
    ref = OracleEngine(**okw)
    for b in range(NB):
        parts = [generate_numpy(spec, (b * world + r) * B, B) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[i] for p in parts]) for i in range(3)])
    ref.flush()
    cat = lambda name, cls, fields: cls(*[np.concatenate([getattr(r[name], f) for r in res]) for f in fields])  # noqa: E731
    assert_fired_equal(cat("fired", capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    assert_events_equal(cat("ealerts", capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    assert_events_equal(cat("calerts", capi.EventRows, ("ts", "key", "val")), ref.poll_consec_alerts())
    assert sum(r["counters"]["late_dropped"] for r in res) == ref.counters()["late_dropped"]


# ---- combiner mode --------------------------------------------------------------------------------
CKW = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000,
           efilter_enabled=1, efilter_threshold=90.0, walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=52.0,
           key_mode=capi.FMQ_KEYS_DENSE, table_capacity=1000, pane_ring=64, max_batch=1 << 16)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("vt", ["f64", "i64"])
@pytest.mark.parametrize("overlap", [0, 1])
@pytest.mark.parametrize("flags", [0, 1])
def test_combiner_mode_emulated_ranks_on_one_gpu(world, vt, overlap, flags):
    """`world` engines in ONE process on one GPU play the ranks (their tables are attached to each other
    as plain device pointers).  flags=0: the two collectives are done by hand between synchronised streams (what a
    caller with its own collective library does); flags=1: the library exchanges the batch scalars and both
    barriers itself through device flags in the peers' control blocks, nothing is synchronised by the host.
    Union of the ranks' outputs == oracle over the whole stream."""
    import torch
    from monitor_systam_flink_quickstart_b200.shard import exchange_tensor
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy, generate_torch
    from oracle_lib import sort_walerts
    kw = dict(CKW, value_type=capi.FMQ_VAL_F64 if vt == "f64" else capi.FMQ_VAL_I64)
    if overlap:
        kw["window_slide_ms"] = kw["window_size_ms"]          # tumbling: fire / retire may run beside the next ingest
    if vt == "i64":
        kw.update(walert_source=capi.FMQ_ALERT_SUM, walert_op=capi.FMQ_CMP_LT, walert_mul=8.0, walert_div0=60.0, walert_div1=1024.0,
                  walert_div2=1024.0, walert_threshold=0.9)
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000, value_type=vt)
    engs = [capi.Engine(**kw) for _ in range(world)]
    for r, e in enumerate(engs):
        e.attach_peers(r, world, [x.table_ptr for x in engs])
        if overlap:
            e.combiner_overlap(True)
    if flags:
        for e in engs:
            e.shard_alloc(world, 4)
        for r, e in enumerate(engs):
            e.shard_attach(r, world, [x.shard_base for x in engs])
            e.combiner_use_flags(True)
    xs = [exchange_tensor(e, "cuda:0") for e in engs]
    ref = OracleEngine(**{k: v for k, v in kw.items()})
    nb, b = 6, 7001

    def collective_max():
        for e in engs:
            e.synchronize()
        m = torch.stack(xs).max(dim=0).values
        for x in xs:
            x.copy_(m)
        torch.cuda.synchronize()

    def fire(wm):
        if flags:                                 # every rank's step is queued before the next step of any rank: no spin waits on unsubmitted work
            for e in engs:
                e.fire_global(wm)
            for e in engs:
                e.retire()
            return
        collective_max()
        for e in engs:
            e.fire_global(wm)
        for e in engs:
            e.synchronize()                       # barrier: everybody finished reading its peers
        for e in engs:
            e.retire()

    for i in range(nb):
        parts = []
        for r, e in enumerate(engs):
            start = (i * world + r) * b
            parts.append(generate_numpy(spec, start, b))
            if i % 3 == 0:                                        # device columns
                cols = generate_torch(spec, start, b, device="cuda:0")
                torch.cuda.synchronize()
                e.push_local(*cols)
            elif i % 3 == 1:                                      # host columns, narrow wire format
                ts_n, key_n, val_n = parts[-1]
                base = int(ts_n.min())
                e.push_local((ts_n - base).astype(np.uint32), key_n.astype(np.uint32), val_n, base)
            else:                                                 # host columns as they are
                e.push_local(*parts[-1])
            if not flags:
                e.synchronize()
        fire(capi.FMQ_WM_AUTO)
        ref.push_batch(*[np.concatenate([p[j] for p in parts]) for j in range(3)])
    if overlap or flags:                          # these modes do not pre-arm the exchange in retire: a flush starts with an empty local push
        for e in engs:
            e.push_local(np.zeros(0, np.int64), np.zeros(0, np.uint64), np.zeros(0, np.float64 if vt == "f64" else np.int64))
    fire(capi.FMQ_WM_FINAL)
    ref.flush()
    res = [dict(fired=e.poll_fired(), wal=e.poll_window_alerts(), eal=e.poll_event_alerts(), c=e.counters()) for e in engs]
    cat = lambda name, cls, fields: cls(*[np.concatenate([getattr(r[name], f) for r in res]) for f in fields])  # noqa: E731
    assert_fired_equal(cat("fired", capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    a, w = sort_walerts(cat("wal", capi.WindowAlerts, ("key", "window_end", "value"))), sort_walerts(ref.poll_window_alerts())
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.window_end, w.window_end)
    np.testing.assert_allclose(a.value, w.value, rtol=1e-9)
    assert_events_equal(cat("eal", capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    rc = ref.counters()
    assert sum(r["c"]["late_dropped"] for r in res) == rc["late_dropped"] > 0
    assert sum(r["c"]["windows_fired"] for r in res) == rc["windows_fired"]
    assert all(r["c"]["watermark"] == rc["watermark"] for r in res)
    # every rank fired only its own slot range
    per = (1000 + world - 1) // world
    for r, x in enumerate(res):
        assert len(x["fired"]) > 0 and x["fired"].key.min() >= per * r and x["fired"].key.max() < per * (r + 1)


def _combiner_worker(rank, world, port, outdir, tumbling=False, flags=True):
    import torch
    import torch.distributed as dist
    from monitor_systam_flink_quickstart_b200.shard import CombinerOperator
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    kw = dict(CKW, window_slide_ms=CKW["window_size_ms"]) if tumbling else CKW
    eng = capi.Engine(device=rank, **kw)
    op = CombinerOperator(eng, rank, world, flags=flags)
    assert op.overlap == tumbling
    cols = [generate_torch(spec, (b * world + rank) * B, B, device=f"cuda:{rank}") for b in range(NBC)]
    torch.cuda.synchronize()
    for b in range(NBC):
        op.process_batch(*cols[b])     # no host synchronisation between batches: the streams order everything
    eng.synchronize()
    op.flush()
    res = dict(fired=eng.poll_fired(), ealerts=eng.poll_event_alerts(), counters=eng.counters())
    pickle.dump(res, open(os.path.join(outdir, f"r{rank}.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("tumbling", [False, True])       # True: fire / retire overlapped with the next batch's ingest
@pytest.mark.parametrize("world,flags", [(2, False), (2, True), (4, True), (8, True)])   # flags: device flags instead of NCCL all-reduces
def test_two_gpu_combiner_over_nvlink_equals_oracle(tumbling, world, flags):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
    port = 29700 + os.getpid() % 2000 + 7 * world + int(flags) + 2 * int(tumbling)
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_combiner_worker, args=(world, port, d, tumbling, flags), nprocs=world, join=True)
        res = [pickle.load(open(os.path.join(d, f"r{r}.pkl"), "rb")) for r in range(world)]
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    ref = OracleEngine(**dict(CKW, window_slide_ms=CKW["window_size_ms"] if tumbling else CKW["window_slide_ms"]))
    for b in range(NBC):
        parts = [generate_numpy(spec, (b * world + r) * B, B) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[i] for p in parts]) for i in range(3)])
    ref.flush()
    cat = lambda name, cls, fields: cls(*[np.concatenate([getattr(r[name], f) for r in res]) for f in fields])  # noqa: E731
    assert_fired_equal(cat("fired", capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    assert_events_equal(cat("ealerts", capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    assert sum(r["counters"]["late_dropped"] for r in res) == ref.counters()["late_dropped"]


# ---- peer-pull keyBy exchange -------------------------------------------------------------------------
@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
def test_peer_pull_exchange_emulated_ranks_on_one_gpu(key_mode):
    """3 engines in one process play the ranks; their send buffers are attached to each other as plain
    device pointers and the size matrix is assembled by hand.  Every operator runs behind the exchange
    (windows, filter, keyed ValueState, median); union of the ranks' outputs == oracle."""
    import torch
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy, generate_torch
    from oracle_lib import sort_walerts
    world = 3
    kw = dict(KW, key_mode=key_mode, rolling_max_enabled=0, median_log_capacity=1 << 18)
    engs = [capi.Engine(**kw) for _ in range(world)]
    for e in engs:
        e.exchange_alloc()
    for r, e in enumerate(engs):
        e.exchange_attach(r, world, [x.exchange_base for x in engs])
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    ref = OracleEngine(**{k: v for k, v in kw.items()})
    cur_max, wm, bound, b = -(2 ** 63) + KW["max_out_of_orderness_ms"], -(2 ** 63), KW["max_out_of_orderness_ms"], 9001
    for i in range(5):
        cols = [generate_torch(spec, (i * world + r) * b, b, device="cuda:0") for r in range(world)]
        torch.cuda.synchronize()
        matrix = np.stack([e.exchange_split(*cols[r], i & 1, world) for r, e in enumerate(engs)])
        cur_max = max(cur_max, max(int(c[0].max()) for c in cols))
        wm = max(wm, cur_max - bound)
        got = [e.exchange_pull_push(i & 1, matrix, wm) for e in engs]
        assert sum(got) == world * b and got == matrix.sum(axis=0).tolist()
        for e in engs:
            e.synchronize()
        parts = [generate_numpy(spec, (i * world + r) * b, b) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[j] for p in parts]) for j in range(3)])
    for e in engs:
        e.flush()
    ref.flush()

    def gk(r, keys):
        return keys * np.uint64(world) + np.uint64(r) if key_mode == capi.FMQ_KEYS_DENSE else keys
    fired, cal, med = [], [], []
    for r, e in enumerate(engs):
        f = e.poll_fired(); f.key = gk(r, f.key); fired.append(f)
        c = e.poll_consec_alerts(); c.key = gk(r, c.key); cal.append(c)
        m = e.poll_medians(); m.key = gk(r, m.key); med.append(m)
    cat = lambda parts, cls, fields: cls(*[np.concatenate([getattr(p, f) for p in parts]) for f in fields])  # noqa: E731
    assert_fired_equal(cat(fired, capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    assert_events_equal(cat(cal, capi.EventRows, ("ts", "key", "val")), ref.poll_consec_alerts())
    a, w = sort_walerts(cat(med, capi.WindowAlerts, ("key", "window_end", "value"))), sort_walerts(ref.poll_medians())
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.value.view(np.uint64), w.value.view(np.uint64))
    assert sum(e.counters()["late_dropped"] for e in engs) == ref.counters()["late_dropped"] > 0


def _peer_pull_worker(rank, world, port, outdir, key_mode):
    import torch
    import torch.distributed as dist
    from monitor_systam_flink_quickstart_b200.shard import PeerPullOperator
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    eng = capi.Engine(device=rank, key_mode=key_mode, **KW)
    op = PeerPullOperator(eng, rank, world, KW["max_out_of_orderness_ms"])
    for b in range(NB):
        ts, key, val = generate_torch(spec, (b * world + rank) * B, B, device=f"cuda:{rank}")
        torch.cuda.synchronize()
        op.process_batch(ts, key, val)
        eng.synchronize()
    op.flush()
    res = dict(fired=eng.poll_fired(), ealerts=eng.poll_event_alerts(), calerts=eng.poll_consec_alerts(), counters=eng.counters())
    for name in ("fired", "ealerts", "calerts"):
        res[name].key = op.global_keys(res[name].key)
    pickle.dump(res, open(os.path.join(outdir, f"r{rank}.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
def test_two_gpu_peer_pull_equals_oracle(key_mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
    world, port = 2, 29800 + os.getpid() % 2000 + key_mode
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_peer_pull_worker, args=(world, port, d, key_mode), nprocs=world, join=True)
        res = [pickle.load(open(os.path.join(d, f"r{r}.pkl"), "rb")) for r in range(world)]
    spec = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
    ref = OracleEngine(**{k: v for k, v in KW.items()})
    for b in range(NB):
        parts = [generate_numpy(spec, (b * world + r) * B, B) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[i] for p in parts]) for i in range(3)])
    ref.flush()
    cat = lambda name, cls, fields: cls(*[np.concatenate([getattr(r[name], f) for r in res]) for f in fields])  # noqa: E731
    assert_fired_equal(cat("fired", capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    assert_events_equal(cat("ealerts", capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    assert_events_equal(cat("calerts", capi.EventRows, ("ts", "key", "val")), ref.poll_consec_alerts())
    assert sum(r["counters"]["late_dropped"] for r in res) == ref.counters()["late_dropped"]
