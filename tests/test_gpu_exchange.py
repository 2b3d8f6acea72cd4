"""The keyBy exchange inside the library (fmq_shard_*): each rank's scatter kernel writes owner runs straight into
the owners' receive segments, counts and the largest timestamp travel behind device flags, the receiver's kernels
wait on the flags — no collective library and no host round trip per batch.  Checked against the oracle fed the
whole batch (rank slices concatenated in rank order: that is the per-key arrival order the exchange preserves):
  * `world` engines of ONE process on one GPU attached to each other by pointer (runs on the 1-GPU test box);
  * one process per GPU on 2 / 4 / 8 GPUs, buffers mapped through CUDA IPC, `gloo` used once to pass the blobs."""
import os
import pickle
import tempfile

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.shard import ExchangeOperator
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
from oracle_lib import OracleEngine, assert_events_equal, assert_fired_equal, sort_walerts

pytestmark = pytest.mark.gpu

KW = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000,
          efilter_enabled=1, efilter_threshold=90.0, walert_source=capi.FMQ_ALERT_AVG, walert_op=capi.FMQ_CMP_GT, walert_threshold=52.0,
          table_capacity=1 << 12, pane_ring=64, max_batch=1 << 17)
KEYED = dict(consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1)
SPEC = StreamSpec(n_keys=900, events_per_ms=4, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)


def _cat(parts, cls, fields):
    return cls(*[np.concatenate([getattr(p, f) for p in parts]) for f in fields])


def _check_union(results, ref, keyed, median=False):
    """results: per rank dict(fired, walerts, ealerts, calerts[, medians], counters) with GLOBAL keys."""
    assert_fired_equal(_cat([r["fired"] for r in results], capi.Fired, ("key", "window_end", "count", "sum", "min", "max")), ref.poll_fired())
    a = sort_walerts(_cat([r["walerts"] for r in results], capi.WindowAlerts, ("key", "window_end", "value"))); w = sort_walerts(ref.poll_window_alerts())
    np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.window_end, w.window_end)
    np.testing.assert_allclose(a.value, w.value, rtol=1e-9, atol=0)
    assert_events_equal(_cat([r["ealerts"] for r in results], capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    if keyed:
        assert_events_equal(_cat([r["calerts"] for r in results], capi.EventRows, ("ts", "key", "val")), ref.poll_consec_alerts())
    if median:
        a = sort_walerts(_cat([r["medians"] for r in results], capi.WindowAlerts, ("key", "window_end", "value"))); w = sort_walerts(ref.poll_medians())
        np.testing.assert_array_equal(a.key, w.key); np.testing.assert_array_equal(a.value.view(np.uint64), w.value.view(np.uint64))
        assert_events_equal(_cat([r["late"] for r in results], capi.EventRows, ("ts", "key", "val")), ref.poll_late())
    rc = ref.counters()
    assert sum(r["counters"]["late_dropped"] for r in results) == rc["late_dropped"] > 0
    assert sum(r["counters"]["events_in"] for r in results) == rc["events_in"]
    assert all(r["counters"]["watermark"] == rc["watermark"] for r in results)       # ONE watermark for the whole job


def _collect(eng, op, keyed, median=False):
    res = dict(fired=eng.poll_fired(), walerts=eng.poll_window_alerts(), ealerts=eng.poll_event_alerts(), counters=eng.counters())
    if keyed:
        res["calerts"] = eng.poll_consec_alerts()
    if median:
        res["medians"] = eng.poll_medians()
        res["late"] = eng.poll_late()
    for name in ("fired", "walerts", "ealerts", "calerts", "medians", "late"):
        if name in res:
            res[name].key = op.global_keys(res[name].key)
    return res


@pytest.mark.parametrize("key_mode", [capi.FMQ_KEYS_HASHED, capi.FMQ_KEYS_DENSE])
@pytest.mark.parametrize("flavour", ["window", "keyed", "median"])
def test_exchange_emulated_ranks_on_one_gpu(key_mode, flavour):
    """3 engines of this process play the ranks.  'window': windows + filter, counts never leave the device;
    'keyed': plus the consecutive-readings rule (the one path that reads the batch size back); 'median': plus the
    value log, whose rows the ingest kernel stages per warp."""
    import torch
    from monitor_systam_flink_quickstart_b200.stream import generate_torch
    world, b, nb = 3, 9001, 6
    keyed, median = flavour == "keyed", flavour == "median"
    kw = dict(KW, key_mode=key_mode, **(KEYED if keyed else {}), **(dict(median_log_capacity=1 << 18, late_capacity=1 << 14) if median else {}))
    engs = [capi.Engine(**kw) for _ in range(world)]
    for e in engs:
        e.shard_alloc(world, 1 << 14)
    ops = [ExchangeOperator(e, r, world, 1 << 14, bases=[x.shard_base for x in engs]) for r, e in enumerate(engs)]
    ref = OracleEngine(**kw)
    for i in range(nb):
        n = 0 if i == 3 else b - 7 * i                                  # an empty batch and ragged sizes on the way
        cols = [generate_torch(SPEC, (i * world + r) * b, n, device="cuda:0") for r in range(world)]
        torch.cuda.synchronize()
        for r, op in enumerate(ops):                                    # every rank's send is queued before any recv: nothing here
            op.send(*cols[r])                                           # ever spins on work that has not been submitted yet
        got = [op.recv(want_count=(i % 2 == 0)) for op in ops]          # odd batches: nothing about the batch is read back
        assert i % 2 or sum(got) == world * n
        for e in engs:
            e.synchronize()
        parts = [generate_numpy(SPEC, (i * world + r) * b, n) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[j] for p in parts]) for j in range(3)])
    for e in engs:
        e.flush()
    ref.flush()
    _check_union([_collect(e, op, keyed, median) for e, op in zip(engs, ops)], ref, keyed, median)


def test_exchange_host_columns_and_explicit_watermarks():
    """Host-resident slices go through the same send; explicit watermarks (the same on every rank) replace the automatic one."""
    world, b = 2, 5000
    kw = dict(KW, **KEYED)
    engs = [capi.Engine(**kw) for _ in range(world)]
    for e in engs:
        e.shard_alloc(world, b)
    ops = [ExchangeOperator(e, r, world, b, bases=[x.shard_base for x in engs]) for r, e in enumerate(engs)]
    ref = OracleEngine(**kw)
    wms = [capi.FMQ_WM_KEEP, SPEC.t0 + 1000, capi.FMQ_WM_AUTO, capi.FMQ_WM_KEEP, SPEC.t0 + 9000]
    for i, wm in enumerate(wms):
        parts = [generate_numpy(SPEC, (i * world + r) * b, b) for r in range(world)]
        for r, op in enumerate(ops):
            op.send(*parts[r])
        for op in ops:
            op.recv(wm)
        ref.push_batch(*[np.concatenate([p[j] for p in parts]) for j in range(3)], wm)
    for e in engs:
        e.flush()
    ref.flush()
    _check_union([_collect(e, op, True) for e, op in zip(engs, ops)], ref, True)


# ---- one process per GPU ------------------------------------------------------------------------------------------
NB, B = 6, 20_000


def _worker(rank, world, port, outdir, key_mode, keyed):
    import torch
    import torch.distributed as dist
    from monitor_systam_flink_quickstart_b200.stream import generate_torch
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)        # set-up only: the data path uses no collective library
    eng = capi.Engine(device=rank, key_mode=key_mode, **dict(KW, **(KEYED if keyed else {})))
    op = ExchangeOperator(eng, rank, world, B)
    for b in range(NB):
        ts, key, val = generate_torch(SPEC, (b * world + rank) * B, B, device=f"cuda:{rank}")
        torch.cuda.synchronize()
        op.process_batch(ts, key, val)                                   # no host synchronisation between batches (window flavour)
    eng.synchronize()
    op.flush()
    pickle.dump(_collect(eng, op, keyed), open(os.path.join(outdir, f"r{rank}.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("key_mode,keyed", [(capi.FMQ_KEYS_HASHED, True), (capi.FMQ_KEYS_DENSE, False)])
def test_exchange_one_process_per_gpu_equals_oracle(world, key_mode, keyed):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    port = 30100 + os.getpid() % 2000 + 10 * world + key_mode
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, port, d, key_mode, keyed), nprocs=world, join=True)
        res = [pickle.load(open(os.path.join(d, f"r{r}.pkl"), "rb")) for r in range(world)]
    ref = OracleEngine(**dict(KW, **(KEYED if keyed else {})))
    for b in range(NB):
        parts = [generate_numpy(SPEC, (b * world + r) * B, B) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[i] for p in parts]) for i in range(3)])
    ref.flush()
    _check_union(res, ref, keyed)
