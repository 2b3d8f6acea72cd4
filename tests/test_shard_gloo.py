"""N>1 host path on CPU: world_size-2 over gloo.  Each rank runs the real ShardedOperator
(bucket-size exchange, variable all-to-all, watermark all-reduce) with CPU test doubles for the
two CUDA pieces (numpy owner split; the CPU oracle as the engine).  The union of the ranks' outputs
must equal a single oracle fed the whole stream."""
import os
import pickle
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.shard import ShardedOperator, owner_of_numpy
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy
from oracle_lib import OracleEngine, assert_events_equal, assert_fired_equal

KW = dict(window_enabled=1, agg_mask=15, window_size_ms=10_000, window_slide_ms=5000, max_out_of_orderness_ms=3000,
          efilter_enabled=1, efilter_threshold=90.0, consec_n=3, consec_threshold=50.0, consec_reset_on_fire=1)
SPEC = StreamSpec(n_keys=300, events_per_ms=2, jitter_ms=3000, late_frac_ppm=30000, late_extra_ms=20000)
NB, B = 5, 6000


def numpy_partition(ts, key, val, world):
    k = key.numpy().view(np.uint64)
    own = owner_of_numpy(k, world)
    order = np.argsort(own, kind="stable")
    counts = np.bincount(own, minlength=world)
    o = torch.from_numpy(order)
    return ts[o], key[o], val[o], counts


class TorchFedOracle(OracleEngine):
    """Accepts the CPU tensors the sharded layer hands over."""
    def push_batch(self, ts, key, val, wm=capi.FMQ_WM_AUTO):
        super().push_batch(ts.numpy(), key.numpy().view(np.uint64), val.numpy(), wm)


def _worker(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    eng = TorchFedOracle(**KW)
    op = ShardedOperator(eng, rank, world, KW["max_out_of_orderness_ms"], partition=numpy_partition)
    owned = 0
    for b in range(NB):
        ts, key, val = generate_numpy(SPEC, (b * world + rank) * B, B)      # this rank's slice of global batch b
        owned += op.process_batch(torch.from_numpy(ts), torch.from_numpy(key.view(np.int64)), torch.from_numpy(val))
    op.flush()
    res = dict(fired=eng.poll_fired(), ealerts=eng.poll_event_alerts(), calerts=eng.poll_consec_alerts(), counters=eng.counters(),
               owned=owned, sent=op.shuffle_bytes_sent)
    pickle.dump(res, open(os.path.join(outdir, f"r{rank}.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_sharded_union_equals_single_instance(world):
    port = 29500 + os.getpid() % 2000
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, port, d), nprocs=world, join=True)
        res = [pickle.load(open(os.path.join(d, f"r{r}.pkl"), "rb")) for r in range(world)]
    # single instance over the same global stream, same batch boundaries (global batch b = slices of all ranks in rank order)
    ref = OracleEngine(**KW)
    for b in range(NB):
        parts = [generate_numpy(SPEC, (b * world + r) * B, B) for r in range(world)]
        ref.push_batch(*[np.concatenate([p[i] for p in parts]) for i in range(3)])
    ref.flush()
    cat = lambda name, cls, fields: cls(*[np.concatenate([getattr(r[name], f) for r in res]) for f in fields])  # noqa: E731
    fired = cat("fired", capi.Fired, ("key", "window_end", "count", "sum", "min", "max"))
    assert_fired_equal(fired, ref.poll_fired(), f64_sum_rtol=0.0)          # same per-key order -> bit-exact even for doubles
    assert_events_equal(cat("ealerts", capi.EventRows, ("ts", "key", "val")), ref.poll_event_alerts())
    assert_events_equal(cat("calerts", capi.EventRows, ("ts", "key", "val")), ref.poll_consec_alerts())
    rc = ref.counters()
    assert sum(r["counters"]["late_dropped"] for r in res) == rc["late_dropped"] > 0
    assert sum(r["counters"]["windows_fired"] for r in res) == rc["windows_fired"]
    assert all(r["counters"]["watermark"] == rc["watermark"] for r in res)  # one watermark for the whole job
    assert sum(r["owned"] for r in res) == NB * B * world
    # every key is owned by exactly one rank
    k0, k1 = set(res[0]["fired"].key.tolist()), set(res[1]["fired"].key.tolist())
    assert not (k0 & k1) and all(owner_of_numpy(np.array([k], np.uint64), world)[0] == 0 for k in list(k0)[:50])
    assert all(r["sent"] > 0 for r in res)
