"""fmq_partition_by_owner_noalloc: the owner split for callers that are not allowed to allocate
device memory while they run (an operator thread whose GPU is shared with other subtasks — a
cudaMalloc or cudaFree would stall everybody's kernels — or a process whose device memory all
belongs to its own arena).  Its scratch is a fixed array inside the library, so this test calls it
through the C ABI twice in a row (the second call finds the first call's leftovers in that array)
and checks both results, bit for bit, against the handle-based fmq_partition_by_owner and against
a numpy stable sort by mix64(key) % world; a third call runs the largest batch it accepts.

Kept in a file pytest collects last: a failure here cannot hide a failure anywhere else."""
import ctypes as C

import numpy as np
import pytest

from monitor_systam_flink_quickstart_b200 import capi
from monitor_systam_flink_quickstart_b200.shard import owner_of_numpy
from monitor_systam_flink_quickstart_b200.stream import StreamSpec, generate_numpy

pytestmark = pytest.mark.gpu


def _device_columns(spec, start, n):
    import torch
    ts, key, val = generate_numpy(spec, start, n)
    return [torch.from_numpy(x.view(np.int64) if x.dtype == np.uint64 else x).cuda() for x in (ts, key, val)]


def _noalloc_split(lib, cols, world):
    import torch
    n = cols[0].numel()
    outs = [torch.empty_like(c) for c in cols]              # the caller's own buffers
    counts = np.zeros(world, np.uint64)
    st = lib.fmq_partition_by_owner_noalloc(torch.cuda.current_device(), None, cols[0].data_ptr(), cols[1].data_ptr(),
                                            cols[2].data_ptr(), n, world, outs[0].data_ptr(), outs[1].data_ptr(),
                                            outs[2].data_ptr(), counts.ctypes.data)
    assert st == 0, lib.fmq_last_error(None)
    return outs, counts


def test_noalloc_split_matches_handle_based_split_twice_in_a_row():
    import torch
    lib = capi.load_library()
    eng = capi.Engine(efilter_enabled=1, max_batch=1 << 20)           # reference entry point needs a handle
    spec = StreamSpec(n_keys=20_000, events_per_ms=7)
    torch.cuda.synchronize()
    # second call is smaller and uses another world size: it must not be fooled by the counters
    # the first call left in the shared scratch
    for start, n, world in [(0, 300_007, 8), (1_000_000, 70_001, 3)]:
        cols = _device_columns(spec, start, n)
        torch.cuda.synchronize()
        got, got_counts = _noalloc_split(lib, cols, world)
        want = [torch.empty_like(c) for c in cols]
        want_counts = eng.partition_by_owner(cols[0], cols[1], cols[2], world, *want)
        torch.cuda.synchronize()
        np.testing.assert_array_equal(got_counts, want_counts)
        assert int(got_counts.sum()) == n
        for g, w in zip(got, want):                                    # bit-exact: integer / byte work
            assert torch.equal(g.view(torch.int64), w.view(torch.int64))
        _check_against_numpy(cols, got, got_counts, world)


def _check_against_numpy(cols, got, got_counts, world):
    """Independent of the library: a stable argsort of the owner column is the expected permutation."""
    owner = owner_of_numpy(cols[1].cpu().numpy().view(np.uint64), world)
    order = np.argsort(owner, kind="stable")
    np.testing.assert_array_equal(got_counts, np.bincount(owner, minlength=world).astype(np.uint64))
    for g, c in zip(got, cols):
        np.testing.assert_array_equal(g.cpu().numpy().view(np.int64), c.cpu().numpy().view(np.int64)[order])


def test_noalloc_split_at_its_largest_batch():
    """n == FMQ_NOALLOC_MAX_EVENTS uses every counter of the built-in scratch."""
    import torch
    lib = capi.load_library()
    n = 4194304
    cols = _device_columns(StreamSpec(n_keys=3_000_000, events_per_ms=500), 12_345, n)
    torch.cuda.synchronize()
    for world in (256, 5):
        got, got_counts = _noalloc_split(lib, cols, world)
        assert int(got_counts.sum()) == n
        _check_against_numpy(cols, got, got_counts, world)


def test_noalloc_split_rejects_batches_beyond_its_scratch():
    lib = capi.load_library()
    counts = np.zeros(2, np.uint64)
    too_many = 4194304 + 1                                              # FMQ_NOALLOC_MAX_EVENTS + 1
    st = lib.fmq_partition_by_owner_noalloc(0, None, 8, 8, 8, too_many, 2, 8, 8, 8, counts.ctypes.data)
    assert st == 1
    assert b"FMQ_NOALLOC_MAX_EVENTS" in lib.fmq_last_error(None)
    st = lib.fmq_partition_by_owner_noalloc(0, None, 8, 8, 8, 10, 0, 8, 8, 8, counts.ctypes.data)
    assert st == 1
