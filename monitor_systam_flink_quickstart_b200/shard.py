"""keyBy across GPUs: one process per GPU, every rank owns the state of the keys
with ``mix64(key) % world == rank``.

Per micro-batch each rank (1) splits its slice of the batch into ``world`` owner buckets with the
CUDA multi-split kernel behind ``fmq_partition_by_owner`` (stable, so per-key arrival order
survives), (2) exchanges the bucket sizes, (3) moves the three 8-byte columns with a
variable-size all-to-all over NCCL (NVLink 5 / NVSwitch), buckets concatenated in source-rank
order, (4) agrees on the watermark with an all-reduce(max) of the per-rank max timestamp — the
reference has ONE watermark assigner in front of the keyBy (BandwidthMonitorWithEventTime.java:30-45),
so every rank must judge lateness against the same value — and (5) pushes what it received into
its own engine.  The union of the ranks' outputs equals the single-GPU output.

The exchange is written against ``torch.distributed`` only, so the same code runs over ``gloo``
on CPU tensors in the world_size-2 tests (with a CPU partitioner and the CPU oracle injected as
the engine — test doubles; the product default is the CUDA engine and fails without a GPU).
"""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

from . import capi

_M1, _M2 = np.uint64(0xBF58476D1CE4E5B9), np.uint64(0x94D049BB133111EB)
I64_MIN = -(2 ** 63)


def mix64_numpy(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30); x *= _M1
    x ^= x >> np.uint64(27); x *= _M2
    x ^= x >> np.uint64(31)
    return x


def owner_of_numpy(key: np.ndarray, world: int) -> np.ndarray:
    """Host restatement of ``fmq_owner_of`` (mix64(key) % world)."""
    return (mix64_numpy(key) % np.uint64(world)).astype(np.int64)


class ShardedOperator:
    """One rank of the keyBy-sharded operator.

    engine      : this rank's engine (capi.Engine on its GPU)
    partition   : callable(ts, key, val, world) -> (ts_b, key_b, val_b, counts[world]) bucketed by owner;
                  default = the CUDA multi-split of the engine
    """

    def __init__(self, engine, rank: int, world: int, bound_ms: int, group=None,
                 partition: Optional[Callable] = None, auto_watermark: bool = True):
        import torch.distributed as dist
        self.dist = dist
        self.engine, self.rank, self.world, self.group = engine, rank, world, group
        self.bound = bound_ms
        self.auto_watermark = auto_watermark
        self._partition = partition or self._cuda_partition
        self.cur_max_ts = I64_MIN + bound_ms
        self.wm = I64_MIN
        self.shuffle_bytes_sent = 0
        self._send = None
        self._recv = None
        cfg = getattr(engine, "cfg", None)
        self.dense = bool(cfg is not None and cfg.key_mode == capi.FMQ_KEYS_DENSE and partition is None)

    def global_keys(self, local_keys: np.ndarray) -> np.ndarray:
        """Output rows of a DENSE engine carry the rank-local id; this maps them back."""
        if not self.dense:
            return local_keys
        return local_keys * np.uint64(self.world) + np.uint64(self.rank)

    # -- step 1: owner buckets on the GPU ------------------------------------------------------
    def _cuda_partition(self, ts, key, val, world):
        import torch
        n = ts.numel()
        # persistent bucket buffers: no allocator traffic (and no cross-stream block recycling) per step
        if self._send is None or self._send[0].numel() < n:
            self._send = [torch.empty(n, dtype=t.dtype, device=t.device) for t in (ts, key, val)]
        outs = [b[:n] for b in self._send]
        counts = self.engine.partition_by_owner(ts, key, val, world, *outs)
        return outs[0], outs[1], outs[2], counts.astype(np.int64)

    def _recv_buffers(self, like, n_recv):
        import torch
        if self._recv is None or self._recv[0].numel() < n_recv:
            cap = max(n_recv, int(getattr(self.engine.cfg, "max_batch", 0) or 0)) if hasattr(self.engine, "cfg") else n_recv
            self._recv = [torch.empty(cap, dtype=t.dtype, device=t.device) for t in like]
        return [b[:n_recv] for b in self._recv]

    def process_batch(self, ts, key, val, watermark_after: Optional[int] = None):
        """ts/key/val: this rank's slice of the micro-batch (torch tensors; int64 bit patterns for
        key, int64 or float64 for val).  Returns the number of records this rank now owns.
        On CUDA everything (multi-split kernels, NCCL, the engine's kernel chain) is ordered on the
        engine's own stream, so no host synchronisation is needed between the steps beyond the
        bucket-size read-back."""
        import contextlib
        import torch
        ctx = contextlib.nullcontext()
        if ts.is_cuda:
            if getattr(self, "_ext", None) is None:
                self._ext = torch.cuda.ExternalStream(self.engine.stream, device=ts.device)
            ctx = torch.cuda.stream(self._ext)
        with ctx:
            return self._process_batch(ts, key, val, watermark_after)

    def _process_batch(self, ts, key, val, watermark_after):
        import torch
        dist, W = self.dist, self.world
        n = int(ts.numel())
        local_max = int(ts.max().item()) if n else I64_MIN
        ts_b, key_b, val_b, counts = self._partition(ts, key, val, W)
        # -- step 2: bucket sizes (W x W matrix, 8 B each) --
        send_counts = torch.as_tensor(np.asarray(counts, dtype=np.int64), device=ts.device)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts, group=self.group)
        rc = recv_counts.cpu().tolist(); sc = [int(c) for c in counts]
        n_recv = sum(rc)
        # -- step 3: variable-size all-to-all of the three columns --
        outs = self._recv_buffers((ts_b, key_b, val_b), n_recv)
        for out, col in zip(outs, (ts_b, key_b, val_b)):
            dist.all_to_all_single(out, col, output_split_sizes=rc, input_split_sizes=sc, group=self.group)
        self.shuffle_bytes_sent += 24 * (n - sc[self.rank])
        # -- step 4: one watermark for the whole job --
        if watermark_after is None:
            m = torch.tensor([local_max], dtype=torch.int64, device=ts.device)
            dist.all_reduce(m, op=dist.ReduceOp.MAX, group=self.group)
            self.cur_max_ts = max(self.cur_max_ts, int(m.item()))
            pot = self.cur_max_ts - self.bound
            if pot >= self.wm:
                self.wm = pot
            watermark_after = self.wm
        # -- step 5: keyed operator on the records this rank owns --
        self.engine.push_batch(outs[0], outs[1], outs[2], watermark_after)
        return n_recv

    def flush(self, wm: int = capi.FMQ_WM_FINAL):
        self.engine.flush(wm)


class _DevicePtr:
    """Lets torch alias a raw device pointer (CUDA array interface): used for the 4 x int64
    exchange buffer of an engine, so that the all-reduce runs on it in place."""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<i8", "data": (ptr, False), "version": 2}


def exchange_tensor(engine, device):
    import torch
    return torch.as_tensor(_DevicePtr(engine.exchange_ptr, 4), device=device)


class CombinerOperator:
    """Multi-GPU window aggregation without an event shuffle (include/fmq_b200.h, "combiner mode").

    Every rank folds its own slice of each micro-batch into a full-size local table; per batch the
    ranks exchange 32 bytes (one all-reduce(MAX): max timestamp and live pane range) instead of
    24 bytes per event, and the fire kernel of the rank that owns a slot range reads that range of
    every rank's table over NVLink peer memory.  Needs dense key ids and only covers the commutative
    window aggregates + the stateless per-event filter; keyed ValueState goes through
    :class:`ShardedOperator`.
    """

    def __init__(self, engine, rank: int, world: int, group=None, overlap=None, flags: bool = True):
        """overlap (default: on for tumbling windows): fire / barrier / retire of batch k run on a second
        stream while batch k+1 is ingested (fmq_combiner_overlap).
        flags (default): the 32 bytes of batch scalars and both barriers travel as peer stores behind device flags
        inside the library (fmq_combiner_use_flags) — no collective is called per batch, `torch.distributed` only
        hands the IPC blobs around at set-up.  flags=False: the round-1 schedule with two NCCL all-reduces per batch."""
        import torch
        import torch.distributed as dist
        self.dist, self.engine, self.rank, self.world, self.group = dist, engine, rank, world, group
        self.flags = bool(flags)
        dev = torch.device("cuda", torch.cuda.current_device())
        blobs = [None] * world
        dist.all_gather_object(blobs, engine.ipc_export(), group=group)
        tables = [engine.table_ptr if r == rank else engine.ipc_open(blobs[r]) for r in range(world)]
        engine.attach_peers(rank, world, tables)
        tumbling = engine.cfg.window_size_ms == engine.cfg.window_slide_ms
        self.overlap = tumbling if overlap is None else bool(overlap)
        if self.flags:
            cblobs = [None] * world
            dist.all_gather_object(cblobs, engine.shard_alloc(world, 4), group=group)
            engine.shard_attach_ipc(rank, world, cblobs)
            if self.overlap:
                engine.combiner_overlap(True)
            engine.combiner_use_flags(True)
            self._empty = (np.zeros(0, np.int64), np.zeros(0, np.uint64), np.zeros(0, np.float64 if engine.cfg.value_type == capi.FMQ_VAL_F64 else np.int64))
            self._ext = self._ext2 = None
            dist.barrier(group=group)                        # every rank has mapped every table and control block
            return
        self._ext = torch.cuda.ExternalStream(engine.stream, device=dev)
        self._ext2, self.group2 = self._ext, group
        if self.overlap:
            engine.combiner_overlap(True)
            self._ext2 = torch.cuda.ExternalStream(engine.fire_stream, device=dev)
            ranks = list(range(world)) if group is None else dist.get_process_group_ranks(group)
            self.group2 = dist.new_group(ranks=ranks, backend=dist.get_backend(group))   # its own communicator: no ordering against the first one
        self._x = exchange_tensor(engine, dev)
        self._token = torch.zeros(1, dtype=torch.int64, device=dev)
        self._empty = (np.zeros(0, np.int64), np.zeros(0, np.uint64), np.zeros(0, np.float64 if engine.cfg.value_type == capi.FMQ_VAL_F64 else np.int64))
        dist.barrier(group=group)

    def _fire(self, watermark_after):
        if self.flags:                                       # kernels only: the device-side flags order the ranks
            self.engine.fire_global(watermark_after)
            self.engine.retire()
            return
        import torch
        dist = self.dist
        with torch.cuda.stream(self._ext):
            dist.all_reduce(self._x, op=dist.ReduceOp.MAX, group=self.group)     # global scalars + "all ingested"
            self.engine.fire_global(watermark_after)
        with torch.cuda.stream(self._ext2):
            dist.all_reduce(self._token, group=self.group2)                        # "all peers done reading"
            self.engine.retire()

    def process_batch(self, ts, key, val, watermark_after: int = capi.FMQ_WM_AUTO, ts_base: int = 0):
        """One micro-batch: everything below is queued on the engine's streams, no host round trip
        (host columns: the call returns once they are copied in)."""
        if self.flags:
            self.engine.push_local(ts, key, val, ts_base)
            self._fire(watermark_after)
            return
        import torch
        with torch.cuda.stream(self._ext):
            self.engine.push_local(ts, key, val, ts_base)
        self._fire(watermark_after)

    def flush(self, wm: int = capi.FMQ_WM_FINAL):
        if self.overlap or self.flags:        # the exchange is not pre-armed by retire in these modes
            self.engine.push_local(*self._empty)
        self._fire(wm)

    def join(self):
        """Order the engine's main stream after everything queued on the fire stream (timing, hand-over)."""
        if self.overlap:
            import torch
            dev = torch.device("cuda", torch.cuda.current_device())
            e1 = torch.cuda.ExternalStream(self.engine.stream, device=dev)
            e2 = torch.cuda.ExternalStream(self.engine.fire_stream, device=dev)
            ev = torch.cuda.Event()
            ev.record(e2)
            e1.wait_event(ev)


class PeerPullOperator:
    """keyBy exchange over NVLink peer memory (include/fmq_b200.h, "peer pull"): same semantics as
    :class:`ShardedOperator` (every operator, both key modes, per-key arrival order), but the events are
    never sent: each rank splits into a buffer its peers have mapped through CUDA IPC, ONE collective
    carries the bucket sizes and the max timestamp, and every rank copies its own buckets straight out
    of the peers' memory before running its operator chain."""

    def __init__(self, engine, rank: int, world: int, bound_ms: int, group=None):
        import torch
        import torch.distributed as dist
        self.dist, self.engine, self.rank, self.world, self.group, self.bound = dist, engine, rank, world, group, bound_ms
        self.dev = torch.device("cuda", torch.cuda.current_device())
        blobs = [None] * world
        dist.all_gather_object(blobs, engine.exchange_alloc(), group=group)
        bases = [engine.exchange_base if r == rank else engine.ipc_open(blobs[r]) for r in range(world)]
        engine.exchange_attach(rank, world, bases)
        self._ext = torch.cuda.ExternalStream(engine.stream, device=self.dev)
        self._mine = torch.zeros(world + 1, dtype=torch.int64, device=self.dev)
        self._all = torch.zeros(world * (world + 1), dtype=torch.int64, device=self.dev)
        self.parity = 0
        self.cur_max_ts = I64_MIN + bound_ms
        self.wm = I64_MIN
        self.dense = engine.cfg.key_mode == capi.FMQ_KEYS_DENSE
        dist.barrier(group=group)

    def global_keys(self, local_keys: np.ndarray) -> np.ndarray:
        return local_keys * np.uint64(self.world) + np.uint64(self.rank) if self.dense else local_keys

    def process_batch(self, ts, key, val, watermark_after: Optional[int] = None) -> int:
        import torch
        dist, W = self.dist, self.world
        with torch.cuda.stream(self._ext):
            counts = self.engine.exchange_split(ts, key, val, self.parity, W)          # synchronises the engine stream
            local_max = int(ts.max().item()) if ts.numel() else I64_MIN
            self._mine.copy_(torch.as_tensor(np.append(counts.astype(np.int64), local_max)), non_blocking=False)
            dist.all_gather_into_tensor(self._all, self._mine, group=self.group)        # sizes + max ts; also "all splits done"
            allv = self._all.cpu().numpy().reshape(W, W + 1)
            matrix = allv[:, :W].astype(np.uint64)
            if watermark_after is None:
                self.cur_max_ts = max(self.cur_max_ts, int(allv[:, W].max()))
                pot = self.cur_max_ts - self.bound
                if pot >= self.wm:
                    self.wm = pot
                watermark_after = self.wm
            n = self.engine.exchange_pull_push(self.parity, matrix, watermark_after)
        self.parity ^= 1
        return n

    def flush(self, wm: int = capi.FMQ_WM_FINAL):
        self.engine.flush(wm)


class ExchangeOperator:
    """keyBy exchange inside the library (include/fmq_b200.h, fmq_shard_*): the reference's structure — a hash
    exchange in front of owner-held keyed state (BandwidthMonitorWithEventTime.java:45-47) — with nothing but kernel
    launches per micro-batch.  Each rank's scatter kernel writes every owner's events straight into that owner's
    receive segment over NVLink, counts and the batch's largest timestamp travel as peer stores behind a flag, and
    the receiving rank's kernels wait on those flags on the device.  Every operator, both key modes, per-key arrival
    order.  `torch.distributed` (or any other transport) is used ONCE, at set-up, to hand the 64-byte IPC blobs
    around; pass `bases` instead to attach engines living in this process."""

    def __init__(self, engine, rank: int, world: int, max_send: int, group=None, bases=None):
        self.engine, self.rank, self.world = engine, rank, world
        self.dense = engine.cfg.key_mode == capi.FMQ_KEYS_DENSE
        if bases is not None:                                # same process: (engines already shard_alloc'ed)
            engine.shard_attach(rank, world, bases)
            return
        import torch.distributed as dist
        blob = engine.shard_alloc(world, max_send)
        blobs = [None] * world
        dist.all_gather_object(blobs, blob, group=group)
        engine.shard_attach_ipc(rank, world, blobs)
        dist.barrier(group=group)                            # every rank has mapped every buffer before the first store

    def global_keys(self, local_keys: np.ndarray) -> np.ndarray:
        return local_keys * np.uint64(self.world) + np.uint64(self.rank) if self.dense else local_keys

    def send(self, ts, key, val):
        self.engine.shard_send(ts, key, val)

    def recv(self, watermark_after: int = capi.FMQ_WM_AUTO, want_count: bool = False):
        return self.engine.shard_recv(watermark_after, want_count)

    def process_batch(self, ts, key, val, watermark_after: int = capi.FMQ_WM_AUTO):
        """send + recv of one micro-batch; the send of the next batch overlaps this one's kernels on its own stream."""
        self.engine.shard_send(ts, key, val)
        self.engine.shard_recv(watermark_after)

    def flush(self, wm: int = capi.FMQ_WM_FINAL):
        self.engine.flush(wm)
