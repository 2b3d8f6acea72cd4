"""ctypes binding of the C ABI declared in ``include/fmq_b200.h``.

This is the thinnest possible layer: POD structs, symbol signatures, and an ``Engine`` class that
turns numpy arrays / CUDA tensors into the pointers the ABI wants.  It is what a Python caller
(and the parity tests) use; a Flink shim would bind the same symbols through JNI or Panama
(see INTEGRATION.md).  There is no CPU fallback: if the CUDA library is missing or no device is
usable, construction fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Optional

import numpy as np

FMQ_VAL_F64, FMQ_VAL_I64 = 0, 1
FMQ_MEM_HOST, FMQ_MEM_DEVICE = 0, 1
FMQ_KEYS_HASHED, FMQ_KEYS_DENSE = 0, 1
FMQ_TEXT_CPU, FMQ_TEXT_FLOW = 0, 1
FMQ_TEXT_OK, FMQ_TEXT_MALFORMED, FMQ_TEXT_VALUE_ON_HOST, FMQ_TEXT_KEY_ON_HOST = 0, 1, 2, 3
FMQ_CMP_GT, FMQ_CMP_LT, FMQ_CMP_GE, FMQ_CMP_LE = 0, 1, 2, 3
FMQ_AGG_COUNT, FMQ_AGG_SUM, FMQ_AGG_MIN, FMQ_AGG_MAX = 1, 2, 4, 8
FMQ_ALERT_NONE, FMQ_ALERT_SUM, FMQ_ALERT_AVG, FMQ_ALERT_MIN, FMQ_ALERT_MAX, FMQ_ALERT_COUNT = range(6)
FMQ_WM_AUTO = -(2 ** 63)
FMQ_WM_KEEP = -(2 ** 63) + 1
FMQ_WM_FINAL = 2 ** 63 - 1

STATUS_NAMES = {0: "OK", 1: "BAD_ARG", 2: "TABLE_FULL", 3: "RING_OVERFLOW", 4: "OUTPUT_OVERFLOW",
                5: "CUDA", 6: "BAD_TS", 7: "NO_DEVICE", 8: "POISONED", 9: "PEER_TIMEOUT"}


class FmqConfig(C.Structure):
    """Mirror of ``struct fmq_config`` (include/fmq_b200.h)."""
    _fields_ = [
        ("struct_size", C.c_uint32), ("device", C.c_int32), ("value_type", C.c_int32), ("key_mode", C.c_int32),
        ("window_enabled", C.c_int32), ("agg_mask", C.c_uint32),
        ("window_size_ms", C.c_int64), ("window_slide_ms", C.c_int64), ("window_offset_ms", C.c_int64),
        ("max_out_of_orderness_ms", C.c_int64), ("emit_fired", C.c_int32),
        ("walert_source", C.c_int32), ("walert_op", C.c_int32), ("reserved1", C.c_int32),
        ("walert_mul", C.c_double), ("walert_div0", C.c_double), ("walert_div1", C.c_double),
        ("walert_div2", C.c_double), ("walert_threshold", C.c_double),
        ("efilter_enabled", C.c_int32), ("efilter_op", C.c_int32), ("efilter_threshold", C.c_double),
        ("rolling_max_enabled", C.c_int32), ("consec_n", C.c_int32), ("consec_op", C.c_int32),
        ("consec_reset_on_fire", C.c_int32), ("consec_threshold", C.c_double),
        ("table_capacity", C.c_uint64), ("pane_ring", C.c_uint32), ("reserved2", C.c_uint32),
        ("max_batch", C.c_uint64), ("fired_capacity", C.c_uint64), ("walert_capacity", C.c_uint64),
        ("ealert_capacity", C.c_uint64), ("calert_capacity", C.c_uint64),
        ("text_max_bytes", C.c_uint64), ("median_log_capacity", C.c_uint64), ("median_capacity", C.c_uint64),
        ("late_capacity", C.c_uint64), ("overlapped_polls", C.c_int32), ("reserved3", C.c_int32),
        ("text_dict_capacity", C.c_uint64), ("text_dict_arena_bytes", C.c_uint64),
        ("allowed_lateness_ms", C.c_int64),
    ]


class FmqFiredCols(C.Structure):
    _fields_ = [("key", C.c_void_p), ("window_end", C.c_void_p), ("count", C.c_void_p),
                ("sum", C.c_void_p), ("min", C.c_void_p), ("max", C.c_void_p)]


class FmqWAlertCols(C.Structure):
    _fields_ = [("key", C.c_void_p), ("window_end", C.c_void_p), ("value", C.c_void_p)]


class FmqEventCols(C.Structure):
    _fields_ = [("ts", C.c_void_p), ("key", C.c_void_p), ("val", C.c_void_p)]


class FmqPackedColumns(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("ts_width", C.c_uint32), ("key_width", C.c_uint32), ("val_width", C.c_uint32),
                ("ts_base", C.c_int64), ("ts", C.c_void_p), ("key", C.c_void_p), ("val", C.c_void_p)]


class FmqCounters(C.Structure):
    _fields_ = [("events_in", C.c_uint64), ("late_dropped", C.c_uint64), ("windows_fired", C.c_uint64),
                ("window_alerts", C.c_uint64), ("event_alerts", C.c_uint64), ("consec_alerts", C.c_uint64),
                ("keys", C.c_uint64), ("watermark", C.c_int64), ("gpu_launches", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def make_config(**kw) -> FmqConfig:
    """Zero-initialised config with ``struct_size`` set; keyword arguments override fields."""
    cfg = FmqConfig()
    cfg.struct_size = C.sizeof(FmqConfig)
    cfg.emit_fired = 1
    for k, v in kw.items():
        if not hasattr(cfg, k):
            raise AttributeError(f"fmq_config has no field {k!r}")
        setattr(cfg, k, v)
    return cfg


class FmqError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"fmq status {status} ({STATUS_NAMES.get(status, '?')}): {msg}")
        self.status = status


# FMQ_LIB_PATH lets a developer A/B-test another build of the SAME CUDA library (never a fallback)
_LIB_PATH = os.environ.get("FMQ_LIB_PATH") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "libfmq_b200.so")
_lib = None


def library_path() -> str:
    return _LIB_PATH


def load_library() -> C.CDLL:
    """Load ``libfmq_b200.so`` (built in-tree by ``__graft_entry__.build()``).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise FileNotFoundError(
            f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "— there is no CPU fallback for the hot path")
    lib = C.CDLL(_LIB_PATH)
    _declare(lib, "fmq_")
    _lib = lib
    return lib


def _declare(lib, p: str, has_mem: bool = True, side_outputs: Optional[bool] = None):
    """Attach argtypes/restype for every entry point; `p` is the symbol prefix.  `has_mem`: the calls
    take the host/device `mem` argument (the product does); `side_outputs`: the library exports
    poll_late / poll_medians (default: same as has_mem)."""
    side_outputs = has_mem if side_outputs is None else side_outputs
    vp, sz, i32, i64 = C.c_void_p, C.c_size_t, C.c_int32, C.c_int64
    psz = C.POINTER(C.c_size_t)
    mem = [i32] if has_mem else []

    def sig(name, args, res=i32):
        f = getattr(lib, p + name)
        f.argtypes, f.restype = args, res
    sig("create", [C.POINTER(FmqConfig), C.POINTER(vp)])
    sig("push_batch", [vp, vp, vp, vp, sz, i64] + mem)
    sig("flush", [vp, i64])
    sig("poll_fired", [vp, C.POINTER(FmqFiredCols), sz] + mem + [psz, psz])
    sig("poll_window_alerts", [vp, C.POINTER(FmqWAlertCols), sz] + mem + [psz, psz])
    sig("poll_event_alerts", [vp, C.POINTER(FmqEventCols), sz] + mem + [psz, psz])
    if side_outputs:
        sig("poll_late", [vp, C.POINTER(FmqEventCols), sz] + mem + [psz, psz])
        sig("poll_medians", [vp, C.POINTER(FmqWAlertCols), sz] + mem + [psz, psz])
    sig("poll_consec_alerts", [vp, C.POINTER(FmqEventCols), sz] + mem + [psz, psz])
    sig("poll_rolling_max", [vp, vp, sz] + mem + [psz])
    sig("get_counters", [vp, C.POINTER(FmqCounters)])
    sig("get_key_state", [vp, vp, sz, vp, vp, vp])
    sig("destroy", [vp])
    if has_mem:   # product-only entry points
        sig("abi_version", [])
        sig("build_info", [], C.c_char_p)
        sig("synchronize", [vp])
        sig("last_device_ms", [vp, C.POINTER(C.c_float)])
        pf = C.POINTER(C.c_float)
        sig("get_push_timing", [vp, C.c_uint32, pf, pf, pf, pf])
        sig("stream", [vp], vp)
        sig("last_error", [vp], C.c_char_p)
        sig("owner_of", [C.c_uint64, C.c_uint32], C.c_uint32)
        sig("java_string_hash", [C.c_char_p, sz])
        sig("flink_key_group", [i32, i32])
        sig("flink_subtask", [i32, i32, i32])
        sig("partition_by_owner", [vp, vp, vp, vp, sz, C.c_uint32, vp, vp, vp, vp])
        sig("text_key_hash", [C.c_char_p, sz], C.c_uint64)
        sig("parse_text", [vp, vp, sz, i32, i32, i32, vp, vp, vp, vp, sz, i32, psz, psz])
        sig("push_text", [vp, vp, sz, i32, i32, i32, i64, psz])
        sig("dict_size", [vp, psz])
        sig("dict_lookup", [vp, vp, sz, vp, sz, vp])
        sig("poll_all_begin", [vp, C.POINTER(FmqFiredCols), sz, C.POINTER(FmqWAlertCols), sz, C.POINTER(FmqEventCols), sz, psz])
        sig("poll_all_end", [vp])
        sig("snapshot_size", [vp, psz])
        sig("snapshot", [vp, vp, sz, psz])
        sig("restore", [vp, vp, sz])
        sig("exchange_alloc", [vp, vp])
        sig("exchange_base", [vp], vp)
        sig("exchange_attach", [vp, i32, i32, C.POINTER(vp)])
        sig("exchange_split", [vp, vp, vp, vp, sz, i32, vp])
        sig("exchange_pull_push", [vp, i32, vp, i64, psz])
        sig("table_ptr", [vp], vp)
        sig("exchange_ptr", [vp], vp)
        sig("ipc_export", [vp, vp])
        sig("ipc_open", [vp, vp, C.POINTER(vp)])
        sig("attach_peers", [vp, i32, i32, C.POINTER(vp)])
        sig("push_local", [vp, vp, vp, vp, sz, i32])
        sig("fire_global", [vp, i64])
        sig("retire", [vp])
        sig("combiner_overlap", [vp, i32])
        sig("fire_stream", [vp], vp)
        sig("push_packed", [vp, C.POINTER(FmqPackedColumns), sz, i64, i32])
        sig("stage_host", [vp, C.POINTER(FmqPackedColumns), sz])
        sig("push_local_packed", [vp, C.POINTER(FmqPackedColumns), sz, i32])
        sig("partition_by_owner_noalloc", [i32, vp, vp, vp, vp, sz, C.c_uint32, vp, vp, vp, vp])
        sig("shard_alloc", [vp, i32, sz, vp])
        sig("shard_base", [vp], vp)
        sig("shard_attach", [vp, i32, i32, C.POINTER(vp)])
        sig("shard_attach_ipc", [vp, i32, i32, vp])
        sig("shard_send", [vp, vp, vp, vp, sz, i32])
        sig("shard_recv", [vp, i64, psz])
        sig("combiner_use_flags", [vp, i32])


def _ptr(a) -> int:
    """Raw address of a numpy array or torch tensor (host or device)."""
    if a is None:
        return 0
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()          # torch.Tensor


def _is_device(a) -> bool:
    return (not isinstance(a, np.ndarray)) and hasattr(a, "is_cuda") and a.is_cuda


@dataclass
class Fired:
    key: np.ndarray
    window_end: np.ndarray
    count: np.ndarray
    sum: np.ndarray
    min: np.ndarray
    max: np.ndarray

    def __len__(self):
        return len(self.key)


@dataclass
class WindowAlerts:
    key: np.ndarray
    window_end: np.ndarray
    value: np.ndarray

    def __len__(self):
        return len(self.key)


@dataclass
class EventRows:
    ts: np.ndarray
    key: np.ndarray
    val: np.ndarray

    def __len__(self):
        return len(self.key)


class EngineBase:
    """Shared driver for anything exporting the fmq-shaped ABI: the product library, and any other
    library with the same call shapes under another symbol prefix (with or without the ``mem``
    argument) that a caller loads itself."""

    def __init__(self, lib, prefix: str, cfg: FmqConfig, has_mem: bool):
        self._lib, self._p, self._has_mem = lib, prefix, has_mem
        self.cfg = cfg
        self._vdtype = np.float64 if cfg.value_type == FMQ_VAL_F64 else np.int64
        h = C.c_void_p()
        st = self._f("create")(C.byref(cfg), C.byref(h))
        if st != 0:
            raise FmqError(st, self._err(None))
        self._h = h

    def _f(self, name):
        return getattr(self._lib, self._p + name)

    def _err(self, h) -> str:
        f = getattr(self._lib, self._p + "last_error", None)
        if f is None:
            return "(the library exports no last_error)"
        m = f(h)
        return m.decode() if m else ""

    def _check(self, st):
        if st != 0:
            raise FmqError(st, self._err(self._h))

    # --- ingest -------------------------------------------------------------------------------
    def push_batch(self, ts, key, val, watermark_after: int = FMQ_WM_AUTO):
        n = len(ts)
        assert len(key) == n and len(val) == n
        dev = _is_device(ts)
        if not dev:
            ts = np.ascontiguousarray(ts, dtype=np.int64)
            key = np.ascontiguousarray(key, dtype=np.uint64)
            val = np.ascontiguousarray(val, dtype=self._vdtype)
        args = [self._h, _ptr(ts), _ptr(key), _ptr(val), n, watermark_after]
        if self._has_mem:
            args.append(FMQ_MEM_DEVICE if dev else FMQ_MEM_HOST)
        self._check(self._f("push_batch")(*args))

    @staticmethod
    def _packed(ts, key, val, ts_base):
        n = len(ts)
        assert len(key) == n and len(val) == n
        dev = _is_device(ts)
        cols = [a if dev else np.ascontiguousarray(a) for a in (ts, key, val)]
        w = [a.element_size() if dev else a.dtype.itemsize for a in cols]
        return FmqPackedColumns(C.sizeof(FmqPackedColumns), w[0], w[1], w[2], ts_base, _ptr(cols[0]), _ptr(cols[1]), _ptr(cols[2])), n, dev, cols

    def push_packed(self, ts, key, val, watermark_after: int = FMQ_WM_AUTO, ts_base: int = 0):
        """fmq_push_packed: the element size of each column picks the wire format (uint32 ts offsets
        from `ts_base`, uint32 key ids, int32 values are widened on the device)."""
        pc, n, dev, _keep = self._packed(ts, key, val, ts_base)
        self._check(self._f("push_packed")(self._h, C.byref(pc), n, watermark_after, FMQ_MEM_DEVICE if dev else FMQ_MEM_HOST))

    def stage_host(self, ts, key, val, ts_base: int = 0):
        """fmq_stage_host: start copying the NEXT host batch in; the caller keeps the arrays alive and
        untouched and passes the very same ones to the next push_batch / push_packed."""
        pc, n, dev, cols = self._packed(ts, key, val, ts_base)
        assert not dev and all(c is a for c, a in zip(cols, (ts, key, val))), "stage_host needs contiguous host arrays (no hidden copies)"
        self._check(self._f("stage_host")(self._h, C.byref(pc), n))

    def flush(self, watermark: int = FMQ_WM_FINAL):
        self._check(self._f("flush")(self._h, watermark))

    # --- outputs ------------------------------------------------------------------------------
    def _poll_loop(self, name, make_cols, chunk):
        parts = []
        while True:
            cols, arrays = make_cols(chunk)
            n_out, n_left = C.c_size_t(0), C.c_size_t(0)
            args = [self._h, C.byref(cols), chunk]
            if self._has_mem:
                args.append(FMQ_MEM_HOST)
            args += [C.byref(n_out), C.byref(n_left)]
            self._check(self._f(name)(*args))
            parts.append([a[: n_out.value] for a in arrays])
            if n_left.value == 0:
                break
        return [np.concatenate([p[i] for p in parts]) for i in range(len(parts[0]))]

    # --- zero-allocation polls: the caller supplies (ideally pinned) column buffers ---------------
    def _poll_into(self, name, cols_cls, arrays, offset_rows=0):
        """Single poll call into caller-owned arrays (numpy or pinned torch tensors, all of one
        capacity).  Returns (rows copied, rows left)."""
        cap = len(arrays[0]) - offset_rows
        ptrs = [(_ptr(a) + offset_rows * 8) if a is not None else 0 for a in arrays]
        cols = cols_cls(*ptrs)
        n_out, n_left = C.c_size_t(0), C.c_size_t(0)
        args = [self._h, C.byref(cols), cap]
        if self._has_mem:
            args.append(FMQ_MEM_DEVICE if _is_device(arrays[0]) else FMQ_MEM_HOST)
        args += [C.byref(n_out), C.byref(n_left)]
        self._check(self._f(name)(*args))
        return n_out.value, n_left.value

    def poll_all_begin(self, fired6, walert3, ealert3):
        """Overlapped poll: queues the D2H copies of all three outputs into the given (pinned) column
        lists and returns (n_fired, n_walert, n_ealert) at once; call poll_all_end() before reading."""
        f = FmqFiredCols(*[_ptr(a) for a in fired6]) if fired6 else None
        w = FmqWAlertCols(*[_ptr(a) for a in walert3]) if walert3 else None
        e = FmqEventCols(*[_ptr(a) for a in ealert3]) if ealert3 else None
        counts = (C.c_size_t * 3)()
        self._check(self._lib.fmq_poll_all_begin(self._h, C.byref(f) if f else None, len(fired6[0]) if fired6 else 0,
                                                 C.byref(w) if w else None, len(walert3[0]) if walert3 else 0,
                                                 C.byref(e) if e else None, len(ealert3[0]) if ealert3 else 0, counts))
        return tuple(counts)

    def poll_all_end(self):
        self._check(self._lib.fmq_poll_all_end(self._h))

    def poll_fired_into(self, key, window_end, count, sum, min, max):
        return self._poll_into("poll_fired", FmqFiredCols, [key, window_end, count, sum, min, max])

    def poll_window_alerts_into(self, key, window_end, value):
        return self._poll_into("poll_window_alerts", FmqWAlertCols, [key, window_end, value])

    def poll_event_alerts_into(self, ts, key, val):
        return self._poll_into("poll_event_alerts", FmqEventCols, [ts, key, val])

    def poll_consec_alerts_into(self, ts, key, val):
        return self._poll_into("poll_consec_alerts", FmqEventCols, [ts, key, val])

    def poll_fired(self, chunk: int = 1 << 20) -> Fired:
        def mk(c):
            arrs = [np.empty(c, np.uint64), np.empty(c, np.int64), np.empty(c, np.int64),
                    np.empty(c, self._vdtype), np.empty(c, self._vdtype), np.empty(c, self._vdtype)]
            return FmqFiredCols(*[a.ctypes.data for a in arrs]), arrs
        return Fired(*self._poll_loop("poll_fired", mk, chunk))

    def poll_window_alerts(self, chunk: int = 1 << 20) -> WindowAlerts:
        def mk(c):
            arrs = [np.empty(c, np.uint64), np.empty(c, np.int64), np.empty(c, np.float64)]
            return FmqWAlertCols(*[a.ctypes.data for a in arrs]), arrs
        return WindowAlerts(*self._poll_loop("poll_window_alerts", mk, chunk))

    def poll_medians(self, chunk: int = 1 << 20) -> WindowAlerts:
        """(key, window_end, median) rows of ComputeCpuMiddle's window function."""
        def mk(c):
            arrs = [np.empty(c, np.uint64), np.empty(c, np.int64), np.empty(c, np.float64)]
            return FmqWAlertCols(*[a.ctypes.data for a in arrs]), arrs
        return WindowAlerts(*self._poll_loop("poll_medians", mk, chunk))

    def _poll_events(self, name, chunk) -> EventRows:
        def mk(c):
            arrs = [np.empty(c, np.int64), np.empty(c, np.uint64), np.empty(c, self._vdtype)]
            return FmqEventCols(*[a.ctypes.data for a in arrs]), arrs
        return EventRows(*self._poll_loop(name, mk, chunk))

    def poll_event_alerts(self, chunk: int = 1 << 20) -> EventRows:
        return self._poll_events("poll_event_alerts", chunk)

    def poll_late(self, chunk: int = 1 << 20) -> EventRows:
        """The late side output (sideOutputLateData)."""
        return self._poll_events("poll_late", chunk)

    def poll_consec_alerts(self, chunk: int = 1 << 20) -> EventRows:
        return self._poll_events("poll_consec_alerts", chunk)

    def poll_rolling_max(self, n: int) -> np.ndarray:
        out = np.empty(n, self._vdtype)
        n_out = C.c_size_t(0)
        args = [self._h, out.ctypes.data, n]
        if self._has_mem:
            args.append(FMQ_MEM_HOST)
        args.append(C.byref(n_out))
        self._check(self._f("poll_rolling_max")(*args))
        return out[: n_out.value]

    def counters(self) -> dict:
        c = FmqCounters()
        self._check(self._f("get_counters")(self._h, C.byref(c)))
        return c.as_dict()

    def key_state(self, keys):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        n = len(keys)
        found, consec, rolling = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, self._vdtype)
        self._check(self._f("get_key_state")(self._h, keys.ctypes.data, n, found.ctypes.data,
                                             consec.ctypes.data, rolling.ctypes.data))
        return found, consec, rolling

    def close(self):
        if getattr(self, "_h", None):
            self._f("destroy")(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Engine(EngineBase):
    """One native handle of libfmq_b200 = one operator subtask on one GPU."""

    def __init__(self, cfg: Optional[FmqConfig] = None, **kw):
        lib = load_library()
        super().__init__(lib, "fmq_", cfg if cfg is not None else make_config(**kw), has_mem=True)

    def synchronize(self):
        self._check(self._lib.fmq_synchronize(self._h))

    def last_device_ms(self) -> float:
        ms = C.c_float(0)
        self._check(self._lib.fmq_last_device_ms(self._h, C.byref(ms)))
        return ms.value

    def push_timing(self, back: int = 0) -> dict:
        """Device ms of the push `back` calls ago, by stage (CUDA events on the handle's stream)."""
        t, i, k, f = C.c_float(0), C.c_float(0), C.c_float(0), C.c_float(0)
        self._check(self._lib.fmq_get_push_timing(self._h, back, C.byref(t), C.byref(i), C.byref(k), C.byref(f)))
        return {"total_ms": t.value, "ingest_ms": i.value, "keyed_ms": k.value, "fire_ms": f.value}

    @property
    def stream(self) -> int:
        return self._lib.fmq_stream(self._h) or 0

    @property
    def handle(self):
        return self._h

    # ---- text ingest on the device ------------------------------------------------------------------
    def parse_text(self, text: bytes, fmt: int, utc_offset_hours: int = 8):
        """-> (ts, key, val, status) numpy arrays, one entry per line."""
        cap = text.count(b"\n") + 1
        ts, key, val, st = np.empty(cap, np.int64), np.empty(cap, np.uint64), np.empty(cap, self._vdtype), np.empty(cap, np.uint32)
        n, bad = C.c_size_t(0), C.c_size_t(0)
        self._check(self._lib.fmq_parse_text(self._h, text, len(text), FMQ_MEM_HOST, fmt, utc_offset_hours, ts.ctypes.data, key.ctypes.data,
                                             val.ctypes.data, st.ctypes.data, cap, FMQ_MEM_HOST, C.byref(n), C.byref(bad)))
        k = n.value
        return ts[:k], key[:k], val[:k], st[:k]

    def push_text(self, text: bytes, fmt: int, watermark_after: int = FMQ_WM_AUTO, utc_offset_hours: int = 8) -> int:
        n = C.c_size_t(0)
        self._check(self._lib.fmq_push_text(self._h, text, len(text), FMQ_MEM_HOST, fmt, utc_offset_hours, watermark_after, C.byref(n)))
        return n.value

    def dict_size(self) -> int:
        n = C.c_size_t(0)
        self._check(self._lib.fmq_dict_size(self._h, C.byref(n)))
        return n.value

    def dict_lookup(self, ids, stride: int = 64):
        """Names behind dictionary ids (text_dict_capacity > 0) -> list of str (None for an id never handed out)."""
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        n = len(ids)
        out = np.zeros((n, stride), np.uint8); lens = np.zeros(n, np.uint32)
        if n:
            self._check(self._lib.fmq_dict_lookup(self._h, ids.ctypes.data, n, out.ctypes.data, stride, lens.ctypes.data))
        return [None if l == 0xFFFFFFFF else bytes(out[i, :min(int(l), stride)]).decode("utf-8", "replace") for i, l in enumerate(lens)]

    # ---- checkpointing (snapshotState / initializeState) ------------------------------------------
    def snapshot(self) -> np.ndarray:
        n = C.c_size_t(0)
        self._check(self._lib.fmq_snapshot_size(self._h, C.byref(n)))
        buf = np.empty(n.value, np.uint8)
        w = C.c_size_t(0)
        self._check(self._lib.fmq_snapshot(self._h, buf.ctypes.data, n.value, C.byref(w)))
        return buf[: w.value]

    def restore(self, snap: np.ndarray):
        snap = np.ascontiguousarray(snap, dtype=np.uint8)
        self._check(self._lib.fmq_restore(self._h, snap.ctypes.data, snap.nbytes))

    # ---- combiner mode (multi-GPU without an event shuffle) -----------------------------------
    @property
    def table_ptr(self) -> int:
        return self._lib.fmq_table_ptr(self._h) or 0

    @property
    def exchange_ptr(self) -> int:
        return self._lib.fmq_exchange_ptr(self._h) or 0

    def ipc_export(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self._lib.fmq_ipc_export(self._h, buf))
        return buf.raw

    def ipc_open(self, blob: bytes) -> int:
        p = C.c_void_p()
        self._check(self._lib.fmq_ipc_open(self._h, blob, C.byref(p)))
        return p.value

    def attach_peers(self, rank: int, world: int, tables):
        arr = (C.c_void_p * world)(*[C.c_void_p(t) for t in tables])
        self._check(self._lib.fmq_attach_peers(self._h, rank, world, arr))

    def push_local(self, ts, key, val, ts_base: int = 0):
        """Combiner mode, stage 1: host or device columns; narrow element sizes pick the packed format."""
        pc, n, dev, _keep = self._packed(ts, key, val, ts_base)
        self._check(self._lib.fmq_push_local_packed(self._h, C.byref(pc), n, FMQ_MEM_DEVICE if dev else FMQ_MEM_HOST))

    def fire_global(self, watermark_after: int = FMQ_WM_AUTO):
        self._check(self._lib.fmq_fire_global(self._h, watermark_after))

    def retire(self):
        self._check(self._lib.fmq_retire(self._h))

    def combiner_overlap(self, enable: bool = True):
        self._check(self._lib.fmq_combiner_overlap(self._h, 1 if enable else 0))

    def combiner_use_flags(self, enable: bool = True):
        self._check(self._lib.fmq_combiner_use_flags(self._h, 1 if enable else 0))

    @property
    def fire_stream(self) -> int:
        return self._lib.fmq_fire_stream(self._h)

    # ---- peer-pull keyBy exchange ---------------------------------------------------------------
    def exchange_alloc(self) -> bytes:
        buf = C.create_string_buffer(64)
        self._check(self._lib.fmq_exchange_alloc(self._h, buf))
        return buf.raw

    @property
    def exchange_base(self) -> int:
        return self._lib.fmq_exchange_base(self._h) or 0

    def exchange_attach(self, rank: int, world: int, bases):
        arr = (C.c_void_p * world)(*[C.c_void_p(b) for b in bases])
        self._check(self._lib.fmq_exchange_attach(self._h, rank, world, arr))

    def exchange_split(self, ts, key, val, parity: int, world: int) -> np.ndarray:
        counts = np.zeros(world, np.uint64)
        self._check(self._lib.fmq_exchange_split(self._h, _ptr(ts), _ptr(key), _ptr(val), len(ts), parity, counts.ctypes.data))
        return counts

    def exchange_pull_push(self, parity: int, matrix: np.ndarray, watermark_after: int = FMQ_WM_AUTO) -> int:
        m = np.ascontiguousarray(matrix, dtype=np.uint64)
        n = C.c_size_t(0)
        self._check(self._lib.fmq_exchange_pull_push(self._h, parity, m.ctypes.data, watermark_after, C.byref(n)))
        return n.value

    # ---- keyBy exchange on the device (fmq_shard_*) ------------------------------------------------
    def shard_alloc(self, world: int, max_send: int) -> bytes:
        """Allocates this rank's receive buffer; returns the 64-byte CUDA IPC blob the peers attach."""
        buf = C.create_string_buffer(64)
        self._check(self._lib.fmq_shard_alloc(self._h, world, max_send, buf))
        return buf.raw

    @property
    def shard_base(self) -> int:
        return self._lib.fmq_shard_base(self._h) or 0

    def shard_attach(self, rank: int, world: int, bases):
        """Peers in this process: their fmq_shard_base pointers."""
        arr = (C.c_void_p * world)(*[C.c_void_p(b) for b in bases])
        self._check(self._lib.fmq_shard_attach(self._h, rank, world, arr))

    def shard_attach_ipc(self, rank: int, world: int, blobs):
        """Peers in other processes: the blobs of every rank's shard_alloc, in rank order."""
        raw = b"".join(blobs)
        assert len(raw) == 64 * world
        self._check(self._lib.fmq_shard_attach_ipc(self._h, rank, world, raw))

    def shard_send(self, ts, key, val):
        n = len(ts)
        dev = _is_device(ts)
        if not dev:
            ts = np.ascontiguousarray(ts, dtype=np.int64); key = np.ascontiguousarray(key, dtype=np.uint64)
            val = np.ascontiguousarray(val, dtype=self._vdtype)
        self._check(self._lib.fmq_shard_send(self._h, _ptr(ts), _ptr(key), _ptr(val), n, FMQ_MEM_DEVICE if dev else FMQ_MEM_HOST))

    def shard_recv(self, watermark_after: int = FMQ_WM_AUTO, want_count: bool = False):
        n = C.c_size_t(0)
        self._check(self._lib.fmq_shard_recv(self._h, watermark_after, C.byref(n) if want_count else None))
        return n.value if want_count else None

    def partition_by_owner(self, ts, key, val, world: int, ts_out, key_out, val_out) -> np.ndarray:
        counts = np.zeros(world, np.uint64)
        self._check(self._lib.fmq_partition_by_owner(self._h, _ptr(ts), _ptr(key), _ptr(val), len(ts), world,
                                                     _ptr(ts_out), _ptr(key_out), _ptr(val_out), counts.ctypes.data))
        return counts
