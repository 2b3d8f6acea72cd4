"""Test-side loader of the CPU oracle (oracle/liboracle_flink.so).  TEST INFRASTRUCTURE ONLY:
nothing under the product package imports this module."""
import ctypes as C
import os
import subprocess

import numpy as np

from monitor_systam_flink_quickstart_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "oracle", "liboracle_flink.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
        _lib = C.CDLL(_SO)
        capi._declare(_lib, "orc_", has_mem=False, side_outputs=True)
        vp, sz, i32, i64 = C.c_void_p, C.c_size_t, C.c_int32, C.c_int64
        psz = C.POINTER(C.c_size_t)
        _lib.orc_java_string_hash.argtypes, _lib.orc_java_string_hash.restype = [C.c_char_p, sz], i32
        _lib.orc_murmur_hash.argtypes, _lib.orc_murmur_hash.restype = [i32], i32
        _lib.orc_key_group.argtypes, _lib.orc_key_group.restype = [i32, i32], i32
        _lib.orc_subtask.argtypes, _lib.orc_subtask.restype = [i32, i32, i32], i32
        # multi-thread baseline
        _lib.orcp_create.argtypes, _lib.orcp_create.restype = [C.POINTER(capi.FmqConfig), i32, C.POINTER(vp)], i32
        _lib.orcp_push_batch.argtypes, _lib.orcp_push_batch.restype = [vp, vp, vp, vp, sz, i64], i32
        _lib.orcp_flush.argtypes, _lib.orcp_flush.restype = [vp, i64], i32
        _lib.orcp_get_counters.argtypes, _lib.orcp_get_counters.restype = [vp, C.POINTER(capi.FmqCounters)], i32
        _lib.orcp_poll_fired.argtypes = [vp, C.POINTER(capi.FmqFiredCols), sz, psz, psz]
        _lib.orcp_poll_window_alerts.argtypes = [vp, C.POINTER(capi.FmqWAlertCols), sz, psz, psz]
        _lib.orcp_poll_event_alerts.argtypes = [vp, C.POINTER(capi.FmqEventCols), sz, psz, psz]
        _lib.orcp_poll_consec_alerts.argtypes = [vp, C.POINTER(capi.FmqEventCols), sz, psz, psz]
        _lib.orcp_get_key_state.argtypes, _lib.orcp_get_key_state.restype = [vp, vp, sz, vp, vp, vp], i32
        _lib.orcp_discard_outputs.argtypes = [vp]
        _lib.orcp_destroy.argtypes = [vp]
    return _lib


class OracleEngine(capi.EngineBase):
    """The literal single-thread Flink restatement behind the same driver as the product."""

    def __init__(self, cfg=None, **kw):
        super().__init__(lib(), "orc_", cfg if cfg is not None else capi.make_config(**kw), has_mem=False)


class ParallelOracle(capi.EngineBase):
    """The key-partitioned multi-thread CPU baseline (oracle/cpu_parallel.cpp)."""

    def __init__(self, cfg=None, threads=4, **kw):
        self._threads = threads
        l = lib()
        cfg = cfg if cfg is not None else capi.make_config(**kw)
        self._lib, self._p, self._has_mem, self.cfg = l, "orcp_", False, cfg
        self._vdtype = np.float64 if cfg.value_type == capi.FMQ_VAL_F64 else np.int64
        h = C.c_void_p()
        st = l.orcp_create(C.byref(cfg), threads, C.byref(h))
        if st != 0:
            raise capi.FmqError(st, "orcp_create")
        self._h = h

    def discard_outputs(self):
        self._lib.orcp_discard_outputs(self._h)


# ---------------------------------------------------------------------------------------------
# helpers shared by the parity tests
def sort_fired(f):
    o = np.lexsort((f.key, f.window_end))
    return capi.Fired(*(getattr(f, n)[o] for n in ("key", "window_end", "count", "sum", "min", "max")))


def sort_walerts(a):
    o = np.lexsort((a.key, a.window_end))
    return capi.WindowAlerts(a.key[o], a.window_end[o], a.value[o])


def sort_events(e):
    o = np.lexsort((e.val.view(np.uint64), e.key, e.ts))
    return capi.EventRows(e.ts[o], e.key[o], e.val[o])


def assert_fired_equal(got, want, f64_sum_rtol=1e-9, agg_mask=0xF):
    """Bit-exact on key / window / count / min / max (and integer sums); double sums within the
    1e-9 relative tolerance BASELINE.json's north_star states."""
    got, want = sort_fired(got), sort_fired(want)
    assert len(got) == len(want), f"fired rows: got {len(got)} want {len(want)}"
    np.testing.assert_array_equal(got.key, want.key)
    np.testing.assert_array_equal(got.window_end, want.window_end)
    np.testing.assert_array_equal(got.count, want.count)
    if agg_mask & capi.FMQ_AGG_SUM:
        if got.sum.dtype == np.float64:
            np.testing.assert_allclose(got.sum, want.sum, rtol=f64_sum_rtol, atol=0.0)
        else:
            np.testing.assert_array_equal(got.sum, want.sum)
    if agg_mask & capi.FMQ_AGG_MIN:
        np.testing.assert_array_equal(got.min.view(np.uint64), want.min.view(np.uint64))
    if agg_mask & capi.FMQ_AGG_MAX:
        np.testing.assert_array_equal(got.max.view(np.uint64), want.max.view(np.uint64))


def assert_events_equal(got, want):
    got, want = sort_events(got), sort_events(want)
    assert len(got) == len(want), f"event rows: got {len(got)} want {len(want)}"
    np.testing.assert_array_equal(got.ts, want.ts)
    np.testing.assert_array_equal(got.key, want.key)
    np.testing.assert_array_equal(got.val.view(np.uint64), want.val.view(np.uint64))
