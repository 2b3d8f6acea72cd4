"""Python mirror of the Flink operator the reference's DSL chain instantiates.

``GpuWindowAggOperator`` has the surface of ``OneInputStreamOperator`` [Flink 1.8.0]:
``open() / process_element(record) / process_watermark(wm) / close()`` — the hooks a
``stream.keyBy(k).transform("gpuWindowAgg", type, operator)`` would drive in place of
``keyBy(1).timeWindow(..).reduce(..).map(..).filter(..)``
(chapter3/.../BandwidthMonitorWithEventTime.java:45-55) or ``keyBy(0).max(2)``
(chapter2/.../ComputeCpuMax.java:26).  Records are buffered into three columns; a watermark (or a
full buffer) hands the micro-batch to the CUDA engine through the C ABI and drains what fired into
the collector, stamped with Flink's record timestamp ``window_end - 1``.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np

from . import capi
from .text import KeyDictionary


class GpuWindowAggOperator:
    """collector(kind, row) receives: ("window", (key, window_end-1, count, sum, min, max)),
    ("window_alert", (key, window_end-1, value)), ("event_alert", (ts, key, val)),
    ("consec_alert", (ts, key, val)), ("rolling_max", (key, value))."""

    def __init__(self, collector: Callable, batch_capacity: int = 1 << 16, engine_factory: Optional[Callable] = None,
                 auto_watermark: bool = False, **config):
        """auto_watermark: False (default) = watermarks come from upstream through process_watermark, like in a
        Flink job whose assigner sits in front of the keyBy; a buffer that fills up between two of them is pushed
        with FMQ_WM_KEEP, so event time only ever moves in process_watermark ([Flink 1.8.0] WindowOperator).
        True = the engine derives the watermark itself at every micro-batch boundary (FMQ_WM_AUTO)."""
        self._collector = collector
        self._between = capi.FMQ_WM_AUTO if auto_watermark else capi.FMQ_WM_KEEP
        self._cap = batch_capacity
        self._config = dict(config)
        self._config.setdefault("max_batch", batch_capacity)
        self._factory = engine_factory or (lambda **kw: capi.Engine(**kw))
        self.keys = KeyDictionary()
        self._engine = None

    # -- OneInputStreamOperator#open -----------------------------------------------------------
    def open(self):
        self._engine = self._factory(**self._config)
        vt = np.float64 if self._config.get("value_type", capi.FMQ_VAL_F64) == capi.FMQ_VAL_F64 else np.int64
        self._ts = np.empty(self._cap, np.int64)
        self._key = np.empty(self._cap, np.uint64)
        self._val = np.empty(self._cap, vt)
        self._n = 0
        return self

    # -- OneInputStreamOperator#processElement ---------------------------------------------------
    def process_element(self, timestamp_ms: int, key, value):
        """`key` may be a string (dictionary-encoded here, like the Java shim) or an integer id."""
        if self._n == self._cap:
            self._push(self._between)
        i = self._n
        self._ts[i] = timestamp_ms
        self._key[i] = self.keys.encode(key) if isinstance(key, str) else key
        self._val[i] = value
        self._n = i + 1

    # -- OneInputStreamOperator#processWatermark -------------------------------------------------
    def process_watermark(self, watermark_ms: int = capi.FMQ_WM_AUTO):
        """FMQ_WM_AUTO derives the watermark from the data with the configured bound
        (BoundedOutOfOrdernessTimestampExtractor, chapter3/README.md:379-396)."""
        self._push(watermark_ms)

    def _push(self, wm):
        n = self._n
        wm = capi.FMQ_WM_AUTO if wm is None else wm
        keys_in_batch = self._key[:n].copy()
        self._engine.push_batch(self._ts[:n], self._key[:n], self._val[:n], wm)
        self._n = 0
        self._drain(n, keys_in_batch)

    def _drain(self, n_last: int, keys_last):
        e, cfg, out = self._engine, self._config, self._collector
        if cfg.get("rolling_max_enabled"):
            for k, v in zip(keys_last, e.poll_rolling_max(n_last)):
                out("rolling_max", (self._name(k), v.item()))
        if cfg.get("efilter_enabled"):
            r = capi_sorted_events(e.poll_event_alerts())
            for t, k, v in zip(r.ts, r.key, r.val):
                out("event_alert", (int(t), self._name(k), v.item()))
        if cfg.get("consec_n", 0) > 0:
            r = capi_sorted_events(e.poll_consec_alerts())
            for t, k, v in zip(r.ts, r.key, r.val):
                out("consec_alert", (int(t), self._name(k), v.item()))
        if cfg.get("window_enabled"):
            if cfg.get("emit_fired", 1):
                f = e.poll_fired()
                o = np.lexsort((f.key, f.window_end))          # timers fire in timestamp order
                for i in o:
                    out("window", (self._name(f.key[i]), int(f.window_end[i]) - 1, int(f.count[i]), f.sum[i].item(), f.min[i].item(), f.max[i].item()))
            if cfg.get("walert_source", 0):
                a = e.poll_window_alerts()
                for i in np.lexsort((a.key, a.window_end)):
                    out("window_alert", (self._name(a.key[i]), int(a.window_end[i]) - 1, float(a.value[i])))

    def _name(self, k):
        k = int(k)
        return self.keys.decode(k) if k < len(self.keys) else k

    def metrics(self) -> dict:
        """numLateRecordsDropped & co."""
        return self._engine.counters()

    # -- OneInputStreamOperator#close: end of input = Watermark(Long.MAX_VALUE) ------------------
    def close(self):
        if self._engine is None:
            return
        if self._n:
            self._push(self._between)
        self._engine.flush(capi.FMQ_WM_FINAL)
        self._drain(0, np.empty(0, np.uint64))
        self._engine.close()
        self._engine = None


def capi_sorted_events(r):
    o = np.lexsort((r.key, r.ts))
    return capi.EventRows(r.ts[o], r.key[o], r.val[o])
