"""TEST INFRASTRUCTURE — NOT PRODUCT CODE.

Second, independent restatement of the Flink 1.8.0 semantics on the hot path, in plain Python
(dicts + heapq), used to cross-check ``oracle/flink_literal.cpp`` on small random streams
(hypothesis) and to replay the reference's README sessions G1-G8.  Written
separately from the C++ oracle on purpose: the two share no code.

Follows, per record: SlidingEventTimeWindows.assignWindows -> WindowOperator.isWindowLate ->
Heap{Reducing,Aggregating}State.add -> EventTimeTrigger timer; per watermark:
InternalTimerServiceImpl.advanceWatermark, with the user
lambdas of chapter3/.../BandwidthMonitorWithEventTime.java:47-55, chapter2/.../ComputeCpuAvg.java:39-50,
chapter2/.../ComputeCpuMax.java:26 and chapter1/.../Main.java:31.
"""
from __future__ import annotations

import heapq
import math
import struct

I64_MIN, I64_MAX = -(2 ** 63), 2 ** 63 - 1


def _wrap64(x: int) -> int:
    x &= (1 << 64) - 1
    return x - (1 << 64) if x >= (1 << 63) else x


def _java_mod(a: int, b: int) -> int:
    """Java's truncating % (sign follows the dividend)."""
    r = abs(a) % abs(b)
    return r if a >= 0 else -r


def _dbits(d: float) -> int:
    if d != d:
        return 0x7FF8000000000000
    return struct.unpack("<q", struct.pack("<d", d))[0]


def _compare_to(a, b) -> int:
    """Double.compareTo / Long.compareTo."""
    if a < b:
        return -1
    if a > b:
        return 1
    if isinstance(a, float) or isinstance(b, float):
        x, y = _dbits(float(a)), _dbits(float(b))
        return 0 if x == y else (-1 if x < y else 1)
    return 0


def _jmax(a, b):
    if isinstance(a, float):
        if a != a:
            return a
        if b != b:
            return b
        if a == 0.0 and b == 0.0:
            return b if math.copysign(1.0, a) < 0 else a
    return a if a >= b else b


def _jmin(a, b):
    if isinstance(a, float):
        if a != a:
            return a
        if b != b:
            return b
        if a == 0.0 and b == 0.0:
            return a if math.copysign(1.0, a) < 0 else b
    return a if a <= b else b


_CMP = {0: lambda v, t: v > t, 1: lambda v, t: v < t, 2: lambda v, t: v >= t, 3: lambda v, t: v <= t}


class NaiveFlinkModel:
    """cfg: dict with the fmq_config field names (missing = 0)."""

    def __init__(self, **cfg):
        g = cfg.get
        self.is_f64 = g("value_type", 0) == 0
        self.window = bool(g("window_enabled", 0))
        self.size, self.slide, self.off = g("window_size_ms", 0), g("window_slide_ms", 0), g("window_offset_ms", 0)
        self.bound = g("max_out_of_orderness_ms", 0)
        self.lateness = g("allowed_lateness_ms", 0)
        self.cleanup = []          # heap of (maxTimestamp + lateness, key, window_start)
        self.refire = set()        # (key, window_start) that took a record after having fired, this batch
        self.emit_fired = g("emit_fired", 1)
        self.wsrc, self.wop, self.wthr = g("walert_source", 0), g("walert_op", 0), g("walert_threshold", 0.0)
        self.scale = [g(k, 0.0) or 1.0 for k in ("walert_mul", "walert_div0", "walert_div1", "walert_div2")]
        self.ef, self.efop, self.efthr = g("efilter_enabled", 0), g("efilter_op", 0), g("efilter_threshold", 0.0)
        self.rmax = g("rolling_max_enabled", 0)
        self.cn, self.cop, self.cthr = g("consec_n", 0), g("consec_op", 0), g("consec_threshold", 0.0)
        self.creset = g("consec_reset_on_fire", 0)
        self.wm = I64_MIN
        self.cur_max_ts = I64_MIN + self.bound
        self.last_wm = I64_MIN
        self.state = {}            # (key, window_start) -> [count, sum, min, max]
        self.timers = []           # heap of (ts, key, window_start)
        self.timer_set = set()
        self.kmax, self.kcons = {}, {}
        self.fired, self.walerts, self.ealerts, self.calerts, self.rolling = [], [], [], [], []
        self.late_dropped = self.events_in = self.windows_fired = 0

    # ---- per element -------------------------------------------------------------------------
    def _element(self, ts, key, val):
        self.events_in += 1
        if self.ef and _CMP[self.efop](float(val), self.efthr):
            self.ealerts.append((ts, key, val))
        if self.rmax:
            cur = self.kmax.get(key)
            if cur is None or not (_compare_to(cur, val) > 0):
                cur = val
            self.kmax[key] = cur
            self.rolling.append(cur)
        if self.cn > 0:
            over = _CMP[self.cop](float(val), self.cthr)
            c = self.kcons.get(key, 0) + 1 if over else 0
            if (c == self.cn) if self.creset else (c >= self.cn):
                self.calerts.append((ts, key, val))
                if self.creset:
                    c = 0
            self.kcons[key] = c
        if not self.window:
            return
        self.cur_max_ts = max(self.cur_max_ts, ts)
        last_start = ts - _java_mod(ts - self.off + self.slide, self.slide)
        skipped = True
        s = last_start
        while s > ts - self.size:
            if not (s + self.size - 1 + self.lateness <= self.wm):
                skipped = False
                a = self.state.get((key, s))
                if a is None:
                    self.state[(key, s)] = [1, val, val, val]
                else:
                    a[0] += 1
                    a[1] = a[1] + val if self.is_f64 else _wrap64(a[1] + val)
                    a[2], a[3] = _jmin(a[2], val), _jmax(a[3], val)
                t = (s + self.size - 1, key, s)
                if t[0] <= self.wm:
                    self.refire.add((key, s))          # fired already, still kept: fires again at the end of this batch
                elif t not in self.timer_set:
                    self.timer_set.add(t)
                    heapq.heappush(self.timers, t)
                    if self.lateness > 0:
                        heapq.heappush(self.cleanup, (t[0] + self.lateness, key, s))
            s -= self.slide
        if skipped and ts + self.lateness <= self.wm:
            self.late_dropped += 1

    def _alert_value(self, a):
        cnt, sm, mn, mx = a
        if self.wsrc == 2:
            return 0.0 if cnt == 0 else float(sm) / float(cnt)
        src = {1: sm, 3: mn, 4: mx, 5: cnt}[self.wsrc]
        v = float(src) * self.scale[0]
        for d in self.scale[1:]:
            v = v / d
        return v

    def _emit(self, key, s, a):
        self.windows_fired += 1
        wend = s + self.size
        if self.emit_fired:
            self.fired.append((key, wend, a[0], a[1], a[2], a[3]))
        if self.wsrc:
            v = self._alert_value(a)
            if _CMP[self.wop](v, self.wthr):
                self.walerts.append((key, wend, v))

    def _refire_pending(self):
        for key, s in sorted(self.refire):
            a = self.state.get((key, s))
            if a is not None:
                self._emit(key, s, a)
        self.refire = set()

    def _advance(self, wm):
        if wm <= self.wm:
            return
        self.wm = wm
        while self.timers and self.timers[0][0] <= wm:
            t = heapq.heappop(self.timers)
            self.timer_set.discard(t)
            _, key, s = t
            a = self.state.get((key, s)) if self.lateness > 0 else self.state.pop((key, s), None)
            if a is None:
                continue
            self._emit(key, s, a)
        while self.cleanup and self.cleanup[0][0] <= wm:
            _, key, s = heapq.heappop(self.cleanup)
            self.state.pop((key, s), None)

    def push_batch(self, ts, key, val, watermark_after=I64_MIN):
        self.rolling = []
        for t, k, v in zip(ts, key, val):
            self._element(int(t), int(k), float(v) if self.is_f64 else int(v))
        if self.window:
            self._refire_pending()
            if watermark_after == I64_MIN + 1:       # FMQ_WM_KEEP: no watermark follows this batch
                return
            if watermark_after == I64_MIN:
                pot = self.cur_max_ts - self.bound
                if pot >= self.last_wm:
                    self.last_wm = pot
                watermark_after = self.last_wm
            self._advance(watermark_after)

    def flush(self, wm=I64_MAX):
        if self.window:
            self._advance(wm)


# ---- keyBy routing (MathUtils.murmurHash, KeyGroupRangeAssignment; pinned by golden G8) --------
def java_string_hash(s: str) -> int:
    h = 0
    for ch in s:
        h = (31 * h + ord(ch)) & 0xFFFFFFFF
    return h - (1 << 32) if h >= (1 << 31) else h


def murmur_hash(code: int) -> int:
    m = 0xFFFFFFFF
    code &= m
    code = (code * 0xCC9E2D51) & m
    code = ((code << 15) | (code >> 17)) & m
    code = (code * 0x1B873593) & m
    code = ((code << 13) | (code >> 19)) & m
    code = (code * 5 + 0xE6546B64) & m
    code ^= 4
    code ^= code >> 16
    code = (code * 0x85EBCA6B) & m
    code ^= code >> 13
    code = (code * 0xC2B2AE35) & m
    code ^= code >> 16
    s = code - (1 << 32) if code >= (1 << 31) else code
    if s >= 0:
        return s
    return -s if s != -(1 << 31) else 0


def key_group(key_hash: int, max_parallelism: int = 128) -> int:
    return murmur_hash(key_hash) % max_parallelism


def subtask(kg: int, parallelism: int, max_parallelism: int = 128) -> int:
    return kg * parallelism // max_parallelism
