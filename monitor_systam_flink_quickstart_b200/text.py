"""Host-side record parsing and key dictionary (the `map` functions in front of the hot path).

Mirrors what the reference does immediately before `keyBy`:
  * ``"<epoch-s> <host> <cpu> <usage>"``  -> (host, cpu, usage)      chapter1/.../Main.java:21-25,
    chapter2/.../ComputeCpuAvg.java:22-25
  * ``"<ISO-8601 local time> <channel> <flow>"`` -> (time, channel, flow) with the timestamp
    ``(int) LocalDateTime.parse(..).toEpochSecond(ZoneOffset.ofHours(8)) * 1000L``
    chapter3/.../BandwidthMonitorWithEventTime.java:33-34,39-44
String keys are dictionary-encoded to the uint64 ids the C ABI carries; the dictionary decodes
them again on emit.  This is plain host code (SURVEY.md §8f N2 moves it onto the device later).
"""
from __future__ import annotations

import calendar
import time
from typing import Dict, List, Tuple


class KeyDictionary:
    """String key <-> dense uint64 id (what the Java shim keeps next to the native handle)."""

    def __init__(self):
        self._ids: Dict[str, int] = {}
        self._names: List[str] = []

    def encode(self, name: str) -> int:
        i = self._ids.get(name)
        if i is None:
            i = len(self._names)
            self._ids[name] = i
            self._names.append(name)
        return i

    def decode(self, i: int) -> str:
        return self._names[int(i)]

    def __len__(self):
        return len(self._names)


def _to_int32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x >= (1 << 31) else x


def iso_local_to_epoch_ms(s: str, utc_offset_hours: int = 8) -> int:
    """``(int) LocalDateTime.parse(s).toEpochSecond(ZoneOffset.ofHours(h)) * 1000L`` including the
    reference's 32-bit cast (wraps in 2038)."""
    t = time.strptime(s, "%Y-%m-%dT%H:%M:%S")
    sec = calendar.timegm(t) - utc_offset_hours * 3600
    return _to_int32(sec) * 1000


def parse_cpu_line(line: str) -> Tuple[int, str, str, float]:
    """-> (epoch seconds, host, cpu, usage).  Main.java:21-25 (the timestamp field is ignored there)."""
    items = line.split(" ")
    return int(items[0]), items[1], items[2], float(items[3])


def parse_flow_line(line: str) -> Tuple[int, str, int]:
    """-> (event time ms, channel, flow bytes).  BandwidthMonitorWithEventTime.java:33-44."""
    items = line.split(" ")
    return iso_local_to_epoch_ms(items[0]), items[1], int(items[2])


def text_key_hash(name: str) -> int:
    """Host restatement of ``fmq_text_key_hash`` (FNV-1a over the UTF-8 bytes, finished with mix64):
    the key id the device text parser assigns to a host / channel string."""
    m = (1 << 64) - 1
    h = 0xCBF29CE484222325
    for b in name.encode("utf-8"):
        h ^= b
        h = (h * 0x100000001B3) & m
    h ^= h >> 30; h = (h * 0xBF58476D1CE4E5B9) & m
    h ^= h >> 27; h = (h * 0x94D049BB133111EB) & m
    h ^= h >> 31
    return 0 if h == m else h
