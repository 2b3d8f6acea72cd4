"""Counter-based synthetic monitoring streams.

Event ``i`` is a pure function of ``(seed, i)`` through splitmix64, so the host (numpy) and the
device (torch int64 arithmetic on ``cuda``) generate bit-identical columns without ever shipping
them: ``ts[i]:int64 ms``, ``key[i]:uint64``, ``val[i]:f64|i64`` — 24 B per event, the record layout
at the C-ABI boundary.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

_M1, _M2, _G = 0xBF58476D1CE4E5B9, 0x94D049BB133111EB, 0x9E3779B97F4A7C15
_U64 = (1 << 64) - 1


def _s64(x: int) -> int:
    """Two's-complement reinterpretation of a u64 constant as int64 (for torch)."""
    x &= _U64
    return x - (1 << 64) if x >= (1 << 63) else x


@dataclass
class StreamSpec:
    """Describes one synthetic workload.  Times are in ms."""
    n_keys: int = 1_000_000
    events_per_ms: int = 1000            # base time of event i is t0 + i // events_per_ms
    t0: int = 1_566_957_600_000          # 2019-08-28T10:00:00+08:00, the reference's example epoch
    jitter_ms: int = 5000                # ts = base - U[0, jitter]
    late_frac_ppm: int = 1000            # this many events per million get an extra delay ...
    late_extra_ms: int = 25000           # ... of U[jitter, jitter + late_extra]
    value_type: str = "f64"              # "f64": U[0,100);  "i64": U[0, 2^20) flow bytes
    key_kind: str = "uniform"            # "uniform" | "host_metric" | "zipf"
    metrics_per_host: int = 8
    zipf_s: float = 1.2
    seed: int = 42
    _zipf_cdf: object = field(default=None, repr=False, compare=False)


def _splitmix_np(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(_G)).astype(np.uint64)
    x ^= x >> np.uint64(30); x *= np.uint64(_M1)
    x ^= x >> np.uint64(27); x *= np.uint64(_M2)
    x ^= x >> np.uint64(31)
    return x


def _lsr_t(x, s):
    """Logical shift right on torch int64."""
    return (x >> s) & ((1 << (64 - s)) - 1)


def _splitmix_t(x):
    x = x + _s64(_G)
    x = x ^ _lsr_t(x, 30); x = x * _s64(_M1)
    x = x ^ _lsr_t(x, 27); x = x * _s64(_M2)
    x = x ^ _lsr_t(x, 31)
    return x


def _zipf_table(spec: StreamSpec, size: int = 1 << 16) -> np.ndarray:
    """Inverse-CDF lookup for Zipf(s) over n_keys ids: ``size`` quantiles -> rank (uint64).
    Coarse by design: both host and device use the same table, so parity is unaffected."""
    if spec._zipf_cdf is None:
        # continuous approximation of the Zipf CDF: F(k) ~ (H(k)) / H(n), H(k) ~ integral form
        s, n = spec.zipf_s, spec.n_keys
        u = (np.arange(size, dtype=np.float64) + 0.5) / size
        # exact head (first 4096 ranks), power-law tail
        head = min(n, 4096)
        w = 1.0 / np.arange(1, head + 1, dtype=np.float64) ** s
        tail_mass = ((head + 0.5) ** (1 - s) - (n + 0.5) ** (1 - s)) / (s - 1) if n > head else 0.0
        total = w.sum() + tail_mass
        cdf_head = np.cumsum(w) / total
        ranks = np.empty(size, np.uint64)
        in_head = u <= cdf_head[-1]
        ranks[in_head] = np.searchsorted(cdf_head, u[in_head]).astype(np.uint64)
        if (~in_head).any():
            rem = (u[~in_head] - cdf_head[-1]) * total           # mass into the tail
            x = ((head + 0.5) ** (1 - s) - rem * (s - 1)) ** (1.0 / (1 - s))
            ranks[~in_head] = np.clip(np.floor(x), head, n - 1).astype(np.uint64)
        spec._zipf_cdf = ranks
    return spec._zipf_cdf


def _zipf_next(spec: StreamSpec) -> np.ndarray:
    """First rank of the following quantile (n_keys after the last one)."""
    tab = _zipf_table(spec)
    return np.append(tab[1:], np.uint64(spec.n_keys))


def generate_numpy(spec: StreamSpec, start: int, n: int):
    """Events [start, start+n) on the host.  Returns (ts:int64, key:uint64, val:f64|i64)."""
    i = np.arange(start, start + n, dtype=np.uint64)
    seed = np.uint64(spec.seed)
    h1 = _splitmix_np(i ^ seed)
    h2 = _splitmix_np(h1)
    h3 = _splitmix_np(h2)
    # --- key ---
    if spec.key_kind == "uniform":
        key = h1 % np.uint64(spec.n_keys)
    elif spec.key_kind == "host_metric":
        hosts = np.uint64(spec.n_keys // spec.metrics_per_host)
        key = ((h1 % hosts) << np.uint64(32)) + ((h1 >> np.uint64(40)) % np.uint64(spec.metrics_per_host))
    elif spec.key_kind == "zipf":
        tab = _zipf_table(spec)
        q = (h1 >> np.uint64(48)).astype(np.int64)
        # a quantile of the tail covers many ranks: spread the draw over all of them, so that a stream
        # over 10^8 ids really touches millions of distinct ids (the head quantiles are single ranks)
        span = np.maximum(_zipf_next(spec)[q] - tab[q], np.uint64(1))
        rank = tab[q] + (h1 & np.uint64((1 << 48) - 1)) % span
        # spread ranks over the id space so hot keys are not adjacent ids
        key = (rank * np.uint64(0x9E3779B1)) % np.uint64(spec.n_keys)
    else:
        raise ValueError(spec.key_kind)
    # --- timestamp: base - jitter, a small tail beyond the out-of-orderness bound ---
    base = np.int64(spec.t0) + (i // np.uint64(spec.events_per_ms)).astype(np.int64)
    jit = (h2 % np.uint64(spec.jitter_ms + 1)).astype(np.int64) if spec.jitter_ms > 0 else np.zeros(n, np.int64)
    if spec.late_frac_ppm > 0:
        is_late = ((h2 >> np.uint64(32)) % np.uint64(1_000_000)) < np.uint64(spec.late_frac_ppm)
        extra = (h3 % np.uint64(spec.late_extra_ms + 1)).astype(np.int64) + np.int64(spec.jitter_ms)
        jit = np.where(is_late, extra, jit)
    ts = np.maximum(base - jit, 0)
    # --- value ---
    if spec.value_type == "f64":
        val = (h3 >> np.uint64(11)).astype(np.float64) * (100.0 / 9007199254740992.0)
    else:
        val = ((h3 >> np.uint64(20)) & np.uint64((1 << 20) - 1)).astype(np.int64)
    return ts, key.astype(np.uint64), val


def generate_torch(spec: StreamSpec, start: int, n: int, device="cuda"):
    """Same events as :func:`generate_numpy`, produced on ``device`` with int64 arithmetic.
    ``key`` is returned as an int64 tensor holding the uint64 bit pattern."""
    import torch
    i = torch.arange(start, start + n, dtype=torch.int64, device=device)
    h1 = _splitmix_t(i ^ spec.seed)
    h2 = _splitmix_t(h1)
    h3 = _splitmix_t(h2)

    def umod(x, m):   # unsigned x mod m for int64-held u64 (m < 2^62)
        q = _lsr_t(x, 1) % m
        return (q * 2 + (x & 1)) % m
    if spec.key_kind == "uniform":
        key = umod(h1, spec.n_keys)
    elif spec.key_kind == "host_metric":
        hosts = spec.n_keys // spec.metrics_per_host
        key = (umod(h1, hosts) << 32) + (_lsr_t(h1, 40) % spec.metrics_per_host)
    elif spec.key_kind == "zipf":
        tab = torch.from_numpy(_zipf_table(spec).astype(np.int64)).to(device)
        nxt = torch.from_numpy(_zipf_next(spec).astype(np.int64)).to(device)
        q = _lsr_t(h1, 48)
        span = torch.clamp(nxt[q] - tab[q], min=1)
        rank = tab[q] + (h1 & ((1 << 48) - 1)) % span
        key = (rank * 0x9E3779B1) % spec.n_keys
    else:
        raise ValueError(spec.key_kind)
    base = spec.t0 + i // spec.events_per_ms
    jit = umod(h2, spec.jitter_ms + 1) if spec.jitter_ms > 0 else torch.zeros_like(i)
    if spec.late_frac_ppm > 0:
        is_late = (_lsr_t(h2, 32) % 1_000_000) < spec.late_frac_ppm
        extra = umod(h3, spec.late_extra_ms + 1) + spec.jitter_ms
        jit = torch.where(is_late, extra, jit)
    ts = torch.clamp(base - jit, min=0)
    if spec.value_type == "f64":
        val = _lsr_t(h3, 11).to(torch.float64) * (100.0 / 9007199254740992.0)
    else:
        val = _lsr_t(h3, 20) & ((1 << 20) - 1)
    return ts, key, val
