// Device-side building blocks shared by all kernels of libfmq_b200 (sm_100a only).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace fmq {

typedef unsigned long long u64;
typedef long long i64;
typedef unsigned int u32;

#define FMQ_MAX_PEERS 8
constexpr u64 KEY_EMPTY = 0xFFFFFFFFFFFFFFFFull;
constexpr u32 SLOT_INVALID = 0xFFFFFFFFu;
constexpr i64 I64_MIN_ = (i64)0x8000000000000000ull;
constexpr i64 I64_MAX_ = (i64)0x7FFFFFFFFFFFFFFFull;

// sticky error bits kept in DevState::err_flags
constexpr u32 ERRB_TABLE_FULL = 1u, ERRB_RING = 2u, ERRB_OUTPUT = 4u, ERRB_BAD_INPUT = 8u;

// One accumulator cell per (ring pane, key slot): 32 B = exactly one L2 sector, so the add pair
// {cnt,sum} and the max pair {mx,mn} of one event each travel as ONE sector request when issued
// from adjacent lanes of the same RED instruction (measured: tools/micro_atomics.cu).
// All-zero bytes are the identity of every field.
struct __align__(32) Acc32 {
    u64 cnt;   // F64 jobs: count kept as a double (exact to 2^53) so cnt and sum share red.add.f64
    u64 sum;   // double bits or int64 (wraps like Java long)
    u64 mx;    // order-preserving encoding, see enc_max
    u64 mn;    // complement of the encoding, so min is also a red.max.u64
};

// Per-handle device-resident scalar state.  Everything a kernel needs to know about "where the
// stream is" lives here, so a push never needs a host round trip.
struct DevState {
    // event-time position
    i64 wm;                 // current watermark (INT64_MIN before the first one)
    i64 cur_max_ts;         // BoundedOutOfOrdernessTimestampExtractor.currentMaxTimestamp
    i64 last_emitted_wm;    // .lastEmittedWatermark
    i64 batch_max_ts;       // max ts of the batch being ingested (reset by plan)
    // ingest parameters derived from wm by the plan kernel
    i64 origin_ts;          // records with ts < origin_ts are late (all their windows have fired)
    i64 origin_pane;        // absolute pane index whose start is origin_ts
    int anchored;           // 0 while wm == INT64_MIN (nothing can be late)
    int pad0;
    i64 min_pane, max_pane; // absolute pane range currently holding data (min > max: none)
    // fire plan for the current watermark advance
    i64 fire_first_end;     // exclusive end of the first window to fire
    i64 fire_count;         // number of consecutive window ends (step = slide)
    i64 expire_lo, expire_hi; // absolute panes to clear after firing (lo > hi: none)
    // counters
    u64 events_in, late_dropped, windows_fired, window_alerts, event_alerts, consec_alerts, keys;
    // output cursors (rows appended since the last full drain)
    u64 n_fired, n_walert, n_ealert, n_calert;
    // skew adaptation: one warp per block samples how many lanes of a warp hit the same cell; the plan
    // kernel turns warp-level pre-aggregation on for the next batch when that is common
    u64 dup_sampled, dup_found; u32 combine_dups;
    u32 pending_clear;               // panes whose clear is still in flight while the next batch is ingested (overlapped combiner mode)
    u64 n_late;                      // rows in the late side output
    u64 n_dirty;                     // allowedLateness: (key, pane) cells that took a record after one of their windows had fired, this batch
    u64 n_mlog, n_mwork, n_median;   // median path: logged records, records selected for the window being fired, emitted rows
    u32 err_flags;
    u32 pad1;
};

__host__ __device__ __forceinline__ u64 mix64(u64 x) {
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31; return x;
}

// ---- order-preserving encodings -------------------------------------------------------------
// f64: total order  -NaN.. < -inf < .. < -0.0 < +0.0 < .. < +inf ; NaN is forced to the top for
// BOTH max and min so that it propagates like java.lang.Math.max/min.  i64: bias to unsigned.
__device__ __forceinline__ u64 order_f64(double v) {
    u64 b = (u64)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double unorder_f64(u64 u) {
    u64 b = (u >> 63) ? (u ^ 0x8000000000000000ull) : ~u;
    return __longlong_as_double((i64)b);
}
template <typename V> struct IsF64 { static constexpr bool value = false; };
template <> struct IsF64<double> { static constexpr bool value = true; };
template <typename V> struct Codec;
template <> struct Codec<double> {
    static __device__ __forceinline__ u64 enc_max(double v) { return (v != v) ? ~0ull : order_f64(v); }
    static __device__ __forceinline__ u64 enc_min(double v) { return (v != v) ? ~0ull : ~order_f64(v); }
    static __device__ __forceinline__ u64 dec_max(u64 u) { return (u64)__double_as_longlong(unorder_f64(u)); }
    static __device__ __forceinline__ u64 dec_min(u64 u) { return u == ~0ull ? 0x7FFFFFFFFFFFFFFFull : (u64)__double_as_longlong(unorder_f64(~u)); }
    static __device__ __forceinline__ double as_double(u64 bits) { return __longlong_as_double((i64)bits); }
    static __device__ __forceinline__ double from_bits(u64 bits) { return __longlong_as_double((i64)bits); }
    // Double.compareTo order (NaN canonical and greatest, -0.0 < +0.0): used by rolling max
    static __device__ __forceinline__ u64 cmp_key(double v) { return (v != v) ? ~0ull : order_f64(v); }
};
template <> struct Codec<i64> {
    static __device__ __forceinline__ u64 enc_max(i64 v) { return (u64)v ^ 0x8000000000000000ull; }
    static __device__ __forceinline__ u64 enc_min(i64 v) { return ~((u64)v ^ 0x8000000000000000ull); }
    static __device__ __forceinline__ u64 dec_max(u64 u) { return u ^ 0x8000000000000000ull; }
    static __device__ __forceinline__ u64 dec_min(u64 u) { return (~u) ^ 0x8000000000000000ull; }
    static __device__ __forceinline__ double as_double(u64 bits) { return (double)(i64)bits; }
    static __device__ __forceinline__ i64 from_bits(u64 bits) { return (i64)bits; }
    static __device__ __forceinline__ u64 cmp_key(i64 v) { return (u64)v ^ 0x8000000000000000ull; }
};

__device__ __forceinline__ bool cmp_op(double v, int op, double thr) {
    switch (op) { case 0: return v > thr; case 1: return v < thr; case 2: return v >= thr; default: return v <= thr; }
}

// ---- streaming loads: the event columns are read exactly once ---------------------------------
// sm_100 has 256-bit global loads; only they accept the direct .L2::evict_first qualifier, which
// keeps the one-pass event stream from pushing the L2-resident accumulator table out.
__device__ __forceinline__ ulonglong4 ldg_stream_v4(const void* p) {
    ulonglong4 r;
    asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ u64 ldg_stream(const void* p) {
    u64 r;
    asm volatile("ld.global.nc.L1::no_allocate.u64 %0, [%1];" : "=l"(r) : "l"(p));
    return r;
}

// ---- TMA bulk copies (cp.async.bulk, 1-D) completing on an mbarrier ---------------------------------
__device__ __forceinline__ u32 smem_u32(const void* p) { return (u32)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64* bar, u32 count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64* bar, u32 bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64* bar, u32 parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}"
        ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA, SASS UBLKCP), byte count signalled on `bar`, L2 evict-first
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, u32 bytes, u64* bar, u64 policy) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}

// ---- accumulator updates: REDs that ask the L2 to keep their sector -------------------------------
// The table is what should stay in the L2 while 400 MB of events stream through it per micro-batch:
// evict_last on the REDs (and on the directory buckets) beside evict_first on the stream took the dense
// ingest kernel from 0.290 to 0.272 ms per 2^24 events (A/B on the GPU, profiles/r01_notes.md).
__device__ __forceinline__ u64 l2_keep_policy() {
    u64 p; asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p)); return p;
}
__device__ __forceinline__ void red_add_keep(double* p, double v) {
    asm volatile("red.global.add.L2::cache_hint.f64 [%0], %1, %2;" :: "l"(p), "d"(v), "l"(l2_keep_policy()) : "memory");
}
__device__ __forceinline__ void red_add_keep(u64* p, u64 v) {
    asm volatile("red.global.add.L2::cache_hint.u64 [%0], %1, %2;" :: "l"(p), "l"(v), "l"(l2_keep_policy()) : "memory");
}
__device__ __forceinline__ void red_max_keep(u64* p, u64 v) {
    asm volatile("red.global.max.L2::cache_hint.u64 [%0], %1, %2;" :: "l"(p), "l"(v), "l"(l2_keep_policy()) : "memory");
}

// ---- key directory: bucketized open addressing, keys are never removed ---------------------------
// The directory is an array of 32-byte buckets of 4 keys; slot = 4 * bucket + position.  One
// 256-bit load fetches a whole bucket in a single L2 sector request, so a lookup is normally ONE
// dependent memory access (plain per-key linear probing at the same load factor cost ~11
// serialised probes for the slowest lane of a warp — ncu, profiles/r01_notes.md).
// home bucket = mulhi(mix64(key), nbuckets); `cap` (slots) is any multiple of 4.
// Stale data cannot mislead the probe: a slot only ever changes EMPTY -> key, and the CAS that
// claims it is resolved in L2 and returns the winner.  Buckets fill left to right and all
// inserters of one key walk the same sequence, so a key can never be entered twice.
__device__ __forceinline__ ulonglong4 ld_bucket(const u64* bucket) {
    ulonglong4 r;
    asm volatile("ld.global.L1::no_allocate.L2::evict_last.v4.b64 {%0, %1, %2, %3}, [%4];"
                 : "=l"(r.x), "=l"(r.y), "=l"(r.z), "=l"(r.w) : "l"(bucket) : "memory");
    return r;
}
__device__ __forceinline__ u64 dir_home_bucket(u64 key, u64 nbuckets) { return __umul64hi(mix64(key), nbuckets); }

// branch-free search of one loaded bucket: position 0..3 of `key`, or -1
__device__ __forceinline__ int bucket_match(const ulonglong4& kv, u64 key) {
    int j = -1;
    j = (kv.w == key) ? 3 : j; j = (kv.z == key) ? 2 : j; j = (kv.y == key) ? 1 : j; j = (kv.x == key) ? 0 : j;
    return j;
}

__device__ __noinline__ u32 dir_resolve(u64* __restrict__ dkeys, u64 nbuckets, u64 key, u64 b, ulonglong4 kv, DevState* st);
// The lookup inlined into the ingest loop: the key is in a bucket of its probe sequence (every event of a running
// job).  A bucket with a free slot means first sight of the key: the claim (a compare-and-swap ladder) lives out of
// line — ncu of the hashed ingest kernel had `no_instruction` as its top stall (3 per issue): four inlined copies of
// that ladder per tile made the streaming loop larger than the instruction cache.
__device__ __forceinline__ u32 dir_resolve_inline(u64* __restrict__ dkeys, u64 nbuckets, u64 key, u64 b, ulonglong4 kv, DevState* st) {
#pragma unroll 1
    for (u64 probe = 0; probe < nbuckets; ++probe) {
        const int j = bucket_match(kv, key);
        if (j >= 0) return (u32)(b * 4 + j);
        if (kv.x == KEY_EMPTY || kv.y == KEY_EMPTY || kv.z == KEY_EMPTY || kv.w == KEY_EMPTY) return dir_resolve(dkeys, nbuckets, key, b, kv, st);
        b = (b + 1 == nbuckets) ? 0 : b + 1;
        kv = ld_bucket(dkeys + b * 4);
    }
    atomicOr(&st->err_flags, ERRB_TABLE_FULL);
    return SLOT_INVALID;
}

// Full lookup-or-insert starting at bucket `b` whose contents `kv` are already loaded (slow path:
// first sight of a key, or a key that overflowed its home bucket).
__device__ __noinline__ u32 dir_resolve(u64* __restrict__ dkeys, u64 nbuckets, u64 key, u64 b, ulonglong4 kv, DevState* st) {
    for (u64 probe = 0; probe < nbuckets; ++probe) {
        const u64 k4[4] = {kv.x, kv.y, kv.z, kv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            u64 k = k4[j];
            if (k == KEY_EMPTY) {
                k = atomicCAS(dkeys + b * 4 + j, KEY_EMPTY, key);
                if (k == KEY_EMPTY) { atomicAdd(&st->keys, 1ull); return (u32)(b * 4 + j); }
            }
            if (k == key) return (u32)(b * 4 + j);
        }
        b = (b + 1 == nbuckets) ? 0 : b + 1;
        kv = ld_bucket(dkeys + b * 4);
    }
    atomicOr(&st->err_flags, ERRB_TABLE_FULL);
    return SLOT_INVALID;
}
__device__ __forceinline__ u32 dir_find_or_insert(u64* __restrict__ dkeys, u64 cap, u64 key, DevState* st) {
    const u64 nb = cap >> 2, b = dir_home_bucket(key, nb);
    return dir_resolve(dkeys, nb, key, b, ld_bucket(dkeys + b * 4), st);
}
__device__ __forceinline__ u32 dir_find(const u64* __restrict__ dkeys, u64 cap, u64 key) {
    const u64 nb = cap >> 2; u64 b = dir_home_bucket(key, nb);
    for (u64 probe = 0; probe < nb; ++probe) {
        const ulonglong4 kv = ld_bucket(dkeys + b * 4);
        const u64 k4[4] = {kv.x, kv.y, kv.z, kv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { if (k4[j] == key) return (u32)(b * 4 + j); if (k4[j] == KEY_EMPTY) return SLOT_INVALID; }
        b = (b + 1 == nb) ? 0 : b + 1;
    }
    return SLOT_INVALID;
}

// ---- warp-aggregated append: one atomic per warp, rows of a warp land contiguously --------------
// Must be called by all 32 lanes.  Returns the row index for lanes with pred, or ~0 when the
// buffer is full (and raises the sticky overflow flag).
__device__ __forceinline__ u64 warp_append(bool pred, u64* counter, u64 cap, DevState* st) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return ~0ull;
    unsigned lane = threadIdx.x & 31;
    int leader = __ffs(m) - 1;
    u64 base = 0;
    if ((int)lane == leader) base = atomicAdd(counter, (u64)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (!pred) return ~0ull;
    u64 row = base + __popc(m & ((1u << lane) - 1));
    if (row >= cap) { atomicOr(&st->err_flags, ERRB_OUTPUT); return ~0ull; }
    return row;
}

// ---- block-aggregated append: ONE atomic per block per call --------------------------------------
// Same-address atomics with a return value serialise in a single L2 slice (a few ns each), so a
// reservation per warp per event caps a streaming kernel at a few Gev/s (measured: 8 Gev/s).  Here
// the block scans its per-thread row counts and reserves once.  Must be called by every thread of
// the block (THREADS = blockDim.x); returns the first row index of this thread's `cnt` rows, which
// are consecutive and in thread order, or ~0 if the buffer cannot hold the block's rows.
template <int THREADS>
__device__ __forceinline__ u64 block_append(u32 cnt, u64* counter, u64 cap, DevState* st) {
    __shared__ u32 wsum[THREADS / 32];
    __shared__ u64 base_s;
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    u32 inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += t; }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    u32 wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) { const u32 x = wsum[i]; if (i < (int)w) wbase += x; total += x; }
    if (threadIdx.x == 0) base_s = total ? atomicAdd(counter, (u64)total) : 0;
    __syncthreads();
    const u64 base = base_s;
    if (total == 0) return ~0ull;
    if (base + total > cap) { if (threadIdx.x == 0) atomicOr(&st->err_flags, ERRB_OUTPUT); return ~0ull; }
    return base + wbase + (inc - cnt);
}

// Two appends (each thread contributes 0 or 1 row to each) behind ONE pair of block barriers: the two
// counts travel through the scan packed in one word, two threads reserve concurrently.
template <int THREADS>
__device__ __forceinline__ void block_append2(bool pa, bool pb, u64* ctr_a, u64 cap_a, u64* ctr_b, u64 cap_b, DevState* st, u64& row_a, u64& row_b) {
    __shared__ u32 wsum2[THREADS / 32];
    __shared__ u64 base2[2];
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const u32 cnt = (pa ? 1u : 0u) | (pb ? 0x10000u : 0u);
    u32 inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += t; }
    if (lane == 31) wsum2[w] = inc;
    __syncthreads();
    u32 wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) { const u32 x = wsum2[i]; if (i < (int)w) wbase += x; total += x; }
    const u32 ta = total & 0xFFFFu, tb = total >> 16;
    if (threadIdx.x == 0) base2[0] = ta ? atomicAdd(ctr_a, (u64)ta) : 0;
    if (threadIdx.x == 32) base2[1] = tb ? atomicAdd(ctr_b, (u64)tb) : 0;
    __syncthreads();
    const u32 ex = wbase + inc - cnt;
    row_a = ~0ull; row_b = ~0ull;
    if (ta) { if (base2[0] + ta > cap_a) { if (threadIdx.x == 0) atomicOr(&st->err_flags, ERRB_OUTPUT); } else if (pa) row_a = base2[0] + (ex & 0xFFFFu); }
    if (tb) { if (base2[1] + tb > cap_b) { if (threadIdx.x == 0) atomicOr(&st->err_flags, ERRB_OUTPUT); } else if (pb) row_b = base2[1] + (ex >> 16); }
}

// Same with small per-thread counts (each <= 255): the thread's rows of list A start at row_a, of list B at row_b.
template <int THREADS>
__device__ __forceinline__ void block_append2n(u32 na, u32 nb, u64* ctr_a, u64 cap_a, u64* ctr_b, u64 cap_b, DevState* st, u64& row_a, u64& row_b) {
    __shared__ u32 wsum2n[THREADS / 32];
    __shared__ u64 base2n[2];
    const u32 lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const u32 cnt = na | (nb << 16);
    u32 inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { u32 t = __shfl_up_sync(0xffffffffu, inc, o); if ((int)lane >= o) inc += t; }
    if (lane == 31) wsum2n[w] = inc;
    __syncthreads();
    u32 wbase = 0, total = 0;
#pragma unroll
    for (int i = 0; i < THREADS / 32; ++i) { const u32 x = wsum2n[i]; if (i < (int)w) wbase += x; total += x; }
    const u32 ta = total & 0xFFFFu, tb = total >> 16;
    if (threadIdx.x == 0) base2n[0] = ta ? atomicAdd(ctr_a, (u64)ta) : 0;
    if (threadIdx.x == 32) base2n[1] = tb ? atomicAdd(ctr_b, (u64)tb) : 0;
    __syncthreads();
    const u32 ex = wbase + inc - cnt;
    row_a = ~0ull; row_b = ~0ull;
    if (ta) { if (base2n[0] + ta > cap_a) { if (threadIdx.x == 0) atomicOr(&st->err_flags, ERRB_OUTPUT); } else if (na) row_a = base2n[0] + (ex & 0xFFFFu); }
    if (tb) { if (base2n[1] + tb > cap_b) { if (threadIdx.x == 0) atomicOr(&st->err_flags, ERRB_OUTPUT); } else if (nb) row_b = base2n[1] + (ex >> 16); }
}

__device__ __forceinline__ i64 warp_max_i64(i64 v) {
    for (int o = 16; o > 0; o >>= 1) { i64 t = __shfl_xor_sync(0xffffffffu, v, o); v = t > v ? t : v; }
    return v;
}
__device__ __forceinline__ i64 warp_min_i64(i64 v) {
    for (int o = 16; o > 0; o >>= 1) { i64 t = __shfl_xor_sync(0xffffffffu, v, o); v = t < v ? t : v; }
    return v;
}
__device__ __forceinline__ u64 warp_sum_u64(u64 v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// floor division for possibly negative numerators, positive divisor
__host__ __device__ __forceinline__ i64 floor_div(i64 a, i64 b) { i64 q = a / b; return (a % b != 0 && a < 0) ? q - 1 : q; }

}  // namespace fmq
