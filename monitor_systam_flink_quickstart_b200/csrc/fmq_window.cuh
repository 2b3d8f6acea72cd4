// Event-time window path: ingest (A1 max-ts, A2 pane assign, A3 late drop, A6-i filter, A5 keyed
// accumulate), plan (A1 watermark, fire/expire schedule) and fire (A5f emit + A6-ii/iii alert +
// cleanup).  Row numbers refer to SURVEY.md §8a.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

struct WindowGeom {
    i64 size, slide, off;       // ms
    u32 ring_mask;              // pane_ring - 1
    u32 agg_mask;
    u64 capacity;               // key slots per pane
    u32 slide32;                // slide if it fits 32 bits, else 0
    u32 panes_per_window;       // size / slide
    i64 lateness;               // allowedLateness in ms (0: a window's state goes when it fires)
    u32* dirty;                 // lateness > 0: one bit per (ring pane, key slot), dirty_words 32-bit words per pane
    u64 dirty_words;
};

// allowedLateness: a record landed in a pane of which at least one window has fired already: that window must fire again.
// Panes with relative index below `fired_panes` (counted from the pane origin) are such panes.
__device__ __forceinline__ void mark_dirty(const WindowGeom& g, DevState* st, u32 ring_idx, u32 slot) {
    const u32 bit = 1u << (slot & 31);
    const u32 old = atomicOr(&g.dirty[(u64)ring_idx * g.dirty_words + (slot >> 5)], bit);
    if (!(old & bit)) atomicAdd(&st->n_dirty, 1ull);
}
// Number of panes, counted from the pane origin, that belong to a window which has fired under the watermark in force
// and is still kept.  Window ends are e = off (mod slide); fired: e <= wm + 1; kept: e > wm + 1 - lateness.  If such ends
// exist, the panes concerned are exactly origin .. (largest fired end's index) - 1: the origin is the first pane whose
// LAST window is still kept, and every pane up to there has a window inside the fired-and-kept range.
__device__ __forceinline__ u32 fired_panes(const WindowGeom& g, const DevState* st) {
    if (g.lateness <= 0 || !st->anchored) return 0u;
    const i64 hi_idx = floor_div(st->wm + 1 - g.off, g.slide);                    // index of the largest end <= wm + 1
    const i64 lo_idx = floor_div(st->wm + 1 - g.lateness - g.off, g.slide) + 1;   // index of the smallest end > wm + 1 - lateness
    if (lo_idx > hi_idx) return 0u;
    const i64 n = (hi_idx - 1) - st->origin_pane + 1;
    return n <= 0 ? 0u : (n > 0x7FFFFFFF ? 0x7FFFFFFFu : (u32)n);
}

struct IngestArgs {
    const i64* ts; const u64* key; const void* val; size_t n;
    // Segmented input (keyBy exchange: one segment per source rank, filled by that rank's scatter kernel):
    // segment s of every column starts seg_stride elements after segment s-1 and holds seg_n[s] events,
    // a count that lives in device memory (n is then the capacity of a segment).  nseg <= 1: one plain batch of n.
    u32 nseg; size_t seg_stride; const u64* seg_n;
    DevState* st;
    u64* dkeys;
    Acc32* acc;
    WindowGeom g;
    int window_enabled;
    int vec32;                  // all three columns are 32-byte aligned
    int efilter_enabled, efilter_op; double efilter_thr;
    i64* ea_ts; u64* ea_key; u64* ea_val; u64 ea_cap;
    // late side output (sideOutputLateData): dropped records are also copied here
    i64* lt_ts; u64* lt_key; u64* lt_val; u64 lt_cap;
    // median path (N3): every accepted record is also appended to a value log (pane, slot, value)
    int mlog_enabled; i64* ml_pane; u32* ml_slot; u64* ml_val; u64 ml_cap;
};

// One accumulator update = (count, sum, max, min) contributions for cell `idx`.
// Two adjacent lanes share one update per round: the even lane writes the first 8 B of a pair, the odd
// lane the second, so each RED instruction carries both halves of a 16 B pair in one 32 B sector
// request.  Round 0 serves the even lane's update, round 1 the odd lane's.  Must be reached by all 32 lanes.
template <typename V>
__device__ __forceinline__ void accumulate_pair_raw(Acc32* __restrict__ acc, bool valid, u64 idx, u64 cntb, u64 sumb, u64 mx, u64 mn, u32 agg_mask) {
    const unsigned lane = threadIdx.x & 31;
    const bool odd = lane & 1;
    // exchange: each lane needs the partner's update as well, to apply the half it is responsible for
    const u64 idx_p = __shfl_xor_sync(0xffffffffu, idx, 1);
    const u64 cnt_p = __shfl_xor_sync(0xffffffffu, cntb, 1), sum_p = __shfl_xor_sync(0xffffffffu, sumb, 1);
    const u64 mx_p = __shfl_xor_sync(0xffffffffu, mx, 1), mn_p = __shfl_xor_sync(0xffffffffu, mn, 1);
    const bool valid_p = __shfl_xor_sync(0xffffffffu, (int)valid, 1) != 0;
    const bool want_mx = agg_mask & 8u, want_mn = agg_mask & 4u;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const bool mine = (r == (int)odd);
        const bool ok = mine ? valid : valid_p;
        if (!ok) continue;
        Acc32* a = acc + (mine ? idx : idx_p);
        const u64 add = odd ? (mine ? sumb : sum_p) : (mine ? cntb : cnt_p);
        if constexpr (IsF64<V>::value) {
            red_add_keep(reinterpret_cast<double*>(odd ? &a->sum : &a->cnt), __longlong_as_double((i64)add));   // red.global.add.f64
        } else {
            red_add_keep(odd ? &a->sum : &a->cnt, add);                                                          // red.global.add.u64 (wraps like a Java long)
        }
        if (odd ? want_mn : want_mx) red_max_keep(odd ? &a->mn : &a->mx, odd ? (mine ? mn : mn_p) : (mine ? mx : mx_p));   // red.global.max.u64
    }
}

// Lean form of the pair-lane update for one event per lane with no combining (uniform keys).
template <typename V>
__device__ __forceinline__ void accumulate_pair_lean(Acc32* __restrict__ acc, bool valid, u64 idx, V val, u32 agg_mask) {
    const unsigned lane = threadIdx.x & 31;
    const bool odd = lane & 1;
    u64 vbits; memcpy(&vbits, &val, 8);
    const u64 idx_p = __shfl_xor_sync(0xffffffffu, idx, 1);
    const u64 vb_p = __shfl_xor_sync(0xffffffffu, vbits, 1);
    const bool valid_p = __shfl_xor_sync(0xffffffffu, (int)valid, 1) != 0;
    const bool want_mx = agg_mask & 8u, want_mn = agg_mask & 4u;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const bool mine = (r == (int)odd);
        const u64 ii = mine ? idx : idx_p;
        const u64 vb = mine ? vbits : vb_p;
        const bool ok = mine ? valid : valid_p;
        if (!ok) continue;
        Acc32* a = acc + ii;
        V v; memcpy(&v, &vb, 8);
        if constexpr (IsF64<V>::value) {
            double add; if (odd) memcpy(&add, &vb, 8); else add = 1.0;
            red_add_keep(reinterpret_cast<double*>(odd ? &a->sum : &a->cnt), add);      // red.global.add.f64
        } else {
            red_add_keep(odd ? &a->sum : &a->cnt, odd ? vb : 1ull);                      // red.global.add.u64 (wraps like a Java long)
        }
        if (odd ? want_mn : want_mx) red_max_keep(odd ? &a->mn : &a->mx, odd ? Codec<V>::enc_min(v) : Codec<V>::enc_max(v));   // red.global.max.u64
    }
}

// Per-block cache of heavy-hitter cells (skewed keys): groups of >= HOT_GROUP lanes of one warp that hit
// the same cell are accumulated in shared memory and written to the global table ONCE per block, at
// the end of the kernel, instead of once per warp round.
constexpr int HOT_SLOTS = 64;
constexpr int HOT_GROUP = 3;
struct HotTable { u64 idx[HOT_SLOTS]; u64 cnt[HOT_SLOTS], sum[HOT_SLOTS], mx[HOT_SLOTS], mn[HOT_SLOTS]; };

__device__ __forceinline__ void hot_init(HotTable* t) {
    for (int i = threadIdx.x; i < HOT_SLOTS; i += blockDim.x) { t->idx[i] = ~0ull; t->cnt[i] = 0; t->sum[i] = 0; t->mx[i] = 0; t->mn[i] = 0; }
}
// true if the contribution was absorbed by the block's table
template <typename V>
__device__ __forceinline__ bool hot_absorb(HotTable* t, u64 idx, u64 cntb, u64 sumb, u64 mx, u64 mn) {
    u32 hsh = (u32)((idx * 0x9E3779B97F4A7C15ull) >> 58);
    for (int probe = 0; probe < 4; ++probe, hsh = (hsh + 1) & (HOT_SLOTS - 1)) {
        u64 k = t->idx[hsh];
        if (k == ~0ull) k = atomicCAS(&t->idx[hsh], ~0ull, idx), k = (k == ~0ull) ? idx : k;
        if (k != idx) continue;
        if constexpr (IsF64<V>::value) {
            atomicAdd(reinterpret_cast<double*>(&t->cnt[hsh]), __longlong_as_double((i64)cntb));
            atomicAdd(reinterpret_cast<double*>(&t->sum[hsh]), __longlong_as_double((i64)sumb));
        } else { atomicAdd(&t->cnt[hsh], cntb); atomicAdd(&t->sum[hsh], sumb); }
        atomicMax(&t->mx[hsh], mx); atomicMax(&t->mn[hsh], mn);
        return true;
    }
    return false;
}
template <typename V>
__device__ __forceinline__ void hot_flush(const HotTable* t, Acc32* __restrict__ acc, u32 agg_mask) {
    for (int i = threadIdx.x; i < HOT_SLOTS; i += blockDim.x) {
        const u64 idx = t->idx[i];
        if (idx == ~0ull) continue;
        Acc32* a = acc + idx;
        if constexpr (IsF64<V>::value) {
            atomicAdd(reinterpret_cast<double*>(&a->cnt), __longlong_as_double((i64)t->cnt[i]));
            atomicAdd(reinterpret_cast<double*>(&a->sum), __longlong_as_double((i64)t->sum[i]));
        } else { atomicAdd(&a->cnt, t->cnt[i]); atomicAdd(&a->sum, t->sum[i]); }
        if (agg_mask & 8u) atomicMax(&a->mx, t->mx[i]);
        if (agg_mask & 4u) atomicMax(&a->mn, t->mn[i]);
    }
}

// A5 for one event per lane.
// combine = true: lanes of a warp that hit the SAME cell are combined first (the group's lowest lane
// carries the whole group's contribution).  With skewed keys (Zipf(1.2): one key gets 18 % of the
// events) same-address REDs otherwise serialise in one L2 slice — measured 5.2 ms instead of 0.28 ms
// per 2^24 events, 1.35 ms with combining.  With uniform keys the match instruction and the wider
// exchange cost +25 %, so the mode is chosen per batch from a sample (`sample` = this warp measures):
// dup[0] += lanes looked at, dup[1] += lanes that share their cell with another lane.
template <typename V>
__device__ __forceinline__ void accumulate_pair(Acc32* __restrict__ acc, bool valid, u64 idx, V val, u32 agg_mask, bool combine, bool sample, u32 (&dup)[2],
                                                HotTable* hot) {
    if (!combine && !sample) { accumulate_pair_lean<V>(acc, valid, idx, val, agg_mask); return; }
    const unsigned lane = threadIdx.x & 31;
    const unsigned m = __match_any_sync(0xffffffffu, valid ? idx : (0xFFFF000000000000ull | lane));
    const unsigned shared_m = __ballot_sync(0xffffffffu, valid && (m & (m - 1)) != 0);
    if (sample) { dup[0] += __popc(__ballot_sync(0xffffffffu, valid)); dup[1] += __popc(shared_m); }
    if (!combine || shared_m == 0) { accumulate_pair_lean<V>(acc, valid, idx, val, agg_mask); return; }
    u64 vbits; memcpy(&vbits, &val, 8);
    u64 cntb, sumb = vbits, mx = Codec<V>::enc_max(val), mn = Codec<V>::enc_min(val);
    const int leader = __ffs(m) - 1;
    u64 c_acc = 0, s_acc = 0, mx_acc = 0, mn_acc = 0; double cd = 0.0, sd = 0.0;
    // every lane walks the members of ITS group (lowest lane first: fixed order, deterministic double sums);
    // the trip count is the largest group of the warp, not 32
    const int gmax = __reduce_max_sync(0xffffffffu, __popc(m));
    unsigned rem = m;
    for (int it = 0; it < gmax; ++it) {
        const bool use = rem != 0;
        const int src = use ? __ffs(rem) - 1 : (int)lane;
        rem &= rem - 1;
        const u64 s_i = __shfl_sync(0xffffffffu, sumb, src), mx_i = __shfl_sync(0xffffffffu, mx, src), mn_i = __shfl_sync(0xffffffffu, mn, src);
        if (use) {
            if constexpr (IsF64<V>::value) { cd += 1.0; sd += __longlong_as_double((i64)s_i); } else { c_acc += 1; s_acc += s_i; }
            mx_acc = mx_i > mx_acc ? mx_i : mx_acc; mn_acc = mn_i > mn_acc ? mn_i : mn_acc;
        }
    }
    if constexpr (IsF64<V>::value) { cntb = (u64)__double_as_longlong(cd); sumb = (u64)__double_as_longlong(sd); } else { cntb = c_acc; sumb = s_acc; }
    bool mine = valid && (int)lane == leader;
    if (mine && __popc(m) >= HOT_GROUP && hot_absorb<V>(hot, idx, cntb, sumb, mx_acc, mn_acc)) mine = false;   // heavy hitter: once per block
    accumulate_pair_raw<V>(acc, mine, idx, cntb, sumb, mx_acc, mn_acc, agg_mask);
}

constexpr int INGEST_THREADS = 256;
constexpr int INGEST_EPT = 4;          // events per thread per tile (one 32 B load per column)
#ifndef FMQ_INGEST_MIN_BLOCKS
#define FMQ_INGEST_MIN_BLOCKS 3
#endif

// Loads the (up to) four consecutive events of this thread.
template <typename V>
__device__ __forceinline__ void load_events(const IngestArgs& a, size_t i, i64 (&ts_)[INGEST_EPT], u64 (&key_)[INGEST_EPT],
                                            u64 (&vb_)[INGEST_EPT], u32& in_mask) {
    const V* __restrict__ val = static_cast<const V*>(a.val);
    if (a.vec32 && i + INGEST_EPT <= a.n) {
        const ulonglong4 x = ldg_stream_v4(a.ts + i), y = ldg_stream_v4(a.key + i), z = ldg_stream_v4(val + i);
        ts_[0] = (i64)x.x; ts_[1] = (i64)x.y; ts_[2] = (i64)x.z; ts_[3] = (i64)x.w;
        key_[0] = y.x; key_[1] = y.y; key_[2] = y.z; key_[3] = y.w;
        vb_[0] = z.x; vb_[1] = z.y; vb_[2] = z.z; vb_[3] = z.w;
        in_mask = 0xFu;
    } else {
        in_mask = 0;
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e) {
            const bool in = i + e < a.n;
            in_mask |= in ? (1u << e) : 0u;
            ts_[e] = in ? (i64)ldg_stream(a.ts + i + e) : 0;
            key_[e] = in ? ldg_stream(a.key + i + e) : 0;
            vb_[e] = in ? ldg_stream(val + i + e) : 0;
        }
    }
}

// Same for the tile that starts `off` elements into the columns and holds `cnt` events; `i` = this thread's first event in it.
template <typename V>
__device__ __forceinline__ void load_events_at(const IngestArgs& a, size_t off, u32 cnt, u32 i, i64 (&ts_)[INGEST_EPT], u64 (&key_)[INGEST_EPT],
                                               u64 (&vb_)[INGEST_EPT], u32& in_mask) {
    const u64* ts = reinterpret_cast<const u64*>(a.ts) + off; const u64* key = a.key + off; const u64* val = static_cast<const u64*>(a.val) + off;
    if (a.vec32 && i + INGEST_EPT <= cnt) {
        const ulonglong4 x = ldg_stream_v4(ts + i), y = ldg_stream_v4(key + i), z = ldg_stream_v4(val + i);
        ts_[0] = (i64)x.x; ts_[1] = (i64)x.y; ts_[2] = (i64)x.z; ts_[3] = (i64)x.w;
        key_[0] = y.x; key_[1] = y.y; key_[2] = y.z; key_[3] = y.w;
        vb_[0] = z.x; vb_[1] = z.y; vb_[2] = z.z; vb_[3] = z.w;
        in_mask = 0xFu;
    } else {
        in_mask = 0;
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e) {
            const bool in = i + e < cnt;
            in_mask |= in ? (1u << e) : 0u;
            ts_[e] = in ? (i64)ldg_stream(ts + i + e) : 0;
            key_[e] = in ? ldg_stream(key + i + e) : 0;
            vb_[e] = in ? ldg_stream(val + i + e) : 0;
        }
    }
}

// A6-i: per-event threshold filter + compaction (Main.java:31); one reservation per block tile,
// rows keep arrival order inside the tile.  Called by every thread of the block.
template <typename V>
__device__ __forceinline__ u32 filter_events(const IngestArgs& a, DevState* st, const i64 (&ts_)[INGEST_EPT], const u64 (&key_)[INGEST_EPT],
                                             const u64 (&vb_)[INGEST_EPT], u32 ok_mask) {
    u32 hits = 0;
#pragma unroll
    for (int e = 0; e < INGEST_EPT; ++e) {
        V v; memcpy(&v, &vb_[e], 8);
        hits |= (cmp_op((double)v, a.efilter_op, a.efilter_thr) ? 1u : 0u) << e;
    }
    hits &= ok_mask;
    const u32 nh = __popc(hits);
    u64 row = block_append<INGEST_THREADS>(nh, &st->n_ealert, a.ea_cap, st);
    if (row != ~0ull) {
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e)
            if (hits & (1u << e)) { a.ea_ts[row] = ts_[e]; a.ea_key[row] = key_[e]; a.ea_val[row] = vb_[e]; ++row; }
    }
    return nh;
}

// Chooses the kernel on the device-resident stream position without a host round trip: the fast
// kernel is launched first and returns immediately unless its preconditions hold, and vice versa.
__device__ __forceinline__ bool ingest_fast_ok(const DevState* st, const WindowGeom& g, int window_enabled = 1) {
    if (!window_enabled) return true;                            // filter-only jobs (chapter1) need no pane origin
    // 32-bit pane offsets and 32-bit cell indices: ring * slide < 2^32 ms and fewer than 2^31 accumulator cells (64 GB)
    return st->anchored && g.slide32 != 0 && (u64)(g.ring_mask + 1) * g.slide32 < 0x100000000ull &&
           (u64)(g.ring_mask + 1) * g.capacity < 0x80000000ull;
}

// ---- segmented input ---------------------------------------------------------------------------------
// Where a batch is and how large it is, as the kernels see it: up to FMQ_MAX_PEERS segments whose sizes may
// live in device memory (so that a batch received from peers is consumed without a host round trip).
struct SegTab { u64 n[FMQ_MAX_PEERS]; u64 tile_end[FMQ_MAX_PEERS]; };

template <int TILE>
__device__ __forceinline__ void seg_tab_fill(const IngestArgs& a, SegTab* t, DevState* st) {      // one thread
    u64 acc = 0; const u32 ns = a.nseg ? a.nseg : 1;
    for (u32 s = 0; s < FMQ_MAX_PEERS; ++s) {
        u64 n = 0;
        if (s < ns) {
            n = a.seg_n ? a.seg_n[s] : a.n;
            if (n > a.n) { n = a.n; atomicOr(&st->err_flags, ERRB_BAD_INPUT); }   // a count beyond the segment's capacity
        }
        t->n[s] = n; acc += (n + TILE - 1) / TILE; t->tile_end[s] = acc;
    }
}
// tile index -> element offset of its first event (same for every column) and number of events in it
template <int TILE>
__device__ __forceinline__ size_t seg_locate(const IngestArgs& a, const SegTab& t, u64 tile, u32& cnt) {
    if (a.nseg <= 1) {                                           // one plain batch
        const u64 first = tile * TILE, rem = t.n[0] - first;
        cnt = rem < (u64)TILE ? (u32)rem : (u32)TILE;
        return (size_t)first;
    }
    u32 s = 0;
    while (tile >= t.tile_end[s]) ++s;                           // <= 8 steps, uniform across the warp
    const u64 first = (tile - (s ? t.tile_end[s - 1] : 0)) * TILE, rem = t.n[s] - first;
    cnt = rem < (u64)TILE ? (u32)rem : (u32)TILE;
    return (size_t)s * a.seg_stride + first;
}

// ---- ingest, fast path ------------------------------------------------------------------------------
// Preconditions (checked on the device): the stream is anchored (a watermark exists, so the pane
// origin is known), pane_ring * slide < 2^32, so the pane of a record is a 32-bit quotient of
// (ts - origin), and the table has fewer than 2^31 cells.  Per event the late test, the pane and the
// accumulate are predicated, not branched.  The event columns are streamed into shared memory by the
// TMA engine (cp.async.bulk 1-D copies completing on an mbarrier, L2 evict-first policy).
struct IngestRun {            // per-thread running scalars of one batch
    u64 late = 0, nea = 0; i64 dmax = I64_MIN_; u32 pmin = 0xFFFFFFFFu, pmax = 0; u32 err = 0;
    u32 dup[2] = {0, 0};
    u32 dumin = 0xFFFFFFFFu, dumax = 0;     // lean core: smallest / largest (ts - origin) among the accumulated events
};
struct IngestConst {
    i64 origin_ts, origin_pane; u32 origin_ring, slide, recip; u64 span, nbuckets; bool combine;
    // lean core
    u32 cap32, m32; bool exact32; u64 keep_policy;
    u32 fired_n;                 // allowedLateness: relative panes below this have a fired window (0 = lateness off)
};
// pane quotient of a 32-bit offset: one multiply when ring * slide^2 <= 2^32 (ceil(2^32/slide) is then exact for every
// offset inside the ring), else the estimate-and-correct form
__device__ __forceinline__ u32 pane_quotient(u32 du, const IngestConst& k) {
    if (k.exact32) return __umulhi(du, k.m32);
    u32 q = __umulhi(du, k.recip);                                 // q <= du / slide <= q + 2
    u32 rem = du - q * k.slide;
    if (rem >= k.slide) { q++; rem -= k.slide; }
    if (rem >= k.slide) { q++; }
    return q;
}

// The window part of one tile (A1 max-ts, A3 late test, A2 pane, A5 accumulate) for the four events held by
// this lane.  Must be reached by all 32 lanes of the warp.  Returns which of the four were late / accumulated,
// their pane quotient and slot (for the side outputs the callers append in their own way).
template <typename V, bool DENSE, int EPT>
__device__ __forceinline__ void ingest_core(const IngestArgs& a, DevState* st, const IngestConst& k, IngestRun& r,
                                            const i64 (&ts_)[EPT], const u64 (&key_)[EPT], const u64 (&vb_)[EPT], u32 ok_mask,
                                            HotTable* hot, bool sample, u32& late_mask, u32& vmask, u32 (&q_)[EPT], u32 (&slot_)[EPT]) {
    // ---- per event: A1 max-ts, A3 late test, A2 pane, and (hashed keys) the home-bucket load:
    //      all four directory loads are in flight before any of them is consumed ----
    u32 ring_[EPT]; u64 hb_[EPT]; ulonglong4 kv_[EPT]; u32 act_mask = 0; late_mask = 0;
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        const bool ok = ok_mask & (1u << e);
        const i64 d = ts_[e] - k.origin_ts;
        r.dmax = (ok && d > r.dmax) ? d : r.dmax;                  // A1 (ts = origin + d)
        const bool is_late = ok && d < 0;                          // A3: every window of this record has fired
        r.late += is_late ? 1 : 0;
        late_mask |= is_late ? (1u << e) : 0u;
        const bool beyond = ok && d >= 0 && (u64)d >= k.span;      // would alias an older pane of the ring
        r.err |= beyond ? ERRB_RING : 0u;
        const u32 du = (u32)d;
        u32 q = __umulhi(du, k.recip);                             // A2: q <= du / slide <= q + 2
        u32 rem = du - q * k.slide;
        if (rem >= k.slide) { q++; rem -= k.slide; }
        if (rem >= k.slide) { q++; }
        bool act = ok && d >= 0 && !beyond;
        if (DENSE) {
            if (act && key_[e] >= a.g.capacity) { r.err |= ERRB_BAD_INPUT; act = false; }   // id outside the table
        } else {
            hb_[e] = dir_home_bucket(key_[e], k.nbuckets);
            if (act) kv_[e] = ld_bucket(a.dkeys + hb_[e] * 4);
        }
        r.pmin = (act && q < r.pmin) ? q : r.pmin; r.pmax = (act && q > r.pmax) ? q : r.pmax;
        q_[e] = q;
        ring_[e] = (k.origin_ring + q) & a.g.ring_mask;
        act_mask |= act ? (1u << e) : 0u;
    }
    // ---- A5: slot -> pair-lane accumulate ----
    vmask = 0;
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        bool valid = act_mask & (1u << e);
        u32 slot = (u32)key_[e];                                   // dense ids index the table directly
        if (!DENSE) {
            if (valid) slot = dir_resolve_inline(a.dkeys, k.nbuckets, key_[e], hb_[e], kv_[e], st);
            valid = valid && slot != SLOT_INVALID;
        }
        const u64 idx = (u64)ring_[e] * a.g.capacity + slot;
        V v; memcpy(&v, &vb_[e], 8);
        if (k.fired_n && valid && q_[e] < k.fired_n) mark_dirty(a.g, st, ring_[e], slot);
        accumulate_pair<V>(a.acc, valid, idx, v, a.g.agg_mask, k.combine, e == 0 && sample, r.dup, hot);
        slot_[e] = slot; vmask |= valid ? (1u << e) : 0u;
    }
}

// ---- lean core (warp-ring kernel, no warp-level combining) ------------------------------------------------
// ncu of the first warp-ring kernel: 264 instructions per event and lane, issue slots 52 % busy with 16 warps per
// SM, lg_throttle 0 — the kernel was bound by its own instruction count, not by the L2 (profiles/r02_notes.md).
// This form does the same work in 32-bit arithmetic wherever the geometry allows it: one unsigned compare is the
// late AND the ring test, the pane quotient is one multiply, a cell is a 32-bit index (tables < 2^31 cells), the
// pair-lane exchange moves three words, and the order-preserving max / min encoding is three logic operations.
template <typename V>
__device__ __forceinline__ void red_half(char* acc_half, u32 c, u32 lo, u32 hi, bool odd, u32 oddmask, bool want, u64 policy) {
    if (c == 0xFFFFFFFFu) return;
    char* p = acc_half + (u64)c * 32u;
    if constexpr (IsF64<V>::value) {
        const double addv = odd ? __hiloint2double((int)hi, (int)lo) : 1.0;
        asm volatile("red.global.add.L2::cache_hint.f64 [%0], %1, %2;" :: "l"(p), "d"(addv), "l"(policy) : "memory");
    } else {
        const u64 addv = odd ? (((u64)hi << 32) | lo) : 1ull;
        asm volatile("red.global.add.L2::cache_hint.u64 [%0], %1, %2;" :: "l"(p), "l"(addv), "l"(policy) : "memory");
    }
    if (want) {
        u32 uh, ul;
        if constexpr (IsF64<V>::value) {
            const u32 m = (u32)((int)hi >> 31);                       // all ones for negative values
            uh = hi ^ (m | 0x80000000u) ^ oddmask; ul = lo ^ m ^ oddmask;   // order_f64, complemented on the min lanes
            const double v = __hiloint2double((int)hi, (int)lo);
            if (v != v) { uh = 0xFFFFFFFFu; ul = 0xFFFFFFFFu; }      // NaN wins both max and min (Math.max / Math.min)
        } else {
            uh = hi ^ 0x80000000u ^ oddmask; ul = lo ^ oddmask;
        }
        const u64 x = ((u64)uh << 32) | ul;
        asm volatile("red.global.max.L2::cache_hint.u64 [%0], %1, %2;" :: "l"(p + 16), "l"(x), "l"(policy) : "memory");
    }
}
// A5 for one event per lane: cw = cell index or 0xFFFFFFFF (nothing to add).  Must be reached by all 32 lanes.
template <typename V>
__device__ __forceinline__ void accumulate_lean32(char* acc_half, u32 cw, u64 vb, bool odd, u32 oddmask, bool want, u64 policy) {
    const u32 lo = (u32)vb, hi = (u32)(vb >> 32);
    const u32 cw_p = __shfl_xor_sync(0xffffffffu, cw, 1), lo_p = __shfl_xor_sync(0xffffffffu, lo, 1), hi_p = __shfl_xor_sync(0xffffffffu, hi, 1);
    // first the even lane's event (both halves in one instruction = one sector request), then the odd lane's
    red_half<V>(acc_half, odd ? cw_p : cw, odd ? lo_p : lo, odd ? hi_p : hi, odd, oddmask, want, policy);
    red_half<V>(acc_half, odd ? cw : cw_p, odd ? lo : lo_p, odd ? hi : hi_p, odd, oddmask, want, policy);
}

template <typename V, bool DENSE, int EPT>
__device__ __forceinline__ void ingest_core_lean(const IngestArgs& a, DevState* st, const IngestConst& k, IngestRun& r,
                                                 const i64 (&ts_)[EPT], const u64 (&key_)[EPT], const u64 (&vb_)[EPT], u32 in_mask,
                                                 u32& late_mask, u32& vmask, u32 (&q_)[EPT], u32 (&slot_)[EPT]) {
    const bool odd = threadIdx.x & 1;
    const u32 oddmask = odd ? 0xFFFFFFFFu : 0u;
    const bool want = odd ? (a.g.agg_mask & 4u) != 0 : (a.g.agg_mask & 8u) != 0;
    char* acc_half = reinterpret_cast<char*>(a.acc) + (odd ? 8 : 0);
    u64 hb_[EPT]; ulonglong4 kv_[EPT]; u32 act_mask = 0;
    late_mask = 0; vmask = 0;
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        const bool in = (in_mask >> e) & 1u;
        const i64 d = ts_[e] - k.origin_ts;
        const bool inr = (u64)d < k.span;                          // A3 + ring test: late (negative) and beyond-the-ring offsets both fail
        bool good = in && ts_[e] >= 0;
        if (DENSE) good = good && key_[e] < (u64)k.cap32; else good = good && key_[e] != KEY_EMPTY;
        const bool act = good && inr;
        const u32 du = (u32)d;
        q_[e] = pane_quotient(du, k);                              // A2
        r.dumin = min(r.dumin, act ? du : 0xFFFFFFFFu); r.dumax = max(r.dumax, act ? du : 0u);   // A1 and the live pane range, divided once at the end
        if (in && !act) {                                          // rare: late record, bad input, or a span the ring cannot hold
            if (!good) r.err |= ERRB_BAD_INPUT;
            else { r.dmax = d > r.dmax ? d : r.dmax; if (d < 0) { r.late++; late_mask |= 1u << e; } else r.err |= ERRB_RING; }
        }
        if (!DENSE) { hb_[e] = dir_home_bucket(key_[e], k.nbuckets); if (act) kv_[e] = ld_bucket(a.dkeys + hb_[e] * 4); }
        act_mask |= act ? (1u << e) : 0u;
    }
#pragma unroll
    for (int e = 0; e < EPT; ++e) {
        bool valid = (act_mask >> e) & 1u;
        u32 slot = (u32)key_[e];
        if (!DENSE) {
            if (valid) slot = dir_resolve_inline(a.dkeys, k.nbuckets, key_[e], hb_[e], kv_[e], st);
            valid = valid && slot != SLOT_INVALID;
        }
        const u32 cell = ((k.origin_ring + q_[e]) & a.g.ring_mask) * k.cap32 + slot;
        slot_[e] = slot; vmask |= valid ? (1u << e) : 0u;
        if (k.fired_n && valid && q_[e] < k.fired_n) mark_dirty(a.g, st, (k.origin_ring + q_[e]) & a.g.ring_mask, slot);
        accumulate_lean32<V>(acc_half, valid ? cell : 0xFFFFFFFFu, vb_[e], odd, oddmask, want, k.keep_policy);
    }
}

// ---- ingest, warp-ring kernel ------------------------------------------------------------------------
// Same preconditions and per-event work as the block-staged kernel above, without a single block-wide
// barrier in the streaming loop: every WARP owns a private ring of IW_STAGES tiles of 128 events (3 x 1 KB
// bulk copies per tile, cp.async.bulk completing on the warp's own mbarrier, L2 evict-first), refilled by
// its lane 0 as soon as the 32 lanes hold the tile in registers (__syncwarp + cross-proxy fence).  Rows of
// the per-event filter (and of the late / median side outputs) are staged per warp in shared memory and
// written out behind ONE reservation per ~100 rows, so warps never wait for each other: what is left to
// stall a warp is the RED / store path itself.  (ncu of the block-staged kernel: `barrier` was the top
// stall, 20 % of the samples sat in the mbarrier spin while the L2 ran at 66 %, profiles/r01_notes.md.)
#ifndef FMQ_IW_WARPS
#define FMQ_IW_WARPS 4
#endif
#ifndef FMQ_IW_STAGES
#define FMQ_IW_STAGES 3
#endif
#ifndef FMQ_IW_BLOCKS
#define FMQ_IW_BLOCKS 4
#endif
#ifndef FMQ_IW_EPT
#define FMQ_IW_EPT 4
#endif
#ifndef FMQ_IW_ROWS
#define FMQ_IW_ROWS 128
#endif
constexpr int IW_WARPS = FMQ_IW_WARPS, IW_STAGES = FMQ_IW_STAGES, IW_THREADS = IW_WARPS * 32;
constexpr int IW_EPT = FMQ_IW_EPT;                          // events per lane per tile
constexpr int IW_TILE = 32 * IW_EPT;                        // events per warp tile
constexpr int IW_STAGE_BYTES = 3 * IW_TILE * 8;
constexpr int IW_ROWS = FMQ_IW_ROWS;                        // rows a warp stages before it reserves output space (>= 64)
constexpr int IW_ROWBUF_BYTES = 3 * IW_ROWS * 8;
__host__ __device__ constexpr int iw_warp_bytes(bool extras) { return IW_STAGES * IW_STAGE_BYTES + (extras ? 3 : 1) * IW_ROWBUF_BYTES; }
__host__ __device__ constexpr int iw_smem_bytes(bool extras) { return IW_WARPS * iw_warp_bytes(extras); }

__device__ __forceinline__ void mbar_arrive(u64* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// rows (3 x 8 B) staged by one warp; n is warp-uniform
struct WarpRows { u64* buf; u32 n; };
__device__ __forceinline__ void wr_push(WarpRows& w, bool pred, u64 c0, u64 c1, u64 c2) {      // all 32 lanes; needs room for 32 rows
    const u32 lane = threadIdx.x & 31;
    const unsigned m = __ballot_sync(0xffffffffu, pred);
    if (pred) { const u32 i = w.n + __popc(m & ((1u << lane) - 1)); w.buf[i] = c0; w.buf[IW_ROWS + i] = c1; w.buf[2 * IW_ROWS + i] = c2; }
    w.n += __popc(m);
}
template <typename T0, typename T1>
__device__ __forceinline__ void wr_flush(WarpRows& w, u64* counter, u64 cap, T0* d0, T1* d1, u64* d2, u32& err) {   // all 32 lanes
    if (w.n == 0) return;
    const u32 lane = threadIdx.x & 31;
    __syncwarp();
    u64 base = 0;
    if (lane == 0) base = atomicAdd(counter, (u64)w.n);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base + w.n > cap) err |= ERRB_OUTPUT;
    else for (u32 i = lane; i < w.n; i += 32) { d0[base + i] = (T0)w.buf[i]; d1[base + i] = (T1)w.buf[IW_ROWS + i]; d2[base + i] = w.buf[2 * IW_ROWS + i]; }
    __syncwarp();
    w.n = 0;
}

// The value log (N3) takes a row for EVERY accepted record: one reservation per 128 staged rows was 1.3 x 10^5 atomics with
// a return value on one address per 2^24 records, which serialise in their L2 slice (+0.43 ms per batch, measured).  A warp
// therefore reserves up to ML_CHUNK rows at a time — never more than it can still produce — and fills what it did not use
// (records dropped as late, ragged tiles) with rows of pane ML_PANE_NONE, which belong to no window and are dropped
// by the next compaction.
constexpr u32 ML_CHUNK = 2048;
constexpr i64 ML_PANE_NONE = -(1ll << 62);
struct LogSpace { u64 base; u32 left; };                    // warp-uniform
__device__ __forceinline__ void ml_flush(WarpRows& w, LogSpace& sp, u64 future, const IngestArgs& a, DevState* st, u32& err) {   // all 32 lanes
    if (w.n == 0) return;
    const u32 lane = threadIdx.x & 31;
    __syncwarp();
    u32 done = 0;
    while (done < w.n) {
        if (sp.left == 0) {
            const u64 most = (u64)(w.n - done) + future;
            const u32 want = most < ML_CHUNK ? (u32)most : ML_CHUNK;
            u64 b = 0;
            if (lane == 0) b = atomicAdd(&st->n_mlog, (u64)want);
            b = __shfl_sync(0xffffffffu, b, 0);
            if (b + want > a.ml_cap) { err |= ERRB_OUTPUT; break; }
            sp.base = b; sp.left = want;
        }
        const u32 m = sp.left < w.n - done ? sp.left : w.n - done;
        for (u32 i = lane; i < m; i += 32) {
            a.ml_pane[sp.base + i] = (i64)w.buf[done + i]; a.ml_slot[sp.base + i] = (u32)w.buf[IW_ROWS + done + i]; a.ml_val[sp.base + i] = w.buf[2 * IW_ROWS + done + i];
        }
        sp.base += m; sp.left -= m; done += m;
    }
    __syncwarp();
    w.n = 0;
}
__device__ __forceinline__ void ml_finish(LogSpace& sp, const IngestArgs& a) {
    for (u32 i = threadIdx.x & 31; i < sp.left; i += 32) { a.ml_pane[sp.base + i] = ML_PANE_NONE; a.ml_slot[sp.base + i] = 0u; a.ml_val[sp.base + i] = 0ull; }
    sp.left = 0;
}

// COMBINE: warp-level pre-aggregation of lanes that hit the same cell (skewed keys).  The host picks the variant
// from a hint the plan kernel leaves in pinned host memory (it may lag by a batch: the choice only changes the
// order of additions, never the result beyond the double-sum tolerance).
template <typename V, bool DENSE, bool EXTRAS, bool COMBINE>
__global__ void __launch_bounds__(IW_THREADS, FMQ_IW_BLOCKS) ingest_ring_kernel(const __grid_constant__ IngestArgs a) {
    extern __shared__ __align__(128) unsigned char iw_smem[];
    __shared__ __align__(8) u64 bars[IW_WARPS][IW_STAGES];
    __shared__ SegTab seg;
    __shared__ HotTable hot_tab;
    DevState* st = a.st;
    if (!ingest_fast_ok(st, a.g, a.window_enabled)) return;      // the generic kernel handles this batch
    const u32 lane = threadIdx.x & 31, wl = threadIdx.x >> 5;
    IngestConst k{};
    if (a.window_enabled) {
        k.origin_ts = st->origin_ts;
        k.origin_pane = st->origin_pane;
        k.combine = COMBINE;
        k.origin_ring = (u32)st->origin_pane & a.g.ring_mask;
        k.slide = a.g.slide32;
        k.recip = 0xFFFFFFFFu / k.slide;
        k.span = (u64)(a.g.ring_mask + 1 - st->pending_clear) * k.slide;
        k.nbuckets = a.g.capacity >> 2;
        k.cap32 = (u32)a.g.capacity;
        k.exact32 = k.slide >= 2 && (u64)(a.g.ring_mask + 1) * k.slide * k.slide <= 0x100000000ull;
        k.m32 = k.exact32 ? (u32)((0x100000000ull + k.slide - 1) / k.slide) : 0u;
        k.keep_policy = l2_keep_policy();
        k.fired_n = fired_panes(a.g, st);
    }
    IngestRun r;
    if (threadIdx.x == 0) seg_tab_fill<IW_TILE>(a, &seg, st);
    if (COMBINE) hot_init(&hot_tab);
    if (lane == 0) {
        for (int s = 0; s < IW_STAGES; ++s) mbar_init(&bars[wl][s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();                                             // the only block barrier before the epilogue
    const u64 ntiles = seg.tile_end[FMQ_MAX_PEERS - 1];
    const u64 gw = (u64)blockIdx.x * IW_WARPS + wl, nw = (u64)gridDim.x * IW_WARPS;
    unsigned char* wbase = iw_smem + (size_t)wl * iw_warp_bytes(EXTRAS);
    WarpRows ea{reinterpret_cast<u64*>(wbase + IW_STAGES * IW_STAGE_BYTES), 0};
    WarpRows lt{reinterpret_cast<u64*>(wbase + IW_STAGES * IW_STAGE_BYTES + IW_ROWBUF_BYTES), 0};
    WarpRows ml{reinterpret_cast<u64*>(wbase + IW_STAGES * IW_STAGE_BYTES + 2 * IW_ROWBUF_BYTES), 0};
    LogSpace ml_space{0, 0};
    const u64* cts = reinterpret_cast<const u64*>(a.ts); const u64* cval = static_cast<const u64*>(a.val);
    u64 policy;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    auto stage_col = [&](int s, int c) { return reinterpret_cast<u64*>(wbase + s * IW_STAGE_BYTES + c * IW_TILE * 8); };
    auto issue = [&](u64 t, int s) {                             // lane 0: start the three copies of tile t, or just arrive
        u32 cnt; const size_t off = seg_locate<IW_TILE>(a, seg, t, cnt);
        if (a.vec32 && cnt == IW_TILE) {
            mbar_expect_tx(&bars[wl][s], 3u * IW_TILE * 8u);
            tma_load_1d(stage_col(s, 0), cts + off, IW_TILE * 8u, &bars[wl][s], policy);
            tma_load_1d(stage_col(s, 1), a.key + off, IW_TILE * 8u, &bars[wl][s], policy);
            tma_load_1d(stage_col(s, 2), cval + off, IW_TILE * 8u, &bars[wl][s], policy);
        } else {
            mbar_arrive(&bars[wl][s]);                            // ragged or unaligned tile: the lanes load it themselves
        }
    };
    if (lane == 0)
        for (int s = 0; s < IW_STAGES; ++s) { const u64 t = gw + (u64)s * nw; if (t < ntiles) issue(t, s); }
    // per-event filter A6-i (Main.java:31) as one comparison: LT / LE compare the negated operands
    const bool f_flip = a.efilter_op == 1 || a.efilter_op == 3, f_strict = a.efilter_op < 2;
    const double f_thr = f_flip ? -a.efilter_thr : a.efilter_thr;
    u32 nea_w = 0;                                               // rows this warp produced (uniform)
    int s = 0; u32 phase = 0;
    for (u64 t = gw; t < ntiles; t += nw) {
        mbar_wait(&bars[wl][s], phase);
        u32 cnt; const size_t off = seg_locate<IW_TILE>(a, seg, t, cnt);
        // lane x takes events x, x+32, x+64, x+96 of the tile: conflict-free 8-byte shared loads
        i64 ts_[IW_EPT]; u64 key_[IW_EPT]; u64 vb_[IW_EPT]; u32 in_mask = (1u << IW_EPT) - 1u;
        if (a.vec32 && cnt == IW_TILE) {
            const u64 *sts = stage_col(s, 0), *skey = stage_col(s, 1), *sval = stage_col(s, 2);
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) { ts_[e] = (i64)sts[lane + 32 * e]; key_[e] = skey[lane + 32 * e]; vb_[e] = sval[lane + 32 * e]; }
        } else {
            in_mask = 0;
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) {
                const u32 i = lane + 32 * e; const bool in = i < cnt;
                in_mask |= in ? (1u << e) : 0u;
                ts_[e] = in ? (i64)ldg_stream(cts + off + i) : 0; key_[e] = in ? ldg_stream(a.key + off + i) : 0; vb_[e] = in ? ldg_stream(cval + off + i) : 0;
            }
        }
        __syncwarp();                                            // every lane holds its events: the stage is free
        if (lane == 0) {
            // generic-proxy reads of the stage (all lanes, ordered by the warp barrier) before the async-proxy refill
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            const u64 tn = t + (u64)IW_STAGES * nw;
            if (tn < ntiles) issue(tn, s);
        }
        if (++s == IW_STAGES) { s = 0; phase ^= 1u; }
        if (a.efilter_enabled) {
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) {
                V v; memcpy(&v, &vb_[e], 8);
                const double x = f_flip ? -(double)v : (double)v;
                const bool hit = ((in_mask >> e) & 1u) && ts_[e] >= 0 && key_[e] != KEY_EMPTY && (f_strict ? x > f_thr : x >= f_thr);
                if (ea.n + 32 > IW_ROWS) wr_flush(ea, &st->n_ealert, a.ea_cap, a.ea_ts, a.ea_key, a.ea_val, r.err);
                const u32 before = ea.n;
                wr_push(ea, hit, (u64)ts_[e], key_[e], vb_[e]);
                nea_w += ea.n - before;
            }
        }
        if (!a.window_enabled) {                                 // filter-only job: input validity is still checked
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) if (((in_mask >> e) & 1u) && (ts_[e] < 0 || key_[e] == KEY_EMPTY)) r.err |= ERRB_BAD_INPUT;
            continue;
        }
        u32 late_mask, vmask, q_[IW_EPT], slot_[IW_EPT];
        if constexpr (!COMBINE) {
            ingest_core_lean<V, DENSE, IW_EPT>(a, st, k, r, ts_, key_, vb_, in_mask, late_mask, vmask, q_, slot_);
            if (wl == 0 && (((t - gw) / nw) & 3u) == 0) {        // skew sample (one warp per block, every 4th tile): lanes sharing their cell with another lane
                const bool v0 = vmask & 1u;
                const u64 c0 = ((u64)((k.origin_ring + q_[0]) & a.g.ring_mask) << 32) | slot_[0];
                const unsigned m = __match_any_sync(0xffffffffu, v0 ? c0 : (0xFFFF000000000000ull | lane));
                r.dup[0] += __popc(__ballot_sync(0xffffffffu, v0)); r.dup[1] += __popc(__ballot_sync(0xffffffffu, v0 && (m & (m - 1)) != 0));
            }
        } else {
            u32 ok_mask = 0;
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) ok_mask |= ((ts_[e] >= 0 && key_[e] != KEY_EMPTY) ? 1u : 0u) << e;
            if ((ok_mask & in_mask) != in_mask) r.err |= ERRB_BAD_INPUT;
            ok_mask &= in_mask;
            ingest_core<V, DENSE, IW_EPT>(a, st, k, r, ts_, key_, vb_, ok_mask, &hot_tab, wl == 0, late_mask, vmask, q_, slot_);
        }
        if (EXTRAS && a.lt_cap) {                                // late side output (chapter3/README.md:213-228, option 3)
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) {
                if (lt.n + 32 > IW_ROWS) wr_flush(lt, &st->n_late, a.lt_cap, a.lt_ts, a.lt_key, a.lt_val, r.err);
                wr_push(lt, (late_mask >> e) & 1u, (u64)ts_[e], key_[e], vb_[e]);
            }
        }
        if (EXTRAS && a.mlog_enabled) {                          // N3: every accepted value is logged (ComputeCpuMiddle.java:36-38)
#pragma unroll
            for (int e = 0; e < IW_EPT; ++e) {
                // rows this warp can still produce: the rest of this tile and its later tiles
                if (ml.n + 32 > IW_ROWS) ml_flush(ml, ml_space, (u64)(IW_EPT - e) * 32u + (u64)IW_TILE * ((ntiles - 1 - t) / nw), a, st, r.err);
                wr_push(ml, (vmask >> e) & 1u, (u64)(k.origin_pane + (i64)q_[e]), (u64)slot_[e], vb_[e]);
            }
        }
    }
    if (a.efilter_enabled) wr_flush(ea, &st->n_ealert, a.ea_cap, a.ea_ts, a.ea_key, a.ea_val, r.err);
    if (EXTRAS && a.lt_cap) wr_flush(lt, &st->n_late, a.lt_cap, a.lt_ts, a.lt_key, a.lt_val, r.err);
    if (EXTRAS && a.mlog_enabled) { ml_flush(ml, ml_space, 0, a, st, r.err); ml_finish(ml_space, a); }
    if (COMBINE && a.window_enabled) { __syncthreads(); hot_flush<V>(&hot_tab, a.acc, a.g.agg_mask); }
    // ---- per-warp reduction of the batch scalars, one atomic each per warp ----
    r.nea = nea_w;                                               // uniform: lane 0 reports it
    for (int o = 16; o > 0; o >>= 1) r.err |= __shfl_xor_sync(0xffffffffu, r.err, o);
    if (a.window_enabled) {
        // lean core: the extreme offsets become the extreme panes and the largest timestamp here, once per lane
        if (r.dumin <= r.dumax) {
            const u32 qlo = pane_quotient(r.dumin, k), qhi = pane_quotient(r.dumax, k);
            r.pmin = qlo < r.pmin ? qlo : r.pmin; r.pmax = qhi > r.pmax ? qhi : r.pmax;
            r.dmax = (i64)r.dumax > r.dmax ? (i64)r.dumax : r.dmax;
        }
        r.late = warp_sum_u64(r.late); r.dmax = warp_max_i64(r.dmax);
        r.pmin = __reduce_min_sync(0xffffffffu, r.pmin); r.pmax = __reduce_max_sync(0xffffffffu, r.pmax);
    }
    if (threadIdx.x == 0 && r.dup[0]) { atomicAdd(&st->dup_sampled, (u64)r.dup[0]); atomicAdd(&st->dup_found, (u64)r.dup[1]); }
    if (lane == 0) {
        if (r.nea) atomicAdd(&st->event_alerts, r.nea);
        if (r.err) atomicOr(&st->err_flags, r.err);
        if (a.window_enabled) {
            if (r.late) atomicAdd(&st->late_dropped, r.late);
            if (r.dmax != I64_MIN_) atomicMax(&st->batch_max_ts, k.origin_ts + r.dmax);
            if (r.pmin <= r.pmax) { atomicMin(&st->min_pane, st->origin_pane + (i64)r.pmin); atomicMax(&st->max_pane, st->origin_pane + (i64)r.pmax); }
        }
    }
}

// ---- ingest, generic kernel -------------------------------------------------------------------------
// Used while the stream is not anchored yet (no watermark: nothing can be late and the pane origin
// is unknown, so pane indices need a 64-bit division) or when the geometry does not fit 32 bits.
template <typename V, bool DENSE>
__global__ void __launch_bounds__(INGEST_THREADS) ingest_generic_kernel(const __grid_constant__ IngestArgs a) {
    DevState* st = a.st;
    __shared__ SegTab seg;
    if (ingest_fast_ok(st, a.g, a.window_enabled)) return;       // the fast kernel handles this batch
    if (threadIdx.x == 0) seg_tab_fill<INGEST_THREADS * INGEST_EPT>(a, &seg, st);
    __syncthreads();
    const int anchored = st->anchored;
    const i64 origin_ts = anchored ? st->origin_ts : (a.g.off - a.g.slide);
    const i64 origin_pane = anchored ? st->origin_pane : -1;
    const u32 R = a.g.ring_mask + 1 - st->pending_clear;
    const u32 gfired = fired_panes(a.g, st);

    u64 late = 0, nea = 0; i64 max_ts = I64_MIN_; i64 pmin = I64_MAX_, pmax = I64_MIN_; u32 err = 0;

    constexpr int tile = INGEST_THREADS * INGEST_EPT;
    const u64 ntiles = seg.tile_end[FMQ_MAX_PEERS - 1];
    for (u64 t = blockIdx.x; t < ntiles; t += gridDim.x) {
        u32 cnt; const size_t off = seg_locate<tile>(a, seg, t, cnt);
        i64 ts_[INGEST_EPT]; u64 key_[INGEST_EPT]; u64 vb_[INGEST_EPT]; u32 in_mask;
        load_events_at<V>(a, off, cnt, threadIdx.x * INGEST_EPT, ts_, key_, vb_, in_mask);
        u32 ok_mask = 0;
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e) ok_mask |= ((ts_[e] >= 0 && key_[e] != KEY_EMPTY) ? 1u : 0u) << e;
        if ((ok_mask & in_mask) != in_mask) err |= ERRB_BAD_INPUT;
        ok_mask &= in_mask;
        if (a.efilter_enabled) nea += filter_events<V>(a, st, ts_, key_, vb_, ok_mask);
        if (!a.window_enabled) continue;
        i64 gp_[INGEST_EPT]; u32 gs_[INGEST_EPT]; u32 gmask = 0, glate = 0;
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e) {
            const i64 ts = ts_[e]; const u64 key = key_[e];
            V v; memcpy(&v, &vb_[e], 8);
            bool valid = false; u64 idx = 0; i64 gpane = 0;
            if (ok_mask & (1u << e)) {
                max_ts = ts > max_ts ? ts : max_ts;                 // A1: currentMaxTimestamp
                const i64 d = ts - origin_ts;
                if (d < 0) {
                    late++;                                          // A3
                    glate |= 1u << e;
                } else {
                    const u64 prel = (u64)d / (u64)a.g.slide;        // A2 (64-bit)
                    if (anchored && prel >= R) {
                        err |= ERRB_RING;
                    } else {
                        const i64 p = origin_pane + (i64)prel;
                        pmin = p < pmin ? p : pmin; pmax = p > pmax ? p : pmax;
                        gpane = p;
                        u32 slot;
                        if (DENSE) { slot = key < a.g.capacity ? (u32)key : SLOT_INVALID; if (slot == SLOT_INVALID) err |= ERRB_BAD_INPUT; }
                        else slot = dir_find_or_insert(a.dkeys, a.g.capacity, key, st);
                        if (slot != SLOT_INVALID) {
                            valid = true; idx = (u64)((u32)p & a.g.ring_mask) * a.g.capacity + slot;
                            if (gfired && prel < gfired) mark_dirty(a.g, st, (u32)p & a.g.ring_mask, slot);
                        }
                    }
                }
            }
            accumulate_pair_lean<V>(a.acc, valid, idx, v, a.g.agg_mask);  // A5
            gp_[e] = valid ? gpane : 0; gs_[e] = valid ? (u32)(idx % a.g.capacity) : 0; gmask |= valid ? (1u << e) : 0u;
        }
        if (a.lt_cap) {
            u64 row = block_append<INGEST_THREADS>(__popc(glate), &st->n_late, a.lt_cap, st);
            if (row != ~0ull) {
#pragma unroll
                for (int e = 0; e < INGEST_EPT; ++e)
                    if (glate & (1u << e)) { a.lt_ts[row] = ts_[e]; a.lt_key[row] = key_[e]; a.lt_val[row] = vb_[e]; ++row; }
            }
        }
        if (a.mlog_enabled) {
            u64 row = block_append<INGEST_THREADS>(__popc(gmask), &st->n_mlog, a.ml_cap, st);
            if (row != ~0ull) {
#pragma unroll
                for (int e = 0; e < INGEST_EPT; ++e)
                    if (gmask & (1u << e)) { a.ml_pane[row] = gp_[e]; a.ml_slot[row] = gs_[e]; a.ml_val[row] = vb_[e]; ++row; }
            }
        }
    }
    nea = warp_sum_u64(nea);
    for (int o = 16; o > 0; o >>= 1) err |= __shfl_xor_sync(0xffffffffu, err, o);
    if (a.window_enabled) { late = warp_sum_u64(late); max_ts = warp_max_i64(max_ts); pmin = warp_min_i64(pmin); pmax = warp_max_i64(pmax); }
    if ((threadIdx.x & 31) == 0) {
        if (nea) atomicAdd(&st->event_alerts, nea);
        if (err) atomicOr(&st->err_flags, err);
        if (a.window_enabled) {
            if (late) atomicAdd(&st->late_dropped, late);
            if (max_ts != I64_MIN_) atomicMax(&st->batch_max_ts, max_ts);
            if (pmin <= pmax) { atomicMin(&st->min_pane, pmin); atomicMax(&st->max_pane, pmax); }
        }
    }
}

// ---- plan: one thread turns "watermark_after" into a fire / expire schedule ---------------------
struct PlanArgs {
    DevState* st; WindowGeom g; i64 bound; i64 watermark_after; u64 n_events; int window_enabled; int is_auto;
    int is_keep;      // FMQ_WM_KEEP: bookkeeping only, the watermark stays where it is
    const u64* seg_n; u32 nseg; u64 seg_cap;   // segmented input: the event count lives in device memory (n_events is then ignored)
    u32* combine_hint;                         // pinned host word: the skew decision, for the host's choice of ingest kernel
    int overlap;      // combiner mode with fire / retire on a second stream: the next ingest starts before the expired panes are cleared
};

__global__ void plan_kernel(const PlanArgs a) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    DevState* st = a.st;
    u64 n_events = a.n_events;
    if (a.seg_n) { n_events = 0; for (u32 s = 0; s < a.nseg; ++s) n_events += a.seg_n[s] < a.seg_cap ? a.seg_n[s] : a.seg_cap; }
    st->events_in += n_events;
    st->fire_count = 0; st->expire_lo = 1; st->expire_hi = 0; st->pending_clear = 0;
    st->n_dirty = 0;                                           // the re-fire pass of this batch has run (it precedes the plan)
    if (!a.window_enabled) return;
    // skew adaptation for the next batch: pre-aggregate inside warps when > 3 % of the sampled lanes
    // shared their cell with another lane of the same warp
    if (st->dup_sampled >= 4096) {
        st->combine_dups = (st->dup_found * 32 > st->dup_sampled) ? 1u : 0u; st->dup_sampled = 0; st->dup_found = 0;
        if (a.combine_hint) *a.combine_hint = st->combine_dups;
    }
    // A1: BoundedOutOfOrdernessTimestampExtractor (chapter3/README.md:379-396)
    if (st->batch_max_ts > st->cur_max_ts) st->cur_max_ts = st->batch_max_ts;
    st->batch_max_ts = I64_MIN_;
    i64 target = a.watermark_after;
    if (a.is_auto) {
        const i64 pot = st->cur_max_ts - a.bound;
        if (pot >= st->last_emitted_wm) st->last_emitted_wm = pot;
        target = st->last_emitted_wm;
    }
    // ring validity: every pane that holds data must fit the ring at once
    const bool has_data = st->min_pane <= st->max_pane;
    if (has_data && (st->max_pane - st->min_pane) >= (i64)(a.g.ring_mask + 1)) st->err_flags |= ERRB_RING;
    if (a.is_keep) return;                                     // no watermark came with this batch
    if (target <= st->wm) return;                              // watermarks never go backwards
    const i64 wm_old = st->wm;
    st->wm = target;
    // a watermark this close to INT64_MIN cannot make anything late or fire anything (ts >= 0) and would
    // overflow the pane arithmetic below: it is recorded and otherwise leaves the stream un-anchored
    const i64 wm_floor = I64_MIN_ + a.g.size + a.g.off + a.g.slide + a.g.lateness + 2;
    if (target < wm_floor) return;
    const i64 size = a.g.size, slide = a.g.slide, off = a.g.off;
    const bool final_wm = (target == I64_MAX_);
    // first pane that stays live: p*slide + off + size - 1 > target
    i64 new_lo;
    if (final_wm) new_lo = has_data ? st->max_pane + 1 : 0;
    else          new_lo = floor_div(target - a.g.lateness + 1 - size - off, slide) + 1;   // allowedLateness keeps a window's panes that much longer
    if (has_data) {
        // windows whose end-1 lies in (wm_old, target], clipped to windows that can hold data
        const i64 e_data_lo = st->min_pane * slide + off + slide;      // window whose last pane is min_pane
        const i64 e_data_hi = st->max_pane * slide + off + size;       // window whose first pane is max_pane
        i64 e_lo = e_data_lo, e_hi = e_data_hi;
        if (wm_old >= wm_floor) {
            const i64 e_wm_lo = (floor_div(wm_old + 1 - off, slide) + 1) * slide + off;   // smallest end > wm_old + 1
            if (e_wm_lo > e_lo) e_lo = e_wm_lo;
        }
        if (!final_wm) {
            const i64 e_wm_hi = floor_div(target + 1 - off, slide) * slide + off;          // largest end <= target + 1
            if (e_wm_hi < e_hi) e_hi = e_wm_hi;
        }
        if (e_hi >= e_lo) { st->fire_first_end = e_lo; st->fire_count = (e_hi - e_lo) / slide + 1; }
        // panes whose last window has now fired
        const i64 x_hi = (new_lo - 1 < st->max_pane) ? new_lo - 1 : st->max_pane;
        if (x_hi >= st->min_pane) {
            st->expire_lo = st->min_pane; st->expire_hi = x_hi;
            if (a.overlap) { const i64 cnt = x_hi - st->min_pane + 1; st->pending_clear = cnt > (i64)a.g.ring_mask ? a.g.ring_mask + 1 : (u32)cnt; }
        }
    }
    st->anchored = 1;
    st->origin_pane = new_lo;
    st->origin_ts = final_wm ? I64_MAX_ : new_lo * slide + off;
}

// after fire: retire the expired panes from the live range (runs as the tail of fire's stream order)
__global__ void commit_kernel(DevState* st) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (st->expire_lo <= st->expire_hi) {
        st->min_pane = st->expire_hi + 1;
        if (st->min_pane > st->max_pane) { st->min_pane = I64_MAX_; st->max_pane = I64_MIN_; }
    }
    st->fire_count = 0; st->expire_lo = 1; st->expire_hi = 0;
}
// Overlapped combiner mode: the next batch's ingest kernel is already updating min_pane / max_pane with
// atomics while this runs, so the live range is only ever moved with an atomic (and never reset to
// "empty": a stale lower bound is corrected by pack_scalars_kernel's clamp to the pane origin); the
// plan of the next batch resets the fire / expire fields.
__global__ void commit_overlapped_kernel(DevState* st) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    if (st->expire_lo <= st->expire_hi) atomicMax(reinterpret_cast<long long*>(&st->min_pane), (long long)(st->expire_hi + 1));
}

// ---- fire: thread per key slot; combine panes, emit, evaluate the alert, clear expired panes -----
struct FireArgs {
    DevState* st; const u64* dkeys; Acc32* acc; WindowGeom g;
    int dense_keys;             // 1: key == slot, no directory
    // combiner mode (multi-GPU): every rank accumulates its share of the events into a
    // full-size local table; a window fires on the rank that owns the slot range [slot_lo, slot_hi)
    // by reading that range of EVERY rank's table over NVLink peer memory and combining the cells.
    int world;                  // 1 = single GPU
    int clear_in_fire;          // 1: the firing thread also clears expired panes (single GPU)
    int refire;                 // allowedLateness pass (runs BEFORE the plan): the windows that had fired under the watermark still
                                // in force and are still kept fire again for the keys whose panes took records in this batch
    u64 slot_lo, slot_hi;
    const Acc32* peer_acc[FMQ_MAX_PEERS];
    int emit_fired; int walert_source, walert_op; double mul, d0, d1, d2, thr;
    u64* f_key; i64* f_wend; i64* f_count; u64* f_sum; u64* f_min; u64* f_max; u64 f_cap;
    u64* wa_key; i64* wa_wend; double* wa_val; u64 wa_cap;
};

#ifndef FMQ_FIRE_THREADS
#define FMQ_FIRE_THREADS 256
#endif
#ifndef FMQ_FIRE_SPT
#define FMQ_FIRE_SPT 2
#endif
#ifndef FMQ_FIRE_MINB
#define FMQ_FIRE_MINB 4
#endif
#ifndef FMQ_FIRE_GRID
#define FMQ_FIRE_GRID 8
#endif
constexpr int FIRE_THREADS = FMQ_FIRE_THREADS;

// [batch max ts, -(min pane), max pane, 0]: the three batch scalars every rank must agree on,
// laid out so that ONE all-reduce(MAX) over int64 makes them global (combiner mode).
__global__ void pack_scalars_kernel(const DevState* st, i64* x) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    x[0] = st->batch_max_ts;
    // panes before the origin have expired (plan_kernel): with fire / retire overlapped, the commit that
    // moves min_pane past them may not have run yet
    i64 lo = st->min_pane;
    if (st->anchored && lo < st->origin_pane) lo = st->origin_pane;
    x[1] = (lo <= st->max_pane) ? -lo : I64_MIN_;
    x[2] = (lo <= st->max_pane) ? st->max_pane : I64_MIN_;
    x[3] = 0;
}
__global__ void unpack_scalars_kernel(DevState* st, const i64* x) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    st->batch_max_ts = x[0];
    if (x[1] != I64_MIN_) { st->min_pane = -x[1]; st->max_pane = x[2]; }
}
// combiner mode: clear the panes whose last window has fired everywhere (runs after the barrier that
// follows the peer-reading fire), then retire them like commit_kernel does
__global__ void __launch_bounds__(256) retire_kernel(DevState* st, Acc32* acc, u64 capacity, u32 ring_mask) {
    const i64 x_lo = st->expire_lo, x_hi = st->expire_hi;
    if (x_lo > x_hi) return;
    const ulonglong4 z = make_ulonglong4(0, 0, 0, 0);
    for (i64 p = x_lo; p <= x_hi; ++p) {
        ulonglong4* pane = reinterpret_cast<ulonglong4*>(acc + (u64)((u32)p & ring_mask) * capacity);
        for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < capacity; i += (u64)gridDim.x * blockDim.x) pane[i] = z;
    }
}

// Geometry of the windows being fired is the same for every key: a few threads work it out once per block
// (64-bit divisions) and leave it in shared memory; ncu of the first version counted 630 instructions per key slot,
// most of them this arithmetic repeated by every thread.
constexpr int FIRE_MAXW = 32;             // windows per pass over the slots
constexpr int FIRE_SPT = FMQ_FIRE_SPT;     // key slots per thread: both slots' cells are requested before either is used
struct FireWin { i64 wend; u32 p0; u32 np; };     // first pane (low 32 bits: only its ring index is needed) and number of panes holding data

template <typename V, bool PEERS>
__global__ void __launch_bounds__(FIRE_THREADS, FMQ_FIRE_MINB) fire_kernel(const __grid_constant__ FireArgs a) {
    __shared__ FireWin wins[FIRE_MAXW];
    DevState* st = a.st;
    const i64 size = a.g.size, slide = a.g.slide, off = a.g.off;
    const i64 pmin = st->min_pane, pmax = st->max_pane;
    i64 nfire = st->fire_count, first_end = st->fire_first_end;
    const i64 x_lo = st->expire_lo, x_hi = st->expire_hi;
    if (a.refire) {
        // window ends e = off (mod slide) with e - 1 <= wm (fired) and e - 1 + lateness > wm (still kept), clipped to windows that can hold data
        if (st->n_dirty == 0 || !st->anchored || pmin > pmax) return;
        const i64 wm = st->wm;
        i64 e_hi = floor_div(wm + 1 - off, slide) * slide + off;
        i64 e_lo = (floor_div(wm + 1 - a.g.lateness - off, slide) + 1) * slide + off;
        const i64 d_lo = pmin * slide + off + slide, d_hi = pmax * slide + off + size;
        if (d_lo > e_lo) e_lo = d_lo;
        if (d_hi < e_hi) e_hi = d_hi;
        if (e_hi < e_lo) return;
        first_end = e_lo; nfire = (e_hi - e_lo) / slide + 1;
    } else if (nfire == 0 && (x_lo > x_hi || !a.clear_in_fire)) return;
    const u64 C = a.g.capacity;
    constexpr bool is_f64 = IsF64<V>::value;
    const bool want_mx = a.g.agg_mask & 8u, want_mn = a.g.agg_mask & 4u;
    u64 fired_local = 0, alerts_local = 0;
    const u64 nslots = a.slot_hi - a.slot_lo;
    const u64 tile = (u64)FIRE_THREADS * FIRE_SPT;
    const u64 ntiles = (nslots + tile - 1) / tile;                   // block-uniform trip count
    for (i64 j0 = 0; j0 < nfire || j0 == 0; j0 += FIRE_MAXW) {
        const int nw = (int)((nfire - j0) < FIRE_MAXW ? (nfire - j0) : FIRE_MAXW);
        const bool last_pass = j0 + FIRE_MAXW >= nfire;
        __syncthreads();
        if ((int)threadIdx.x < nw) {
            const i64 wend = first_end + (j0 + threadIdx.x) * slide;
            // panes of [wend - size, wend): absolute indices, clipped to the panes holding data
            i64 p0 = (wend - size - off) / slide, p1 = (wend - off) / slide - 1;   // exact: wend == off (mod slide)
            if (p0 < pmin) p0 = pmin;
            if (p1 > pmax) p1 = pmax;
            wins[threadIdx.x] = FireWin{wend, (u32)p0, p1 >= p0 ? (u32)(p1 - p0 + 1) : 0u};
        }
        __syncthreads();
        for (u64 t = blockIdx.x; t < ntiles; t += gridDim.x) {
            u64 slot_[FIRE_SPT], key_[FIRE_SPT]; bool live_[FIRE_SPT];
#pragma unroll
            for (int q = 0; q < FIRE_SPT; ++q) {
                slot_[q] = a.slot_lo + t * tile + (u64)q * FIRE_THREADS + threadIdx.x;
                key_[q] = slot_[q] < a.slot_hi ? (a.dense_keys ? slot_[q] : a.dkeys[slot_[q]]) : KEY_EMPTY;
                live_[q] = key_[q] != KEY_EMPTY;
            }
            for (int j = 0; j < nw; ++j) {
                const FireWin w = wins[j];
                i64 count_[FIRE_SPT]; u64 sum_[FIRE_SPT], mx_[FIRE_SPT], mn_[FIRE_SPT]; double v_[FIRE_SPT]; bool fired_[FIRE_SPT], hit_[FIRE_SPT];
                if (!PEERS && w.np == 1) {
                    // tumbling windows (and the common sliding case of one pane with data): one cell per slot, all requested first
                    ulonglong4 c_[FIRE_SPT];
#pragma unroll
                    for (int q = 0; q < FIRE_SPT; ++q)
                        c_[q] = live_[q] ? *reinterpret_cast<const ulonglong4*>(a.acc + (u64)(w.p0 & a.g.ring_mask) * C + slot_[q]) : make_ulonglong4(0, 0, 0, 0);
#pragma unroll
                    for (int q = 0; q < FIRE_SPT; ++q) {
                        count_[q] = is_f64 ? (i64)__longlong_as_double((i64)c_[q].x) : (i64)c_[q].x;
                        sum_[q] = c_[q].y; mx_[q] = c_[q].z; mn_[q] = c_[q].w;
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < FIRE_SPT; ++q) {
                        u64 cnt_i = 0; double cnt_d = 0.0; u64 sum_i = 0; double sum_d = 0.0; u64 mx = 0, mn = 0;
                        if (live_[q]) {
                            for (u32 pi = 0; pi < w.np; ++pi) {
                                const u64 cell = (u64)((w.p0 + pi) & a.g.ring_mask) * C + slot_[q];
                                if constexpr (!PEERS) {
                                    const ulonglong4 c = *reinterpret_cast<const ulonglong4*>(a.acc + cell);
                                    if (is_f64) { cnt_d += __longlong_as_double((i64)c.x); sum_d += __longlong_as_double((i64)c.y); }
                                    else        { cnt_i += c.x; sum_i += c.y; }
                                    mx = c.z > mx ? c.z : mx; mn = c.w > mn ? c.w : mn;
                                } else {
                                    // every peer's cell is requested before the first one is used: over NVLink a load takes
                                    // microseconds, and a dependent chain of `world` of them per thread was what made the
                                    // fire chain grow with the GPU count
                                    ulonglong4 c[FMQ_MAX_PEERS];
#pragma unroll
                                    for (int r = 0; r < FMQ_MAX_PEERS; ++r)
                                        c[r] = r < a.world ? *reinterpret_cast<const ulonglong4*>(a.peer_acc[r] + cell) : make_ulonglong4(0, 0, 0, 0);
#pragma unroll
                                    for (int r = 0; r < FMQ_MAX_PEERS; ++r) {          // fixed rank order: deterministic double sums
                                        if (r >= a.world) break;
                                        if (is_f64) { cnt_d += __longlong_as_double((i64)c[r].x); sum_d += __longlong_as_double((i64)c[r].y); }
                                        else        { cnt_i += c[r].x; sum_i += c[r].y; }
                                        mx = c[r].z > mx ? c[r].z : mx; mn = c[r].w > mn ? c[r].w : mn;
                                    }
                                }
                            }
                        }
                        count_[q] = is_f64 ? (i64)cnt_d : (i64)cnt_i;
                        sum_[q] = is_f64 ? (u64)__double_as_longlong(sum_d) : sum_i; mx_[q] = mx; mn_[q] = mn;
                    }
                }
                u32 nf = 0, na = 0;
#pragma unroll
                for (int q = 0; q < FIRE_SPT; ++q) {
                    fired_[q] = live_[q] && count_[q] > 0;
                    if (a.refire && fired_[q]) {                  // only the keys whose panes of this window took records in this batch
                        bool any = false;
                        for (u32 pi = 0; pi < w.np; ++pi)
                            any = any || ((a.g.dirty[(u64)((w.p0 + pi) & a.g.ring_mask) * a.g.dirty_words + (slot_[q] >> 5)] >> (slot_[q] & 31)) & 1u);
                        fired_[q] = any;
                    }
                    // A6-ii/iii: ((double)sum * 8.0 / 60 / 1024 / 1024) < 100, evaluated left to right
                    double v = 0.0; bool hit = false;
                    if (a.walert_source != 0 && fired_[q]) {
                        if (a.walert_source == 2) {
                            v = Codec<V>::as_double(sum_[q]) / (double)count_[q];            // ComputeCpuAvg.java:49
                        } else {
                            double src;
                            if (a.walert_source == 1) src = Codec<V>::as_double(sum_[q]);
                            else if (a.walert_source == 3) src = Codec<V>::as_double(Codec<V>::dec_min(mn_[q]));
                            else if (a.walert_source == 4) src = Codec<V>::as_double(Codec<V>::dec_max(mx_[q]));
                            else src = (double)count_[q];
                            v = src * a.mul / a.d0 / a.d1 / a.d2;
                        }
                        hit = cmp_op(v, a.walert_op, a.thr);
                    }
                    v_[q] = v; hit_[q] = hit;
                    nf += fired_[q] ? 1u : 0u; na += hit ? 1u : 0u;
                }
                fired_local += nf; alerts_local += na;
                // A5f: compaction of the fired rows and of the alert rows behind one pair of block barriers
                u64 row_f = ~0ull, row_a = ~0ull;
                block_append2n<FIRE_THREADS>(a.emit_fired ? nf : 0u, a.walert_source != 0 ? na : 0u, &st->n_fired, a.f_cap, &st->n_walert, a.wa_cap, st, row_f, row_a);
#pragma unroll
                for (int q = 0; q < FIRE_SPT; ++q) {
                    if (fired_[q] && row_f != ~0ull) {
                        a.f_key[row_f] = key_[q]; a.f_wend[row_f] = w.wend; a.f_count[row_f] = count_[q]; a.f_sum[row_f] = sum_[q];
                        a.f_min[row_f] = want_mn ? Codec<V>::dec_min(mn_[q]) : 0;
                        a.f_max[row_f] = want_mx ? Codec<V>::dec_max(mx_[q]) : 0;
                        ++row_f;
                    }
                    if (hit_[q] && row_a != ~0ull) { a.wa_key[row_a] = key_[q]; a.wa_wend[row_a] = w.wend; a.wa_val[row_a] = v_[q]; ++row_a; }
                }
            }
            if (a.refire && last_pass) {                          // the marks of this batch are spent
                const i64 r0 = (first_end - size - off) / slide, r1 = (first_end + (nfire - 1) * slide - off) / slide - 1;
                const i64 c0 = r0 < pmin ? pmin : r0, c1 = r1 > pmax ? pmax : r1;
#pragma unroll
                for (int q = 0; q < FIRE_SPT; ++q)
                    if (live_[q]) for (i64 p = c0; p <= c1; ++p) {
                        u32* wd = &a.g.dirty[(u64)((u32)p & a.g.ring_mask) * a.g.dirty_words + (slot_[q] >> 5)];
                        const u32 bit = 1u << (slot_[q] & 31);
                        if (*wd & bit) atomicAnd(wd, ~bit);
                    }
            }
            // cleanup: WindowOperator.clearAllState for panes whose last window has fired
            if (a.clear_in_fire && !a.refire && last_pass) {
                const ulonglong4 z = make_ulonglong4(0, 0, 0, 0);
#pragma unroll
                for (int q = 0; q < FIRE_SPT; ++q)
                    if (live_[q]) for (i64 p = x_lo; p <= x_hi; ++p)
                        *reinterpret_cast<ulonglong4*>(&a.acc[(u64)((u32)p & a.g.ring_mask) * C + slot_[q]]) = z;
            }
        }
    }
    fired_local = warp_sum_u64(fired_local); alerts_local = warp_sum_u64(alerts_local);
    if ((threadIdx.x & 31) == 0) {
        if (fired_local) atomicAdd(&st->windows_fired, fired_local);
        if (alerts_local) atomicAdd(&st->window_alerts, alerts_local);
    }
}

}  // namespace fmq
