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
};

struct IngestArgs {
    const i64* ts; const u64* key; const void* val; size_t n;
    DevState* st;
    u64* dkeys;
    Acc32* acc;
    WindowGeom g;
    int window_enabled;
    int vec32;                  // all three columns are 32-byte aligned
    int efilter_enabled, efilter_op; double efilter_thr;
    i64* ea_ts; u64* ea_key; u64* ea_val; u64 ea_cap;
};

// Two adjacent lanes share one event per round: the even lane updates the first 8 B of a pair,
// the odd lane the second, so each RED instruction carries both halves of a 16 B pair in one
// 32 B sector.  Round 0 serves the even lane's event, round 1 the odd lane's.
// Must be reached by all 32 lanes.
template <typename V>
__device__ __forceinline__ void accumulate_pair(Acc32* __restrict__ acc, bool valid, u64 idx, V val, u32 agg_mask) {
    const unsigned lane = threadIdx.x & 31;
    const bool odd = lane & 1;
    u64 vbits; memcpy(&vbits, &val, 8);
    u64 idx_p = __shfl_xor_sync(0xffffffffu, idx, 1);
    u64 vb_p = __shfl_xor_sync(0xffffffffu, vbits, 1);
    bool valid_p = __shfl_xor_sync(0xffffffffu, (int)valid, 1) != 0;
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
            double* p = reinterpret_cast<double*>(odd ? &a->sum : &a->cnt);
            double add; if (odd) memcpy(&add, &vb, 8); else add = 1.0;
            atomicAdd(p, add);                                // red.global.add.f64
        } else {
            u64* p = odd ? &a->sum : &a->cnt;
            atomicAdd(p, odd ? vb : 1ull);                    // red.global.add.u64 (wraps like a Java long)
        }
        if (odd ? want_mn : want_mx) {
            u64* p = odd ? &a->mn : &a->mx;
            atomicMax(p, odd ? Codec<V>::enc_min(v) : Codec<V>::enc_max(v));   // red.global.max.u64
        }
    }
}

constexpr int INGEST_THREADS = 256;
constexpr int INGEST_EPT = 4;          // events per thread per tile (one 32 B load per column)

template <typename V>
__global__ void __launch_bounds__(INGEST_THREADS) ingest_kernel(const __grid_constant__ IngestArgs a) {
    DevState* st = a.st;
    const int anchored = st->anchored;
    const i64 origin_ts = anchored ? st->origin_ts : (a.g.off - a.g.slide);
    const i64 origin_pane = anchored ? st->origin_pane : -1;
    const u32 R = a.g.ring_mask + 1;
    const V* __restrict__ val = static_cast<const V*>(a.val);

    u64 late = 0, nea = 0; i64 max_ts = I64_MIN_; i64 pmin = I64_MAX_, pmax = I64_MIN_; u32 err = 0;

    const size_t tile = (size_t)INGEST_THREADS * INGEST_EPT;
    const size_t ntiles = (a.n + tile - 1) / tile;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        // each thread owns 4 consecutive events: one 256-bit load per column (3 x 32 B in flight)
        const size_t i = t * tile + (size_t)threadIdx.x * INGEST_EPT;
        i64 ts_[INGEST_EPT]; u64 key_[INGEST_EPT]; u64 vb_[INGEST_EPT]; bool in_[INGEST_EPT];
        if (a.vec32 && i + INGEST_EPT <= a.n) {
            const ulonglong4 x = ldg_stream_v4(a.ts + i), y = ldg_stream_v4(a.key + i), z = ldg_stream_v4(val + i);
            ts_[0] = (i64)x.x; ts_[1] = (i64)x.y; ts_[2] = (i64)x.z; ts_[3] = (i64)x.w;
            key_[0] = y.x; key_[1] = y.y; key_[2] = y.z; key_[3] = y.w;
            vb_[0] = z.x; vb_[1] = z.y; vb_[2] = z.z; vb_[3] = z.w;
            in_[0] = in_[1] = in_[2] = in_[3] = true;
        } else {
#pragma unroll
            for (int e = 0; e < INGEST_EPT; ++e) {
                in_[e] = i + e < a.n;
                ts_[e] = in_[e] ? (i64)ldg_stream(a.ts + i + e) : 0;
                key_[e] = in_[e] ? ldg_stream(a.key + i + e) : 0;
                vb_[e] = in_[e] ? ldg_stream(val + i + e) : 0;
            }
        }
        // ---- A6-i: per-event threshold filter + compaction (Main.java:31), one reservation per tile;
        //      rows keep arrival order inside the tile ----
        if (a.efilter_enabled) {
            u32 hits = 0;
#pragma unroll
            for (int e = 0; e < INGEST_EPT; ++e) {
                V v; memcpy(&v, &vb_[e], 8);
                const bool ok = in_[e] && ts_[e] >= 0 && key_[e] != KEY_EMPTY;
                if (ok && cmp_op((double)v, a.efilter_op, a.efilter_thr)) hits |= 1u << e;
            }
            const u32 nh = __popc(hits);
            nea += nh;
            u64 row = block_append<INGEST_THREADS>(nh, &st->n_ealert, a.ea_cap, st);
            if (row != ~0ull) {
#pragma unroll
                for (int e = 0; e < INGEST_EPT; ++e)
                    if (hits & (1u << e)) { a.ea_ts[row] = ts_[e]; a.ea_key[row] = key_[e]; a.ea_val[row] = vb_[e]; ++row; }
            }
        }
#pragma unroll
        for (int e = 0; e < INGEST_EPT; ++e) {
            const i64 ts = ts_[e]; const u64 key = key_[e];
            V v; memcpy(&v, &vb_[e], 8);
            bool ok = in_[e];
            if (ok && (ts < 0 || key == KEY_EMPTY)) { err |= ERRB_BAD_INPUT; ok = false; }
            if (!a.window_enabled) continue;
            bool valid = false; u64 idx = 0;
            if (ok) {
                max_ts = ts > max_ts ? ts : max_ts;                 // A1: currentMaxTimestamp
                const i64 d = ts - origin_ts;
                if (d < 0) {
                    late++;                                          // A3: every window of this record has fired
                } else {
                    // A2: pane index relative to the origin; 32-bit division on the common path
                    const u64 prel = (a.g.slide32 && (u64)d < 0x100000000ull) ? (u64)((u32)d / a.g.slide32)
                                                                              : (u64)d / (u64)a.g.slide;
                    if (anchored && prel >= R) {
                        err |= ERRB_RING;
                    } else {
                        const i64 p = origin_pane + (i64)prel;
                        pmin = p < pmin ? p : pmin; pmax = p > pmax ? p : pmax;
                        const u32 slot = dir_find_or_insert(a.dkeys, a.g.capacity, key, st);
                        if (slot != SLOT_INVALID) { valid = true; idx = (u64)((u32)p & a.g.ring_mask) * a.g.capacity + slot; }
                    }
                }
            }
            accumulate_pair<V>(a.acc, valid, idx, v, a.g.agg_mask);  // A5
        }
    }
    // ---- per-warp reduction of the batch scalars, one atomic each per warp ----
    if (a.window_enabled) {
        late = warp_sum_u64(late); max_ts = warp_max_i64(max_ts); pmin = warp_min_i64(pmin); pmax = warp_max_i64(pmax);
    }
    for (int o = 16; o > 0; o >>= 1) err |= __shfl_xor_sync(0xffffffffu, err, o);
    nea = warp_sum_u64(nea);
    if ((threadIdx.x & 31) == 0) {
        if (nea) atomicAdd(&st->event_alerts, nea);
        if (late) atomicAdd(&st->late_dropped, late);
        if (max_ts != I64_MIN_) atomicMax(&st->batch_max_ts, max_ts);
        if (pmin <= pmax) { atomicMin(&st->min_pane, pmin); atomicMax(&st->max_pane, pmax); }
        if (err) atomicOr(&st->err_flags, err);
    }
}

// ---- plan: one thread turns "watermark_after" into a fire / expire schedule ---------------------
struct PlanArgs {
    DevState* st; WindowGeom g; i64 bound; i64 watermark_after; u64 n_events; int window_enabled; int is_auto;
};

__global__ void plan_kernel(const PlanArgs a) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    DevState* st = a.st;
    st->events_in += a.n_events;
    st->fire_count = 0; st->expire_lo = 1; st->expire_hi = 0;
    if (!a.window_enabled) return;
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
    if (target <= st->wm) return;                              // watermarks never go backwards
    const i64 wm_old = st->wm;
    st->wm = target;
    const i64 size = a.g.size, slide = a.g.slide, off = a.g.off;
    const bool final_wm = (target == I64_MAX_);
    // first pane that stays live: p*slide + off + size - 1 > target
    i64 new_lo;
    if (final_wm) new_lo = has_data ? st->max_pane + 1 : 0;
    else          new_lo = floor_div(target + 1 - size - off, slide) + 1;
    if (has_data) {
        // windows whose end-1 lies in (wm_old, target], clipped to windows that can hold data
        const i64 e_data_lo = st->min_pane * slide + off + slide;      // window whose last pane is min_pane
        const i64 e_data_hi = st->max_pane * slide + off + size;       // window whose first pane is max_pane
        i64 e_lo = e_data_lo, e_hi = e_data_hi;
        if (wm_old != I64_MIN_) {
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
        if (x_hi >= st->min_pane) { st->expire_lo = st->min_pane; st->expire_hi = x_hi; }
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

// ---- fire: thread per key slot; combine panes, emit, evaluate the alert, clear expired panes -----
struct FireArgs {
    DevState* st; const u64* dkeys; Acc32* acc; WindowGeom g;
    int emit_fired; int walert_source, walert_op; double mul, d0, d1, d2, thr;
    u64* f_key; i64* f_wend; i64* f_count; u64* f_sum; u64* f_min; u64* f_max; u64 f_cap;
    u64* wa_key; i64* wa_wend; double* wa_val; u64 wa_cap;
};

constexpr int FIRE_THREADS = 256;

template <typename V>
__global__ void __launch_bounds__(FIRE_THREADS) fire_kernel(const __grid_constant__ FireArgs a) {
    DevState* st = a.st;
    const i64 nfire = st->fire_count;
    const i64 x_lo = st->expire_lo, x_hi = st->expire_hi;
    if (nfire == 0 && x_lo > x_hi) return;
    const i64 first_end = st->fire_first_end;
    const i64 pmin = st->min_pane, pmax = st->max_pane;
    const i64 size = a.g.size, slide = a.g.slide, off = a.g.off;
    const u64 C = a.g.capacity;
    constexpr bool is_f64 = IsF64<V>::value;
    u64 fired_local = 0, alerts_local = 0;
    const u64 nslots_rounded = (C + FIRE_THREADS - 1) / FIRE_THREADS * FIRE_THREADS;   // block-uniform trip count
    for (u64 slot = (u64)blockIdx.x * blockDim.x + threadIdx.x; slot < nslots_rounded; slot += (u64)gridDim.x * blockDim.x) {
        const bool live = slot < C && a.dkeys[slot] != KEY_EMPTY;
        const u64 key = live ? a.dkeys[slot] : 0;
        for (i64 j = 0; j < nfire; ++j) {
            const i64 wend = first_end + j * slide;
            // panes of [wend - size, wend): absolute indices, clipped to the panes holding data
            i64 p0 = (wend - size - off) / slide, p1 = (wend - off) / slide - 1;   // exact: wend == off (mod slide)
            if (p0 < pmin) p0 = pmin;
            if (p1 > pmax) p1 = pmax;
            u64 cnt_i = 0; double cnt_d = 0.0; u64 sum_i = 0; double sum_d = 0.0; u64 mx = 0, mn = 0;
            if (live) {
                for (i64 p = p0; p <= p1; ++p) {
                    const Acc32 c = a.acc[(u64)((u32)p & a.g.ring_mask) * C + slot];
                    if (is_f64) { cnt_d += __longlong_as_double((i64)c.cnt); sum_d += __longlong_as_double((i64)c.sum); }
                    else        { cnt_i += c.cnt; sum_i += c.sum; }
                    mx = c.mx > mx ? c.mx : mx; mn = c.mn > mn ? c.mn : mn;
                }
            }
            const i64 count = is_f64 ? (i64)cnt_d : (i64)cnt_i;
            const bool fired = live && count > 0;
            const u64 sum_bits = is_f64 ? (u64)__double_as_longlong(sum_d) : sum_i;
            if (fired) fired_local++;
            if (a.emit_fired) {
                const u64 row = block_append<FIRE_THREADS>(fired ? 1u : 0u, &st->n_fired, a.f_cap, st);
                if (fired && row != ~0ull) {
                    a.f_key[row] = key; a.f_wend[row] = wend; a.f_count[row] = count; a.f_sum[row] = sum_bits;
                    a.f_min[row] = (a.g.agg_mask & 4u) ? Codec<V>::dec_min(mn) : 0;
                    a.f_max[row] = (a.g.agg_mask & 8u) ? Codec<V>::dec_max(mx) : 0;
                }
            }
            if (a.walert_source != 0) {
                // A6-ii/iii: ((double)sum * 8.0 / 60 / 1024 / 1024) < 100, evaluated left to right
                double v = 0.0; bool hit = false;
                if (fired) {
                    if (a.walert_source == 2) {
                        v = Codec<V>::as_double(sum_bits) / (double)count;            // ComputeCpuAvg.java:49
                    } else {
                        double src;
                        if (a.walert_source == 1) src = Codec<V>::as_double(sum_bits);
                        else if (a.walert_source == 3) src = Codec<V>::as_double(Codec<V>::dec_min(mn));
                        else if (a.walert_source == 4) src = Codec<V>::as_double(Codec<V>::dec_max(mx));
                        else src = (double)count;
                        v = src * a.mul / a.d0 / a.d1 / a.d2;
                    }
                    hit = cmp_op(v, a.walert_op, a.thr);
                }
                if (hit) alerts_local++;
                const u64 row = block_append<FIRE_THREADS>(hit ? 1u : 0u, &st->n_walert, a.wa_cap, st);
                if (hit && row != ~0ull) { a.wa_key[row] = key; a.wa_wend[row] = wend; a.wa_val[row] = v; }
            }
        }
        // cleanup: WindowOperator.clearAllState for panes whose last window has fired
        if (live) {
            const ulonglong4 z = make_ulonglong4(0, 0, 0, 0);
            for (i64 p = x_lo; p <= x_hi; ++p)
                *reinterpret_cast<ulonglong4*>(&a.acc[(u64)((u32)p & a.g.ring_mask) * C + slot]) = z;
        }
    }
    fired_local = warp_sum_u64(fired_local); alerts_local = warp_sum_u64(alerts_local);
    if ((threadIdx.x & 31) == 0) {
        if (fired_local) atomicAdd(&st->windows_fired, fired_local);
        if (alerts_local) atomicAdd(&st->window_alerts, alerts_local);
    }
}

}  // namespace fmq
