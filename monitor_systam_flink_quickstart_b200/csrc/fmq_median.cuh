// Non-incremental window function: per-(key, window) median,
// chapter2/.../ComputeCpuMiddle.java:34-47 — "buffer every value of the window, sort, take the middle".
// The tutorial itself warns that this path is slow (chapter2/README.md:231); it is built from the
// pieces that already exist: the ingest kernels log (pane, slot, value) for every accepted record;
// when windows fire, the records of ALL of them are selected from the log in one pass, sorted by
// ((window, slot), value) with the packed radix passes (value in Double.compareTo order, like
// Collections.sort), and one thread per (window, key) picks the middle element(s).
#pragma once
#include "fmq_device.cuh"

namespace fmq {

constexpr int MED_THREADS = 256;

// All windows [j0, j0 + nj) of the fire plan at once: a record of pane p belongs to the windows j with
// p - P0 - (k - 1) <= j <= p - P0 (P0 = first pane of window 0, k = panes per window); every such (record, window)
// pair becomes one item ((j - j0) << slot_bits | slot) << 32 | index, vals[index] = the value.  No host round trip
// per window: the plan is read on the device.
__global__ void __launch_bounds__(MED_THREADS) median_select_all_kernel(const i64* __restrict__ ml_pane, const u32* __restrict__ ml_slot, const u64* __restrict__ ml_val,
                                                                       DevState* st, i64 size, i64 slide, i64 off, i64 j0, i64 nj, u32 slot_bits,
                                                                       u64* __restrict__ items, u64* __restrict__ vals, u64 cap) {
    const u64 n = st->n_mlog;
    const i64 nfire = st->fire_count, k = size / slide;
    const i64 P0 = (st->fire_first_end - size - off) / slide;             // exact: window ends are off (mod slide)
    const i64 jhi_all = (j0 + nj < nfire ? j0 + nj : nfire) - 1;
    const u64 nr = (n + MED_THREADS - 1) / MED_THREADS * MED_THREADS;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < nr; i += (u64)gridDim.x * blockDim.x) {
        i64 ja = 1, jb = 0;
        if (i < n) {
            const i64 r = ml_pane[i] - P0;
            ja = r - (k - 1) > j0 ? r - (k - 1) : j0; jb = r < jhi_all ? r : jhi_all;
        }
        const u32 cnt = jb >= ja ? (u32)(jb - ja + 1) : 0u;
        u64 row = block_append<MED_THREADS>(cnt, &st->n_mwork, cap, st);
        if (cnt && row != ~0ull) {
            const u64 slot = ml_slot[i], v = ml_val[i];
            for (i64 j = ja; j <= jb; ++j, ++row) { items[row] = ((((u64)(j - j0) << slot_bits) | slot) << 32) | (u32)row; vals[row] = v; }
        }
    }
}

// keep only the records of panes > x_hi (their windows have not all fired yet)
__global__ void __launch_bounds__(MED_THREADS) median_compact_kernel(const i64* __restrict__ ip, const u32* __restrict__ is, const u64* __restrict__ iv, u64 n,
                                                                    DevState* st, i64 x_hi, i64* __restrict__ op, u32* __restrict__ os, u64* __restrict__ ov, u64 cap) {
    const u64 nr = (n + MED_THREADS - 1) / MED_THREADS * MED_THREADS;
    for (u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x; i < nr; i += (u64)gridDim.x * blockDim.x) {
        const bool keep = i < n && ip[i] > x_hi;
        const u64 row = block_append<MED_THREADS>(keep ? 1u : 0u, &st->n_mwork, cap, st);
        if (keep && row != ~0ull) { op[row] = ip[i]; os[row] = is[i]; ov[row] = iv[i]; }
    }
}

struct MedianEmitArgs {
    const u64* items; const u64* vals; u64 n;          // sorted by ((window, slot), value); unused tail items are all ones
    const u64* dkeys; int dense_keys; i64 window_end; DevState* st;
    u64* m_key; i64* m_wend; double* m_val; u64 m_cap;
    u32 slot_bits; i64 slide;                          // window of an item = window_end + (group key >> slot_bits) * slide
};

// thread t handles the run that starts at t: its end by binary search on the slot, then
// odd -> v[n/2]; even -> (v[n/2] + v[n/2 - 1]) / 2   (ComputeCpuMiddle.java:43-47)
__global__ void __launch_bounds__(MED_THREADS) median_emit_kernel(const MedianEmitArgs a) {
    const u64 nr = (a.n + MED_THREADS - 1) / MED_THREADS * MED_THREADS;
    for (u64 t = (u64)blockIdx.x * blockDim.x + threadIdx.x; t < nr; t += (u64)gridDim.x * blockDim.x) {
        bool head = false; u32 slot = 0; double med = 0.0;
        if (t < a.n && (u32)(a.items[t] >> 32) != 0xFFFFFFFFu) {
            slot = (u32)(a.items[t] >> 32);
            head = (t == 0) || ((u32)(a.items[t - 1] >> 32) != slot);
            if (head) {
                u64 lo = t + 1, hi = a.n;                               // first position whose slot differs
                while (lo < hi) { const u64 mid = (lo + hi) >> 1; if ((u32)(a.items[mid] >> 32) == slot) lo = mid + 1; else hi = mid; }
                const u64 len = lo - t;
                const double hi_v = __longlong_as_double((i64)a.vals[(u32)a.items[t + len / 2]]);
                if (len & 1) med = hi_v;
                else med = (hi_v + __longlong_as_double((i64)a.vals[(u32)a.items[t + len / 2 - 1]])) / 2;
            }
        }
        const u64 row = block_append<MED_THREADS>(head ? 1u : 0u, &a.st->n_median, a.m_cap, a.st);
        if (head && row != ~0ull) {
            const u32 s = slot & ((1u << a.slot_bits) - 1u);
            a.m_key[row] = a.dense_keys ? (u64)s : a.dkeys[s]; a.m_wend[row] = a.window_end + (i64)(slot >> a.slot_bits) * a.slide; a.m_val[row] = med;
        }
    }
}

}  // namespace fmq
