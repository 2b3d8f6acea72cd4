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

constexpr int MED_RPT = 8;          // records per thread per reservation: one block-wide scan and ONE atomic per 2048 records

// All windows [j0, j0 + nj) of the fire plan at once: a record of pane p belongs to the windows j with
// p - P0 - (k - 1) <= j <= p - P0 (P0 = first pane of window 0, k = panes per window); every such (record, window)
// pair becomes one item ((j - j0) << slot_bits | slot) << 32 | index, vals[index] = the value.  No host round trip
// per window: the plan is read on the device.  (Rows of one thread are consecutive; the order of items is free.)
__global__ void __launch_bounds__(MED_THREADS) median_select_all_kernel(const i64* __restrict__ ml_pane, const u32* __restrict__ ml_slot, const u64* __restrict__ ml_val,
                                                                       DevState* st, i64 size, i64 slide, i64 off, i64 j0, i64 nj, u32 slot_bits,
                                                                       u64* __restrict__ items, u64* __restrict__ vals, u64 cap) {
    const u64 n = st->n_mlog;
    const i64 nfire = st->fire_count, k = size / slide;
    const i64 P0 = (st->fire_first_end - size - off) / slide;             // exact: window ends are off (mod slide)
    const i64 jhi_all = (j0 + nj < nfire ? j0 + nj : nfire) - 1;
    constexpr u64 PER = (u64)MED_THREADS * MED_RPT;
    for (u64 base = (u64)blockIdx.x * PER; base < n; base += (u64)gridDim.x * PER) {
        i64 ja[MED_RPT]; u32 cnt[MED_RPT], tot = 0;
#pragma unroll
        for (int r = 0; r < MED_RPT; ++r) {
            const u64 i = base + (u64)r * MED_THREADS + threadIdx.x;
            i64 a = 1, b = 0;
            if (i < n) {
                const i64 q = ml_pane[i] - P0;                            // (rows of pane ML_PANE_NONE lie before every window)
                a = q - (k - 1) > j0 ? q - (k - 1) : j0; b = q < jhi_all ? q : jhi_all;
            }
            ja[r] = a; cnt[r] = b >= a ? (u32)(b - a + 1) : 0u; tot += cnt[r];
        }
        u64 row = block_append<MED_THREADS>(tot, &st->n_mwork, cap, st);
        if (tot && row != ~0ull) {
#pragma unroll
            for (int r = 0; r < MED_RPT; ++r) {
                if (!cnt[r]) continue;
                const u64 i = base + (u64)r * MED_THREADS + threadIdx.x;
                const u64 slot = ml_slot[i], v = ml_val[i];
                for (u32 c = 0; c < cnt[r]; ++c, ++row) { items[row] = ((((u64)(ja[r] + c - j0) << slot_bits) | slot) << 32) | (u32)row; vals[row] = v; }
            }
        }
    }
}

// keep only the records of panes > x_hi (their windows have not all fired yet)
__global__ void __launch_bounds__(MED_THREADS) median_compact_kernel(const i64* __restrict__ ip, const u32* __restrict__ is, const u64* __restrict__ iv, u64 n,
                                                                    DevState* st, i64 x_hi, i64* __restrict__ op, u32* __restrict__ os, u64* __restrict__ ov, u64 cap) {
    constexpr u64 PER = (u64)MED_THREADS * MED_RPT;
    for (u64 base = (u64)blockIdx.x * PER; base < n; base += (u64)gridDim.x * PER) {
        i64 pane[MED_RPT]; u32 keep = 0;
#pragma unroll
        for (int r = 0; r < MED_RPT; ++r) {
            const u64 i = base + (u64)r * MED_THREADS + threadIdx.x;
            pane[r] = i < n ? ip[i] : x_hi;
            keep |= (pane[r] > x_hi ? 1u : 0u) << r;
        }
        u64 row = block_append<MED_THREADS>(__popc(keep), &st->n_mwork, cap, st);
        if (keep && row != ~0ull) {
#pragma unroll
            for (int r = 0; r < MED_RPT; ++r) {
                if (!((keep >> r) & 1u)) continue;
                const u64 i = base + (u64)r * MED_THREADS + threadIdx.x;
                op[row] = pane[r]; os[row] = is[i]; ov[row] = iv[i]; ++row;
            }
        }
    }
}

// ---- the middle of every (window, key) run -----------------------------------------------------------------------
// Items arrive sorted by group = (window, slot) only (two 11-bit passes for a 2^20-slot table); the values of a run are
// put in order where they lie: a run of up to MED_SMALL values by ONE thread in local memory (1 M keys x 10 values
// per window is the tutorial's shape), a longer one by a block that radix-selects the middle rank, most significant byte
// first.  The eight passes over the 64-bit values that a full (group, value) sort would need are gone.
constexpr u32 MED_SMALL = 32;

struct MedianPickArgs {
    const u64* items; u64* keys;                       // items sorted by group; keys[t] = value of item t in Double.compareTo order (written for long runs only)
    const u64* vals; u64 n;                            // unused tail items are all ones
    const u64* dkeys; int dense_keys; i64 window_end; DevState* st;
    u64* m_key; i64* m_wend; double* m_val; u64 m_cap;
    u32 slot_bits; i64 slide;                          // window of an item = window_end + (group >> slot_bits) * slide
    u64* big; u64 big_cap;                             // big[0] = number of long runs, then (start, length) pairs
};

__device__ __forceinline__ double median_value(u64 hi_key, u64 lo_key, bool odd) {      // ComputeCpuMiddle.java:43-47
    const double hi = hi_key == ~0ull ? __longlong_as_double(0x7FF8000000000000ll) : unorder_f64(hi_key);
    if (odd) return hi;
    const double lo = lo_key == ~0ull ? __longlong_as_double(0x7FF8000000000000ll) : unorder_f64(lo_key);
    return (hi + lo) / 2;
}
__device__ __forceinline__ void median_row(const MedianPickArgs& a, u64 row, u32 grp, double med) {
    const u32 s = grp & ((1u << a.slot_bits) - 1u);
    a.m_key[row] = a.dense_keys ? (u64)s : a.dkeys[s]; a.m_wend[row] = a.window_end + (i64)(grp >> a.slot_bits) * a.slide; a.m_val[row] = med;
}

// Short runs, every lane at work: the block holds MED_THREADS positions and MED_SMALL neighbours on either side in shared
// memory; the thread at position p finds its run's ends, counts the values that precede its own in (value, position)
// order and remembers the largest of them — the one thread per run whose count is len/2 holds the upper middle value and,
// in that maximum, the lower one.  The first thread of a run that is too long hands it to median_long_kernel.
constexpr u32 MED_OUT = 1024;        // rows a block stages before it reserves output space (one atomic per ~35 tiles, not one per tile)
__global__ void __launch_bounds__(MED_THREADS) median_pick_kernel(const MedianPickArgs a) {
    constexpr u32 H = MED_SMALL, TILE = MED_THREADS + 2 * H;
    static_assert(2 * H <= MED_THREADS, "two positions per thread cover the tile");
    __shared__ u32 sg[TILE];
    __shared__ u64 sk[TILE];
    __shared__ u32 og[MED_OUT];
    __shared__ double ov[MED_OUT];
    __shared__ u32 on;
    __shared__ u64 obase;
    if (threadIdx.x == 0) on = 0;
    auto flush = [&]() {                                             // whole block; `on` is stable between the barriers around it
        const u32 m = on;
        if (threadIdx.x == 0) obase = m ? atomicAdd((unsigned long long*)&a.st->n_median, (unsigned long long)m) : 0ull;
        __syncthreads();
        if (m && obase + m > a.m_cap) { if (threadIdx.x == 0) atomicOr(&a.st->err_flags, ERRB_OUTPUT); }
        else for (u32 i = threadIdx.x; i < m; i += MED_THREADS) median_row(a, obase + i, og[i], ov[i]);
        __syncthreads();
        if (threadIdx.x == 0) on = 0;
    };
    // position base + x - H of the tile at `base`: its group and its value in comparison order (the one random access per item)
    auto fetch = [&](u64 base, u32 x, u32& g, u64& key) {
        const u64 p = base + x;
        g = 0xFFFFFFFFu; key = 0;
        if (x < TILE && base < a.n && p >= H && p - H < a.n) {
            const u64 it = a.items[p - H];
            g = (u32)(it >> 32);
            if (g != 0xFFFFFFFFu) key = Codec<double>::cmp_key(__longlong_as_double((i64)a.vals[(u32)it]));
        }
    };
    const u64 step = (u64)gridDim.x * MED_THREADS;
    u32 g0, g1; u64 k0, k1;                                          // the next tile, in flight while this one is worked on
    fetch((u64)blockIdx.x * MED_THREADS, threadIdx.x, g0, k0); fetch((u64)blockIdx.x * MED_THREADS, threadIdx.x + MED_THREADS, g1, k1);
    for (u64 base = (u64)blockIdx.x * MED_THREADS; base < a.n; base += step) {
        __syncthreads();                                             // the previous tile has been read, its rows are staged
        if (on + MED_THREADS > MED_OUT) flush();
        sg[threadIdx.x] = g0; sk[threadIdx.x] = k0;
        if (threadIdx.x < 2 * H) { sg[threadIdx.x + MED_THREADS] = g1; sk[threadIdx.x + MED_THREADS] = k1; }
        __syncthreads();
        fetch(base + step, threadIdx.x, g0, k0); fetch(base + step, threadIdx.x + MED_THREADS, g1, k1);
        const u32 me = threadIdx.x + H;
        const u32 grp = sg[me];
        if (grp != 0xFFFFFFFFu) {
            u32 l = 0, r = 0;
            while (l < H && sg[me - l - 1] == grp) ++l;
            while (r < H && sg[me + r + 1] == grp) ++r;
            const u32 len = l + r + 1;
            const u64 mine = sk[me];
            if (len <= MED_SMALL) {
                u32 rank = 0; u64 below = 0;
                for (u32 x = me - l; x <= me + r; ++x) {
                    const u64 o = sk[x];
                    if (o < mine || (o == mine && x < me)) { ++rank; if (o > below) below = o; }
                }
                if (rank == len / 2) { const u32 i = atomicAdd(&on, 1u); og[i] = grp; ov[i] = median_value(mine, below, len & 1); }
            } else {
                const u64 t = base + threadIdx.x;
                a.keys[t] = mine;                                    // long runs are finished from global memory
                if (l == 0) {
                    u64 lo = t + r + 1, hi = a.n;                    // first position of another group
                    while (lo < hi) { const u64 mid = (lo + hi) >> 1; if ((u32)(a.items[mid] >> 32) == grp) lo = mid + 1; else hi = mid; }
                    const u64 e = atomicAdd((unsigned long long*)&a.big[0], 1ull);
                    if (e < a.big_cap) { a.big[2 + 2 * e] = t; a.big[3 + 2 * e] = lo - t; } else atomicOr(&a.st->err_flags, ERRB_OUTPUT);
                }
            }
        }
    }
    __syncthreads();
    flush();
}

// long runs, one block each: the key of rank len/2 byte by byte from the top (a 256-bin count of the keys that share the
// bytes chosen so far), then for an even length the rank below it = the same key again if it repeats, else the largest smaller one
__global__ void __launch_bounds__(MED_THREADS) median_long_kernel(const MedianPickArgs a) {
    __shared__ u32 hist[256];
    __shared__ u64 sh_prefix, sh_rank, sh_below, sh_nbelow;
    const u64 nbig = a.big[0] < a.big_cap ? a.big[0] : a.big_cap;
    for (u64 e = blockIdx.x; e < nbig; e += gridDim.x) {
        const u64 start = a.big[2 + 2 * e], len = a.big[3 + 2 * e];
        const u64* k = a.keys + start;
        if (threadIdx.x == 0) { sh_prefix = 0; sh_rank = len / 2; sh_below = 0; sh_nbelow = 0; }
        for (int shift = 56; shift >= 0; shift -= 8) {
            hist[threadIdx.x] = 0;                                   // MED_THREADS == 256 bins
            __syncthreads();
            const u64 mask = shift == 56 ? 0ull : ~0ull << (shift + 8), prefix = sh_prefix;
            for (u64 i = threadIdx.x; i < len; i += MED_THREADS) { const u64 x = k[i]; if ((x & mask) == prefix) atomicAdd(&hist[(x >> shift) & 255u], 1u); }
            __syncthreads();
            if (threadIdx.x < 32) {                                  // the bin that holds the wanted rank
                u32 c[8], sum = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) { c[j] = hist[threadIdx.x * 8 + j]; sum += c[j]; }
                u32 inc = sum;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const u32 v = __shfl_up_sync(0xffffffffu, inc, o); if ((int)threadIdx.x >= o) inc += v; }
                const u64 r = sh_rank; u64 before = inc - sum;
                __syncwarp();
                if (r >= before && r < (u64)inc) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (r >= before && r < before + c[j]) { sh_prefix = prefix | ((u64)(threadIdx.x * 8 + j) << shift); sh_rank = r - before; }
                        before += c[j];
                    }
                }
            }
            __syncthreads();
        }
        const u64 hi_key = sh_prefix;
        u64 lo_key = hi_key;
        if (!(len & 1)) {
            u64 mx = 0, cnt = 0;
            for (u64 i = threadIdx.x; i < len; i += MED_THREADS) { const u64 x = k[i]; if (x < hi_key) { ++cnt; if (x > mx) mx = x; } }
            if (cnt) { atomicMax((unsigned long long*)&sh_below, mx); atomicAdd((unsigned long long*)&sh_nbelow, cnt); }
            __syncthreads();
            if (sh_nbelow > len / 2 - 1) lo_key = sh_below;          // ranks 0 .. nbelow-1 are smaller keys: rank len/2-1 is the largest of them
        }
        if (threadIdx.x == 0) {
            const u64 row = atomicAdd((unsigned long long*)&a.st->n_median, 1ull);
            if (row < a.m_cap) median_row(a, row, (u32)(a.items[start] >> 32), median_value(hi_key, lo_key, len & 1));
            else atomicOr(&a.st->err_flags, ERRB_OUTPUT);
        }
        __syncthreads();
    }
}

}  // namespace fmq
