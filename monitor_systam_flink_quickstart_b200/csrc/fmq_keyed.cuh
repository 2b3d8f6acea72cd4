// Keyed ValueState operators in per-key ARRIVAL order:
//   (i)  rolling max  = keyBy(0).max(2), chapter2/.../ComputeCpuMax.java:26
//        (StreamGroupedReduce + ComparableAggregator: one output per input, Double.compareTo order)
//   (ii) N consecutive over-threshold readings (BASELINE.json north_star; spec in SURVEY §8 A7-ii)
// Pipeline per batch: resolve key -> directory slot, stable LSD radix sort of (slot, arrival idx),
// then every run of equal slots is walked in order, seeded from / written back to the per-key state.
#pragma once
#include "fmq_device.cuh"

namespace fmq {

// per-key state, indexed by directory slot
//   kflags[slot] : bit 31 = rolling max present, bits 0..30 = consecutive-breach counter
//   kmax[slot]   : rolling max value bits
constexpr u32 KF_HAS_MAX = 0x80000000u;
constexpr u32 KF_SEEN = 0x40000000u;       // dense-id mode: this id has been seen by a keyed operator
constexpr u32 KF_COUNT_MASK = 0x3FFFFFFFu;

// ---- step 1: key -> (slot << 32 | arrival index), one packed 8-byte item per event ----------------
__global__ void __launch_bounds__(256) resolve_pack_kernel(const u64* __restrict__ key, size_t n, u64* dkeys, u64 dcap, int dense,
                                                           DevState* st, u64* __restrict__ packed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const u64 k = ldg_stream(key + i);
        u32 s = SLOT_INVALID;
        if (k == KEY_EMPTY || (dense && k >= dcap)) atomicOr(&st->err_flags, ERRB_BAD_INPUT);
        else s = dense ? (u32)k : dir_find_or_insert(dkeys, dcap, k, st);
        packed[i] = ((u64)s << 32) | (u32)i;
    }
}

// ---- step 2: stable LSD radix sort of the packed items by slot, BITS (<= 11) bits per pass -------
// Same scheme as the owner split (a warp owns a contiguous chunk, counts[digit][chunk] scanned
// digit-major), tuned for this job: 8-byte items, up to 2048 bins so that 22 slot bits need two
// passes, and four rounds of loads issued ahead of the ranking so that a warp is not one
// dependent load per 32 items.
constexpr u32 PS_CHUNK = 4096;
constexpr int PS_AHEAD = 4;
__host__ __device__ inline u32 ps_num_chunks(size_t n) { return (u32)((n + PS_CHUNK - 1) / PS_CHUNK); }

// KeyFn: u64 operator()(u64 item) -> the sort key of an item; a pass ranks by (key >> shift) & (2^BITS - 1)
struct SlotKey {            // item = (slot << 32 | index): sort by slot
    __device__ __forceinline__ u64 operator()(u64 item) const { return item >> 32; }
};
struct F64ValueKey {        // item = (slot << 32 | index into vals): sort by vals[index] in Double.compareTo order
    const u64* vals;
    __device__ __forceinline__ u64 operator()(u64 item) const { return Codec<double>::cmp_key(__longlong_as_double((i64)vals[(u32)item])); }
};

template <int BITS, bool SCATTER, typename KeyFn>
__global__ void __launch_bounds__(256) ps_pass_kernel(const u64* __restrict__ in, u64* __restrict__ out, size_t n, u32 nchunks, u32 shift,
                                                      u32* __restrict__ counts /* hist: written; scatter: scanned offsets */, KeyFn keyfn) {
    extern __shared__ u32 ps_smem[];
    constexpr u32 NB = 1u << BITS;
    const u32 wl = threadIdx.x >> 5, lane = threadIdx.x & 31;
    u32* bins = ps_smem + wl * NB;
    const u32 chunk = blockIdx.x * 8 + wl;
    if (chunk >= nchunks) return;
    for (u32 d = lane; d < NB; d += 32) bins[d] = SCATTER ? counts[(size_t)d * nchunks + chunk] : 0u;
    __syncwarp();
    const size_t lo = (size_t)chunk * PS_CHUNK;
    const size_t hi = (lo + PS_CHUNK < n) ? lo + PS_CHUNK : n;
    for (size_t i0 = lo; i0 < hi; i0 += 32 * PS_AHEAD) {
        u64 it[PS_AHEAD];
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r) { const size_t i = i0 + r * 32 + lane; it[r] = i < hi ? in[i] : ~0ull; }
        if constexpr (!SCATTER) {
            // counting needs no order: one shared-memory atomic per item on the warp's private bins
#pragma unroll
            for (int r = 0; r < PS_AHEAD; ++r)
                if (i0 + r * 32 + lane < hi) atomicAdd(&bins[(u32)(keyfn(it[r]) >> shift) & (NB - 1)], 1u);
        } else {
#pragma unroll
        for (int r = 0; r < PS_AHEAD; ++r) {
            const bool act = i0 + r * 32 + lane < hi;
            const unsigned am = __ballot_sync(0xffffffffu, act);
            if (act) {
                const u32 d = (u32)(keyfn(it[r]) >> shift) & (NB - 1);
                const unsigned m = __match_any_sync(am, d);
                const int leader = __ffs(m) - 1;
                u32 base = 0;
                if ((int)lane == leader) { base = bins[d]; bins[d] = base + __popc(m); }
                base = __shfl_sync(m, base, leader);
                out[(size_t)base + __popc(m & ((1u << lane) - 1))] = it[r];
            }
            __syncwarp();
        }
        }
    }
    __syncwarp();
    if (!SCATTER) for (u32 d = lane; d < NB; d += 32) counts[(size_t)d * nchunks + chunk] = bins[d];
}

// ---- step 3: apply the operators along every key's run, in arrival order ---------------------------
struct KeyedArgs {
    const u64* sorted; size_t n;                    // (slot << 32 | arrival index), sorted by slot, stable
    const i64* ts; const u64* key; const void* val;
    DevState* st; u32* kflags; u64* kmax;
    int rolling_enabled; u64* rolling_out;          // [n], input order
    int consec_n, consec_op, consec_reset; double consec_thr;
    i64* ca_ts; u64* ca_key; u64* ca_val; u64 ca_cap;
};

constexpr u32 KA_RANGE = 512;                       // nominal sorted positions per warp

// A run (= all records of one key in this batch) belongs to the warp whose nominal range contains
// its first position; the warp follows its last run past the end of the range, so no state ever
// crosses warps.  Inside a warp 32 consecutive positions are handled per round: the per-record
// transition functions are composed with a segmented warp scan, seeded from the key's stored state
// at run heads and from the previous round otherwise.
//   consecutive-breach counter:  over -> c+1 (alert when it reaches N, then 0 if reset-on-fire), else -> 0;
//     composite of a sub-run = (has_reset, k): c -> has_reset ? k : c + k
//   rolling max (ComparableAggregator): keep cur iff cur.compareTo(v) > 0, else take v = max-scan
//     on the compareTo order with ties going to the later record
template <typename V>
__global__ void __launch_bounds__(256) keyed_scan_kernel(const __grid_constant__ KeyedArgs a) {
    const V* __restrict__ val = static_cast<const V*>(a.val);
    const u32 lane = threadIdx.x & 31;
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t r_lo = warp * KA_RANGE;
    if (r_lo >= a.n) return;
    const size_t r_hi = (r_lo + KA_RANGE < a.n) ? r_lo + KA_RANGE : a.n;
    const u32 N = (u32)a.consec_n;
    // ---- find the first run head at or after r_lo ----
    size_t pos = r_lo;
    {
        bool found = false;
        while (pos < r_hi) {
            const size_t p = pos + lane;
            bool head = false;
            if (p < r_hi) { const u32 s = (u32)(a.sorted[p] >> 32); head = (p == 0) || ((u32)(a.sorted[p - 1] >> 32) != s); }
            const unsigned hm = __ballot_sync(0xffffffffu, head);
            if (hm) { pos += __ffs(hm) - 1; found = true; break; }
            pos += 32;
        }
        if (!found) return;                            // the whole range continues a run owned by an earlier warp
    }
    // carried from the previous round (lane-uniform)
    u32 c_state = 0; u64 c_mkey = 0, c_mbits = 0; bool c_hasmax = false; u32 c_slot = SLOT_INVALID;
    u64 nalert = 0;
    bool done = false;
    while (!done && pos < a.n) {
        const size_t p = pos + lane;
        const bool in = p < a.n;
        const u64 item = in ? a.sorted[p] : ~0ull;
        const u32 slot = (u32)(item >> 32), idx = (u32)item;
        const u32 prev_slot = __shfl_up_sync(0xffffffffu, slot, 1);
        bool head = in && (lane == 0 ? (slot != c_slot) : (slot != prev_slot));
        // the warp stops at the first head at or beyond the end of its nominal range
        const unsigned stopm = __ballot_sync(0xffffffffu, in && head && p >= r_hi);
        const u32 nact = stopm ? (u32)(__ffs(stopm) - 1) : 32u;
        const bool act = in && lane < nact && slot != SLOT_INVALID;
        // warp-uniform exit: a later run head past the range, the end of the array, or the invalid slots (sorted last)
        if (stopm || __ballot_sync(0xffffffffu, !in || slot == SLOT_INVALID)) done = true;
        if (!act) head = false;
        // gathers: value of this record, stored state at run heads
        V v = act ? val[idx] : V(0);
        u64 vb; memcpy(&vb, &v, 8);
        u32 fl = 0; u64 smax = 0;
        if (head) { fl = a.kflags[slot]; smax = a.kmax[slot]; }
        // ---- consecutive-breach: element (hr = result independent of what precedes, k) ----
        bool hr = false; u32 k = 0;
        if (N > 0 && act) {
            const bool over = cmp_op((double)v, a.consec_op, a.consec_thr);
            hr = !over; k = over ? 1u : 0u;
            if (head) { const u32 seed = fl & KF_COUNT_MASK; k = over ? seed + 1 : 0; hr = true; }
        }
        // ---- rolling max: element (mkey, mbits, has) ----
        u64 mkey = act ? Codec<V>::cmp_key(v) : 0, mbits = vb; bool seg = head;     // seg = "a head lies at or before me in this round"
        if (a.rolling_enabled && head && (fl & KF_HAS_MAX)) {
            const u64 sk = Codec<V>::cmp_key(Codec<V>::from_bits(smax));
            if (sk > mkey) { mkey = sk; mbits = smax; }
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const bool hr_p = __shfl_up_sync(0xffffffffu, (int)hr, o) != 0; const u32 k_p = __shfl_up_sync(0xffffffffu, k, o);
            const u64 mk_p = __shfl_up_sync(0xffffffffu, mkey, o), mb_p = __shfl_up_sync(0xffffffffu, mbits, o);
            const bool seg_p = __shfl_up_sync(0xffffffffu, (int)seg, o) != 0;
            if ((int)lane >= o && !seg) {                          // my sub-run does not start with a head: prepend the left neighbour's
                if (!hr) { k += k_p; hr = hr_p; }
                if (mk_p > mkey) { mkey = mk_p; mbits = mb_p; }
                seg = seg_p;
            }
        }
        // prepend the carry of the previous round for lanes whose run began before this round
        if (act && !seg) {
            if (!hr) k += c_state;
            if (c_hasmax && c_mkey > mkey) { mkey = c_mkey; mbits = c_mbits; }
        }
        u32 state = k; bool alert = false;
        if (N > 0 && act) {
            if (a.consec_reset) {
                state = k % N;
                alert = cmp_op((double)v, a.consec_op, a.consec_thr) && state == 0;     // the counter just reached N (and went back to 0)
            } else {
                if (state > KF_COUNT_MASK) state = KF_COUNT_MASK;
                alert = state >= N;
            }
        }
        if (N > 0) {
            const u64 row = warp_append(alert, &a.st->n_calert, a.ca_cap, a.st);         // one reservation per warp round
            if (row != ~0ull) { a.ca_ts[row] = a.ts[idx]; a.ca_key[row] = a.key[idx]; a.ca_val[row] = vb; }
            nalert += alert ? 1 : 0;
        }
        if (a.rolling_enabled && act) a.rolling_out[idx] = mbits;
        // run tails write the key's state back
        const u32 next_slot = __shfl_down_sync(0xffffffffu, slot, 1);
        const u64 peek = (p + 1 < a.n) ? a.sorted[p + 1] : ~0ull;
        const bool tail = act && ((lane == 31 || lane + 1 >= nact) ? ((u32)(peek >> 32) != slot) : (next_slot != slot));
        if (tail) {
            a.kflags[slot] = (a.rolling_enabled ? KF_HAS_MAX : 0u) | KF_SEEN | (state & KF_COUNT_MASK);
            if (a.rolling_enabled) a.kmax[slot] = mbits;
        }
        // carry = last active lane
        const int last = (int)(nact < 32 ? nact : 32) - 1;
        if (last >= 0) {
            c_state = __shfl_sync(0xffffffffu, state, last); c_mkey = __shfl_sync(0xffffffffu, mkey, last);
            c_mbits = __shfl_sync(0xffffffffu, mbits, last); c_slot = __shfl_sync(0xffffffffu, slot, last);
            c_hasmax = a.rolling_enabled != 0;
        }
        pos += 32;
    }
    nalert = warp_sum_u64(nalert);
    if (lane == 0 && nalert) atomicAdd(&a.st->consec_alerts, nalert);
}

__global__ void key_state_lookup_kernel(const u64* keys, size_t n, const u64* dkeys, u64 dcap, int dense, const u32* kflags, const u64* kmax,
                                        int* found, int* consec, u64* rolling) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 s = dense ? (keys[i] < dcap ? (u32)keys[i] : SLOT_INVALID) : dir_find(dkeys, dcap, keys[i]);
    if (s == SLOT_INVALID || (dense && !(kflags[s] & KF_SEEN))) { found[i] = 0; consec[i] = 0; rolling[i] = 0; return; }
    const u32 fl = kflags[s];
    found[i] = 1; consec[i] = (int)(fl & KF_COUNT_MASK); rolling[i] = (fl & KF_HAS_MAX) ? kmax[s] : 0;
}

}  // namespace fmq
